"""Turn ncu output into the small text summaries committed under profiles/.

  python tools/ncu_summary.py launches gpurun_out/launches.csv  "<command profiled>"  > profiles/rNN_launches_summary.csv
  ncu -i gpurun_out/prof.ncu-rep --page raw --csv > gpurun_out/prof_raw.csv
  python tools/ncu_summary.py full gpurun_out/prof_raw.csv "<command profiled>"       > profiles/rNN_ncu_full_summary.txt

`launches`: per-kernel totals of `ncu --metrics gpu__time_duration.sum` (cold-cache, serialised: compare shares).
`full`: the first launch of every kernel name in an `ncu --set full` report: duration, DRAM traffic (per-column unit
conversion), achieved GB/s and the utilisation counters DESIGN.md quotes.
"""
import csv
import re
import sys
from collections import OrderedDict

UNIT = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12,
        "ns": 1e-9, "us": 1e-6, "ms": 1e-3, "s": 1.0, "usecond": 1e-6, "nsecond": 1e-9, "msecond": 1e-3, "second": 1.0}


def short(name):
    name = re.sub(r"^void ", "", name).replace("<unnamed>::", "").replace("(anonymous namespace)::", "")
    name = name.split("(")[0]
    return re.sub(r"<.*", "", name) if name.startswith(("cub::", "at::")) else name


def rows_of(path):
    with open(path, newline="") as f:
        lines = [l for l in f if l.startswith('"')]
    return list(csv.reader(lines))


def launches(path, command):
    rows = rows_of(path)
    hdr = rows[0]
    k, v = hdr.index("Kernel Name"), hdr.index("Metric Value")
    agg = OrderedDict()
    for r in rows[1:]:
        n = re.sub(r"<.*", "", short(r[k]))
        t = float(r[v].replace(",", "")) * 1e-6
        a = agg.setdefault(n, [0.0, 0])
        a[0] += t
        a[1] += 1
    total = sum(a[0] for a in agg.values())
    n_l = sum(a[1] for a in agg.values())
    print("# ncu launch list summary: %s" % command)
    print("# ncu --metrics gpu__time_duration.sum --clock-control none (first %d launches; cold-cache and serialised: "
          "compare SHARES, not absolutes)" % n_l)
    print("# total %.2f ms over %d launches" % (total, n_l))
    print("ms_total,share_pct,launches,avg_ms,kernel")
    for n, (t, c) in sorted(agg.items(), key=lambda kv: -kv[1][0]):
        print("%.3f,%.1f,%d,%.4f,%s" % (t, 100 * t / total, c, t / c, n))


WANT = ["gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread",
        "launch__grid_size", "launch__block_size", "smsp__inst_executed.sum"]


def full(path, command):
    rows = rows_of(path)
    hdr, units = rows[0], rows[1]
    col = {h: i for i, h in enumerate(hdr)}

    def get(r, name):
        i = col.get(name)
        if i is None or r[i] == "":
            return None, ""
        return float(r[i].replace(",", "")), units[i]

    print("# ncu --set full --clock-control none --import-source on: %s" % command)
    print("# first captured launch of each kernel; traffic = dram__bytes_read.sum + dram__bytes_write.sum per launch; "
          "achieved = traffic / duration")
    seen = set()
    for r in rows[2:]:
        n = short(r[col["Kernel Name"]])
        if n in seen:
            continue
        seen.add(n)
        d, du = get(r, "gpu__time_duration.sum")
        rd, ru = get(r, "dram__bytes_read.sum")
        wr, wu = get(r, "dram__bytes_write.sum")
        secs = d * UNIT[du]
        rb, wb = rd * UNIT[ru], wr * UNIT[wu]
        print("\nkernel: %s" % n)
        print("  %-62s %.4f ms" % ("gpu__time_duration.sum", secs * 1e3))
        print("  %-62s %.1f MB read + %.1f MB write = %.1f MB  -> %.0f GB/s" %
              ("dram traffic", rb / 1e6, wb / 1e6, (rb + wb) / 1e6, (rb + wb) / secs / 1e9))
        for w in WANT:
            x, u = get(r, w)
            if x is not None:
                print("  %-62s %s %s" % (w, ("%f" % x).rstrip("0").rstrip("."), u))


if __name__ == "__main__":
    {"launches": launches, "full": full}[sys.argv[1]](sys.argv[2], sys.argv[3] if len(sys.argv) > 3 else "")
