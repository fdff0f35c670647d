"""Print the measured host-link / FP32 / host-DRAM peaks of this box (clmgs.peaks) + box facts."""
import json, os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "clm-gs_b200")]
import torch
from clmgs import peaks
out = peaks.measure_all("cuda:0")
out["cpu_count"] = os.cpu_count()
try:
    out["mem_total_gb"] = round(int([l for l in open("/proc/meminfo") if l.startswith("MemTotal")][0].split()[1]) / 1e6, 1)
    out["lscpu"] = [l.strip() for l in subprocess.run(["lscpu"], capture_output=True, text=True).stdout.splitlines()
                    if any(k in l for k in ("Model name", "Socket", "NUMA node", "Thread(s)", "Core(s)", "Flags"))][:10]
    out["lscpu"] = [l if not l.startswith("Flags") else "avx512f=%s" % ("avx512f" in l) for l in out["lscpu"]]
except Exception as e:
    out["err"] = str(e)
print(json.dumps(out, indent=1))
