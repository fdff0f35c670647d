"""Does backing the pinned tables with 2 MiB pages pay?  Host Adam on a sparse row list and the zero-copy row gather,
tables from torch's pinned allocator vs clmgs.hostmem (mmap + MADV_HUGEPAGE + cudaHostRegister)."""
import json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "clm-gs_b200")]
import numpy as np, torch
from clmgs import ops, hostmem
dev = "cuda:0"
M = int(os.environ.get("ROWS", 40_000_000))
rng = np.random.default_rng(0)
out = {}
for mode in ("torch_pinned", "huge_pages"):
    t0 = time.time()
    if mode == "torch_pinned":
        tabs = [torch.zeros(M, 48, pin_memory=True) for _ in range(4)]
    else:
        tabs = [hostmem.pinned_tensor((M, 48), zero=True) for _ in range(4)]
    alloc_s = time.time() - t0
    p, g, m, v = tabs
    res = dict(alloc_s=round(alloc_s, 1), hp=hostmem.huge_page_stats())
    col_lr = np.full(48, 1e-3, np.float32); sig = np.zeros(1, np.int32)
    for density in (1.0, 0.45, 0.1):
        rows = np.arange(M, dtype=np.int32) if density == 1.0 else np.sort(rng.choice(M, int(M * density), replace=False)).astype(np.int32)
        ts = []
        for rep in range(3):
            t0 = time.perf_counter()
            ops.cpu_adam_batched_sparse_step(p, g, m, v, col_lr, rows, [0], [len(rows)], [-1], sig, 0.9, 0.999, 1e-15, rep + 1, 1.0, 16)
            ts.append(time.perf_counter() - t0)
        res["adam_gbs_density_%.2f" % density] = round(len(rows) * 1344 / min(ts) / 1e9, 1)
        ids = torch.from_numpy(rows[: min(len(rows), 8_000_000)].astype(np.int64)).to(dev)
        d = torch.empty(ids.shape[0], 48, device=dev)
        ops.rows_gather(d, p, ids); torch.cuda.synchronize()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record(); ops.rows_gather(d, p, ids); ops.rows_gather(d, p, ids); e.record(); torch.cuda.synchronize()
        res["gather_gbs_density_%.2f" % density] = round(2 * ids.shape[0] * 192 / s.elapsed_time(e) / 1e6, 1)
    out[mode] = res
    del tabs, p, g, m, v
    import gc; gc.collect()
    try: torch._C._host_emptyCache()
    except Exception: pass
print(json.dumps(out, indent=1))
