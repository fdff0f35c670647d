"""Debug helper: zero-copy SH-row gather/scatter throughput vs row density, alone and with the host Adam running."""
import os, sys, time, threading
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "clm-gs_b200")]
import numpy as np, torch
from clmgs import ops
dev = "cuda:0"
N = 40_000_000
h = torch.empty(N, 48).pin_memory(); h.fill_(1.0)
hg = torch.zeros(N, 48).pin_memory()
def timed(fn, reps=3):
    fn(); torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(reps): fn()
    e.record(); torch.cuda.synchronize()
    return s.elapsed_time(e) / reps
rng = np.random.default_rng(0)
# a second set of host tables for the Adam thread to hammer
NA = 20_000_000
p, g, m, v = (torch.zeros(NA, 48) for _ in range(4))
rowsA = np.arange(NA, dtype=np.int32); col_lr = np.full(48, 1e-3, np.float32); sig = np.zeros(1, np.int32)
stop = [False]
def hammer(threads):
    k = 1
    while not stop[0]:
        ops.cpu_adam_batched_sparse_step(p, g, m, v, col_lr, rowsA, [0], [NA], [-1], sig, 0.9, 0.999, 1e-15, k, 1.0, threads); k += 1
for density in (0.9, 0.2, 0.02):
    n = int(N * density) if density < 0.5 else int(N * density) // 4
    base = 0
    ids = np.sort(rng.choice(N if density < 0.5 else N // 4, n, replace=False)).astype(np.int64)
    t_ids = torch.from_numpy(ids).to(dev)
    d = torch.empty(n, 48, device=dev); dg = torch.ones(n, 48, device=dev)
    for adam_threads in (0, 8, 16):
        stop[0] = False
        th = None
        if adam_threads:
            th = threading.Thread(target=hammer, args=(adam_threads,)); th.start(); time.sleep(0.3)
        t_in = timed(lambda: ops.rows_gather(d, h, t_ids))
        t_out = timed(lambda: ops.rows_scatter_grad(dg, hg, t_ids, accum=False))
        stop[0] = True
        if th: th.join()
        print("density %.2f rows %8d  adam_threads %2d : gather %.1f GB/s  scatter %.1f GB/s" % (density, n, adam_threads, n * 192 / t_in / 1e6, n * 192 / t_out / 1e6))
