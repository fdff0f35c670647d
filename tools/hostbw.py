"""Debug helper: host DRAM bandwidth on the GPU box (STREAM-like, torch CPU threads) and host Adam alone."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "clm-gs_b200")]
import numpy as np, torch
from clmgs import ops
torch.set_num_threads(os.cpu_count())
n = 256 * 1024 * 1024
a = torch.ones(n); b = torch.ones(n); c = torch.empty(n)
for name, fn, bytes_ in (("copy", lambda: c.copy_(a), 8 * n), ("add", lambda: torch.add(a, b, out=c), 12 * n), ("scale_inplace", lambda: a.mul_(1.0001), 8 * n)):
    fn(); t0 = time.perf_counter()
    for _ in range(5): fn()
    dt = (time.perf_counter() - t0) / 5
    print("%-14s %.1f GB/s" % (name, bytes_ / dt / 1e9))
# host Adam alone over 5M rows (pinned vs pageable)
N = 5_000_000
for pinned in (False, True):
    mk = (lambda *s: torch.zeros(*s).pin_memory()) if pinned and torch.cuda.is_available() else torch.zeros
    p, g, m, v = mk(N, 48), mk(N, 48), mk(N, 48), mk(N, 48)
    g.fill_(0.01)
    rows = np.arange(N, dtype=np.int32)
    col_lr = np.full(48, 1e-3, np.float32)
    sig = np.zeros(1, np.int32)
    for thr in (8, 16, 32):
        ts = []
        for rep in range(3):
            g.fill_(0.01)
            t0 = time.perf_counter()
            ops.cpu_adam_batched_sparse_step(p, g, m, v, col_lr, rows, [0], [N], [-1], sig, 0.9, 0.999, 1e-15, rep + 1, 1.0, thr)
            ts.append(time.perf_counter() - t0)
        dt = min(ts)
        print("host adam pinned=%s threads=%d: %.1f ms, %.1f GB/s (1536 B/row)" % (pinned, thr, dt * 1e3, N * 1536 / dt / 1e9))
