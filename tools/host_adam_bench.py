"""Host Adam throughput on this machine's cores (no GPU needed): dense vs sparse row lists.

    python tools/host_adam_bench.py [n_rows=20000000] [threads=0] [density=0.45]

Measured on the GPU box's host (16 vCPU, 15 threads, 30 M rows, pageable memory): dense 190 GB/s,
45 % of the rows 136-142 GB/s, 10 % of the rows 114-120 GB/s (counting 1536 B per updated row).
"""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "clm-gs_b200"))
from clmgs import _lib  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 20_000_000
threads = int(sys.argv[2]) if len(sys.argv) > 2 else 0
density = float(sys.argv[3]) if len(sys.argv) > 3 else 0.45
L = _lib.lib()
tabs = [np.zeros((n, 48), np.float32) for _ in range(4)]  # param, grad, m, v
for t in tabs:
    t[:] = 0.001
lr = np.full(48, 1e-3, np.float32)
rng = np.random.default_rng(0)
sig = np.ones(4, np.int32)


def run(rows, zero_grad, label):
    rows = np.ascontiguousarray(rows, np.int32)
    sb, se = np.array([0], np.int64), np.array([len(rows)], np.int64)
    ss, sz = np.array([-1], np.int32), np.array([1 if zero_grad else 0], np.int32)
    best = 1e9
    for it in range(3):
        t0 = time.perf_counter()
        L.call("clmgs_cpu_adam_batched_sparse_step", tabs[0], tabs[1], tabs[2], tabs[3], 48, lr, rows, sb, se, ss, sz, 1,
               sig, 0.9, 0.999, 1e-8, it + 1, 1.0, threads, None)
        best = min(best, time.perf_counter() - t0)
    b = len(rows) * 192 * (6 if zero_grad else 8)
    print("%-34s %9d rows  %7.1f ms  %6.1f GB/s" % (label, len(rows), best * 1e3, b / best / 1e9))


run(np.arange(n), False, "dense, contiguous")
run(np.arange(n), True, "dense, contiguous, zero-grad rows")
run(np.nonzero(rng.random(n) < density)[0], False, "sparse %.0f%%, ascending" % (100 * density))
run(np.nonzero(rng.random(n) < 0.1)[0], False, "sparse 10%, ascending")
