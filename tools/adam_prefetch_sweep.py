"""Host Adam, sparse row lists: sweep of the software-prefetch distance / cache level (CLMGS_CPU_ADAM_PREFETCH).

    python tools/adam_prefetch_sweep.py [n_rows=10000000] [threads=3] [density=0.6]

One child process per setting (the knob is read once per process).  Counts 1344 B per updated row (packed gradients).
"""
import os
import subprocess
import sys

CHILD = r'''
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.join(%r, "clm-gs_b200"))
from clmgs import _lib
n, threads, density = int(sys.argv[1]), int(sys.argv[2]), float(sys.argv[3])
L = _lib.lib()
tabs = [np.full((n, 48), 0.001, np.float32) for _ in range(3)]
rng = np.random.default_rng(0)
rows = np.ascontiguousarray(np.nonzero(rng.random(n) < density)[0], np.int32)
packed = np.full((len(rows), 48), 0.001, np.float32)
lr = np.full(48, 1e-3, np.float32)
sig = np.ones(4, np.int32)
sb, se = np.array([0], np.int64), np.array([len(rows)], np.int64)
ss, sz = np.array([-1], np.int32), np.array([0], np.int32)
best = 1e9
for it in range(4):
    t0 = time.perf_counter()
    L.call("clmgs_cpu_adam_batched_sparse_step_packed", tabs[0], packed, 0, tabs[1], tabs[2], 48, lr, rows, sb, se, ss, sz, 1,
           sig, 0.9, 0.999, 1e-8, it + 1, 1.0, threads, None)
    best = min(best, time.perf_counter() - t0)
print("%%7.1f ms  %%6.1f GB/s  (%%.2f GB/s per thread)" %% (best * 1e3, len(rows) * 1344 / best / 1e9, len(rows) * 1344 / best / 1e9 / threads))
'''

root = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
args = sys.argv[1:] + ["10000000", "3", "0.6"][len(sys.argv) - 1:]
for cfg in ("0", "6,0", "12,0", "6,1", "12,1", "24,1", "48,1"):
    env = dict(os.environ, CLMGS_CPU_ADAM_PREFETCH=cfg)
    out = subprocess.run([sys.executable, "-c", CHILD % root] + args, env=env, capture_output=True, text=True)
    print("prefetch %-5s" % cfg, (out.stdout.strip().splitlines() or [out.stderr[-300:]])[-1])
