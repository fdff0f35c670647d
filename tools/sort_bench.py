"""Time csrc/radix.cu alone: python tools/sort_bench.py [n] [end_bit] [key_bits]"""
import os
import sys

import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "clm-gs_b200"))
from clmgs import ops  # noqa: E402
from clmgs._lib import lib  # noqa: E402
import numpy as np  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 46_000_000
end_bit = int(sys.argv[2]) if len(sys.argv) > 2 else 13
key_bits = int(sys.argv[3]) if len(sys.argv) > 3 else end_bit
dev = "cuda:0"
g = torch.Generator(device=dev).manual_seed(0)
keys = torch.randint(0, 1 << min(key_bits, 31), (n,), device=dev, generator=g, dtype=torch.int64).to(torch.int32)
vals = torch.arange(n, device=dev, dtype=torch.int32)
k1, v1 = torch.empty_like(keys), torch.empty_like(vals)
sc = ops._scratch(lib().call("clmgs_radix_sort_pairs_u32_scratch_bytes", n), dev)
flag = np.zeros(1, np.int32)
st = torch.cuda.current_stream().cuda_stream
ts = []
for it in range(6):
    k0, v0 = keys.clone(), vals.clone()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    lib().call("clmgs_radix_sort_pairs_u32", n, k0, k1, v0, v1, end_bit, sc, sc.numel(), flag, st)
    e1.record()
    torch.cuda.synchronize()
    ts.append(e0.elapsed_time(e1))
passes = (end_bit + 7) // 8
t = min(ts[1:])
print("n=%d end_bit=%d passes=%d: %.3f ms  (%.3f ms/pass, %.0f GB/s at 20 B/pair/pass)" %
      (n, end_bit, passes, t, t / passes, 20.0 * n * passes / t / 1e6))
