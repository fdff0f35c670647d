"""Time the batch cull alone on a prefix of a BASELINE.json scene: python tools/cull_bench.py [config] [n] [cameras]"""
import os
import sys

import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "clm-gs_b200"))
from clmgs import ops, scene as S  # noqa: E402

cfg = sys.argv[1] if len(sys.argv) > 1 else "c4_bigcity_102m"
n = int(sys.argv[2]) if len(sys.argv) > 2 else 10_000_000
B = int(sys.argv[3]) if len(sys.argv) > 3 else 64
spec = S.CONFIGS[cfg]
g = S.make_gaussians(spec, n=n, want_sh=False)
V, K, P = S.make_cameras(spec, B, g["extent"])
dev = "cuda:0"
xyz, ls, rq = (torch.from_numpy(g[k]).to(dev) for k in ("xyz", "log_scale", "raw_quat"))
Vd, Kd = torch.from_numpy(V).to(dev), torch.from_numpy(K).to(dev)
ts = []
for it in range(5):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    out = ops.cull(xyz, ls, rq, Vd, Kd, spec.width, spec.height, activated=False)
    e1.record()
    torch.cuda.synchronize()
    ts.append(e0.elapsed_time(e1))
vis = float(sum(int(c) for c in out["counts"])) / (n * B)
print("%s: n=%d cameras=%d visible=%.2f%%  cull %.3f ms  (%.2f G pair-tests/s)" % (cfg, n, B, 100 * vis, min(ts[1:]), n * B / min(ts[1:]) / 1e6))
