"""Rasterise fwd / bwd alone on one camera of a named config (kernel time, CUDA events, 10 repetitions).
   CLMGS_RASTER_BWD_TILE_ACC=0|1 python tools/raster_bench.py [config] [n_gaussians]   (0, the default: per-warp global reductions)"""
import json, math, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "clm-gs_b200")]
import numpy as np, torch
from clmgs import ops, scene as S
cfg = sys.argv[1] if len(sys.argv) > 1 else "c2_bicycle_6m"
base = S.CONFIGS[cfg]
n = int(sys.argv[2]) if len(sys.argv) > 2 else min(base.n_gaussians, 30_000_000)
spec = S.SceneSpec(base.name, base.n_gaussians, base.width, base.height, 4, base.shape, base.fovx_deg, base.index)
g = S.make_gaussians(spec, n=n)
V, K, P = S.make_cameras(spec, 4, g["extent"])
dev = "cuda:0"
t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
xyz, ls, rq, ol = t(g["xyz"]), t(g["log_scale"]), t(g["raw_quat"]), t(g["opacity_logit"])
W, H = spec.width, spec.height
cull = ops.cull(xyz, ls, rq, t(V), t(K), W, H, activated=False)
cam = int(np.argmax(cull["counts"])); st = sum(cull["counts"][:cam]); ids = cull["gaussian_ids"][st:st + cull["counts"][cam]]
m, s_, q_, o_ = ops.gather_activate(xyz, ls, rq, ol, ids)
rad, m2, dep, con, _ = ops.projection_fwd(m, q_, s_, t(V[cam:cam + 1]), t(K[cam:cam + 1]), W, H)
tw, th = math.ceil(W / 16.0), math.ceil(H / 16.0)
flat, off, n_is = ops.isect_sorted_offsets(m2, rad, dep, 16, tw, th)
col = torch.rand(1, ids.shape[0], 3, device=dev); opac = o_[None].contiguous()
rc, ra, last = ops.rasterize_fwd(m2, con, col, opac, W, H, 16, off, flat)
v = torch.randn_like(rc)
def timeit(fn, reps=10):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(reps): fn()
    e.record(); torch.cuda.synchronize()
    return s.elapsed_time(e) / reps
# pre-allocated outputs are not possible through ops (it zero-fills per call): subtract the zero-fill separately
zf = timeit(lambda: [torch.zeros_like(m2), torch.zeros_like(con), torch.zeros_like(col), torch.zeros_like(opac)])
out = dict(config=cfg, n_visible=int(ids.shape[0]), n_isects=int(n_is), tile_acc=os.environ.get("CLMGS_RASTER_BWD_TILE_ACC", "0"),
           fwd_ms=round(timeit(lambda: ops.rasterize_fwd(m2, con, col, opac, W, H, 16, off, flat)), 4),
           bwd_ms_incl_zero_fill=round(timeit(lambda: ops.rasterize_bwd(m2, con, col, opac, W, H, 16, off, flat, ra, last, v, None)), 4),
           zero_fill_ms=round(zf, 4),
           tile_stage_ms=round(timeit(lambda: ops.isect_sorted_offsets(m2, rad, dep, 16, tw, th)), 4))
print(json.dumps(out))
