"""How much do the zero-copy row-gather CTAs (host link, side stream) slow the compute kernels they share SMs with?
Times the tile stage / rasterise fwd / bwd / SH bwd of one camera alone and while a row gather loops on another stream.
   CLMGS_HOST_LOAD_CTAS=<n> python tools/interference_bench.py [config] [n_gaussians]"""
import json, math, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "clm-gs_b200")]
import numpy as np, torch
from clmgs import ops, scene as S, hostmem
cfg = sys.argv[1] if len(sys.argv) > 1 else "c4_bigcity_102m"
base = S.CONFIGS[cfg]
n = int(sys.argv[2]) if len(sys.argv) > 2 else base.n_gaussians
spec = S.SceneSpec(base.name, base.n_gaussians, base.width, base.height, 4, base.shape, base.fovx_deg, base.index)
g = S.make_gaussians(spec, n=n, want_sh=False)
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
nv = ids.shape[0]
col = torch.rand(1, nv, 3, device=dev); opac = o_[None].contiguous()
rc, ra, last = ops.rasterize_fwd(m2, con, col, opac, W, H, 16, off, flat)
v = torch.randn_like(rc)
rows = torch.rand(nv, 48, device=dev); rows_g = torch.zeros(nv, 48, device=dev); vcol = torch.randn(nv, 3, device=dev)
vm = torch.zeros(nv, 3, device=dev); campos = t(P[cam])
# host table + a long gather on a side stream
M = 20_000_000
h = hostmem.pinned_tensor((M, 48)); h[:1000].fill_(1.0)
gid = torch.from_numpy(np.sort(np.random.default_rng(0).choice(M, 6_000_000, replace=False)).astype(np.int64)).to(dev)
dst = torch.empty(gid.shape[0], 48, device=dev)
side = torch.cuda.Stream(priority=-1)
stages = dict(tile_stage=lambda: ops.isect_sorted_offsets(m2, rad, dep, 16, tw, th),
              raster_fwd=lambda: ops.rasterize_fwd(m2, con, col, opac, W, H, 16, off, flat),
              raster_bwd=lambda: ops.rasterize_bwd(m2, con, col, opac, W, H, 16, off, flat, ra, last, v, None),
              sh_bwd=lambda: ops.sh_colors_bwd(3, m, campos, rows, vcol, rows_g, vm))
def timeit(fn, reps=5):
    fn(); torch.cuda.current_stream().synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(reps): fn()
    e.record(); e.synchronize()
    return s.elapsed_time(e) / reps
out = dict(config=cfg, n_visible=int(nv), n_isects=int(n_is), load_ctas=os.environ.get("CLMGS_HOST_LOAD_CTAS", "48"))
for name, fn in stages.items():
    alone = timeit(fn)
    with torch.cuda.stream(side):
        gs, ge = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        gs.record(side)
        for _ in range(12): ops.rows_gather(dst, h, gid)
        ge.record(side)
    busy = timeit(fn)
    torch.cuda.synchronize()
    out[name] = dict(alone_ms=round(alone, 3), with_gather_ms=round(busy, 3), gather_gbs=round(12 * gid.shape[0] * 192 / gs.elapsed_time(ge) / 1e6, 1))
print(json.dumps(out))
