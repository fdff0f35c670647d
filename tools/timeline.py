"""Device timeline of one training iteration (debug helper, not part of the product).
   python tools/timeline.py [config] [first_camera] [n_cameras]
Prints, for a few cameras in the middle of the step, every stage's start / end on its stream (CUDA events, ms from the
start of the step) and the idle gaps of the compute stream between consecutive stages."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "clm-gs_b200")]
import numpy as np, torch
from clmgs import engine, hostmem, model, scene as S

cfg = sys.argv[1] if len(sys.argv) > 1 else "c2_bicycle_6m"
first = int(sys.argv[2]) if len(sys.argv) > 2 else 1
ncam = int(sys.argv[3]) if len(sys.argv) > 3 else 2
spec = S.CONFIGS[cfg]
N, bsz = spec.n_gaussians, spec.batch
dev = "cuda:0"
sh_host = hostmem.pinned_tensor((N, 48))
g = S.make_gaussians(spec, sh_out=sh_host.numpy())
args = model.TrainArgs(bsz=bsz, sparse_adam=cfg.startswith(("c4", "c5")), cpu_adam_threads=max(1, os.cpu_count() - 1))
gm = model.GaussianModelCLMOffload(3, dev).create_from_arrays(g["xyz"], g["log_scale"], g["raw_quat"], g["opacity_logit"],
                                                             sh_host, grad_table=False).training_setup(args)
nb = 4
V, K, P = S.make_cameras(spec, nb * bsz, g["extent"])
gts = torch.from_numpy(S.make_gt_images(spec, 4)).to(dev)
comm = torch.cuda.Stream(priority=-1)
gen = torch.Generator().manual_seed(0)
COMPUTE = ("cull", "order", "gather_activate", "projection_fwd", "isect_count", "isect_sort", "sh_fwd", "raster_fwd", "loss",
           "raster_bwd", "projection_bwd", "sh_bwd", "activate_bwd_scatter", "gpu_adam")
for b in range(nb):
    cams = [model.Camera(V[b * bsz + i], K[b * bsz + i], gts[i % 4], dev) for i in range(bsz)]
    timer = engine.StageTimer(b == nb - 1)
    torch.cuda.synchronize(); t0 = time.perf_counter()
    engine.clm_offload_train_one_batch(gm, None, cams, gm.parameters_grad_buffer, None, None, comm, gen, timer=timer)
    print("iter", b, "wall %.1f ms" % ((time.perf_counter() - t0) * 1e3), {k: round(v, 1) for k, v in gm.last_batch_stats.items() if "ms" in k})
tl = timer.timeline
comp = [(n, s, e) for n, s, e in tl if n in COMPUTE]
busy = sum(e - s for n, s, e in comp)
print("compute-stream stages: %d, busy %.1f ms, span %.1f ms" % (len(comp), busy, comp[-1][2] - comp[0][1]))
# camera index of a stage = number of raster_fwd stages seen before it
k, prev_end = 0, None
for n, s, e in comp:
    if first <= k < first + ncam:
        gap = (s - prev_end) if prev_end is not None else 0.0
        print("cam %2d  %-22s %9.2f -> %9.2f  (%6.2f ms)  gap before %6.2f" % (k, n, s, e, e - s, gap))
    prev_end = e
    if n == "activate_bwd_scatter":
        k += 1
print("per-stage statistics over the step (ms): n, total, min, median, p90, max")
import collections
by = collections.defaultdict(list)
for n, s_, e_ in tl:
    by[n].append(e_ - s_)
for n, v in sorted(by.items(), key=lambda kv: -sum(kv[1])):
    v = np.array(v)
    print("   %-22s %4d %9.2f %8.3f %8.3f %8.3f %8.3f" % (n, len(v), v.sum(), v.min(), np.median(v), np.percentile(v, 90), v.max()))
slow = sorted([(e_ - s_, n, s_) for n, s_, e_ in tl if n == "raster_fwd"], reverse=True)[:8]
print("slowest raster_fwd instances (ms, start):", [(round(a, 2), round(c, 1)) for a, _, c in slow])
print("copy streams:")
lo = min(s for (n, s, e), kk in zip(comp, range(len(comp))))
for n, s, e in tl:
    if n in ("rows_load_first", "rows_prefetch", "grad_handoff"):
        print("   %-16s %9.2f -> %9.2f (%6.2f ms)" % (n, s, e, e - s))
