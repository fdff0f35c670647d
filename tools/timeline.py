"""Debug helper: print the device timeline of one C2 training iteration (not part of the product)."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "clm-gs_b200")]
import numpy as np, torch
from clmgs import engine, model, scene as S

spec = S.CONFIGS["c2_bicycle_6m"]
N, bsz = spec.n_gaussians, spec.batch
dev = "cuda:0"
sh_host = torch.empty(N, 48).pin_memory()
g = S.make_gaussians(spec, sh_out=sh_host.numpy())
args = model.TrainArgs(bsz=bsz)
gm = model.GaussianModelCLMOffload(3, dev).create_from_arrays(g["xyz"], g["log_scale"], g["raw_quat"], g["opacity_logit"], sh_host).training_setup(args)
V, K, P = S.make_cameras(spec, 5 * bsz, g["extent"])
gts = S.make_gt_images(spec, 4)
comm = torch.cuda.Stream(priority=-1)
gen = torch.Generator().manual_seed(0)
from clmgs import ops, _lib
_orig_call = ops.call
_calls = []
def _traced(name, *a):
    if not _calls_enabled[0] or name.endswith("_bytes") or "cpu_adam" in name or "tsp" in name:
        return _orig_call(name, *a)
    st = torch.cuda.current_stream()
    s_, e_ = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter(); s_.record(st)
    r = _orig_call(name, *a)
    e_.record(st)
    _calls.append((name, s_, e_, t0))
    return r
_calls_enabled = [False]
ops.call = _traced
for b in range(5):
    cams = [model.Camera(V[b * bsz + i], K[b * bsz + i], gts[i % 4], dev) for i in range(bsz)]
    timer = engine.StageTimer(False)
    _calls_enabled[0] = (b == 4)
    base_evt = torch.cuda.Event(enable_timing=True); base_evt.record()
    torch.cuda.synchronize(); t0 = time.perf_counter()
    engine.clm_offload_train_one_batch(gm, None, cams, gm.parameters_grad_buffer, None, None, comm, gen, timer=timer)
    dt = (time.perf_counter() - t0) * 1e3
    print("iter", b, "wall %.1f ms" % dt, {k: round(v, 1) for k, v in gm.last_batch_stats.items() if "ms" in k})
for name, s_, e_, th in _calls[:60]:
    print("%8.2f -> %8.2f  (%6.2f)  host@%7.2f  %s" % (base_evt.elapsed_time(s_), base_evt.elapsed_time(e_), s_.elapsed_time(e_), (th - t0) * 1e3, name))
