"""bench.py -- one CLM-offload training iteration per "step" on synthetic scenes.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--config NAME]

Metric (BASELINE.json): train iters/sec (one iteration = one optimiser step over a batch of cameras: cull -> order ->
per camera {SH gather, project, SH colour, tile sort, rasterise, loss, backward, gradient offload} -> GPU Adam + host
Adam) and the rasterise fwd+bwd ms.  Default workload = the north-star target, BASELINE.json configs[3]: 102 M
Gaussians, 1920x1080, batch 64, sparse Adam, no densification (release_scripts/bigcity.sh:73-92).

`value`  : iterations/s with the camera batches (matrices + uint8 GT images) already in HBM.
`e2e`    : same loop, but every step first copies its camera batch from pinned host memory and ends with a
           device->host read of the losses (the caller-visible API cost).
`roofline`: the stage with the largest device time of the step -- chosen from the measured stage times -- against the
           roof that bounds it: HBM (MEASURED_PEAKS.json), host link (cudaMemcpyAsync pinned<->device measured in this
           run), FP32 FMA (measured in this run).  `stages` gives the same for every stage, `peaks` the denominators.
`--gpus N`: configs with a batch of >= 4 cameras per GPU keep the GLOBAL batch fixed and split the cameras
           (strong scaling: 64 -> 32/16/8 per GPU for configs[3]); smaller batches scale weakly.
`other_configs` (N = 1, default run): configs[1] and configs[2] (clm vs naive) measured in child processes.
Opt-in modes, never part of the default line: `--pipeline` / `--lazy-adam` (relaxed call contract),
           `--spatial-order` (the scene's Gaussians stored in Morton order; noted in `config.spatial_order`).
`--impl reference`: the reference's CPU path (oracle port: C restatement of gsplat's kernels + cpu-adam, all host
           threads) on a bounded sample of the same workload, extrapolated.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for p in (ROOT, os.path.join(ROOT, "clm-gs_b200")):
    if p not in sys.path:
        sys.path.insert(0, p)

import numpy as np  # noqa: E402

METRIC = "train_iters_per_sec"
UNIT = "iter/s"
DEFAULT_CONFIG = "c4_bigcity_102m"
SPARSE_BY_DEFAULT = ("c4_bigcity_102m", "c5_stress_500m")   # release_scripts/bigcity.sh:89-92 (--sparse_adam)


# ------------------------------------------------------------------------------------------
def load_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        d = json.load(open(path))
        return dict(hbm_gbs=float(d["hbm_gbs"]), hbm_source="measured (MEASURED_PEAKS.json)",
                    sm_max_mhz=float(d.get("sm_max_mhz", 1965.0)))
    return dict(hbm_gbs=6650.0, hbm_source="fallback (B200_PROFILING.md)", sm_max_mhz=1965.0)


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md)."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "200"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.proc is None:
            return dict(sm_mhz=None, sm_max_mhz=None, reasons=["nvidia-smi unavailable"])
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 6:
                continue
            try:
                sm.append(float(f[0])); mx.append(float(f[1]))
            except ValueError:
                continue
            for nm, v in zip(names, f[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        return dict(sm_mhz=float(np.median(sm)) if sm else None, sm_max_mhz=max(mx) if mx else None,
                    reasons=sorted(reasons), samples=len(sm))


# kernels of OUR library per entry point (all hand-written; clmgs_isect_emit_sort is set per config in run_ours():
# depth keys + 4 radix passes x 3 kernels + count/scan/emit + 3 kernels per tile-key radix pass)
LAUNCHES = dict(clmgs_cull_count=2, clmgs_cull_emit=1, clmgs_gather_activate=1, clmgs_projection_fwd=1,
                clmgs_sh_colors_fwd=1, clmgs_isect_count=3, clmgs_isect_emit_sort=5, clmgs_isect_sorted_offsets=5, clmgs_isect_offset_encode=1,
                clmgs_rasterize_fwd=1, clmgs_rasterize_fwd_counted=1, clmgs_loss_fwd_bwd=3, clmgs_rasterize_bwd=1, clmgs_projection_bwd=1,
                clmgs_sh_colors_bwd=1, clmgs_activate_bwd_scatter=1, clmgs_rows_gather_i64=1,
                clmgs_rows_prefetch_next=2, clmgs_rows_grad_handoff=2, clmgs_set_signal=1,
                clmgs_bitmap_distance=2, clmgs_bitmap_permute=1, clmgs_extract_ffs=1, clmgs_group_by_ffs=3,
                clmgs_selective_adam=1, clmgs_adam_dense=1, clmgs_extract_first_use=1, clmgs_slot_map=1,
                clmgs_rows_gather_i32=1, clmgs_rows_flush_slots=1, clmgs_sh_colors_fwd_slots=1,
                clmgs_sh_colors_bwd_slots=1, clmgs_densification_stats=1, clmgs_rows_accumulate=1,
                clmgs_rows_scatter_grad_i64=1, clmgs_rows_gather_to_slots=1, clmgs_store_i64_to_host=1)


def count_launches(lib):
    """Wrap _Lib.call to count kernel launches of OUR library."""
    counter = dict(n=0, calls=0)
    orig = lib.call

    def call(name, *a):
        k = LAUNCHES.get(name, 0)
        counter["n"] += k
        counter["calls"] += 1 if k else 0
        return orig(name, *a)

    lib.call = call
    from clmgs import _lib, ops
    _lib.call = call
    ops.call = call
    return counter


# ------------------------------------------------------------------------------------------
def make_host_batches(spec, extent, n_batches, bsz):
    from clmgs import scene as S
    V, K, P = S.make_cameras(spec, n_batches * bsz, extent)
    n_img = min(n_batches * bsz, 16)  # distinct GT images; reused cyclically (content does not affect cost)
    gts = S.make_gt_images(spec, n_img)
    return V, K, P, gts


def algorithmic_work(stats, N, B, W, H, sparse_adam):
    """SURVEY.md 8d per-unit figures x the units of this batch (DESIGN.md 'Roofline arithmetic').
    -> {stage: (bound, amount)}; amount in bytes (hbm / host_link / host_dram) or FLOP (fp32)."""
    tiles = ((W + 15) // 16) * ((H + 15) // 16)
    nv, ni = stats["n_visible"], stats["n_isects"]
    n_union = stats.get("n_union", stats.get("n_need", 0))
    n_first = stats.get("n_first", nv[0] if nv else 0)
    out = dict(
        cull=("hbm", 40 * N + 8 * sum(nv) + N * (1 if B <= 8 else 2 if B <= 16 else 4 if B <= 32 else 8)),
        gather_activate=("hbm", sum(88 * n + 8 * n for n in nv)),
        projection_fwd=("hbm", sum((40 + 28) * n for n in nv)),
        sh_fwd=("hbm", sum((192 + 12 + 12) * n for n in nv)),
        isect_sort=("hbm", sum(36 * i + 4 * tiles + 16 * n for n, i in zip(nv, ni))),
        projection_bwd=("hbm", sum((40 + 12 + 20 + 40) * n for n in nv)),
        sh_bwd=("hbm", sum((192 + 384 + 12 + 12 + 24) * n for n in nv)),
        activate_bwd_scatter=("hbm", sum((44 + 40 + 88 + 8) * n for n in nv)),
        loss=("hbm", (3 * 4 + 3 * 1 + 3 * 4) * W * H * B),
        gpu_adam=("hbm", (308 * n_union + N) if sparse_adam else 308 * N),
        rows_load_first=("host_link_h2d", 192 * n_first),
        rows_prefetch=("host_link_h2d", 192 * max(0, n_union - n_first)),
        grad_handoff=("host_link_d2h", 192 * n_union),
        bulk_load=("host_link_h2d", 236 * N),    # naive_offload: all 59 floats per Gaussian, both ways
        bulk_store=("host_link_d2h", 236 * N),
    )
    work = stats.get("raster_work")
    if work:
        out["raster_fwd"] = ("fp32", 27 * work[0])
        out["raster_bwd"] = ("fp32", 75 * work[2])
    return out


def run_ours(a):
    import torch
    import torch.distributed as dist

    from clmgs import _lib, engine, engine_dist, hostmem, model, peaks as PK, scene as S

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- the product has no CPU fallback (use --impl reference)")
    torch.cuda.set_device(local_rank)
    dev = "cuda:%d" % local_rank
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device(dev))
    assert a.gpus == world, "--gpus %d but WORLD_SIZE=%d (launch with torch.distributed.run)" % (a.gpus, world)

    spec = S.CONFIGS[a.config]
    N = a.n_gaussians or spec.n_gaussians
    batch = a.bsz or spec.batch                      # cameras per optimiser step of the named config
    strong = world > 1 and batch % world == 0 and batch // world >= 4 and a.scaling != "weak"
    bsz = batch // world if strong else batch        # cameras per GPU per step
    global_batch = bsz * world
    sparse_adam = a.sparse_adam if a.sparse_adam is not None else (a.config in SPARSE_BY_DEFAULT)
    spec = S.SceneSpec(spec.name, N, spec.width, spec.height, bsz, spec.shape, spec.fovx_deg, spec.index)
    W, H = spec.width, spec.height
    _tile_bits = (((W + 15) // 16) * ((H + 15) // 16)).bit_length() + 1  # tile_n_bits + cam_n_bits (C = 1)
    LAUNCHES["clmgs_isect_emit_sort"] = 1 + 3 * 4 + 3 + 3 * ((_tile_bits + 7) // 8)
    LAUNCHES["clmgs_isect_sorted_offsets"] = LAUNCHES["clmgs_isect_emit_sort"] + 2  # + offsets fill and suffix-min
    # host Adam threads per rank: this rank's share of the cores minus one for the thread that launches kernels
    host_threads = a.cpu_threads or max(1, (os.cpu_count() or 1) // world - 1)
    if world > 1:
        os.environ.setdefault("CLMGS_CPU_ADAM_PIN", str(local_rank * host_threads))  # each rank its own cores
    t0 = time.time()
    # N > 1: one global batch of world x bsz cameras per step, cameras data-parallel, SH rows + host Adam state and
    # the Adam moments of the GPU-resident attributes sharded by Gaussian id (clmgs/engine_dist.py).
    ctx = engine_dist.DistContext(N) if world > 1 else None
    lo, hi = (ctx.lo, ctx.hi) if ctx else (0, N)
    if ctx is not None:
        # the SH parameter table is ONE node-shared host mapping that every rank's GPU reads (clmgs/shm.py); each
        # rank generates and later updates only its own rows
        sh_host = ctx.create_shared_table().tensor[lo:hi]
    else:
        sh_host = hostmem.pinned_tensor((hi - lo, 48), torch.float32)
    g = S.make_gaussians(spec, sh_out=sh_host.numpy(), sh_rows=(lo, hi),
                         workers=max(1, (os.cpu_count() or 1) // world))  # each rank generates only its SH shard
    if a.spatial_order:
        # opt-in, reported separately: the same Gaussians stored in Morton order.  The synthetic SH rows are i.i.d., so
        # only the geometry is permuted (every rank computes the same permutation).
        perm = S.morton_order(g["xyz"], device=dev)
        for key in ("xyz", "log_scale", "raw_quat", "opacity_logit"):
            g[key] = np.ascontiguousarray(g[key][perm])
        del perm
    args = model.TrainArgs(bsz=global_batch, sparse_adam=sparse_adam, cpu_adam_threads=host_threads)
    gm = model.GaussianModelCLMOffload(3, dev).create_from_arrays(g["xyz"], g["log_scale"], g["raw_quat"],
                                                                 g["opacity_logit"], sh_host,
                                                                 grad_table=(a.strategy == "naive"))
    extent = g["extent"]
    del g
    if ctx is not None:
        engine_dist.training_setup_dist(gm, args, ctx)
    else:
        gm.training_setup(args)

    def train_step(cams, timer=None):
        if a.strategy == "naive":
            return engine.naive_offload_train_one_batch(gm, None, cams, gm.parameters_grad_buffer, None, None, comm, gen,
                                                        timer=timer, offload_all=True)  # all 59 floats, as the reference
        if ctx is None:
            return engine.clm_offload_train_one_batch(gm, None, cams, gm.parameters_grad_buffer, None, None, comm, gen,
                                                      timer=timer)
        return engine_dist.clm_offload_train_one_batch_dist(gm, None, cams, gm.parameters_grad_buffer, None, None,
                                                            comm, gen, ctx, timer=timer)

    n_batches = a.warmup + a.steps
    spec_r = S.SceneSpec(spec.name, N, W, H, bsz, spec.shape, spec.fovx_deg, spec.index + 100 * rank)
    V, K, P, gts = make_host_batches(spec_r, extent, n_batches, bsz)
    gts_pinned = torch.from_numpy(gts).pin_memory()
    setup_s = time.time() - t0
    comm = torch.cuda.Stream(priority=-1)
    gen = torch.Generator().manual_seed(1234)
    counter = count_launches(_lib.lib())

    # GT images live in HBM once (16 distinct ones); a resident camera batch is matrices + views of them
    gts_dev = torch.from_numpy(gts).to(dev)

    def cams_resident(b):
        return [model.Camera(V[b * bsz + i], K[b * bsz + i], gts_dev[(b * bsz + i) % len(gts)], dev, uid=b * bsz + i)
                for i in range(bsz)]

    from clmgs import feed as feed_mod
    cam_feed = feed_mod.CameraFeed(V, K, gts_pinned, bsz, dev)

    def step_e2e(b):
        """host camera batch -> device (clmgs.feed: this call waits for batch b's upload and issues the upload of
        batch b+1 from pinned host memory), iteration, losses -> host.  One batch upload per step."""
        cams = cam_feed.get(b)
        losses, order, sp = train_step(cams)
        return torch.stack(losses).cpu()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, first_batch, nsteps):
        barrier()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t_wall = time.time()
        s.record()
        for b in range(first_batch, first_batch + nsteps):
            fn(b)
        e.record()
        barrier()
        ms = s.elapsed_time(e)
        wall = (time.time() - t_wall) * 1e3
        ms = max(ms, wall) if abs(wall - ms) > 0.05 * wall else ms  # the step ends with a host join
        if world > 1:
            t = torch.tensor([ms], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms

    # ---- roofline denominators of THIS box, measured before the timed region ------------------------
    pk = load_peaks()
    if rank == 0 and not a.no_peaks:
        pk.update(PK.measure_all(dev, host_threads))
    barrier()

    # ---- value: camera batches resident in HBM ----------------------------------------------
    resident = [cams_resident(b) for b in range(n_batches)]
    timer = engine.StageTimer(False)

    def step_resident(b):
        train_step(resident[b], timer=timer)

    for b in range(a.warmup):
        step_resident(b)
    clocks = ClockSampler(local_rank)
    clocks.start()
    counter["n"] = 0
    ms_total = timed(step_resident, a.warmup, a.steps)
    launches = counter["n"]
    clk = clocks.stop()
    ms_per_step = ms_total / a.steps
    # strong scaling: an iteration is one optimiser step over the config's batch whatever N is.
    # weak scaling: a global step over world x bsz cameras counts as `world` iterations of `bsz` cameras.
    it_per_step = 1.0 if (strong or world == 1) else float(world)
    value = it_per_step * 1000.0 / ms_per_step

    # ---- per-stage device times (separate pass: event pairs around every stage, raster work counters) -------
    timer = engine.StageTimer(True)
    nprof = min(a.steps, 3)
    host_stats = []

    def step_profiled(b):
        train_step(resident[b], timer=timer)
        st4 = np.zeros(4)
        _lib.call("clmgs_cpu_adam_last_stats", st4)
        host_stats.append(st4)

    timed(step_profiled, a.warmup, nprof)
    # rasteriser work counters (P_f, evaluated, P_b): one more, untimed step with the counting variant of the forward
    # kernel -- it is slower than the product kernel, so it must not run inside the stage-timing pass
    gm.raster_work_counters = torch.zeros(64, 3, dtype=torch.int64, device=dev)
    train_step(resident[a.warmup + nprof - 1])
    barrier()
    stats = dict(gm.last_batch_stats)
    stats["raster_work"] = [int(x) for x in gm.raster_work_counters.sum(0).tolist()]
    gm.raster_work_counters = None
    stage_ms = {k: v / nprof for k, v in timer.totals.items()}
    stage_n = {k: v / nprof for k, v in timer.counts.items()}
    if "isect_count" in stage_ms:   # S4 = tile counts + emission + sort + offsets (SURVEY.md 8d counts them together)
        stage_ms["isect_sort"] = stage_ms.get("isect_sort", 0.0) + stage_ms.pop("isect_count")
        stage_n.pop("isect_count", None)
    hs = np.mean(host_stats, axis=0) if host_stats else np.zeros(4)
    timer = engine.StageTimer(False)

    # ---- optional: pipelined host Adam (deviates from the reference's "everything done on return") -------
    pipelined = None
    if a.pipeline and ctx is None:
        args.pipeline_host_adam = True

        def run_pipe(first, n):
            for b in range(first, first + n):
                train_step(resident[b])
            engine.wait_host_adam(gm)   # the join of the last batch is inside the timed region

        run_pipe(0, a.warmup)
        barrier()
        t0 = time.time()
        run_pipe(a.warmup, a.steps)
        torch.cuda.synchronize()
        ms_pipe = (time.time() - t0) * 1e3 / a.steps
        pipelined = dict(value=round(1000.0 / ms_pipe, 4), unit=UNIT, ms_per_step=round(ms_pipe, 3),
                         note="TrainArgs.pipeline_host_adam=True: the call returns before the host Adam finished and the "
                              "next call overlaps it; wall clock over K steps incl. the final join")
        args.pipeline_host_adam = False

    # ---- optional: lazy dense Adam (same results after materialisation; see model.TrainArgs.lazy_dense_adam) ---
    lazy = None
    if a.lazy_adam and ctx is None and not sparse_adam and a.strategy == "clm":
        args.lazy_dense_adam = True

        def run_lazy(first, n):
            for b in range(first, first + n):
                train_step(resident[b])
            gm.materialize_host_rows()   # every postponed update is applied inside the timed region

        run_lazy(0, a.warmup)
        barrier()
        t0 = time.time()
        run_lazy(a.warmup, a.steps)
        torch.cuda.synchronize()
        ms_lazy = (time.time() - t0) * 1e3 / a.steps
        lazy = dict(value=round(1000.0 / ms_lazy, 4), unit=UNIT, ms_per_step=round(ms_lazy, 3),
                    note="TrainArgs.lazy_dense_adam=True: zero-gradient host updates of untouched rows are postponed and "
                         "replayed bit-identically when a row is next needed; wall clock over K steps incl. one full "
                         "materialisation of the table at the end")
        args.lazy_dense_adam = False

    # ---- e2e: host buffers in, losses out ----------------------------------------------------------
    step_e2e(a.warmup - 1 if a.warmup else 0)
    ms_e2e = timed(step_e2e, a.warmup, a.steps) / a.steps
    h2d = cam_feed.h2d_bytes_per_batch
    d2h = bsz * 4

    out = None
    if rank == 0:
        work = algorithmic_work(stats, N, bsz, W, H, sparse_adam)
        peak_of = dict(hbm=(pk["hbm_gbs"], "GB/s"), host_link_h2d=(pk.get("h2d_gbs"), "GB/s"),
                       host_link_d2h=(pk.get("d2h_gbs"), "GB/s"), fp32=(pk.get("fma_tflops"), "TFLOP/s"),
                       host_dram=(pk.get("host_triad_gbs"), "GB/s"))
        stages = {}
        for k, ms in sorted(stage_ms.items(), key=lambda kv: -kv[1]):
            ent = dict(ms_per_step=round(ms, 3), launches_per_step=round(stage_n.get(k, 0), 1))
            if k in work and ms > 0:
                bound, amount = work[k]
                peak, unit = peak_of[bound]
                ach = amount / (ms * 1e-3) / (1e12 if bound == "fp32" else 1e9)
                ent.update(bound=bound, algorithmic=int(amount), achieved=round(ach, 2), unit=unit)
                if peak:
                    ent.update(peak=peak, frac=round(ach / peak, 4))
            stages[k] = ent
        if "cull" in stages and bsz >= 4:
            stages["cull"]["note"] = ("FP32-issue bound for >= 4 cameras per pass (ncu: SM busy 89 %, DRAM 11 %): the HBM "
                                      "fraction is not this kernel's roof")
        # host Adam (runs on the host threads while the GPU works): busy time = call time - time waiting for flags
        n_upd, n_zero = float(hs[2]), float(hs[3])
        busy_ms = max(1e-6, float(hs[0] - hs[1]))
        adam_bytes = 1344.0 * n_upd - 384.0 * n_zero
        host_adam = dict(ms_per_step=round(float(hs[0]), 2), wait_for_gradients_ms=round(float(hs[1]), 2),
                         busy_ms=round(busy_ms, 2), rows=int(n_upd), bound="host_dram", algorithmic=int(adam_bytes),
                         achieved=round(adam_bytes / busy_ms / 1e6, 2), unit="GB/s", threads=host_threads,
                         trailing_ms=round(stats.get("cpu_adam_trailing_ms", 0.0), 2))
        if pk.get("host_triad_gbs"):
            host_adam.update(peak=pk["host_triad_gbs"], frac=round(host_adam["achieved"] / pk["host_triad_gbs"], 4))
        stages["host_adam"] = host_adam
        # `roofline`: the device stage with the largest measured time of the step, against ITS roof
        cand = [(v["ms_per_step"], k) for k, v in stages.items() if "frac" in v and k != "host_adam"]
        top = max(cand)[1] if cand else None
        roof = None
        if top:
            t = stages[top]
            nl = max(1.0, t["launches_per_step"])
            roof = dict(bound={"hbm": "hbm", "fp32": "fp32_simt"}.get(t["bound"], "host_link"), kernel=top,
                        achieved=t["achieved"], peak=t["peak"], unit=t["unit"], frac=t["frac"], traffic=None,
                        algorithmic_per_launch=int(t["algorithmic"] / nl), launch_ms=round(t["ms_per_step"] / nl, 4),
                        launches_per_step=nl,
                        peak_source=(pk["hbm_source"] if t["bound"] == "hbm" else "measured in this run (clmgs.peaks)"),
                        note="stage chosen by measured device time; in-stream time, i.e. with the copy streams and the "
                             "host Adam running; `traffic` (dram bytes) is not measurable without ncu -- see profiles/")
            hb = [(v["ms_per_step"], k) for k, v in stages.items() if v.get("bound") == "hbm" and "frac" in v]
            if hb:
                k2 = max(hb)[1]
                roof["dominant_hbm_stage"] = dict(kernel=k2, achieved=stages[k2]["achieved"], peak=stages[k2]["peak"],
                                                  frac=stages[k2]["frac"], ms_per_step=stages[k2]["ms_per_step"])
        raster = dict(fwd_ms_per_camera=round(stage_ms.get("raster_fwd", 0.0) / max(1, bsz), 4),
                      bwd_ms_per_camera=round(stage_ms.get("raster_bwd", 0.0) / max(1, bsz), 4),
                      examined_per_step=stats["raster_work"][0], evaluated_after_pruning_per_step=stats["raster_work"][1],
                      backward_walk_per_step=stats["raster_work"][2])
        out = dict(metric=METRIC, value=round(value, 4), unit=UNIT, n_gpus=world, steps=a.steps, warmup=a.warmup,
                   ms_per_step=round(ms_per_step, 3), higher_is_better=True,
                   scaling="strong" if strong else "weak", vs_baseline=None, dtype="f32", data="synthetic",
                   images_per_sec=round(global_batch * 1000.0 / ms_per_step, 2),
                   config=dict(workload="%s: %d Gaussians, %dx%d, batch %d cameras/iteration, SH degree 3, "
                               "%s_offload (48-float SH rows + Adam state in pinned host memory), %s Adam"
                               % (spec.name, N, W, H, batch, a.strategy, "sparse" if sparse_adam else "dense"),
                               parallelism=("dp%d over cameras, %s: %d cameras per GPU, global batch %d per optimiser step%s; SH "
                                            "table node-shared in host memory (every GPU reads its rows over its own host "
                                            "link); host Adam + the Adam moments of the GPU-resident attributes sharded by "
                                            "Gaussian id; SH gradients: accumulate into the owner's HBM over NVLink (peer "
                                            "memory, chunked); GPU-resident attributes: NCCL reduce-scatter + all-gather"
                                            % (world, "strong scaling" if strong else "weak scaling", bsz, global_batch,
                                               "" if strong else " (counted as %d iterations of %d)" % (world, bsz)))
                               if world > 1 else "1 GPU",
                               l2="inputs larger than L2 (Gaussian tables %.0f MB >> 126 MB); no explicit flush" % (N * 44 / 1e6),
                               sparse_adam=bool(sparse_adam), cpu_adam_threads_per_rank=host_threads,
                               host_cores=os.cpu_count(),
                               visible_per_camera=_summ(stats["n_visible"]), isects_per_camera=_summ(stats["n_isects"]),
                               visible_fraction=round(float(np.mean(stats["n_visible"])) / N, 5),
                               touched_rows_per_step=stats.get("n_union", stats.get("n_need")),
                               setup_s=round(setup_s, 1)),
                   e2e=dict(value=round(it_per_step * 1000.0 / ms_e2e, 4), unit=UNIT,
                            h2d_bytes_per_step=h2d, d2h_bytes_per_step=d2h, ms_per_step=round(ms_e2e, 3)),
                   gpu_launches=launches, clocks=clk, roofline=roof, raster=raster, stages=stages,
                   peaks={k: v for k, v in pk.items()})
        out["config"]["hbm_allocated_gb"] = round(torch.cuda.max_memory_allocated() / 1e9, 2)
        if "grad_exchange" in stats:
            out["config"]["grad_exchange"] = stats["grad_exchange"]
            out["config"]["grad_chunks"] = stats.get("grad_chunks")
        out["config"]["host_huge_pages"] = hostmem.huge_page_stats()
        if a.spatial_order:
            out["config"]["spatial_order"] = "Gaussians stored in Morton order (opt-in; default is the generator's i.i.d. order)"
        if pipelined:
            out["pipelined"] = pipelined
        if lazy:
            out["lazy_dense_adam"] = lazy
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    # free the big tables before anything else runs on this box (child processes, the CPU baseline)
    del gm, resident, cam_feed, gts_dev, sh_host, gts_pinned, train_step, step_resident, step_e2e, step_profiled
    import gc
    gc.collect()
    torch.cuda.empty_cache()
    try:
        torch._C._host_emptyCache()
    except Exception:
        pass
    if out is not None and world == 1:
        if not a.no_cpu_baseline:
            out["cpu_baseline"] = cpu_baseline(a, S.CONFIGS[a.config], N, batch, sparse_adam)
        if a.other_configs:
            out["other_configs"] = run_other_configs(a)
    return out


def _summ(xs):
    xs = list(xs)
    if len(xs) <= 8:
        return xs
    return dict(n=len(xs), min=int(min(xs)), mean=int(np.mean(xs)), max=int(max(xs)))


def run_other_configs(a):
    """configs[1] (bicycle 6 M) and configs[2] (rubble-4K 28 M: clm_offload vs naive_offload) in child processes, a few
    steps each, so that the driver's default run also carries them.  Same code path, same JSON contract per child."""
    legs = [("c2_bicycle_6m", "clm"), ("c3_rubble4k_28m", "clm"), ("c3_rubble4k_28m", "naive")]
    res = {}
    for cfg, strat in legs:
        cmd = [sys.executable, os.path.abspath(__file__), "--config", cfg, "--strategy", strat, "--steps", str(a.other_steps),
               "--warmup", "3", "--no-cpu-baseline", "--no-other-configs", "--no-peaks"]
        key = cfg + ("" if strat == "clm" else "_" + strat)
        try:
            r = subprocess.run(cmd, capture_output=True, text=True, timeout=900)
            line = [ln for ln in r.stdout.splitlines() if ln.startswith("{")]
            if r.returncode != 0 or not line:
                res[key] = dict(error=(r.stderr or r.stdout)[-400:])
                continue
            d = json.loads(line[-1])
            res[key] = dict(value=d["value"], unit=d["unit"], ms_per_step=d["ms_per_step"], e2e=d["e2e"]["value"],
                            steps=d["steps"], warmup=d["warmup"], workload=d["config"]["workload"],
                            visible_fraction=d["config"]["visible_fraction"],
                            touched_rows_per_step=d["config"]["touched_rows_per_step"], clocks=d["clocks"],
                            raster=d.get("raster"),
                            stages={k: {kk: v[kk] for kk in ("ms_per_step", "frac", "bound") if kk in v}
                                    for k, v in d["stages"].items()})
        except Exception as e:  # a child that fails must not take the headline with it
            res[key] = dict(error=str(e)[-400:])
    c, n = res.get("c3_rubble4k_28m", {}), res.get("c3_rubble4k_28m_naive", {})
    if "value" in c and "value" in n:
        res["c3_clm_vs_naive"] = round(c["value"] / n["value"], 3)
    return res


# ------------------------------------------------------------------------------------------
def cpu_baseline(a, spec, N, bsz, sparse_adam):
    """The reference's CPU path restated (oracle port) on a bounded sample, all host threads.
    Sample: ONE camera of the batch at full resolution over the first n Gaussians of the scene (ids are
    i.i.d., so a prefix is a uniform subsample), measured at n = n_s/2 and n = n_s; the per-camera time is
    extrapolated with the linear fit t(n) = t_fixed + t_per_gaussian * n (the loss / per-pixel work does not
    grow with n) to the full N and multiplied by the cameras per iteration."""
    cores = os.cpu_count() or 1
    os.environ["OMP_NUM_THREADS"] = str(cores)   # torchrun exports OMP_NUM_THREADS=1; set before libgomp loads
    import oracle  # noqa: F401
    from oracle import train_step as TS
    from clmgs import model, scene as S

    n_s = min(N, a.cpu_sample_gaussians)
    W, H = spec.width, spec.height
    sspec = S.SceneSpec(spec.name, N, W, H, 1, spec.shape, spec.fovx_deg, spec.index)
    args = model.TrainArgs(bsz=1, sparse_adam=sparse_adam)
    gts = S.make_gt_images(sspec, 1)
    times = {}
    for n in sorted({max(1000, n_s // 2), n_s}):
        g = S.make_gaussians(sspec, n=n)
        V, K, P = S.make_cameras(sspec, 1, g["extent"])
        st = TS.new_state(g)
        t0 = time.time()
        TS.train_one_batch(st, V, K, P, gts, W, H, args, [0])
        times[n] = time.time() - t0
    ns = sorted(times)
    if len(ns) == 2 and times[ns[1]] > times[ns[0]]:
        per_g = (times[ns[1]] - times[ns[0]]) / float(ns[1] - ns[0])
        fixed = max(0.0, times[ns[0]] - per_g * ns[0])
    else:
        per_g, fixed = times[ns[-1]] / float(ns[-1]), 0.0
    full_s = bsz * (fixed + per_g * N)
    return dict(value=round(1.0 / full_s, 6), unit=UNIT, cores=cores, kind="port",
                sample="oracle port (C restatement of gsplat kernels + cpu-adam; OpenMP on the parallel loops, %d threads), 1 "
                       "of %d cameras at %dx%d over the first n Gaussians of %d: %s s; fit t = %.2f s + %.3g s/Gaussian, "
                       "extrapolated to N and x%d cameras = %.0f s/iteration"
                       % (cores, bsz, W, H, N, ", ".join("n=%d: %.2f" % (n, times[n]) for n in ns), fixed, per_g, bsz, full_s),
                sample_seconds=round(sum(times.values()), 3))


def run_reference(a):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return None
    from clmgs import scene as S
    spec = S.CONFIGS[a.config]
    N = a.n_gaussians or spec.n_gaussians
    bsz = a.bsz or spec.batch
    sparse_adam = a.sparse_adam if a.sparse_adam is not None else (a.config in SPARSE_BY_DEFAULT)
    # warm-up + steps, each a bounded sample
    vals = []
    for i in range(a.warmup + a.steps):
        cb = cpu_baseline(a, spec, N, bsz, sparse_adam)
        if i >= a.warmup:
            vals.append(cb)
    v = float(np.median([c["value"] for c in vals]))
    cb = vals[-1]
    cb["value"] = v
    return dict(impl="reference", metric=METRIC, value=v, unit=UNIT, n_gpus=a.gpus, steps=a.steps, warmup=a.warmup,
                ms_per_step=round(1000.0 / v, 1), higher_is_better=True,
                scaling="strong" if (a.gpus > 1 and bsz % a.gpus == 0 and bsz // a.gpus >= 4) else "weak",
                vs_baseline=None, dtype="f32", data="synthetic",
                config=dict(workload="%s: %d Gaussians, %dx%d, batch %d cameras/iteration, SH degree 3, "
                                     "clm_offload (48-float SH rows + Adam state in pinned host memory), %s Adam"
                                     % (spec.name, N, spec.width, spec.height, bsz, "sparse" if sparse_adam else "dense"),
                            note="reference arm: CPU port on the host cores, bounded sample extrapolated to this workload "
                                 "(see cpu_baseline.sample)"),
                cpu_baseline=cb, e2e=dict(value=v, unit=UNIT, h2d_bytes_per_step=0, d2h_bytes_per_step=0))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=8)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", default=DEFAULT_CONFIG)
    ap.add_argument("--n-gaussians", type=int, default=0, help="override the config's Gaussian count (debug)")
    ap.add_argument("--bsz", type=int, default=0)
    ap.add_argument("--scaling", default="auto", choices=["auto", "weak"],
                    help="auto: fixed global batch split over the GPUs where >= 4 cameras per GPU remain (strong scaling, "
                         "SURVEY 8d); weak: the config's batch PER GPU, one optimiser step over N x batch cameras")
    ap.add_argument("--sparse-adam", dest="sparse_adam", action="store_true", default=None)
    ap.add_argument("--dense-adam", dest="sparse_adam", action="store_false")
    ap.add_argument("--cpu-threads", type=int, default=0)
    ap.add_argument("--strategy", default="clm", choices=["clm", "naive"],
                    help="naive = bulk-transfer baseline built from the same kernels (BASELINE.json configs[2] comparison)")
    ap.add_argument("--lazy-adam", action="store_true",
                    help="opt-in (dense Adam): postpone zero-gradient host updates of untouched rows; the timed region "
                         "ends with a full materialisation so that no update is left out of it")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-other-configs", dest="other_configs", action="store_false", default=None)
    ap.add_argument("--other-steps", type=int, default=6)
    ap.add_argument("--no-peaks", action="store_true", help="skip measuring host-link / FMA / host-DRAM peaks")
    ap.add_argument("--pipeline", action="store_true",
                    help="opt-in engine mode: defer the host-Adam join to the next call (reported in the extra "
                         "`pipelined` object; the headline keeps the reference's call contract)")
    ap.add_argument("--cpu-sample-gaussians", type=int, default=8_000_000)
    ap.add_argument("--spatial-order", action="store_true",
                    help="store the scene's Gaussians in Morton order (opt-in; changes the memory access pattern of the "
                         "SH row transfers, not the training math; noted in config)")
    a = ap.parse_args()
    a.warmup = max(a.warmup, 0)
    if a.other_configs is None:   # the default run (north-star config, 1 GPU) also carries configs[1] and configs[2]
        a.other_configs = (a.config == DEFAULT_CONFIG and a.gpus == 1 and a.strategy == "clm" and not a.n_gaussians
                           and a.impl == "ours")
    out = run_reference(a) if a.impl == "reference" else run_ours(a)
    if out is not None:
        print(json.dumps(out))


if __name__ == "__main__":
    main()
