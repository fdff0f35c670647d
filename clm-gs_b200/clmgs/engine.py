"""Host orchestration of one CLM-offload training batch on B200 kernels.

Keeps the contract of ``clm_offload_train_one_batch`` (strategies/clm_offload/engine.py:338-347,925):
same arguments, returns ``(losses, ordered_cams, sparsity)``, same side effects on the model
(GPU-resident attributes + their Adam state, pinned SH table + host Adam state, pinned gradient
table left zeroed).  What changed underneath:

  reference                                             here
  --------------------------------------------------   ----------------------------------------------
  packed projection x2 + cumsum + unique_consecutive   clmgs_cull_count/_emit: one projection pass for all
                                                       cameras, bitmap + ballot compaction
  bsz scatter_to_bit + randperm + (B,B,n) XOR tensor   bitmap comes from the cull; sampled distance kernel
  torch.sort(N bytes) + unique_consecutive + split      counting sort by last-use key (group_by_ffs)
  ~8 torch ops + 2 nonzero_static per camera,           batch-resident SH rows: one device row ("slot") per touched
  retained rows re-packed for every camera              Gaussian, loaded at first use, flushed after last use
  4 gathers + 3 activations + autograd                  clmgs_gather_activate / clmgs_activate_bwd_scatter
  autograd graph per camera                             explicit forward + backward kernel calls
  fused_ssim + torch.compile'd L1                       clmgs_loss_fwd_bwd (writes v_render_colors directly)

Streams: compute = torch's current stream, SH-row traffic on ``comm_stream`` with events, host Adam
on a python thread that drops the GIL inside the C call -- the reference's three-way overlap
(engine.py:423-437, 464-641, 767-826).
"""
import math
import os
import threading

import numpy as np
import torch

from . import hostmem, ops

TILE_SIZE = 16  # strategies/base_engine.py:15


_NVTX = os.environ.get("CLMGS_NVTX", "1") != "0"


class nvtx_range:
    """NVTX range around a host-side phase (the reference marks its phases the same way, engine.py:149-853, so an
    nsys timeline of this engine reads like one of the reference).  CLMGS_NVTX=0 turns them off."""

    def __init__(self, name):
        self.name = name

    def __enter__(self):
        if _NVTX:
            torch.cuda.nvtx.range_push(self.name)
        return self

    def __exit__(self, *a):
        if _NVTX:
            torch.cuda.nvtx.range_pop()


class StageTimer:
    """CUDA-event stage timing on the stream each stage is launched on; every stage is also an
    NVTX range."""

    def __init__(self, enabled=False):
        self.enabled = enabled
        self.pairs = []
        self.totals = {}
        self.counts = {}
        self.base = None
        self.timeline = []   # (name, start_ms, end_ms) relative to the first recorded event of the batch

    class _Ctx:
        def __init__(self, owner, name, stream):
            self.o, self.name, self.stream = owner, name, stream

        def __enter__(self):
            if _NVTX:
                torch.cuda.nvtx.range_push(self.name)
            if self.o.enabled:
                self.s = torch.cuda.Event(enable_timing=True)
                self.e = torch.cuda.Event(enable_timing=True)
                self.s.record(self.stream or torch.cuda.current_stream())
            return self

        def __exit__(self, *a):
            if _NVTX:
                torch.cuda.nvtx.range_pop()
            if self.o.enabled:
                self.e.record(self.stream or torch.cuda.current_stream())
                self.o.pairs.append((self.name, self.s, self.e))

    def __call__(self, name, stream=None):
        return StageTimer._Ctx(self, name, stream)

    def collect(self):
        """Call after a device synchronize."""
        self.timeline = []
        base = self.pairs[0][1] if self.pairs else None
        for name, s, e in self.pairs:
            self.totals[name] = self.totals.get(name, 0.0) + s.elapsed_time(e)
            self.counts[name] = self.counts.get(name, 0) + 1
            self.timeline.append((name, base.elapsed_time(s), base.elapsed_time(e)))
        self.pairs = []


_NULL_TIMER = StageTimer(False)


def calculate_filters(batched_cameras, gaussians, timer=_NULL_TIMER):
    """strategies/base_engine.py:18-76.  Returns (filters, cull) where filters[c] is the ascending
    int64 id list of camera c (views into one flat tensor) and cull carries the visibility bitmap."""
    args = gaussians.args
    viewmats = torch.stack([c.viewmat for c in batched_cameras])
    Ks = torch.stack([c.create_k_on_gpu() for c in batched_cameras])
    W, H = batched_cameras[0].image_width, batched_cameras[0].image_height
    with timer("cull"):
        cull = ops.cull(gaussians._xyz, gaussians._scaling, gaussians._rotation, viewmats, Ks, W, H,
                        activated=False, radius_clip=args.radius_clip)
    assert all(c > 0 for c in cull["counts"]), \
        "every camera must see at least one Gaussian (strategies/base_engine.py:65-67)"
    filters = list(torch.split(cull["gaussian_ids"], cull["counts"]))
    return filters, cull


_PINNED_ROWS = {}


def _pinned_rows_buffer(n, parity=0):
    """Pinned int32 [n] staging buffer for the host-Adam row lists, allocated once (cudaHostAlloc of
    tens of MB costs milliseconds and was on the critical path of every batch).  Two of them, used
    alternately: with pipeline_host_adam the previous batch's host Adam may still be reading its list."""
    key = (n, parity)
    buf = _PINNED_ROWS.get(key)
    if buf is None:
        for k in [k for k in _PINNED_ROWS if k[0] != n]:
            del _PINNED_ROWS[k]
        buf = torch.empty(n, dtype=torch.int32).pin_memory()
        _PINNED_ROWS[key] = buf
    return buf


def _packed_staging(gaussians, n_rows, which=0, cap=None):
    """Pinned [rows, 48] staging buffer the gradient flush DMAs into and the host Adam reads from; kept across
    batches and grown with 25 % headroom (pinning gigabytes takes seconds: it must not happen in steady state), never
    beyond `cap` rows (two of them only in pipelined mode, where the previous batch's host Adam may still be reading
    its own)."""
    bufs = getattr(gaussians, "_packed_grad_staging", None)
    if bufs is None:
        bufs = gaussians._packed_grad_staging = {}
    buf = bufs.get(which)
    if buf is None or buf.shape[0] < n_rows:
        bufs[which] = None
        want = ops._bucket(max(int(n_rows * 1.25), 1))
        if cap is not None:
            want = max(n_rows, min(want, int(cap)))
        buf = bufs[which] = hostmem.pinned_tensor((want, 48), torch.float32)
    return buf


def wait_host_adam(gaussians):
    """Join a host Adam left running by a pipeline_host_adam call (no-op otherwise).  Call it before
    reading the SH table or the host optimiser state."""
    pend = getattr(gaussians, "_pending_adam", None)
    if pend is not None:
        pend["worker"].join()
        gaussians._pending_adam = None
        torch.cuda.synchronize()
        if pend["err"]:
            raise pend["err"][0]


_SIDE_STREAMS = {}


def _side_stream(dev):
    """One extra stream per device for copies nobody on the compute stream waits for."""
    key = torch.device(dev).index if torch.device(dev).index is not None else torch.cuda.current_device()
    st = _SIDE_STREAMS.get(key)
    if st is None:
        st = _SIDE_STREAMS[key] = torch.cuda.Stream(device=dev)
    return st


def order_calculation(filters, cull, n_gaussians, bsz, perm_generator, args, timer=_NULL_TIMER, parity=0,
                      host_rows=True):
    """engine.py:135-298: camera order (TSP over sampled visibility distance), ordered bitmap,
    per-Gaussian last-use groups for the host Adam.  host_rows=False (multi-GPU engine: the host Adam's row lists
    come from the gradient exchange) skips the last-use grouping and its N x 4-byte device->host copy.
    perm_generator: a CPU or a CUDA torch.Generator (the reference passes a CUDA one, train.py:191), or None."""
    bitmap = cull["bitmap"]
    dev = bitmap.device
    with timer("order"):
        n_sampled = max(1, n_gaussians // (bsz * bsz) if bsz >= 32 else n_gaussians // 32)
        stride = max(1, n_gaussians // n_sampled)
        off = 0
        if perm_generator is not None:  # drawn on the generator's own device
            off = int(torch.randint(0, stride, (1,), generator=perm_generator, device=perm_generator.device).item())
        sample_ids = torch.arange(n_sampled, device=dev, dtype=torch.int64) * stride + off
        sample_ids.clamp_(max=n_gaussians - 1)
        dist = ops.bitmap_distance(bitmap, bsz, sample_ids).cpu().numpy()  # D2H sync (reference: .tolist())
        ordered_cams = ops.tsp_tour(dist) if bsz > 1 else [0]
        if args.reorder_by_min_sparsity_at_end and bsz > 1:
            lens = [filters[c].shape[0] for c in ordered_cams]
            mi = bsz - 1
            for k in range(bsz - 1):
                if lens[k] < lens[mi]:
                    mi = k
            ordered_cams = ordered_cams[mi + 1:] + ordered_cams[:mi + 1]
        filters = [filters[i] for i in ordered_cams]
        sparsity = [f.shape[0] / float(n_gaussians) for f in filters]
        ordered_bitmap = ops.bitmap_permute(bitmap, bsz, ordered_cams)
        ffs = ops.extract_ffs(ordered_bitmap)
        rows_dev = counts = rows_host = rows_ready = None
        if host_rows:
            rows_dev, counts = ops.group_by_ffs(ffs, bsz)
            rows_host = _pinned_rows_buffer(n_gaussians, parity)
            # N x 4 bytes to the host (7 ms at 102 M): on a side stream, so that neither the read-backs below nor the
            # first cameras wait for it; its readers (the host Adam thread first of all) wait on rows_ready
            side = _side_stream(dev)
            grouped = torch.cuda.Event()
            grouped.record()
            side.wait_event(grouped)
            with torch.cuda.stream(side):
                rows_host.copy_(rows_dev, non_blocking=True)
                rows_ready = torch.cuda.Event()
                rows_ready.record(side)
            rows_dev.record_stream(side)
        # batch-resident rows: Gaussians grouped by the first camera that needs them; a Gaussian's slot in the
        # device row table is its position in that grouping (so each first-use group is a contiguous slot range)
        first = ops.extract_first_use(ordered_bitmap, bsz)
        first_rows_dev, first_counts = ops.group_by_ffs(first, bsz)
        counts_h = counts.cpu().numpy() if host_rows else None
        first_counts_h = first_counts.cpu().numpy()
        n_untouched = int(first_counts_h[0])
        union_ids = first_rows_dev[n_untouched:]
        # A Gaussian's slot in the batch-resident device tables = its position in the LAST-use grouping: the rows
        # whose last user is camera k are one contiguous slot range, so their gradients leave with one plain DMA copy
        # (copy engine, no SMs) into a packed pinned staging buffer laid out like the host Adam's row list.  Loads
        # scatter into the slots (they are zero-copy gather kernels either way).  Without the last-use grouping
        # (multi-GPU engine, which makes its own slots) fall back to the first-use order.
        slot_of = torch.empty(n_gaussians, dtype=torch.int32, device=dev)
        ops.slot_map(rows_dev[int(counts_h[0]):] if host_rows else union_ids, slot_of)
    row_begin = row_end = None
    if host_rows:
        offs = np.concatenate([[0], np.cumsum(counts_h)]).astype(np.int64)
        # group m of the host Adam = rows whose last user is camera m-1 <-> ffs key bsz+1-m (engine.py:213)
        keys = [0] + [bsz + 1 - m for m in range(1, bsz + 1)]
        row_begin = np.ascontiguousarray(offs[keys], dtype=np.int64)
        row_end = np.ascontiguousarray(offs[[k + 1 for k in keys]], dtype=np.int64)
    first_edges = np.concatenate([[0], np.cumsum(first_counts_h[1:])]).astype(np.int64)  # slots of first-use group k
    return dict(ordered_cams=ordered_cams, filters=filters, sparsity=sparsity, ordered_bitmap=ordered_bitmap,
                visibility=ffs, rows_host=rows_host, rows_ready=rows_ready, rows_dev=rows_dev, row_begin=row_begin,
                row_end=row_end,
                union_ids=union_ids, slot_of=slot_of, first_edges=first_edges)


def geometry_begin(gaussians, camera, ids, timer=_NULL_TIMER):
    """First part of a camera's geometry (engine.py:51-63,85-97): gather + activations, projection, tile counts.
    Nothing here waits for the device; the tile total travels to the host behind an event."""
    W, H = camera.image_width, camera.image_height
    tw, th = math.ceil(W / float(TILE_SIZE)), math.ceil(H / float(TILE_SIZE))
    with timer("gather_activate"):
        means, scales, quats, opac = ops.gather_activate(gaussians._xyz, gaussians._scaling, gaussians._rotation,
                                                         gaussians._opacity, ids)
    with timer("projection_fwd"):
        radii, means2d, depths, conics, _ = ops.projection_fwd(means, quats, scales, camera.viewmat[None],
                                                               camera.K[None], W, H)
    with timer("isect_count"):
        cnt = ops.isect_count_async(means2d, radii, TILE_SIZE, tw, th)
    return dict(means=means, scales=scales, quats=quats, opac=opac, radii=radii, means2d=means2d, conics=conics,
                depths=depths, count_state=cnt, tw=tw, th=th)


def geometry_finish(fw, timer=_NULL_TIMER):
    """Second part (engine.py:89-100): emission + tile sort + offsets, sized by the total that geometry_begin sent to
    the host (it has normally arrived long before: this is issued at least one camera later)."""
    if "flatten_ids" in fw:
        return fw
    with timer("isect_sort"):
        flatten_ids, offsets, n_isects = ops.isect_sort_counted(fw.pop("count_state"), fw.pop("depths"), TILE_SIZE,
                                                                fw["tw"], fw["th"])
    fw.update(offsets=offsets, flatten_ids=flatten_ids, n_isects=n_isects)
    return fw


def geometry_one_camera(gaussians, camera, ids, timer=_NULL_TIMER):
    """The part of the forward that does not touch the offloaded SH rows (engine.py:51-63,85-100):
    gather + activations, projection, tile intersection + sort.  Runs while SH rows are still in
    flight on the copy stream."""
    return geometry_finish(geometry_begin(gaussians, camera, ids, timer), timer)


def shade_one_camera(gaussians, camera, fw, shs, background, timer=_NULL_TIMER, rows_ready=None, ids=None,
                     slot_of=None):
    """SH colour + rasterisation (engine.py:67-78,106-119); needs the SH rows.  With `slot_of`, `shs` is
    the batch-resident row table and row g is slot_of[ids[g]]."""
    W, H = camera.image_width, camera.image_height
    if rows_ready is not None:  # the SH rows arrive on the copy stream
        torch.cuda.current_stream().wait_event(rows_ready)
    with timer("sh_fwd"):
        if slot_of is not None:
            fw["colors"] = ops.sh_colors_fwd_slots(gaussians.active_sh_degree, fw["means"], camera.campos, shs, ids,
                                                   slot_of)
        else:
            fw["colors"] = ops.sh_colors_fwd(gaussians.active_sh_degree, fw["means"], camera.campos, shs)
    fw["bg"] = None if background is None else background.reshape(1, 3)
    with timer("raster_fwd"):
        fw["render"], fw["alphas"], fw["last"] = ops.rasterize_fwd(
            fw["means2d"], fw["conics"], fw["colors"][None], fw["opac"][None], W, H, TILE_SIZE, fw["offsets"],
            fw["flatten_ids"], fw["bg"], work_counters=getattr(gaussians, "raster_work_counters", None))
    return fw


def forward_one_camera(gaussians, camera, ids, shs, background, timer=_NULL_TIMER, rows_ready=None):
    """pipeline_forward_one_step_shs_inplace (engine.py:30-127) without autograd.  Returns a dict
    with everything the backward needs."""
    fw = geometry_one_camera(gaussians, camera, ids, timer)
    return shade_one_camera(gaussians, camera, fw, shs, background, timer, rows_ready)


def backward_one_camera(gaussians, camera, ids, shs, shs_grad, fw, v_render, grads, timer=_NULL_TIMER, slot_of=None):
    """loss.backward() + spherical_harmonics_bwd_inplace + scatter_add_ (engine.py:703-742)."""
    W, H = camera.image_width, camera.image_height
    with timer("raster_bwd"):
        v_means2d, v_conics, v_colors, v_opac, _ = ops.rasterize_bwd(
            fw["means2d"], fw["conics"], fw["colors"][None], fw["opac"][None], W, H, TILE_SIZE, fw["offsets"],
            fw["flatten_ids"], fw["alphas"], fw["last"], v_render, None, fw["bg"])
    with timer("projection_bwd"):
        v_means, v_quats, v_scales = ops.projection_bwd(fw["means"], fw["quats"], fw["scales"], camera.viewmat[None],
                                                        camera.K[None], W, H, 0.3, fw["radii"], fw["conics"],
                                                        v_means2d, None, v_conics)
    with timer("sh_bwd"):
        if slot_of is not None:
            ops.sh_colors_bwd_slots(gaussians.active_sh_degree, fw["means"], camera.campos, shs, v_colors[0], shs_grad,
                                    ids, slot_of, v_means)
        else:
            ops.sh_colors_bwd(gaussians.active_sh_degree, fw["means"], camera.campos, shs, v_colors[0], shs_grad,
                              v_means)
    with timer("activate_bwd_scatter"):
        ops.activate_bwd_scatter(gaussians._scaling, gaussians._rotation, gaussians._opacity, ids, v_means, v_scales,
                                 v_quats, v_opac[0], grads["xyz"], grads["scaling"], grads["rotation"],
                                 grads["opacity"])
    return v_means2d


def clm_offload_train_one_batch(gaussians, scene, batched_cameras, parameters_grad_buffer, background, pipe_args,
                                comm_stream, perm_generator, timer=None):
    """One optimiser step over ``len(batched_cameras)`` cameras.  See the module docstring."""
    timer = timer or _NULL_TIMER
    args = gaussians.args
    bsz = len(batched_cameras)
    N = gaussians.n_gaussians
    dev = gaussians._xyz.device
    compute = torch.cuda.current_stream()
    h_params = gaussians._parameters
    h_grad = parameters_grad_buffer[:N]

    pipe = bool(getattr(args, "pipeline_host_adam", False))
    prev = getattr(gaussians, "_pending_adam", None)
    if prev is not None and not pipe:
        wait_host_adam(gaussians)
        prev = None
    parity = gaussians._batch_parity = 1 - getattr(gaussians, "_batch_parity", 0)

    # ---- stage 1: visibility + order ------------------------------------------------------------
    with torch.no_grad():
        with nvtx_range("calculate_filters"):        # engine.py:376
            filters, cull = calculate_filters(batched_cameras, gaussians, timer)
        with nvtx_range("sort cameras"):             # engine.py:389
            oc = order_calculation(filters, cull, N, bsz, perm_generator, args, timer, parity)
    ordered_cams, filters, sparsity = oc["ordered_cams"], oc["filters"], oc["sparsity"]
    cams = [batched_cameras[i] for i in ordered_cams]
    bitmap = oc["ordered_bitmap"]

    # Lazy dense Adam (opt-in): rows no camera touches skip this step's zero-gradient update; the rows this batch
    # does touch first replay whatever they skipped earlier -- before their coefficients are read by the GPU.
    lazy = bool(getattr(args, "lazy_dense_adam", False)) and not args.sparse_adam
    if not lazy and gaussians.optimizer.cpu_adam.lazy_pending:
        # the flag was switched off with postponed updates outstanding: replay them before this step's update
        gaussians.optimizer.cpu_adam.catch_up(None)
    if lazy:
        assert not pipe, "lazy_dense_adam and pipeline_host_adam are not combined"
        ca = gaussians.optimizer.cpu_adam
        if not ca.lazy_pending:
            ca.last_step[:N] = ca.step_count
        oc["rows_ready"].synchronize()
        ca.catch_up(oc["rows_host"][int(oc["row_end"][0]):N], include_this_step=True)
        ca.lazy_pending = True

    # ---- stage 2: host Adam thread --------------------------------------------------------------
    # flags 0..bsz-2: cameras 0..bsz-2 done; flags bsz-1..: pieces of the last camera's flush
    n_last = filters[-1].shape[0]
    J = max(1, min(int(getattr(args, "final_flush_chunks", 1)), max(1, n_last // 65536)))
    n_flags = bsz - 1 + J
    ring = getattr(gaussians, "_signal_ring", None)
    if ring is None:
        ring = gaussians._signal_ring = [None, None]
    if ring[parity] is None or ring[parity].shape[0] != n_flags:
        ring[parity] = torch.zeros(n_flags, dtype=torch.int32).pin_memory()
    else:
        ring[parity].zero_()  # its last user (two batches ago) has joined: the previous worker joined it
    signal = gaussians.signal_tensor_pinned = ring[parity]
    cpu_adam = gaussians.optimizer.cpu_adam
    worker_err = []

    worker_times = {}
    # host-Adam segments (see clmgs.h): group 0 = untouched rows (skipped when sparse_adam), group m = rows
    # last used by camera m-1; the last group is cut into J pieces that match the J pieces of the final flush
    # (both lists are the last camera's ascending ids).
    rb, re_ = oc["row_begin"], oc["row_end"]
    assert int(re_[bsz] - rb[bsz]) == n_last, "last-use group of the final camera must equal its filter"
    chunk_edges = [n_last * j // J for j in range(J + 1)]
    seg_begin, seg_end, seg_signal = [], [], []
    if not args.sparse_adam and not lazy:
        # rows no camera touched: dense Adam still decays their moments; their gradient rows are zero by the
        # table's invariant, so the update skips reading / re-zeroing them (seg_zero_grad)
        seg_begin.append(int(rb[0])); seg_end.append(int(re_[0])); seg_signal.append(-1)
    for m in range(1, bsz):
        seg_begin.append(int(rb[m])); seg_end.append(int(re_[m])); seg_signal.append(m - 1)
    for j in range(J):
        seg_begin.append(int(rb[bsz]) + chunk_edges[j]); seg_end.append(int(rb[bsz]) + chunk_edges[j + 1])
        seg_signal.append(bsz - 1 + j)

    progress = np.zeros(1, np.int64)
    prev_worker = prev["worker"] if prev is not None else None
    # packed pinned staging for this batch's SH gradients: row s = slot s = list entry n_unt + s (see order_calculation)
    n_unt = int(re_[0])
    h_packed = _packed_staging(gaussians, N - n_unt, parity if pipe else 0, cap=N)

    def cpuadam_thread():
        import time as _t
        if prev_worker is not None:
            prev_worker.join()  # one host Adam at a time, in batch order
        t0 = _t.perf_counter()
        try:
            if prev is not None and prev["err"]:
                raise prev["err"][0]   # the previous batch's host update failed: do not update on top of it
            oc["rows_ready"].synchronize()   # the row list is in pinned memory
            with nvtx_range("cpu_adam.sparse_step()"):  # engine.py:319
                cpu_adam.batched_sparse_step(h_grad, oc["rows_host"], seg_begin, seg_end, seg_signal, signal, 1.0 / bsz,
                                             progress, [1 if (s_ == -1 and not args.sparse_adam and i_ == 0) else 0
                                                        for i_, s_ in enumerate(seg_signal)],
                                             grad_packed=h_packed, packed_base=n_unt)
        except Exception as e:  # surfaced after join
            worker_err.append(e)
        worker_times["cpu_adam_thread_ms"] = (_t.perf_counter() - t0) * 1e3

    worker = threading.Thread(target=cpuadam_thread, daemon=True)
    worker.start()
    try:
        out = _train_cameras(gaussians, batched_cameras, oc, cull, filters, h_params, h_packed, background, comm_stream,
                             signal, timer, worker, worker_err, chunk_edges, prev, pipe)
        if pipe:
            # ids at the ends of the final-flush pieces: the next call loads rows up to hi[c] once piece c is updated
            oc["rows_ready"].synchronize()
            rows_host, g0 = oc["rows_host"], int(rb[bsz])
            hi = [int(rows_host[g0 + chunk_edges[j + 1] - 1]) for j in range(J)]
            gaussians._pending_adam = dict(worker=worker, progress=progress, chunk_hi=hi, n_seg=len(seg_begin), err=worker_err,
                                           keep=[cull, oc, out])
        else:
            gaussians.last_batch_stats.update(worker_times)
        return out
    except BaseException:
        # never leave the host-Adam thread spinning on flags that will not come
        torch.cuda.synchronize()
        h_packed.zero_()          # whatever has not been flushed counts as a zero gradient
        gaussians.signal_tensor_pinned.fill_(1)
        worker.join()
        raise


def _wait_progress(pend, need):
    """Spin (politely) until the previous batch's host Adam has finished `need` segments."""
    import time as _t
    prog, w = pend["progress"], pend["worker"]
    while prog[0] < need and w.is_alive():
        _t.sleep(5e-5)
    if pend["err"]:   # the previous host update died: its rows are only partly updated, do not train on them
        raise pend["err"][0]


def _train_cameras(gaussians, batched_cameras, oc, cull, filters, h_params, h_packed, background, comm_stream, signal,
                   timer, worker, worker_err, chunk_edges, prev=None, pipe=False):
    args = gaussians.args
    bsz = len(batched_cameras)
    N = gaussians.n_gaussians
    dev = gaussians._xyz.device
    compute = torch.cuda.current_stream()
    ordered_cams, sparsity = oc["ordered_cams"], oc["sparsity"]
    cams = [batched_cameras[i] for i in ordered_cams]
    bitmap = oc["ordered_bitmap"]

    # ---- stage 3: state ---------------------------------------------------------------------------
    grads = dict(xyz=torch.zeros_like(gaussians._xyz), opacity=torch.zeros_like(gaussians._opacity),
                 scaling=torch.zeros_like(gaussians._scaling), rotation=torch.zeros_like(gaussians._rotation))
    # Two copy streams: `comm_stream` (the caller's) carries host->device SH rows, `flush_stream` the
    # device->host gradient rows.  PCIe is full duplex: kept on one stream the two directions serialise
    # (measured at C4: 8.1 ms prefetch + 8.8 ms hand-off per camera back to back).
    if getattr(gaussians, "_flush_stream", None) is None or gaussians._flush_stream.device != dev:
        gaussians._flush_stream = torch.cuda.Stream(device=dev, priority=-1)
    flush_stream = gaussians._flush_stream
    start_evt = torch.cuda.Event()
    start_evt.record(compute)
    comm_stream.wait_event(start_evt)  # filters / bitmap are ready for the copy streams
    flush_stream.wait_event(start_evt)
    losses, keep, n_isects_log = [], [cull, oc], []
    lookahead = max(0, int(getattr(args, "geometry_lookahead", 3)))
    geoms = [None] * bsz

    with torch.no_grad():
        # Batch-resident SH rows: one device row per touched Gaussian for the whole batch (slot_of).  The copy
        # stream loads first-use group k (rows first needed by camera k) into its contiguous slot range, in order
        # and ahead of the compute stream; the flush stream stores last-use group k to the pinned gradient table
        # once camera k's backward is done.  Every row crosses PCIe once each way; no re-packing between cameras.
        slot_of, union_ids, fe = oc["slot_of"], oc["union_ids"], oc["first_edges"]
        n_union = int(fe[-1])
        with torch.cuda.stream(comm_stream):
            d_rows = ops._empty((n_union, 48), torch.float32, dev)
            d_rows.record_stream(compute)
            union_ids.record_stream(comm_stream)
            slot_of.record_stream(comm_stream)
            rows_evt = []
            for k in range(bsz):
                a0, a1 = int(fe[k]), int(fe[k + 1])
                if k == 0 and prev is not None:
                    # pipelined: the previous batch's host Adam is still updating rows.  It finishes its last group
                    # in pieces of ascending ids; load our first group in matching id ranges, each as soon as the
                    # piece covering it is done (everything before the last group is done by then as well).
                    hi = torch.tensor(prev["chunk_hi"], dtype=union_ids.dtype, device=dev)
                    cuts = torch.searchsorted(union_ids[a0:a1].contiguous(), hi, right=True).tolist()
                    edges = [0] + cuts[:-1] + [a1 - a0]
                    base = prev["n_seg"] - len(cuts)
                    with timer("rows_load_first", comm_stream):
                        for c in range(len(cuts)):
                            _wait_progress(prev, prev["n_seg"] if c == len(cuts) - 1 else base + c + 1)
                            if edges[c + 1] > edges[c]:
                                ops.rows_gather_to_slots(d_rows, h_params, union_ids[a0 + edges[c]:a0 + edges[c + 1]],
                                                         slot_of)
                    prev["keep"] = None  # the previous batch is completely done: release its buffers
                else:
                    with timer("rows_load_first" if k == 0 else "rows_prefetch", comm_stream):
                        ops.rows_gather_to_slots(d_rows, h_params, union_ids[a0:a1], slot_of)
                e = torch.cuda.Event()
                e.record(comm_stream)
                rows_evt.append(e)
        d_grad = ops._zeros((n_union, 48), torch.float32, dev)   # compute stream; the flush stream waits on bwd events
        d_grad.record_stream(flush_stream)
        rb, re_ = oc["row_begin"], oc["row_end"]
        n_unt = int(re_[0])

        for k in range(bsz):
            ids = filters[k]
            cam = cams[k]
            if _NVTX:
                torch.cuda.nvtx.range_push("micro_batch_idx: " + str(k))   # engine.py:485
            # geometry of this and the next few cameras does not need SH rows: issue it before waiting on the
            # copy stream so that it hides the host-link latency of the first load
            # The first part (projection, tile counts) is issued `lookahead` cameras ahead, the tile sort one camera
            # ahead: its size comes from a count that the first part sent to the host two launches earlier, so
            # reading it does not stall the host (the reference blocks on .item() for every camera).
            for kk in range(k, min(bsz, k + 1 + lookahead)):
                if geoms[kk] is None:
                    geoms[kk] = geometry_begin(gaussians, cams[kk], filters[kk], timer)
            for kk in range(k, min(bsz, k + 2)):
                geometry_finish(geoms[kk], timer)
            fw, geoms[k] = geoms[k], None
            fw = shade_one_camera(gaussians, cam, fw, d_rows, background, timer, rows_ready=rows_evt[k], ids=ids,
                                  slot_of=slot_of)
            with timer("loss"):
                loss_vec, v_render = ops.loss_fwd_bwd(fw["render"][0], cam.original_image, args.lambda_dssim,
                                                      hwc=True)
            losses.append(loss_vec[0])
            v_means2d = backward_one_camera(gaussians, cam, ids, d_rows, d_grad, fw, v_render[None], grads, timer,
                                            slot_of=slot_of)
            if args.densification_stats:  # engine.py:833-844
                with timer("densification_stats"):
                    ops.densification_stats(ids, v_means2d[0], fw["radii"][0], cam.image_width, cam.image_height,
                                            gaussians.max_radii2D, gaussians.xyz_gradient_accum, gaussians.denom)
            bwd_done = torch.cuda.Event()
            bwd_done.record(compute)

            # gradients of the rows whose last user is this camera -> pinned table, then the flag (flush stream)
            # slots are positions in the last-use grouping, so this group is the contiguous range [g0, g1) - n_unt of
            # d_grad and of the packed staging buffer: one DMA copy, no kernel
            m = k + 1
            g0, g1 = int(rb[m]) - n_unt, int(re_[m]) - n_unt
            with torch.cuda.stream(flush_stream):
                flush_stream.wait_event(bwd_done)
                with timer("grad_handoff", flush_stream):
                    if k < bsz - 1:
                        if g1 > g0:
                            h_packed[g0:g1].copy_(d_grad[g0:g1], non_blocking=True)
                        ops.set_signal(signal, k, 1)
                    else:
                        # last camera: in pieces, each announced on its own flag, so the host Adam starts on piece j
                        # while piece j+1 is still crossing the link
                        for j in range(len(chunk_edges) - 1):
                            c0, c1 = g0 + chunk_edges[j], g0 + chunk_edges[j + 1]
                            if c1 > c0:
                                h_packed[c0:c1].copy_(d_grad[c0:c1], non_blocking=True)
                            ops.set_signal(signal, bsz - 1 + j, 1)
            n_isects_log.append(fw["n_isects"])
            del fw, v_render, loss_vec
            if _NVTX:
                torch.cuda.nvtx.range_pop()

        # ---- stage 5: GPU Adam (grads / bsz fused as grad_scale) ------------------------------------
        with timer("gpu_adam"):
            gaussians.optimizer.gpu_adam.step(grads, visibility=oc["visibility"] if args.sparse_adam else None,
                                              grad_scale=1.0 / bsz)

    if pipe:
        # return with the host Adam (and the tail of the GPU work) still running; the next call or
        # wait_host_adam() picks it up.  Everything that crosses streams carries record_stream marks.
        gaussians.last_batch_stats = dict(n_union=n_union, n_first=int(fe[1]), n_visible=[f.shape[0] for f in filters],
                                          n_isects=n_isects_log)
        return losses, ordered_cams, sparsity
    import time as _t
    comm_stream.synchronize()
    flush_stream.synchronize()
    compute.synchronize()
    t_join = _t.perf_counter()
    worker.join()  # "trailing overhead": host Adam still running after the GPU finished (engine.py:888-917)
    join_ms = (_t.perf_counter() - t_join) * 1e3
    if worker_err:
        raise worker_err[0]
    torch.cuda.synchronize()
    if timer.enabled:
        timer.collect()
    gaussians.last_batch_stats = dict(cpu_adam_trailing_ms=join_ms, n_union=n_union, n_first=int(fe[1]),
                                      n_visible=[f.shape[0] for f in filters],
                                      n_isects=n_isects_log)
    return losses, ordered_cams, sparsity


class _NaiveHostState:
    """naive_offload keeps EVERY parameter and its optimiser state on the host (strategies/naive_offload/
    gaussian_model.py:88-147: one cpu_adam.CPUAdam over xyz / f_dc / f_rest / opacity / scaling / rotation): pinned
    copies of the four GPU-resident tensors, their gradient tables and a host Adam each (same batch-scaled
    hyper-parameters as the GPU Adam they replace)."""

    def __init__(self, gaussians):
        from .model import CpuAdam
        self.param, self.grad, self.adam = {}, {}, {}
        for g in gaussians.optimizer.gpu_adam.param_groups:
            p = g["params"][0]
            h = hostmem.pinned_tensor(tuple(p.shape), torch.float32)
            h.copy_(p)
            self.param[g["name"]] = h
            self.grad[g["name"]] = hostmem.pinned_tensor(tuple(p.shape), torch.float32, zero=True)
            self.adam[g["name"]] = CpuAdam(h, [p.shape[1]], [g["lr"]], g["betas"], g["eps"],
                                           gaussians.optimizer.cpu_adam.n_threads)
        self.groups = gaussians.optimizer.gpu_adam.param_groups


def naive_sync_to_gpu(gaussians):
    """After naive_offload_train_one_batch(offload_all=True) the host copies are the up-to-date parameters; bring
    them to the GPU tensors (the next naive step does this itself)."""
    st = getattr(gaussians, "_naive_host", None)
    if st is not None:
        for g in st.groups:
            g["params"][0].copy_(st.param[g["name"]], non_blocking=True)
        torch.cuda.synchronize()


@torch.no_grad()
def naive_offload_train_one_batch(gaussians, scene, batched_cameras, parameters_grad_buffer, background, pipe_args,
                                  comm_stream=None, perm_generator=None, timer=None, offload_all=False):
    """Minimal ZeRO-offload-style baseline for BASELINE.json's "clm_offload vs naive_offload" comparison
    (strategies/naive_offload/engine.py: load ALL offloaded parameters to the GPU in one transfer, render the
    cameras, send ALL gradients back in one transfer, then update on the CPU -- no selection, no overlap).
    Built from the same kernels.  offload_all=False offloads only the 48 SH floats (the other 11 stay on the GPU with
    the GPU Adam, as in clm_offload); offload_all=True is the reference's naive mode: all 59 floats per Gaussian live
    on the host, travel both ways every step and are updated by host Adams (what bench.py --strategy naive runs).
    Same math as clm_offload_train_one_batch; cameras are processed in input order."""
    timer = timer or _NULL_TIMER
    args = gaussians.args
    bsz = len(batched_cameras)
    N = gaussians.n_gaussians
    dev = gaussians._xyz.device
    h_params, h_grad = gaussians._parameters, parameters_grad_buffer[:N]
    wait_host_adam(gaussians)
    if gaussians.optimizer.cpu_adam.lazy_pending:
        gaussians.optimizer.cpu_adam.catch_up(None)
    host = None
    if offload_all:
        host = getattr(gaussians, "_naive_host", None)
        if host is None or host.param["xyz"].shape[0] != N:
            host = gaussians._naive_host = _NaiveHostState(gaussians)
        with timer("bulk_load"):   # the 11 "selection-critical" floats come from the host as well
            for g in host.groups:
                g["params"][0].copy_(host.param[g["name"]], non_blocking=True)
    filters, cull = calculate_filters(batched_cameras, gaussians, timer)
    sparsity = [f.shape[0] / float(N) for f in filters]
    with timer("bulk_load"):
        d_rows = ops._empty((N, 48), torch.float32, dev)
        d_rows.copy_(h_params, non_blocking=True)
    d_grad = ops._zeros((N, 48), torch.float32, dev)
    ident = getattr(gaussians, "_identity_slots", None)
    if ident is None or ident.shape[0] != N:
        ident = gaussians._identity_slots = torch.arange(N, dtype=torch.int32, device=dev)
    grads = dict(xyz=torch.zeros_like(gaussians._xyz), opacity=torch.zeros_like(gaussians._opacity),
                 scaling=torch.zeros_like(gaussians._scaling), rotation=torch.zeros_like(gaussians._rotation))
    losses, n_isects = [], []
    for k in range(bsz):
        cam, ids = batched_cameras[k], filters[k]
        fw = geometry_one_camera(gaussians, cam, ids, timer)
        fw = shade_one_camera(gaussians, cam, fw, d_rows, background, timer, ids=ids, slot_of=ident)
        with timer("loss"):
            loss_vec, v_render = ops.loss_fwd_bwd(fw["render"][0], cam.original_image, args.lambda_dssim, hwc=True)
        losses.append(loss_vec[0])
        backward_one_camera(gaussians, cam, ids, d_rows, d_grad, fw, v_render[None], grads, timer, slot_of=ident)
        n_isects.append(fw["n_isects"])
        del fw, v_render
    if not offload_all:
        with timer("gpu_adam"):
            vis = (cull["bitmap"] != 0).to(torch.uint8) if args.sparse_adam else None
            gaussians.optimizer.gpu_adam.step(grads, visibility=vis, grad_scale=1.0 / bsz)
    with timer("bulk_store"):
        h_grad.copy_(d_grad, non_blocking=True)
        if offload_all:
            for name in host.grad:
                host.grad[name].copy_(grads[name], non_blocking=True)
    if args.sparse_adam:
        rows = torch.nonzero(cull["bitmap"]).flatten().to(torch.int32).cpu().numpy()
    else:
        rows = np.arange(N, dtype=np.int32)
    torch.cuda.synchronize()
    nosig = np.zeros(1, np.int32)
    gaussians.optimizer.cpu_adam.batched_sparse_step(h_grad, rows, [0], [len(rows)], [-1], nosig, 1.0 / bsz)
    if offload_all:
        for g in host.groups:   # the xyz learning rate follows its schedule (update_learning_rate)
            ad = host.adam[g["name"]]
            ad.columns_lr[:] = g["lr"]
            ad.batched_sparse_step(host.grad[g["name"]], rows, [0], [len(rows)], [-1], nosig, 1.0 / bsz)
    if args.sparse_adam:
        h_grad.zero_()  # rows outside `rows` hold zeros already; keep the table's invariant explicit
    if timer.enabled:
        timer.collect()
    gaussians.last_batch_stats = dict(n_visible=[f.shape[0] for f in filters], n_isects=n_isects, n_union=N)
    return losses, list(range(bsz)), sparsity


@torch.no_grad()
def clm_offload_eval_one_cam(camera, gaussians, background, scene=None):
    """engine.py:928-979, with the SH rows fetched by the zero-copy gather kernel instead of a
    CPU-side torch.gather + .to('cuda').  Returns the [3,H,W] image."""
    if hasattr(gaussians, "materialize_host_rows"):
        gaussians.materialize_host_rows()  # postponed (lazy) host updates, a pipelined host Adam
    filters, _ = calculate_filters([camera], gaussians)
    ids = filters[0]
    shs = torch.empty(ids.shape[0], 48, device=ids.device)
    ops.rows_gather(shs, gaussians._parameters, ids)
    fw = forward_one_camera(gaussians, camera, ids, shs, background)
    return fw["render"][0].permute(2, 0, 1).contiguous()
