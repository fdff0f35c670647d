"""Multi-GPU training iteration on one node: cameras data-parallel, Gaussian state sharded by id.

New design -- the reference is single-GPU.  One process per GPU, global batch = all ranks' cameras,
ONE optimiser step per call.  Rank g owns the Gaussian range [g S, (g+1) S), S = ceil(N / G):

  replicated : the 11 GPU-resident floats per Gaussian (every rank culls all N for its cameras)
  sharded    : Adam moments of those 11 floats (ShardedGpuAdam), pinned gradient table, host-Adam moments and the
               responsibility for updating the SH rows of the range
  shared     : the SH parameter table itself -- one node-shared host mapping registered with every rank's GPU
               (clmgs/shm.py), so a rank reads the rows its cameras need from host DRAM over its OWN PCIe link with
               the single-GPU zero-copy kernels and the single-GPU first-use grouping; there is no owner->renderer
               hop and the eight host links add up

  per step   : 1. cull + TSP order of the LOCAL cameras; need_ids = ascending union of their Gaussians
               2. plan: the local camera sequence is cut into K chunks; ids + "chunk of my last use" all-to-all ->
                  every owner knows which of its rows will receive gradients, from whom and in which chunk's
                  exchange the last contribution arrives; it merges the (already sorted) lists with flag tables
                  (no sort) and starts its host Adam thread, which waits on per-chunk flags
               3. copy stream loads first-use group k of the local cameras into the batch-resident rows (slots in
                  ascending-id = owner-major order); compute stream renders camera k, accumulates SH gradients
                  into a slot-shaped buffer and the dense gradients of the replicated attributes
               4. after each chunk of cameras, on a side stream: sparse reduce-scatter of the rows whose gradient
                  is final on this rank (slots are in (last chunk, id) order, so a chunk is one owner-major slice):
                  "peer" (default) -- the rows are added straight into the owners' accumulators, which live in
                  symmetric memory mapped by every rank, with 16-byte vector atomics over NVLink, then a signal per
                  owner; "nccl" -- all-to-all-v, then the owner accumulates out of the receive buffer.  Rows whose
                  last contribution this was leave by DMA into a packed pinned staging buffer + one flag -> the
                  owner's host Adam updates them while later cameras render
               5. NCCL reduce-scatter of the dense gradients -> owner Adam on its slice (moments sharded) ->
                  all-gather of the updated slices
               6. join host Adam; barrier (the next step reads rows other ranks just updated)
  math       : identical to one GPU running the global batch, up to fp32 summation order.
"""
import threading

import numpy as np
import torch
import torch.distributed as dist

from . import ops
from .dist_rows import RowExchange, ShardMap, merge_requests, plan_chunked
from .engine import (StageTimer, _NULL_TIMER, backward_one_camera, calculate_filters, geometry_begin,
                     geometry_finish, order_calculation, shade_one_camera)

_GPU_NAMES = ("xyz", "opacity", "scaling", "rotation")
_SIGNAL_TIMEOUT_MS = 120_000   # a peer that never signals traps the wait instead of hanging the GPU


class DistContext:
    def __init__(self, n_total, group=None):
        self.rank, self.world = dist.get_rank(group), dist.get_world_size(group)
        self.group = group
        self.shard = ShardMap(n_total, self.world)
        self.exchange = RowExchange(self.shard, self.rank, group)
        self.lo, self.hi = self.shard.range_of(self.rank)
        self.n_total = int(n_total)
        self.table = None            # SharedHostTable of the SH parameters (create_shared_table)

    def create_shared_table(self, cols=48):
        """The node-shared SH parameter table [N, cols]; each rank fills rows [lo, hi) and then calls
        training_setup_dist (which registers the mapping with CUDA and synchronises the ranks)."""
        from .shm import SharedHostTable
        self.table = SharedHostTable(self.n_total, cols, self.group)
        return self.table


def training_setup_dist(gaussians, args, ctx):
    """Optimiser set-up of a rank.  args.bsz = GLOBAL batch.  gaussians was built by create_from_arrays with
    sh = ctx.table.tensor[lo:hi] (this rank's rows of the shared table).  Pads the replicated tensors to world x S
    rows (equal counts for reduce-scatter / all-gather) and shards the Adam moments."""
    assert ctx.table is not None, "call ctx.create_shared_table() and fill this rank's rows first"
    dev = gaussians._xyz.device
    N, S, G = ctx.n_total, ctx.shard.shard, ctx.world
    for name in ("_xyz", "_opacity", "_scaling", "_rotation"):
        t = getattr(gaussians, name)
        buf = torch.zeros((G * S,) + tuple(t.shape[1:]), dtype=t.dtype, device=dev)
        buf[:N].copy_(t)
        setattr(gaussians, name + "_padded", buf)
        setattr(gaussians, name, buf[:N])
    gaussians.training_setup(args, gpu_shard=(ctx.lo, ctx.hi))
    gaussians.dist_ctx = ctx
    ctx.symm = None
    if G > 1 and getattr(args, "dist_grad_exchange", "peer") == "peer":
        # Owner-side gradient accumulators in SYMMETRIC memory: every rank maps every rank's buffer, so a renderer
        # adds its gradient rows straight into the owner's HBM over NVLink (vector atomics) -- the sparse
        # reduce-scatter is ONE kernel per (chunk, owner) with no send / receive buffers and no collective call.
        ok, err = 1, None
        try:
            import torch.distributed._symmetric_memory as symm_mem
            ctx.acc_sym = symm_mem.empty(S * 48, dtype=torch.float32, device=dev)
            ctx.acc_sym.zero_()
            symm = symm_mem.rendezvous(ctx.acc_sym, ctx.group if ctx.group is not None else dist.group.WORLD)
            peer_acc = [symm.get_buffer(r, (S, 48), torch.float32) for r in range(G)]
        except Exception as e:   # no peer access on this box / torch build without symmetric memory
            ok, err = 0, e
        flag = torch.tensor([ok], device=dev)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN, group=ctx.group)   # all ranks take the same path
        if int(flag.item()) == 1:
            ctx.symm, ctx.peer_acc = symm, peer_acc
        else:
            import warnings
            warnings.warn("clmgs.engine_dist: symmetric memory is not available (%r); the SH gradients return through "
                          "NCCL all-to-all-v instead of peer-memory accumulation" % (err,))
    ctx.table.register_cuda()
    gaussians.shared_parameters = ctx.table.tensor
    torch.cuda.synchronize()
    dist.barrier(group=ctx.group)   # every rank's rows are in the table
    return gaussians


def clm_offload_train_one_batch_dist(gaussians, scene, local_cameras, parameters_grad_buffer, background, pipe_args,
                                     comm_stream, perm_generator, ctx: DistContext, timer=None):
    """See the module docstring.  gaussians._parameters / parameters_grad_buffer / host-Adam state hold this rank's
    SHARD ([hi-lo, 48]); gaussians.shared_parameters is the whole table; the GPU-resident tensors hold all N.
    Returns (losses of the local cameras in processed order, local order, local sparsity)."""
    timer = timer or _NULL_TIMER
    args = gaussians.args
    bsz_local = len(local_cameras)
    bsz_global = bsz_local * ctx.world
    assert args.bsz == bsz_global, "TrainArgs.bsz must be the global batch size (%d)" % bsz_global
    N, S, G = gaussians.n_gaussians, ctx.shard.shard, ctx.world
    n_local = ctx.hi - ctx.lo
    dev = gaussians._xyz.device
    h_table = gaussians.shared_parameters
    h_grad = parameters_grad_buffer[:n_local]
    assert gaussians._parameters.shape[0] == n_local, "SH table view must be this rank's shard"
    compute = torch.cuda.current_stream()
    if getattr(gaussians, "_flush_stream", None) is None or gaussians._flush_stream.device != dev:
        gaussians._flush_stream = torch.cuda.Stream(device=dev, priority=-1)
    flush_stream = gaussians._flush_stream
    cpu_adam = gaussians.optimizer.cpu_adam
    if getattr(cpu_adam, "lazy_pending", False):
        cpu_adam.catch_up(None)

    with torch.no_grad():
        # ---- 1. visibility + local order ---------------------------------------------------------------
        filters, cull = calculate_filters(local_cameras, gaussians, timer)
        oc = order_calculation(filters, cull, N, bsz_local, perm_generator, args, timer, host_rows=False)
        ordered, filters, sparsity = oc["ordered_cams"], oc["filters"], oc["sparsity"]
        cams = [local_cameras[i] for i in ordered]
        union_ids, fe = oc["union_ids"], oc["first_edges"]     # grouped by first use (int32), id ranges per camera
        n_need = int(fe[-1])
        K = max(1, min(int(getattr(args, "dist_grad_chunks", 8)), bsz_local // 2 if bsz_local >= 2 else 1))
        chunk_of = [p * K // bsz_local for p in range(bsz_local)]            # camera position -> chunk
        with timer("need_ids"):
            need_ids = torch.nonzero(oc["ordered_bitmap"]).flatten()          # ascending union = owner-major order
            assert need_ids.shape[0] == n_need
            # oc["visibility"] = ffs of the ordered bitmap: key k > 0 <-> last used by camera position bsz - k
            last_pos = bsz_local - oc["visibility"][need_ids].to(torch.int64)
            last_chunk = torch.div(last_pos * K, bsz_local, rounding_mode="floor")

        # ---- 2. plan + owner-side merge + host Adam thread ---------------------------------------------
        with timer("plan"):
            plan = plan_chunked(ctx.exchange, need_ids, last_chunk, K, n_local, want_untouched=not args.sparse_adam)
            slot_of = torch.empty(N, dtype=torch.int32, device=dev)
            ops.slot_map(plan.send_order.to(torch.int32), slot_of)            # slots in (last chunk, id) order
            n_uniq = int(plan.acc_order.shape[0])
            n_unt = 0 if plan.untouched is None else int(plan.untouched.shape[0])
            rows_host = _pinned_i32(gaussians, n_local)
            if n_unt:
                rows_host[:n_unt].copy_(plan.untouched.to(torch.int32), non_blocking=True)
            rows_host[n_unt:n_unt + n_uniq].copy_(plan.acc_order.to(torch.int32), non_blocking=True)
            torch.cuda.current_stream().synchronize()
        # host-Adam segments: [untouched (dense Adam only)] + rows whose sum is final after exchange c, c = 0..K-1;
        # the last group leaves in J pieces so that the host works on piece j while piece j+1 crosses the link
        f_edges = np.concatenate([[0], np.cumsum(plan.final_counts)]).astype(np.int64)
        n_lastg = int(plan.final_counts[K - 1])
        J = max(1, min(int(getattr(args, "final_flush_chunks", 8)), max(1, n_lastg // 65536)))
        piece = [n_lastg * j // J for j in range(J + 1)]
        seg_b, seg_e, seg_sig, seg_zg = [], [], [], []
        if n_unt:  # dense Adam: untouched rows decay their moments; gradient rows are zero -> shortcut, no wait
            seg_b.append(0); seg_e.append(n_unt); seg_sig.append(-1); seg_zg.append(1)
        for c in range(K - 1):
            seg_b.append(n_unt + int(f_edges[c])); seg_e.append(n_unt + int(f_edges[c + 1])); seg_sig.append(c); seg_zg.append(0)
        for j in range(J):
            seg_b.append(n_unt + int(f_edges[K - 1]) + piece[j]); seg_e.append(n_unt + int(f_edges[K - 1]) + piece[j + 1])
            seg_sig.append(K - 1 + j); seg_zg.append(0)
        signal = _signal_flags(gaussians, K - 1 + J)
        worker_err, worker_times = [], {}
        from .engine import _packed_staging
        h_packed = _packed_staging(gaussians, n_uniq, cap=n_local)   # row i = summed gradient of acc_order[i] = list entry n_unt + i

        def cpuadam_thread():
            import time as _t
            t0 = _t.perf_counter()
            try:
                cpu_adam.batched_sparse_step(h_grad, rows_host, seg_b, seg_e, seg_sig, signal, 1.0 / bsz_global, None,
                                             seg_zg, grad_packed=h_packed, packed_base=n_unt)
            except Exception as e:  # surfaced after join
                worker_err.append(e)
            worker_times["cpu_adam_thread_ms"] = (_t.perf_counter() - t0) * 1e3

        worker = threading.Thread(target=cpuadam_thread, daemon=True)
        worker.start()
        try:
            out = _render_and_exchange(gaussians, ctx, cams, filters, oc, slot_of, plan, chunk_of, f_edges, piece,
                                       signal, h_table, h_packed, background, comm_stream, flush_stream, timer)
        except BaseException:
            torch.cuda.synchronize()
            h_packed.zero_()
            signal.fill_(1)       # never leave the host-Adam thread spinning on flags that will not come
            worker.join()
            raise
    import time as _t
    comm_stream.synchronize()
    flush_stream.synchronize()
    compute.synchronize()
    t_join = _t.perf_counter()
    worker.join()
    join_ms = (_t.perf_counter() - t_join) * 1e3
    if worker_err:
        raise worker_err[0]
    # the next step's row loads read rows that OTHER ranks' host Adams just updated
    dist.barrier(group=ctx.group)
    if timer.enabled:
        timer.collect()
    gaussians.last_batch_stats = dict(n_visible=[f.shape[0] for f in filters], n_isects=out["n_isects"], n_union=n_need,
                                      n_first=int(fe[1]), n_owner_rows=n_uniq, n_adam_rows=n_uniq,
                                      rows_from_peers=plan.rows_from_peers, grad_chunks=K,
                                      grad_exchange="peer" if ctx.symm is not None else "nccl",
                                      cpu_adam_trailing_ms=join_ms, **worker_times)
    return out["losses"], ordered, sparsity


def _pinned_i32(gaussians, n):
    buf = getattr(gaussians, "_dist_rows_host", None)
    if buf is None or buf.shape[0] < n:
        buf = gaussians._dist_rows_host = torch.empty(max(n, 1), dtype=torch.int32, pin_memory=True)
    return buf


def _signal_flags(gaussians, n):
    buf = getattr(gaussians, "_dist_signal", None)
    if buf is None or buf.shape[0] != n:
        buf = gaussians._dist_signal = torch.zeros(n, dtype=torch.int32, pin_memory=True)
    else:
        buf.zero_()
    gaussians.signal_tensor_pinned = buf
    return buf


def _render_and_exchange(gaussians, ctx, cams, filters, oc, slot_of, plan, chunk_of, f_edges, piece, signal, h_table,
                         h_packed, background, comm_stream, flush_stream, timer):
    args = gaussians.args
    dev = gaussians._xyz.device
    bsz_local, G, S, K = len(cams), ctx.world, ctx.shard.shard, plan.K
    bsz_global = bsz_local * G
    compute = torch.cuda.current_stream()
    union_ids, fe = oc["union_ids"], oc["first_edges"]
    n_need = int(fe[-1])
    lookahead = max(0, int(getattr(args, "geometry_lookahead", 3)))
    if getattr(gaussians, "_xchg_stream", None) is None or gaussians._xchg_stream.device != dev:
        gaussians._xchg_stream = torch.cuda.Stream(device=dev, priority=-1)
    xchg_stream = gaussians._xchg_stream

    # ---- 3. rows in (copy stream, first-use groups) + cameras (compute stream) --------------------------
    start_evt = torch.cuda.Event()
    start_evt.record(compute)
    comm_stream.wait_event(start_evt)
    with torch.cuda.stream(comm_stream):
        d_rows = ops._empty((n_need, 48), torch.float32, dev)
        d_rows.record_stream(compute)
        union_ids.record_stream(comm_stream)
        slot_of.record_stream(comm_stream)
        rows_evt = []
        for k in range(bsz_local):
            a0, a1 = int(fe[k]), int(fe[k + 1])
            with timer("rows_load_first" if k == 0 else "rows_prefetch", comm_stream):
                ops.rows_gather_to_slots(d_rows, h_table, union_ids[a0:a1], slot_of)
            e = torch.cuda.Event()
            e.record(comm_stream)
            rows_evt.append(e)
    d_grad = ops._zeros((n_need, 48), torch.float32, dev)
    n_uniq = int(plan.acc_order.shape[0])
    # owner-side sums, rows in (final chunk, local row) order; peer mode: this rank's symmetric buffer (zeroed at the
    # end of the previous step, before the barrier)
    acc = ctx.peer_acc[ctx.rank][:n_uniq] if ctx.symm is not None else ops._zeros((n_uniq, 48), torch.float32, dev)
    zero_evt = torch.cuda.Event()
    zero_evt.record(compute)
    xchg_stream.wait_event(zero_evt)
    d_grad.record_stream(xchg_stream)
    if ctx.symm is None:
        acc.record_stream(xchg_stream)
        acc.record_stream(flush_stream)
    s_edges = np.concatenate([[0], np.cumsum([sum(c) for c in plan.send_counts])]).astype(np.int64)
    # dense gradients of the replicated attributes, padded to G x S rows for the reduce-scatter
    grads = {n: torch.zeros_like(getattr(gaussians, "_" + n + "_padded")) for n in _GPU_NAMES}
    gview = {n: grads[n][: gaussians.n_gaussians] for n in _GPU_NAMES}
    losses, n_isects, geoms = [], [], [None] * bsz_local
    for k in range(bsz_local):
        cam, ids = cams[k], filters[k]
        for kk in range(k, min(bsz_local, k + 1 + lookahead)):
            if geoms[kk] is None:
                geoms[kk] = geometry_begin(gaussians, cams[kk], filters[kk], timer)
        for kk in range(k, min(bsz_local, k + 2)):
            geometry_finish(geoms[kk], timer)
        fw, geoms[k] = geoms[k], None
        fw = shade_one_camera(gaussians, cam, fw, d_rows, background, timer, rows_ready=rows_evt[k], ids=ids,
                              slot_of=slot_of)
        with timer("loss"):
            loss_vec, v_render = ops.loss_fwd_bwd(fw["render"][0], cam.original_image, args.lambda_dssim, hwc=True)
        losses.append(loss_vec[0])
        v_means2d = backward_one_camera(gaussians, cam, ids, d_rows, d_grad, fw, v_render[None], gview, timer,
                                        slot_of=slot_of)
        if args.densification_stats:
            with timer("densification_stats"):
                ops.densification_stats(ids, v_means2d[0], fw["radii"][0], cam.image_width, cam.image_height,
                                        gaussians.max_radii2D, gaussians.xyz_gradient_accum, gaussians.denom)
        n_isects.append(fw["n_isects"])
        del fw, v_render, loss_vec

        # ---- 4. after the last camera of chunk c: its rows' gradients are final on this rank -> sparse
        # reduce-scatter to the owners (NCCL all-to-all-v straight out of the slot buffer), owners accumulate, and
        # the rows whose LAST contribution this was go to the host (DMA into the packed staging buffer + flag)
        c = chunk_of[k]
        if k == bsz_local - 1 or chunk_of[k + 1] != c:
            done = torch.cuda.Event()
            done.record(compute)
            with torch.cuda.stream(xchg_stream):
                xchg_stream.wait_event(done)
                with timer("grad_exchange", xchg_stream):
                    if ctx.symm is not None:
                        # peer push: add this chunk's rows into each owner's accumulator (its HBM, over NVLink), then
                        # tell the owners; the chunk's sums are complete here once every peer has signalled
                        off = int(s_edges[c])
                        for o in range(G):
                            n_o = plan.send_counts[c][o]
                            if n_o:
                                ops.rows_accumulate(ctx.peer_acc[o], d_grad[off:off + n_o], plan.remote_slot[off:off + n_o])
                            off += n_o
                        for o in range(G):
                            if o != ctx.rank:
                                ctx.symm.put_signal(o, c, _SIGNAL_TIMEOUT_MS)
                        for o in range(G):
                            if o != ctx.rank:
                                ctx.symm.wait_signal(o, c, _SIGNAL_TIMEOUT_MS)
                    else:
                        recv = ops._empty((sum(plan.recv_counts[c]), 48), torch.float32, dev)
                        dist.all_to_all_single(recv, d_grad[int(s_edges[c]):int(s_edges[c + 1])],
                                               output_split_sizes=plan.recv_counts[c],
                                               input_split_sizes=plan.send_counts[c], group=ctx.group)
                        ops.rows_accumulate(acc, recv, plan.dst_index[c])
                summed = torch.cuda.Event()
                summed.record(xchg_stream)
            with torch.cuda.stream(flush_stream):
                flush_stream.wait_event(summed)
                with timer("grad_handoff", flush_stream):   # contiguous on both sides: DMA copies (copy engine, no SMs)
                    f0 = int(f_edges[c])
                    if c < K - 1:
                        if f_edges[c + 1] > f0:
                            h_packed[f0:int(f_edges[c + 1])].copy_(acc[f0:int(f_edges[c + 1])], non_blocking=True)
                        ops.set_signal(signal, c, 1)
                    else:
                        for j in range(len(piece) - 1):
                            c0, c1 = f0 + piece[j], f0 + piece[j + 1]
                            if c1 > c0:
                                h_packed[c0:c1].copy_(acc[c0:c1], non_blocking=True)
                            ops.set_signal(signal, K - 1 + j, 1)

    # ---- 5. GPU-resident attributes: reduce-scatter -> owner Adam on its slice -> all-gather ---------------
    shards = {}
    with timer("reduce_scatter_dense"):
        for n in _GPU_NAMES:
            full = grads[n]
            shards[n] = torch.empty((S,) + tuple(full.shape[1:]), dtype=full.dtype, device=dev)
            dist.reduce_scatter_tensor(shards[n], full, group=ctx.group)
    with timer("gpu_adam"):
        vis = plan.touched_mask.view(torch.uint8) if args.sparse_adam else None
        gaussians.optimizer.gpu_adam.step(shards, visibility=vis, grad_scale=1.0 / bsz_global)
    with timer("all_gather_params"):
        for n in _GPU_NAMES:
            full = getattr(gaussians, "_" + n + "_padded")
            dist.all_gather_into_tensor(full, full[ctx.rank * S:(ctx.rank + 1) * S].clone(), group=ctx.group)
    xchg_stream.synchronize()
    if ctx.symm is not None:
        flush_stream.synchronize()       # the DMA copies out of the accumulator are done
        acc.zero_()                      # ready for the next step's pushes (the caller's barrier orders it before them)
    return dict(losses=losses, n_isects=n_isects)
