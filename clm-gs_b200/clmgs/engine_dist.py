"""Multi-GPU training iteration: cameras data-parallel, offloaded SH state sharded by Gaussian id.

New design -- the reference is single-GPU.  One process per GPU:

  replicated : the 11 GPU-resident floats per Gaussian + their Adam moments (every rank culls all N)
  sharded    : rank g owns SH rows [N g/G, N (g+1)/G): pinned table, gradient table, host-Adam moments
  per batch  : 1. each rank culls / orders ITS cameras (global batch = all ranks' cameras)
               2. SH rows of the union of its cameras come from their owners: ids all-to-all, owners gather
                  from pinned host with the zero-copy kernel, rows all-to-all-v over NCCL/NVLink into a
                  per-batch device cache (HBM is plentiful on B200: no per-camera retention needed)
               3. cameras render from the cache; SH gradients accumulate into a cache-shaped buffer
               4. dense gradients of the replicated attributes: NCCL all-reduce, identical Adam on all ranks
               5. SH gradients go back to owners (all-to-all-v), are summed there (vector atomics), stored to
                  the pinned gradient shard and consumed by the owner's host Adam
  math       : identical to one GPU running the global batch, up to fp32 summation order.
"""
import threading

import numpy as np
import torch
import torch.distributed as dist

from . import ops
from .dist_rows import RowExchange, ShardMap
from .engine import (StageTimer, _NULL_TIMER, backward_one_camera, calculate_filters, geometry_one_camera,
                     shade_one_camera)


class DistContext:
    def __init__(self, n_total, group=None):
        self.rank, self.world = dist.get_rank(group), dist.get_world_size(group)
        self.group = group
        self.shard = ShardMap(n_total, self.world)
        self.exchange = RowExchange(self.shard, self.rank, group)
        self.lo, self.hi = self.shard.range_of(self.rank)


def training_setup_dist(gaussians, args, ctx):
    """Optimiser set-up of a rank (TrainArgs.bsz = GLOBAL batch)."""
    return gaussians.training_setup(args)


def clm_offload_train_one_batch_dist(gaussians, scene, local_cameras, parameters_grad_buffer, background, pipe_args,
                                     comm_stream, perm_generator, ctx: DistContext, timer=None):
    """gaussians._parameters / parameters_grad_buffer / host-Adam state hold this rank's SHARD
    ([hi-lo, 48]); the GPU-resident tensors hold all N.  gaussians.args.bsz must be the GLOBAL batch
    size (it scales lr / betas / eps and the gradient average, GM:180-211, engine.py:870-872).
    Returns (losses of the local cameras, local order, local sparsity)."""
    timer = timer or _NULL_TIMER
    args = gaussians.args
    bsz_local = len(local_cameras)
    bsz_global = bsz_local * ctx.world
    assert args.bsz == bsz_global, "TrainArgs.bsz must be the global batch size (%d)" % bsz_global
    N = gaussians.n_gaussians
    dev = gaussians._xyz.device
    h_params, h_grad = gaussians._parameters, parameters_grad_buffer[: ctx.hi - ctx.lo]
    assert h_params.shape[0] == ctx.hi - ctx.lo, "SH table must be this rank's shard"

    compute = torch.cuda.current_stream()
    with torch.no_grad():
        filters, cull = calculate_filters(local_cameras, gaussians, timer)
        # local order: keep input order (TSP only matters for per-camera retention, which the cache replaces)
        ordered = list(range(bsz_local))
        sparsity = [f.shape[0] / float(N) for f in filters]
        with timer("need_ids"):
            need_ids = torch.nonzero(cull["bitmap"]).flatten()  # union of this rank's cameras, ascending
        # geometry of the local cameras needs no SH rows: queue it on the compute stream, then run the row
        # exchange on the copy stream so that NVLink / PCIe traffic overlaps projection + tile sort
        geoms = [geometry_one_camera(gaussians, local_cameras[k], filters[k], timer) for k in ordered]
        ready = torch.cuda.Event()
        ready.record(compute)
        with torch.cuda.stream(comm_stream):
            comm_stream.wait_event(ready)
            with timer("rows_exchange_in", comm_stream):
                plan = ctx.exchange.plan(need_ids)

                def gather_local(local_idx):
                    # several ranks usually ask for the same row: read each requested row from pinned host ONCE
                    # (PCIe), then serve every requester from device memory
                    uniq, inv = torch.unique(local_idx, return_inverse=True)
                    rows_u = ops._empty((uniq.shape[0], 48), torch.float32, dev)
                    ops.rows_gather(rows_u, h_params, uniq)           # zero-copy read of my pinned shard
                    out = ops._empty((local_idx.shape[0], 48), torch.float32, dev)
                    ops.rows_gather(out, rows_u, inv)                 # device -> device fan-out
                    return out

                cache = ctx.exchange.fetch(plan, gather_local)
            cache.record_stream(compute)
            cache_ready = torch.cuda.Event()
            cache_ready.record(comm_stream)
        need_ids.record_stream(comm_stream)
        # the cache is a slot table: row of Gaussian id = its rank in need_ids (same kernels as the single-GPU engine)
        slot_of = torch.empty(N, dtype=torch.int32, device=dev)
        ops.slot_map(need_ids.to(torch.int32), slot_of)
        cache_grad = ops._zeros((need_ids.shape[0], 48), torch.float32, dev)
        grads = dict(xyz=torch.zeros_like(gaussians._xyz), opacity=torch.zeros_like(gaussians._opacity),
                     scaling=torch.zeros_like(gaussians._scaling), rotation=torch.zeros_like(gaussians._rotation))
        losses, n_isects = [], []
        compute.wait_event(cache_ready)
        for k in ordered:
            cam, ids = local_cameras[k], filters[k]
            fw = shade_one_camera(gaussians, cam, geoms[k], cache, background, timer, ids=ids, slot_of=slot_of)
            geoms[k] = None
            with timer("loss"):
                loss_vec, v_render = ops.loss_fwd_bwd(fw["render"][0], cam.original_image, args.lambda_dssim, hwc=True)
            losses.append(loss_vec[0])
            backward_one_camera(gaussians, cam, ids, cache, cache_grad, fw, v_render[None], grads, timer, slot_of=slot_of)
            n_isects.append(fw["n_isects"])
            del fw, v_render

        # replicated attributes: sum gradients over ranks, same Adam everywhere
        with timer("allreduce_dense"):
            for g in grads.values():
                dist.all_reduce(g, group=ctx.group)
        visibility = None
        if args.sparse_adam:
            visibility = (cull["bitmap"] != 0).to(torch.uint8)
            dist.all_reduce(visibility, op=dist.ReduceOp.MAX, group=ctx.group)
        with timer("gpu_adam"):
            gaussians.optimizer.gpu_adam.step(grads, visibility=visibility, grad_scale=1.0 / bsz_global)

        # SH gradients back to their owners, summed there, then the owner's host Adam
        touched = {}

        def accumulate_local(local_idx, rows):
            uniq, inv = torch.unique(local_idx, return_inverse=True)
            acc = ops._zeros((uniq.shape[0], 48), torch.float32, dev)
            ops.rows_accumulate(acc, rows, inv)
            ops.rows_scatter_grad(acc, h_grad, uniq, accum=False)   # table rows are zero: plain store
            touched["rows"] = uniq

        with timer("rows_exchange_out"):
            ctx.exchange.return_grads(plan, cache_grad, accumulate_local)
        torch.cuda.synchronize()
    n_local = ctx.hi - ctx.lo
    t_rows = touched["rows"]
    if args.sparse_adam:
        rows = t_rows.to(torch.int32).cpu().numpy()
        seg_b, seg_e, seg_zg = [0], [len(rows)], [0]
    else:
        # dense Adam also decays the moments of the rows nobody touched; their gradient rows are zero, so that
        # segment skips the gradient table
        mask = torch.ones(n_local, dtype=torch.bool, device=dev)
        mask[t_rows] = False
        unt = torch.nonzero(mask).flatten()
        rows = torch.cat([unt, t_rows]).to(torch.int32).cpu().numpy()
        seg_b, seg_e, seg_zg = [0, int(unt.shape[0])], [int(unt.shape[0]), len(rows)], [1, 0]
    sig = np.zeros(1, np.int32)
    gaussians.optimizer.cpu_adam.batched_sparse_step(h_grad, rows, seg_b, seg_e, [-1] * len(seg_b), sig, 1.0 / bsz_global,
                                                     None, seg_zg)
    if timer.enabled:
        timer.collect()
    gaussians.last_batch_stats = dict(n_visible=[f.shape[0] for f in filters], n_isects=n_isects,
                                      n_need=int(need_ids.shape[0]), rows_from_peers=int(need_ids.shape[0] - plan.send_counts[ctx.rank]))
    return losses, ordered, sparsity
