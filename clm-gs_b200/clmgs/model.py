"""State the hot path touches: the split GPU / pinned-host Gaussian model and
its optimiser hyper-parameters.

Mirrors ``GaussianModelCLMOffload`` (strategies/clm_offload/gaussian_model.py:18,
tensors :88-111, pinned buffers :33-44) and ``training_setup`` (:122-234) /
``UnifiedAdam`` (optimizer.py:91-184) as far as one training iteration needs
them.  Densification, ply/pt IO and the COLMAP readers are out of scope
(SURVEY.md 2.1 rows 3, 15).

Layout
  GPU   _xyz [N,3], _opacity [N,1] (logit), _scaling [N,3] (log), _rotation [N,4] (raw wxyz)
        + Adam moments for each (44 B + 88 B per Gaussian)
  host  _parameters [N,48] = cat(f_dc[N,3], f_rest[N,45]) viewed as [16 bases][3 rgb]
        (engine.py:71), parameters_grad_buffer [N,48], exp_avg / exp_avg_sq [N,48];
        all pinned and mapped into the GPU's address space (768 B per Gaussian)
"""
from dataclasses import dataclass, field
import math
import os

import numpy as np
import torch

from . import hostmem, ops


@dataclass
class TrainArgs:
    """The handful of reference flags the hot path reads (arguments/__init__.py:84-100,182-229)."""
    bsz: int = 4
    sparse_adam: bool = False
    reorder_by_min_sparsity_at_end: bool = True
    radius_clip: float = 0.0
    sh_degree: int = 3
    lr_scale_mode: str = "sqrt"
    position_lr_init: float = 0.00016
    position_lr_final: float = 0.0000016
    position_lr_delay_mult: float = 0.01
    position_lr_max_steps: int = 30_000
    feature_lr: float = 0.0025
    opacity_lr: float = 0.05
    scaling_lr: float = 0.005
    rotation_lr: float = 0.001
    lambda_dssim: float = 0.2
    spatial_lr_scale: float = 1.0
    densification_stats: bool = False  # accumulate max_radii2D / xyz_gradient_accum / denom per camera (A17)
    cpu_adam_threads: int = 0  # 0 = all hardware threads
    final_flush_chunks: int = 8  # last camera: flush + host Adam pipelined in this many pieces
    dist_grad_chunks: int = 8  # multi-GPU: the local cameras' SH gradients return to their owners in this many exchanges
    dist_grad_exchange: str = "peer"  # "peer": accumulate straight into the owner's HBM over NVLink; "nccl": all-to-all-v
    pipeline_host_adam: bool = False  # opt-in: return before the host Adam finished; the next call overlaps it
    lazy_dense_adam: bool = False  # opt-in (dense Adam only): postpone the zero-gradient updates of untouched rows
    rows_lookahead: int = 2  # SH-row prefetches issued this many cameras ahead on the copy stream
    geometry_lookahead: int = 3  # cameras whose projection/tile sort is issued ahead of their SH rows
    # densification schedule (arguments/__init__.py:216-225)
    percent_dense: float = 0.01
    densification_interval: int = 100
    opacity_reset_interval: int = 3000
    densify_from_iter: int = 500
    densify_until_iter: int = 15_000
    densify_grad_threshold: float = 0.0002
    min_opacity: float = 0.005
    disable_auto_densification: bool = False


class _Group(dict):
    pass


class GpuAdam:
    """Dense (torch fused-Adam rule, bias-corrected) or selective (masked, no bias
    correction) Adam over the four GPU-resident tensors -- optimizer.py:47-88,124-129."""

    def __init__(self, groups, sparse):
        self.param_groups = groups
        self.sparse = sparse
        self.state = {}
        self.step_count = 0
        for g in groups:
            p = g["params"][0]
            self.state[g["name"]] = dict(exp_avg=torch.zeros_like(p), exp_avg_sq=torch.zeros_like(p))

    @torch.no_grad()
    def step(self, grads, visibility=None, grad_scale=1.0):
        """grads: dict name -> dense grad tensor.  visibility: uint8/bool [N] (selective only)."""
        self.step_count += 1
        for g in self.param_groups:
            p, st = g["params"][0], self.state[g["name"]]
            b1, b2 = g["betas"]
            if self.sparse:
                ops.selective_adam(p, grads[g["name"]], st["exp_avg"], st["exp_avg_sq"], visibility, g["lr"], b1, b2,
                                   g["eps"], grad_scale)
            else:
                ops.adam_dense(p, grads[g["name"]], st["exp_avg"], st["exp_avg_sq"], g["lr"], b1, b2, g["eps"],
                               self.step_count, grad_scale)


class ShardedGpuAdam:
    """Multi-GPU: the Adam moments of the GPU-resident attributes exist only for this rank's Gaussian range
    [lo, hi) (44 x 2 / world bytes per Gaussian instead of 88): gradients arrive by reduce-scatter, the owner
    updates its slice of the replicated parameters, an all-gather publishes it.  Same update rules
    as GpuAdam."""

    def __init__(self, groups, sparse, lo, hi):
        self.param_groups, self.sparse, self.lo, self.hi = groups, sparse, lo, hi
        self.state, self.step_count = {}, 0
        for g in groups:
            p = g["params"][0]
            shape = (hi - lo,) + tuple(p.shape[1:])
            self.state[g["name"]] = dict(exp_avg=torch.zeros(shape, dtype=p.dtype, device=p.device),
                                         exp_avg_sq=torch.zeros(shape, dtype=p.dtype, device=p.device))

    @torch.no_grad()
    def step(self, grad_shards, visibility=None, grad_scale=1.0):
        """grad_shards: name -> summed gradient of rows [lo, hi); visibility: uint8/bool [hi - lo] (selective)."""
        self.step_count += 1
        n = self.hi - self.lo
        for g in self.param_groups:
            p = g["params"][0][self.lo:self.hi]
            st, gr = self.state[g["name"]], grad_shards[g["name"]][:n]
            b1, b2 = g["betas"]
            if self.sparse:
                ops.selective_adam(p, gr, st["exp_avg"], st["exp_avg_sq"], visibility, g["lr"], b1, b2, g["eps"],
                                   grad_scale)
            else:
                ops.adam_dense(p, gr, st["exp_avg"], st["exp_avg_sq"], g["lr"], b1, b2, g["eps"], self.step_count,
                               grad_scale)


class CpuAdam:
    """Host Adam over the pinned [N,48] SH table (the reference's cpu_adam.FusedCPUAdam,
    optimizer.py:130-142): per-column learning rates, bias correction on."""

    def __init__(self, param, columns_sizes, columns_lr, betas, eps, n_threads=0, capacity=None):
        self.param = param
        self.columns_lr = np.repeat(np.asarray(columns_lr, np.float32), columns_sizes).astype(np.float32)
        self.betas = list(betas)
        self.eps = eps
        # moments are sized like the parameter buffer (capacity rows) so that densification can grow in place
        rows = max(capacity or 0, param.shape[0])
        self.exp_avg_buffer = _pinned_zeros((rows,) + tuple(param.shape[1:]))
        self.exp_avg_sq_buffer = _pinned_zeros((rows,) + tuple(param.shape[1:]))
        self.exp_avg = self.exp_avg_buffer[:param.shape[0]]
        self.exp_avg_sq = self.exp_avg_sq_buffer[:param.shape[0]]
        # lazy dense Adam: step through which each row is up to date (all rows current while lazy mode is unused)
        self.last_step = np.zeros(rows, np.int32)
        self.lazy_pending = False
        self.step_count = 0
        # default: all cores but one (the thread that launches kernels must not be time-sliced against
        # 16 spinning workers)
        self.n_threads = n_threads if n_threads > 0 else max(1, (os.cpu_count() or 2) - 1)

    def batched_sparse_step(self, grad, rows, seg_begin, seg_end, seg_signal, signal_tensor_pinned, scale,
                            progress=None, seg_zero_grad=None, grad_packed=None, packed_base=0):
        """engine.py:320-327.  Blocking; run it on a worker thread (ctypes drops the GIL)."""
        self.step_count += 1
        ops.cpu_adam_batched_sparse_step(self.param, grad, self.exp_avg, self.exp_avg_sq, self.columns_lr, rows,
                                         seg_begin, seg_end, seg_signal, signal_tensor_pinned, self.betas[0],
                                         self.betas[1], self.eps, self.step_count, scale, self.n_threads, progress,
                                         seg_zero_grad, grad_packed, packed_base)

    def catch_up(self, rows=None, include_this_step=False):
        """Lazy dense Adam: replay the postponed zero-gradient steps of `rows` (int32 ids; None = every row) through
        the last finished step.  include_this_step: the caller is about to apply the next real step to exactly these
        rows, so they are marked as current through it."""
        n = self.param.shape[0]
        ops.cpu_adam_lazy_catch_up(self.param, self.exp_avg, self.exp_avg_sq, self.columns_lr, rows,
                                   self.last_step[:n], self.step_count, self.step_count + (1 if include_this_step else 0),
                                   self.betas[0], self.betas[1], self.eps, self.n_threads)
        if rows is None:
            self.lazy_pending = False

    def resize(self, param):
        """After densify / prune: the parameter view changed length; the moment buffers were compacted in place."""
        self.param = param
        self.exp_avg = self.exp_avg_buffer[:param.shape[0]]
        self.exp_avg_sq = self.exp_avg_sq_buffer[:param.shape[0]]
        self.last_step[:] = self.step_count  # callers materialise before restructuring: every row is current


class UnifiedAdam:
    def __init__(self, gpu_adam, cpu_adam):
        self.gpu_adam, self.cpu_adam = gpu_adam, cpu_adam


def _pinned_zeros(shape):
    # allocated pinned directly (Tensor.pin_memory() would allocate pageable memory first and copy: at C4 that is
    # 78 GB of extra host traffic and, transiently, twice the memory), on 2 MiB pages where available (hostmem.py)
    return hostmem.pinned_tensor(shape, torch.float32, zero=True)


def _pinned_empty(shape):
    return hostmem.pinned_tensor(shape, torch.float32)


def get_expon_lr_func(lr_init, lr_final, lr_delay_steps=0, lr_delay_mult=1.0, max_steps=1000000):
    """utils/general_utils.py get_expon_lr_func (log-linear interpolation with optional delay)."""

    def helper(step):
        if step < 0 or (lr_init == 0.0 and lr_final == 0.0):
            return 0.0
        if lr_delay_steps > 0:
            delay_rate = lr_delay_mult + (1 - lr_delay_mult) * math.sin(0.5 * math.pi * min(max(step / lr_delay_steps, 0), 1))
        else:
            delay_rate = 1.0
        t = min(max(step / max_steps, 0), 1)
        return delay_rate * math.exp(math.log(lr_init) * (1 - t) + math.log(lr_final) * t)

    return helper


class GaussianModelCLMOffload:
    def __init__(self, sh_degree=3, device="cuda:0"):
        self.max_sh_degree = sh_degree
        self.active_sh_degree = sh_degree
        self.device = device
        self.optimizer = None
        self.args = None

    # -- construction ----------------------------------------------------------------
    def create_from_arrays(self, xyz, log_scale, raw_quat, opacity_logit, sh, capacity=None, grad_table=True):
        """sh may already be a pinned [capacity,48] torch tensor (large scenes) or a numpy array.
        grad_table=False: do not allocate the [capacity,48] pinned gradient table (parameters_grad_buffer becomes a
        one-row placeholder).  The clm / multi-GPU engines move gradients through a packed staging buffer and never
        touch the table -- it exists for the reference's call signature and for naive_offload -- so at 102 M
        Gaussians this saves 19.6 GB of pinned memory (96 GB over 8 ranks at 500 M)."""
        dev = self.device
        N = xyz.shape[0]
        self._xyz = torch.from_numpy(np.ascontiguousarray(xyz)).to(dev)
        self._opacity = torch.from_numpy(np.ascontiguousarray(opacity_logit).reshape(N, 1)).to(dev)
        self._scaling = torch.from_numpy(np.ascontiguousarray(log_scale)).to(dev)
        self._rotation = torch.from_numpy(np.ascontiguousarray(raw_quat)).to(dev)
        # the SH table may be a shard of the N Gaussians (multi-GPU: this rank's contiguous id range)
        n_rows = sh.shape[0]
        if isinstance(sh, torch.Tensor):
            self.parameters_buffer = sh
        else:
            self.parameters_buffer = _pinned_empty((capacity or n_rows, 48))
            self.parameters_buffer[:n_rows].copy_(torch.from_numpy(np.ascontiguousarray(sh)))
        self._parameters = self.parameters_buffer[:n_rows]
        self.parameters_grad_buffer = (_pinned_zeros(tuple(self.parameters_buffer.shape)) if grad_table
                                       else torch.zeros((1, 48), dtype=torch.float32, pin_memory=torch.cuda.is_available()))
        return self

    @property
    def n_gaussians(self):
        return self._xyz.shape[0]

    def all_parameters(self):
        return [self._xyz, self._opacity, self._scaling, self._rotation, self._parameters]

    # -- optimiser (GM:122-234) --------------------------------------------------------
    def training_setup(self, args: TrainArgs, gpu_shard=None):
        """gpu_shard = (lo, hi): multi-GPU, keep the moments of the GPU-resident attributes for that id range only
        (the pinned SH table / host Adam state passed to create_from_arrays is then that range's shard, too)."""
        self.args = args
        # densification statistics (GM:122-127, base_gaussian_model max_radii2D)
        self.max_radii2D = torch.zeros(self.n_gaussians, device=self.device)
        self.xyz_gradient_accum = torch.zeros(self.n_gaussians, 1, device=self.device)
        self.denom = torch.zeros(self.n_gaussians, 1, device=self.device)
        bsz = args.bsz
        groups = [
            _Group(name="xyz", params=[self._xyz], lr=args.position_lr_init * args.spatial_lr_scale),
            _Group(name="opacity", params=[self._opacity], lr=args.opacity_lr),
            _Group(name="scaling", params=[self._scaling], lr=args.scaling_lr),
            _Group(name="rotation", params=[self._rotation], lr=args.rotation_lr),
        ]
        columns_lr = np.array([args.feature_lr, args.feature_lr / 20.0], np.float64)
        betas, eps = [0.9, 0.999], 1e-15
        assert args.lr_scale_mode == "sqrt", "overlapped host Adam only supports sqrt lr scaling (engine.py:861)"
        lr_scale = math.sqrt(bsz)
        for g in groups:
            g["lr"] *= lr_scale
            g["eps"] = eps / lr_scale
            g["betas"] = [b ** bsz for b in betas]
        cpu = CpuAdam(self._parameters, [3, 45], columns_lr * lr_scale, [b ** bsz for b in betas], eps / lr_scale,
                      args.cpu_adam_threads, capacity=self.parameters_buffer.shape[0])
        self.percent_dense = args.percent_dense
        gpu = (GpuAdam(groups, args.sparse_adam) if gpu_shard is None
               else ShardedGpuAdam(groups, args.sparse_adam, int(gpu_shard[0]), int(gpu_shard[1])))
        self.optimizer = UnifiedAdam(gpu, cpu)
        self.xyz_scheduler_args = get_expon_lr_func(
            lr_init=args.position_lr_init * args.spatial_lr_scale * lr_scale,
            lr_final=args.position_lr_final * args.spatial_lr_scale * lr_scale,
            lr_delay_mult=args.position_lr_delay_mult, max_steps=args.position_lr_max_steps)
        return self

    # -- densify / prune / opacity reset ----------------------------
    _GPU_NAMES = ("xyz", "opacity", "scaling", "rotation")

    def _gpu_tensors(self):
        st = self.optimizer.gpu_adam.state
        p = [self._xyz, self._opacity, self._scaling, self._rotation]
        return p + [st[n]["exp_avg"] for n in self._GPU_NAMES] + [st[n]["exp_avg_sq"] for n in self._GPU_NAMES]

    def materialize_host_rows(self):
        """Lazy dense Adam (TrainArgs.lazy_dense_adam): bring every SH row and its moments up to the last finished
        step.  Call before reading `_parameters` / the host moments outside the training iteration (evaluation,
        densification and saving do it themselves).  No-op when nothing is postponed."""
        from . import engine
        engine.wait_host_adam(self)
        cpu = self.optimizer.cpu_adam if self.optimizer is not None else None
        if cpu is not None and cpu.lazy_pending:
            cpu.catch_up(None)

    @torch.no_grad()
    def _restructure(self, cls, noise=None, generator=None):
        """classify bytes -> new state.  One plan, one gather kernel for the GPU tensors and their moments, one in-place
        compaction of the pinned tables (strategies/clm_offload/gaussian_model.py:522-739 re-concatenates every
        tensor three times and compacts the pinned table with a temporary of the whole kept set)."""
        if getattr(self, "dist_ctx", None) is not None:
            # the sharded multi-GPU state cannot be restructured in place -- rows would have to migrate
            # between the ranks' host shards and every rank would need the same noise; refuse instead of corrupting
            raise RuntimeError("densify / prune is not supported on the Gaussian-sharded multi-GPU state "
                               "(clmgs.engine_dist); restructure on one GPU and re-shard")
        self.materialize_host_rows()  # joins a pipelined host update and replays postponed (lazy) updates
        dev = self._xyz.device
        n_old = self.n_gaussians
        src, noise_row, (nA, nB, nS, nC) = ops.densify_plan(cls)
        n_new = nA + nB + 2 * nC
        cap = self.parameters_buffer.shape[0]
        if n_new > cap:
            raise RuntimeError("densification needs %d rows but the pinned buffers hold %d "
                               "(the reference pre-allocates them too: gaussian_model.py:33-44)" % (n_new, cap))
        if nC > 0:
            if noise is None:
                noise = torch.randn((2 * nS, 3), device=dev, generator=generator)
            noise = noise.to(dev, torch.float32).contiguous()
            assert tuple(noise.shape) == (2 * nS, 3), (noise.shape, nS)
        widths = (3, 1, 3, 4) * 3
        outs = [torch.empty((n_new, w), dtype=torch.float32, device=dev) for w in widths]
        ops.densify_apply((nA, nB, nS, nC), src, noise_row, noise, self._gpu_tensors(), outs)
        src_host = src.cpu().numpy()
        ops.host_rows_densify(self.parameters_buffer, self.optimizer.cpu_adam.exp_avg_buffer,
                              self.optimizer.cpu_adam.exp_avg_sq_buffer, n_old, src_host, nA, nB + 2 * nC,
                              self.optimizer.cpu_adam.n_threads)
        self._xyz, self._opacity, self._scaling, self._rotation = outs[:4]
        for g, p in zip(self.optimizer.gpu_adam.param_groups, outs[:4]):
            g["params"][0] = p
        for k, n in enumerate(self._GPU_NAMES):
            self.optimizer.gpu_adam.state[n] = dict(exp_avg=outs[4 + k], exp_avg_sq=outs[8 + k])
        self._parameters = self.parameters_buffer[:n_new]
        self.optimizer.cpu_adam.resize(self._parameters)
        # densification_postfix (GM:728-738): the statistics restart from zero
        self.max_radii2D = torch.zeros(n_new, device=dev)
        self.xyz_gradient_accum = torch.zeros(n_new, 1, device=dev)
        self.denom = torch.zeros(n_new, 1, device=dev)
        self._identity_slots = None
        return dict(n_before=n_old, n_after=n_new, n_cloned=nB, n_split_parents=nS, n_children=2 * nC,
                    n_pruned=n_old - nA - nS)

    def densify_and_prune(self, max_grad, min_opacity, extent, max_screen_size, noise=None, generator=None):
        """strategies/base_gaussian_model.py:364-388 (clone + split + prune as one restructuring).
        ``noise`` [2 * n_split_parents, 3] standard normal replaces the internal torch.randn draw (tests)."""
        cls = ops.densify_classify(self.xyz_gradient_accum, self.denom, self._scaling, self._opacity, max_grad,
                                   self.percent_dense * extent, min_opacity,
                                   0.1 * extent if max_screen_size else None)
        return self._restructure(cls, noise, generator)

    def count_split_parents(self, max_grad, extent):
        """How many rows densify_and_prune would split now (sizes an externally supplied noise tensor)."""
        cls = ops.densify_classify(self.xyz_gradient_accum, self.denom, self._scaling, self._opacity, max_grad,
                                   self.percent_dense * extent, 0.0, None)
        return int(((cls & 2) != 0).sum().item())

    def prune_points(self, mask):
        """GM:578-611."""
        return self._restructure(ops.prune_mask_to_class(mask.to(self._xyz.device)))

    @torch.no_grad()
    def reset_opacity(self):
        """GM:488-520."""
        st = self.optimizer.gpu_adam.state["opacity"]
        ops.reset_opacity(self._opacity, st["exp_avg"], st["exp_avg_sq"], 0.01)

    def update_learning_rate(self, iteration):
        for g in self.optimizer.gpu_adam.param_groups:
            if g["name"] == "xyz":
                g["lr"] = self.xyz_scheduler_args(iteration)
                return g["lr"]


class Camera:
    """Fields of scene/cameras.py the hot path reads."""

    def __init__(self, viewmat, K, gt_image_u8, device="cuda:0", uid=0):
        vm = torch.as_tensor(viewmat, dtype=torch.float32)
        self.world_view_transform = vm.t().contiguous().to(device)  # transposed w2c, as the reference stores it
        self.viewmat = vm.contiguous().to(device)
        self.K = torch.as_tensor(K, dtype=torch.float32).contiguous().to(device)
        c2w = torch.linalg.inv(vm.double()).float()
        self.camtoworlds = c2w.unsqueeze(0).contiguous().to(device)
        self.campos = c2w[:3, 3].contiguous().to(device)
        self.original_image = torch.as_tensor(gt_image_u8).contiguous().to(device)  # uint8 [3,H,W]
        self.image_height, self.image_width = int(self.original_image.shape[1]), int(self.original_image.shape[2])
        self.uid = uid

    def create_k_on_gpu(self):
        return self.K
