"""clm_kernels-shaped operators over libclmgs.so: the names and arguments of
submodules/clm_kernels/clm_kernels/__init__.py:4-100 (pybind table ext.cpp:7-23).
The grid/block arguments are accepted and ignored (the library sizes its own launches)."""
import torch

from . import ops


def send_shs2gpu_stream(d_parameters, h_parameters, mask_indices, grid_size=None, block_size=None):
    ops.rows_gather(d_parameters, h_parameters, mask_indices)


def send_shs2gpu_stream_retention(d_parameters, h_parameters, r_parameters, host_indices, rtnt_indices,
                                  param_indices_from_host, param_indices_from_rtnt, grid_size_H=None, block_size_H=None,
                                  grid_size_D=None, block_size_D=None):
    ops.rows_gather_retention(d_parameters, h_parameters, r_parameters, host_indices, rtnt_indices,
                              param_indices_from_host, param_indices_from_rtnt)


def send_shs2cpu_grad_buffer_stream(d_parameters, h_parameters, mask_indices, accum, grid_size=None, block_size=None):
    ops.rows_scatter_grad(d_parameters, h_parameters, mask_indices, accum)


def send_shs2cpu_grad_buffer_stream_retention(d_parameters, h_parameters, r_parameters, host_indices, rtnt_indices,
                                              grad_indices_to_host, grad_indices_to_rtnt, accum, grid_size_H=None,
                                              block_size_H=None, grid_size_D=None, block_size_D=None):
    ops.rows_scatter_grad_retention(d_parameters, h_parameters, r_parameters, host_indices, rtnt_indices,
                                    grad_indices_to_host, grad_indices_to_rtnt, accum)


def set_signal(signal_tensor, microbatch_idx, signal):
    ops.set_signal(signal_tensor, microbatch_idx, signal)


def scatter_to_bit(bitmap, filter, bit):
    ops.scatter_to_bit(bitmap, filter, bit)


def extract_ffs(input, output):
    ops.extract_ffs(input, output)


def compute_cnt_h(bitmap, tmp_buffer, grid_size=None, blk_size=None):
    """Reference: per-thread partials in tmp_buffer [(bsz-1), grid*blk], summed by the caller
    (engine.py:231-232).  Here column 0 receives the totals and the rest is zeroed, so the caller's
    torch.sum(tmp_buffer, dim=1) gives the same cnt_d."""
    bsz = tmp_buffer.shape[0] + 1
    cnt = ops.compute_cnt_d(bitmap, bsz)
    tmp_buffer.zero_()
    tmp_buffer[:, 0] = cnt


class _FusedSSIM(torch.autograd.Function):
    @staticmethod
    def forward(ctx, img1, img2):
        s, v = ops.fused_ssim_fwd_bwd(img1, img2, want_grad=img1.requires_grad)
        ctx.save_for_backward(v)
        return s

    @staticmethod
    def backward(ctx, g):
        (v,) = ctx.saved_tensors
        return (None if v is None else v * g), None


def fused_ssim(img1, img2, padding="same", train=True):
    """clm_kernels.fused_ssim (__init__.py:54-61): mean of the SSIM map, differentiable w.r.t. img1."""
    if padding != "same":
        raise NotImplementedError("only padding='same' is on the CLM path (base_engine.py:93)")
    return _FusedSSIM.apply(img1.contiguous(), img2.contiguous())


def selective_adam_update(param, param_grad, exp_avg, exp_avg_sq, tiles_touched, lr, b1, b2, eps, N, M):
    ops.selective_adam(param, param_grad, exp_avg, exp_avg_sq, tiles_touched, lr, b1, b2, eps)


@torch.no_grad()
def spherical_harmonics_bwd_inplace(degrees_to_use, dirs, coeffs, v_coeffs, v_colors, masks=None):
    """__init__.py:64-85: v_coeffs += ..., returns v_dirs."""
    _, v_dirs = ops.sh_bwd(degrees_to_use, dirs, coeffs, v_colors, v_coeffs=v_coeffs, masks=masks)
    return v_dirs
