"""Thin tensor-level front-end over the C ABI (include/clmgs.h).

Each function allocates its outputs with torch (device memory + the current
stream are the only things torch provides here) and calls exactly one or two
``clmgs_*`` entry points.  No math happens in python.
"""
import math

import torch

from . import _lib
from ._lib import call


def _st():
    return torch.cuda.current_stream().cuda_stream


def _bucket(n):
    """Round a row count up to 1/8-octave steps.  Per-camera sizes (visible Gaussians, tile
    intersections) differ a little from camera to camera; asking torch's caching allocator for the
    exact size makes it cudaMalloc fresh segments in steady state, and cudaMalloc waits for running
    kernels (measured: a 3 ms sort stalled 20 ms behind the SH-row transfer).  Bucketed sizes hit the
    cache."""
    n = int(n)
    if n <= 4096:
        return max(n, 1)
    step = 1 << (n.bit_length() - 4)
    return (n + step - 1) // step * step


def _empty(shape, dtype, device):
    shape = (shape,) if isinstance(shape, int) else tuple(shape)
    n = shape[0]
    return torch.empty((_bucket(n),) + shape[1:], dtype=dtype, device=device)[:n]


def _zeros(shape, dtype, device):
    t = _empty(shape, dtype, device)
    t.zero_()
    return t


def _scratch(nbytes, device):
    return torch.empty(_bucket(max(int(nbytes), 16)), dtype=torch.uint8, device=device)


_WORD_DTYPE = {1: torch.uint8, 2: torch.int16, 4: torch.int32, 8: torch.int64}


def word_bytes_for(B):
    return 1 if B <= 8 else 2 if B <= 16 else 4 if B <= 32 else 8


def _f32c(t):
    assert t.dtype == torch.float32, t.dtype
    return t.contiguous()


# --------------------------------------------------------------------------- S1
def cull(means, scales, quats, viewmats, Ks, width, height, activated=True, eps2d=0.3,
         near_plane=0.01, far_plane=1e10, radius_clip=0.0, packed_outputs=False):
    """Frustum cull for B cameras.  Returns dict(bitmap, counts (list[int]), gaussian_ids
    [nnz] i64 (camera-major, ascending), and with packed_outputs also camera_ids, radii,
    means2d, depths, conics).  One host sync (reading the B counts)."""
    means, scales, quats = _f32c(means), _f32c(scales), _f32c(quats)
    viewmats, Ks = _f32c(viewmats), _f32c(Ks)
    dev = means.device
    N, B = means.shape[0], viewmats.shape[0]
    wb = word_bytes_for(B)
    bitmap = torch.empty(N, dtype=_WORD_DTYPE[wb], device=dev)
    counts = torch.empty(B, dtype=torch.int64, device=dev)
    sbytes = call("clmgs_cull_scratch_bytes", N, B)
    scratch = _scratch(sbytes, dev)
    args = (means, scales, quats, int(activated), N, viewmats, Ks, B, int(width), int(height),
            float(eps2d), float(near_plane), float(far_plane), float(radius_clip))
    call("clmgs_cull_count", *args, bitmap, wb, counts, scratch, scratch.numel(), _st())
    counts_h = counts.tolist()  # host sync (the reference has the same one: packed_fwd.cu:236)
    nnz = sum(counts_h)
    gids = torch.empty(nnz, dtype=torch.int64, device=dev)
    out = dict(bitmap=bitmap, counts=counts_h, counts_dev=counts, gaussian_ids=gids, word_bytes=wb)
    cam = rad = m2 = dep = con = None
    if packed_outputs:
        cam = torch.empty(nnz, dtype=torch.int64, device=dev)
        rad = torch.empty(nnz, dtype=torch.int32, device=dev)
        m2 = torch.empty(nnz, 2, dtype=torch.float32, device=dev)
        dep = torch.empty(nnz, dtype=torch.float32, device=dev)
        con = torch.empty(nnz, 3, dtype=torch.float32, device=dev)
        out.update(camera_ids=cam, radii=rad, means2d=m2, depths=dep, conics=con)
    call("clmgs_cull_emit", *args, bitmap, wb, scratch, gids, cam, rad, m2, dep, con, _st())
    return out


# --------------------------------------------------------------------------- A4
def gather_activate(xyz, log_scale, raw_quat, opacity_logit, ids):
    n, dev = ids.shape[0], xyz.device
    f32 = torch.float32
    means = _empty((n, 3), f32, dev)
    scales = _empty((n, 3), f32, dev)
    quats = _empty((n, 4), f32, dev)
    opac = _empty((n,), f32, dev)
    call("clmgs_gather_activate", _f32c(xyz), _f32c(log_scale), _f32c(raw_quat), _f32c(opacity_logit),
         ids.contiguous(), n, means, scales, quats, opac, _st())
    return means, scales, quats, opac


def activate_bwd_scatter(log_scale, raw_quat, opacity_logit, ids, v_means, v_scales, v_quats,
                         v_opacities, g_xyz, g_log_scale, g_raw_quat, g_opacity_logit):
    call("clmgs_activate_bwd_scatter", log_scale, raw_quat, opacity_logit, ids, ids.shape[0],
         _f32c(v_means), _f32c(v_scales), _f32c(v_quats), _f32c(v_opacities), g_xyz, g_log_scale,
         g_raw_quat, g_opacity_logit, _st())


def densification_stats(ids, v_means2d, radii, width, height, max_radii2d, xyz_gradient_accum, denom):
    call("clmgs_densification_stats", ids, ids.shape[0], _f32c(v_means2d), radii.contiguous(), int(width), int(height),
         max_radii2d, xyz_gradient_accum, denom, _st())


# --------------------------------------------------------------------------- S3
def projection_fwd(means, quats, scales, viewmats, Ks, width, height, eps2d=0.3, near_plane=0.01,
                   far_plane=1e10, radius_clip=0.0, calc_compensations=False):
    means, quats, scales = _f32c(means), _f32c(quats), _f32c(scales)
    viewmats, Ks = _f32c(viewmats), _f32c(Ks)
    dev = means.device
    N, C = means.shape[0], viewmats.shape[0]
    CN = C * N
    radii = _empty((CN,), torch.int32, dev).view(C, N)
    means2d = _empty((CN, 2), torch.float32, dev).view(C, N, 2)
    depths = _empty((CN,), torch.float32, dev).view(C, N)
    conics = _empty((CN, 3), torch.float32, dev).view(C, N, 3)
    comp = _empty((CN,), torch.float32, dev).view(C, N) if calc_compensations else None
    call("clmgs_projection_fwd", C, N, means, quats, scales, viewmats, Ks, int(width), int(height),
         float(eps2d), float(near_plane), float(far_plane), float(radius_clip), radii, means2d, depths,
         conics, comp, _st())
    return radii, means2d, depths, conics, comp


def projection_bwd(means, quats, scales, viewmats, Ks, width, height, eps2d, radii, conics,
                   v_means2d, v_depths, v_conics):
    dev = means.device
    N, C = means.shape[0], viewmats.shape[0]
    v_means = _zeros((N, 3), torch.float32, dev)
    v_quats = _zeros((N, 4), torch.float32, dev)
    v_scales = _zeros((N, 3), torch.float32, dev)
    call("clmgs_projection_bwd", C, N, _f32c(means), _f32c(quats), _f32c(scales), _f32c(viewmats),
         _f32c(Ks), int(width), int(height), float(eps2d), radii.contiguous(), _f32c(conics),
         _f32c(v_means2d), None if v_depths is None else _f32c(v_depths), _f32c(v_conics), v_means, v_quats,
         v_scales, _st())
    return v_means, v_quats, v_scales


# --------------------------------------------------------------------------- SH
def sh_fwd(degree, dirs, coeffs, masks=None):
    dirs, coeffs = _f32c(dirs), _f32c(coeffs)
    K = coeffs.shape[-2]
    N = dirs.numel() // 3
    colors = torch.zeros_like(dirs) if masks is not None else torch.empty_like(dirs)
    m = None if masks is None else masks.contiguous().view(torch.uint8)
    call("clmgs_sh_fwd", N, K, int(degree), dirs, coeffs, m, colors, _st())
    return colors


def sh_bwd(degree, dirs, coeffs, v_colors, v_coeffs=None, masks=None, compute_v_dirs=True):
    """v_coeffs given -> accumulate in place (clm_kernels semantics); else fresh tensor."""
    dirs, coeffs, v_colors = _f32c(dirs), _f32c(coeffs), _f32c(v_colors)
    K = coeffs.shape[-2]
    N = dirs.numel() // 3
    accumulate = v_coeffs is not None
    if v_coeffs is None:
        v_coeffs = torch.zeros_like(coeffs)
    assert v_coeffs.is_contiguous()
    v_dirs = torch.empty_like(dirs) if compute_v_dirs else None
    m = None if masks is None else masks.contiguous().view(torch.uint8)
    call("clmgs_sh_bwd", N, K, int(degree), dirs, coeffs, m, v_colors, v_coeffs, int(accumulate), v_dirs, _st())
    return v_coeffs, v_dirs


# --------------------------------------------------------------------------- S4
def isect_tiles(means2d, radii, depths, tile_size, tile_width, tile_height, sort=True,
                n_cameras=None, camera_ids=None):
    """-> tiles_per_gauss i32 (same shape as radii), isect_ids i64 [n], flatten_ids i32 [n].
    One host sync to read n_isects (reference: isect_tiles.cu:166)."""
    means2d, depths = _f32c(means2d), _f32c(depths)
    radii = radii.contiguous()
    dev = means2d.device
    packed = camera_ids is not None
    if packed:
        C, N = int(n_cameras), radii.shape[0]
        total = N
    else:
        C, N = radii.shape
        total = C * N
    tpg = _empty((max(total, 1),), torch.int32, dev)[:total].view(radii.shape)
    cum = _empty((max(total, 1),), torch.int64, dev)
    sc = _scratch(call("clmgs_isect_count_scratch_bytes", total), dev)
    call("clmgs_isect_count", total, means2d, radii, int(tile_size), int(tile_width), int(tile_height),
         tpg, cum, sc, sc.numel(), _st())
    n_isects = int(cum[total - 1].item()) if total > 0 else 0
    ids = _empty((n_isects,), torch.int64, dev)
    flat = _empty((n_isects,), torch.int32, dev)
    sc2 = _scratch(call("clmgs_isect_sort_scratch_bytes", total, n_isects) if sort else 0, dev)
    call("clmgs_isect_emit_sort", C, N, camera_ids, means2d, radii, depths, cum, int(tile_size),
         int(tile_width), int(tile_height), n_isects, int(bool(sort)), ids, flat, sc2, sc2.numel(), _st())
    return tpg, ids, flat


def isect_sorted_offsets(means2d, radii, depths, tile_size, tile_width, tile_height):
    """The engine's tile stage: depth-sorted flatten_ids per tile + tile offsets, without materialising the
    64-bit isect_ids (same flatten_ids / offsets as isect_tiles + isect_offset_encode).
    -> (flatten_ids [n_isects] int32, offsets [C, th, tw] int32, n_isects)."""
    means2d, depths = _f32c(means2d), _f32c(depths)
    radii = radii.contiguous()
    dev = means2d.device
    C, N = radii.shape
    total = C * N
    tpg = _empty((max(total, 1),), torch.int32, dev)
    cum = _empty((max(total, 1),), torch.int64, dev)
    sc = _scratch(call("clmgs_isect_count_scratch_bytes", total), dev)
    call("clmgs_isect_count", total, means2d, radii, int(tile_size), int(tile_width), int(tile_height),
         tpg, cum, sc, sc.numel(), _st())
    n_isects = int(cum[total - 1].item()) if total > 0 else 0
    flat = _empty((n_isects,), torch.int32, dev)
    off = torch.empty(C, tile_height, tile_width, dtype=torch.int32, device=dev)
    sc2 = _scratch(call("clmgs_isect_sort_scratch_bytes", total, n_isects), dev)
    call("clmgs_isect_sorted_offsets", C, N, None, means2d, radii, depths, cum, int(tile_size), int(tile_width),
         int(tile_height), n_isects, flat, off, sc2, sc2.numel(), _st())
    return flat, off, n_isects


def isect_count_async(means2d, radii, tile_size, tile_width, tile_height):
    """First half of the engine's tile stage without a host sync: tile counts + their inclusive prefix on the
    device, and the total copied to a pinned word behind an event.  Pass the result to isect_sort_counted() a few
    launches later -- by then the event has fired and reading the total costs nothing (the reference blocks on
    cum[-1].item() right here, isect_tiles.cu:165-166)."""
    means2d = _f32c(means2d)
    radii = radii.contiguous()
    dev = means2d.device
    C, N = radii.shape
    total = C * N
    tpg = _empty((max(total, 1),), torch.int32, dev)
    cum = _empty((max(total, 1),), torch.int64, dev)
    count = torch.zeros(1, dtype=torch.int64, pin_memory=True)
    ev = torch.cuda.Event()
    if total > 0:
        sc = _scratch(call("clmgs_isect_count_scratch_bytes", total), dev)
        call("clmgs_isect_count", total, means2d, radii, int(tile_size), int(tile_width), int(tile_height),
             tpg, cum, sc, sc.numel(), _st())
        # by a kernel, not a memcpy: the D2H copy engine is busy with the gradient flush's bulk DMA
        call("clmgs_store_i64_to_host", count, cum[total - 1:total], _st())
    ev.record(torch.cuda.current_stream())
    return dict(means2d=means2d, radii=radii, cum=cum, tpg=tpg, count=count, event=ev, C=C, N=N)


def isect_sort_counted(state, depths, tile_size, tile_width, tile_height):
    """Second half: depth-sorted flatten_ids per tile + tile offsets (see isect_sorted_offsets)."""
    state["event"].synchronize()
    n_isects = int(state["count"][0])
    depths = _f32c(depths)
    dev = depths.device
    C, N = state["C"], state["N"]
    flat = _empty((n_isects,), torch.int32, dev)
    off = torch.empty(C, tile_height, tile_width, dtype=torch.int32, device=dev)
    sc2 = _scratch(call("clmgs_isect_sort_scratch_bytes", C * N, n_isects), dev)
    call("clmgs_isect_sorted_offsets", C, N, None, state["means2d"], state["radii"], depths, state["cum"], int(tile_size),
         int(tile_width), int(tile_height), n_isects, flat, off, sc2, sc2.numel(), _st())
    return flat, off, n_isects


def isect_offset_encode(isect_ids, n_cameras, tile_width, tile_height):
    off = torch.empty(n_cameras, tile_height, tile_width, dtype=torch.int32, device=isect_ids.device)
    call("clmgs_isect_offset_encode", isect_ids.shape[0], isect_ids.contiguous(), int(n_cameras),
         int(tile_width), int(tile_height), off, _st())
    return off


# --------------------------------------------------------------------------- S5
def rasterize_fwd(means2d, conics, colors, opacities, width, height, tile_size, isect_offsets,
                  flatten_ids, backgrounds=None, work_counters=None):
    """work_counters: optional int64 [64, 3] device tensor whose column sums P_f / evaluated / P_b are ACCUMULATED
    into (clmgs_rasterize_fwd_counted; measurement only)."""
    means2d, conics, colors, opacities = map(_f32c, (means2d, conics, colors, opacities))
    dev = means2d.device
    C, th, tw = isect_offsets.shape
    N = opacities.shape[-1]
    rc = torch.empty(C, height, width, 3, device=dev)
    ra = torch.empty(C, height, width, 1, device=dev)
    last = torch.empty(C, height, width, dtype=torch.int32, device=dev)
    bg = None if backgrounds is None else _f32c(backgrounds)
    if work_counters is not None:
        assert work_counters.dtype == torch.int64 and work_counters.numel() == 192 and work_counters.is_contiguous()
        call("clmgs_rasterize_fwd_counted", C, N, flatten_ids.shape[0], means2d, conics, colors, opacities, bg,
             int(width), int(height), int(tile_size), tw, th, isect_offsets.contiguous(),
             flatten_ids.contiguous(), rc, ra, last, work_counters, _st())
        return rc, ra, last
    call("clmgs_rasterize_fwd", C, N, flatten_ids.shape[0], means2d, conics, colors, opacities, bg,
         int(width), int(height), int(tile_size), tw, th, isect_offsets.contiguous(),
         flatten_ids.contiguous(), rc, ra, last, _st())
    return rc, ra, last


def rasterize_bwd(means2d, conics, colors, opacities, width, height, tile_size, isect_offsets,
                  flatten_ids, render_alphas, last_ids, v_render_colors, v_render_alphas=None,
                  backgrounds=None, absgrad=False):
    means2d, conics, colors, opacities = map(_f32c, (means2d, conics, colors, opacities))
    dev = means2d.device
    C, th, tw = isect_offsets.shape
    N = opacities.shape[-1]
    CN = opacities.numel()   # dense [C, N] or packed [nnz]: the kernels index the flattened arrays either way
    v_means2d = _zeros((CN, 2), torch.float32, dev).view(means2d.shape)
    v_conics = _zeros((CN, 3), torch.float32, dev).view(conics.shape)
    v_colors = _zeros((CN, 3), torch.float32, dev).view(colors.shape)
    v_opac = _zeros((CN,), torch.float32, dev).view(opacities.shape)
    v_abs = _zeros((CN, 2), torch.float32, dev).view(means2d.shape) if absgrad else None
    bg = None if backgrounds is None else _f32c(backgrounds)
    vra = None if v_render_alphas is None else _f32c(v_render_alphas)
    call("clmgs_rasterize_bwd", C, N, flatten_ids.shape[0], means2d, conics, colors, opacities, bg,
         int(width), int(height), int(tile_size), tw, th, isect_offsets.contiguous(),
         flatten_ids.contiguous(), _f32c(render_alphas), last_ids.contiguous(), _f32c(v_render_colors),
         vra, v_abs, v_means2d, v_conics, v_colors, v_opac, _st())
    return v_means2d, v_conics, v_colors, v_opac, v_abs


# --------------------------------------------------------------------------- loss
def loss_fwd_bwd(image, gt_u8, lambda_dssim=0.2, hwc=True):
    """image [H,W,3] (hwc) or [3,H,W]; gt_u8 uint8 [3,H,W] -> (loss_vec[3] device: loss,l1,ssim; v_image)."""
    image = _f32c(image)
    if hwc:
        H, W, _ = image.shape
    else:
        _, H, W = image.shape
    out = torch.empty(3, device=image.device)
    v = torch.empty_like(image)
    sc = _scratch(call("clmgs_loss_scratch_bytes", H, W), image.device)
    call("clmgs_loss_fwd_bwd", H, W, image, int(hwc), gt_u8.contiguous(), float(lambda_dssim), out, v, sc,
         sc.numel(), _st())
    return out, v


def fused_ssim_fwd_bwd(img1, img2, want_grad=True):
    """clm_kernels.fused_ssim on float [B,CH,H,W]: -> (ssim scalar tensor, d mean/d img1 or None)."""
    img1, img2 = _f32c(img1), _f32c(img2)
    B, CH, H, W = img1.shape
    out = torch.empty(1, device=img1.device)
    v = torch.empty_like(img1) if want_grad else None
    sc = _scratch(256 + (B * CH * 3 * H * W * 4 if want_grad else 0), img1.device)
    call("clmgs_fused_ssim", B, CH, H, W, img1, img2, out, v, sc, sc.numel(), _st())
    return out[0], v


# --------------------------------------------------------------------------- S2 / S6 rows
def rows_gather(d_rows, h_rows, ids):
    call("clmgs_rows_gather_i64", d_rows, h_rows, ids.contiguous(), ids.shape[0], _st())


def rows_gather_retention(d_rows, h_rows, d_prev, host_ids, prev_src_rows, dst_from_host, dst_from_prev):
    call("clmgs_rows_gather_retention", d_rows, h_rows, d_prev, host_ids, prev_src_rows, dst_from_host,
         dst_from_prev, host_ids.shape[0], prev_src_rows.shape[0], _st())


def rows_scatter_grad(d_grad, h_grad, ids, accum):
    call("clmgs_rows_scatter_grad_i64", d_grad, h_grad, ids.contiguous(), ids.shape[0], int(accum), _st())


def rows_scatter_grad_retention(d_grad, h_grad, d_next_grad, host_ids, next_dst_rows, src_to_host, src_to_next,
                                accum):
    call("clmgs_rows_scatter_grad_retention", d_grad, h_grad, d_next_grad, host_ids, next_dst_rows, src_to_host,
         src_to_next, host_ids.shape[0], next_dst_rows.shape[0], int(accum), _st())


def rows_accumulate(d_dst_rows, d_src_rows, dst_index):
    call("clmgs_rows_accumulate", d_dst_rows, d_src_rows, dst_index.contiguous(), dst_index.shape[0], _st())


def set_signal(h_flags, index, value):
    call("clmgs_set_signal", h_flags, int(index), int(value), _st())


def rows_prefetch_next(d_next, h_rows, d_this, ids_this, ids_next, bitmap, bit_this):
    n_next = ids_next.shape[0]
    src_map = _empty((max(n_next, 1),), torch.int32, d_next.device)
    call("clmgs_rows_prefetch_next", d_next, h_rows, d_this, ids_this, 0 if ids_this is None else ids_this.shape[0],
         ids_next, n_next, bitmap, bitmap.element_size(), int(bit_this), src_map, _st())
    return src_map


def rows_grad_handoff(d_grad, d_next_grad, h_grad, ids_this, n_next, src_map, bitmap, bit_this, bit_next, accum=2):
    call("clmgs_rows_grad_handoff", d_grad, d_next_grad, h_grad, ids_this, ids_this.shape[0], int(n_next), src_map,
         bitmap, bitmap.element_size(), int(bit_this), int(bit_next), int(accum), _st())


# --------------------------------------------------------------------------- A2 bitmaps
def scatter_to_bit(bitmap, ids, bit):
    call("clmgs_scatter_to_bit", bitmap, bitmap.element_size(), ids.contiguous(), ids.shape[0], int(bit), _st())


def extract_ffs(bitmap, out=None):
    if out is None:
        out = torch.empty(bitmap.shape[0], dtype=torch.uint8, device=bitmap.device)
    call("clmgs_extract_ffs", bitmap, bitmap.element_size(), out, bitmap.shape[0], _st())
    return out


def compute_cnt_d(bitmap, bsz):
    cnt = torch.empty(max(bsz - 1, 1), dtype=torch.int32, device=bitmap.device)
    call("clmgs_compute_cnt_d", bitmap, bitmap.element_size(), int(bsz), bitmap.shape[0], cnt, _st())
    return cnt[: bsz - 1]


def bitmap_permute(bitmap, bsz, order):
    import numpy as np
    out = torch.empty_like(bitmap)
    order = np.ascontiguousarray(order, dtype=np.int32)
    call("clmgs_bitmap_permute", bitmap, out, bitmap.element_size(), int(bsz), order, bitmap.shape[0], _st())
    return out


def bitmap_distance(bitmap, bsz, sample_ids):
    buf = torch.empty(2 * bsz * bsz, dtype=torch.int32, device=bitmap.device)
    call("clmgs_bitmap_distance", bitmap, bitmap.element_size(), int(bsz), sample_ids.contiguous(),
         sample_ids.shape[0], buf, _st())
    return buf[: bsz * bsz].view(bsz, bsz)


def group_by_ffs(ffs, bsz):
    N, dev = ffs.shape[0], ffs.device
    out_ids = torch.empty(N, dtype=torch.int32, device=dev)
    counts = torch.empty(bsz + 1, dtype=torch.int64, device=dev)
    sc = _scratch(call("clmgs_group_by_ffs_scratch_bytes", N), dev)
    call("clmgs_group_by_ffs", ffs, N, int(bsz), out_ids, counts, sc, sc.numel(), _st())
    return out_ids, counts


# --------------------------------------------------------------------------- A16 / A15 optimisers
def selective_adam(param, grad, exp_avg, exp_avg_sq, mask, lr, b1, b2, eps, grad_scale=1.0):
    N = mask.shape[0]
    M = param.numel() // N
    m = mask if mask.dtype == torch.uint8 else mask.view(torch.uint8)
    call("clmgs_selective_adam", param, grad, exp_avg, exp_avg_sq, m, float(lr), float(b1), float(b2), float(eps),
         float(grad_scale), N, M, _st())


def adam_dense(param, grad, exp_avg, exp_avg_sq, lr, b1, b2, eps, step, grad_scale=1.0):
    call("clmgs_adam_dense", param, grad, exp_avg, exp_avg_sq, float(lr), float(b1), float(b2), float(eps),
         float(grad_scale), int(step), param.numel(), _st())


def cpu_adam_batched_sparse_step(param, grad, exp_avg, exp_avg_sq, col_lr, rows, seg_begin, seg_end, seg_signal,
                                 signal, b1, b2, eps, step, scale, n_threads=0, progress=None, seg_zero_grad=None,
                                 grad_packed=None, packed_base=0):
    """All arguments are HOST tensors/arrays (pinned where the GPU also touches them).  Blocking.
    Segment s = rows[seg_begin[s]:seg_end[s]], processed after signal[seg_signal[s]] != 0 (no wait if < 0).
    grad_packed: gradients come from this packed staging buffer (row r - packed_base for list entry r) instead of
    the table `grad` (which is then not touched)."""
    import numpy as np
    seg_begin = np.ascontiguousarray(seg_begin, np.int64)
    seg_end = np.ascontiguousarray(seg_end, np.int64)
    seg_signal = np.ascontiguousarray(seg_signal, np.int32)
    zg = None if seg_zero_grad is None else np.ascontiguousarray(seg_zero_grad, np.int32)
    if grad_packed is not None:
        call("clmgs_cpu_adam_batched_sparse_step_packed", param, grad_packed, int(packed_base), exp_avg, exp_avg_sq,
             int(param.shape[1]), col_lr, rows, seg_begin, seg_end, seg_signal, zg, len(seg_begin), signal, float(b1),
             float(b2), float(eps), int(step), float(scale), int(n_threads), progress)
        return
    call("clmgs_cpu_adam_batched_sparse_step", param, grad, exp_avg, exp_avg_sq, int(param.shape[1]), col_lr, rows,
         seg_begin, seg_end, seg_signal, zg, len(seg_begin), signal, float(b1), float(b2), float(eps), int(step),
         float(scale), int(n_threads), progress)


def sh_colors_fwd(degree, means, campos, coeffs):
    """Engine-fused: colours = clamp_min(SH(means - campos) + 0.5, 0).  coeffs [n,48] or [n,K,3]."""
    n = means.shape[0]
    K = coeffs.numel() // (3 * n) if n else 16
    colors = _empty((n, 3), torch.float32, means.device)
    call("clmgs_sh_colors_fwd", n, K, int(degree), means, campos, coeffs, colors, _st())
    return colors


def sh_colors_bwd(degree, means, campos, coeffs, v_colors, v_coeffs, v_means):
    n = means.shape[0]
    K = coeffs.numel() // (3 * n) if n else 16
    call("clmgs_sh_colors_bwd", n, K, int(degree), means, campos, coeffs, _f32c(v_colors), v_coeffs, v_means, _st())


def tsp_tour(dist):
    import numpy as np
    d = np.ascontiguousarray(dist, dtype=np.int32)
    n = d.shape[0]
    tour = np.empty(n, np.int32)
    call("clmgs_tsp_tour", d, n, tour)
    return tour.tolist()


# --------------------------------------------------------------------------- batch-resident rows (slots)
def extract_first_use(bitmap, bsz):
    out = torch.empty(bitmap.shape[0], dtype=torch.uint8, device=bitmap.device)
    call("clmgs_extract_first_use", bitmap, bitmap.element_size(), int(bsz), out, bitmap.shape[0], _st())
    return out


def slot_map(ids32, slot_of):
    call("clmgs_slot_map", ids32, ids32.shape[0], slot_of, _st())


def rows_gather_i32(d_rows, h_rows, ids32):
    call("clmgs_rows_gather_i32", d_rows, h_rows, ids32, ids32.shape[0], _st())


def rows_gather_to_slots(d_rows, h_rows, ids32, slot_of):
    call("clmgs_rows_gather_to_slots", d_rows, h_rows, ids32, ids32.shape[0], slot_of, _st())


def rows_flush_slots(d_grad_rows, slot_of, ids32, h_grad):
    call("clmgs_rows_flush_slots", d_grad_rows, slot_of, ids32, ids32.shape[0], h_grad, _st())


def sh_colors_fwd_slots(degree, means, campos, d_rows, ids, slot_of):
    n = means.shape[0]
    colors = _empty((n, 3), torch.float32, means.device)
    call("clmgs_sh_colors_fwd_slots", n, int(degree), means, campos, d_rows, ids, slot_of, colors, _st())
    return colors


def sh_colors_bwd_slots(degree, means, campos, d_rows, v_colors, d_grad_rows, ids, slot_of, v_means):
    call("clmgs_sh_colors_bwd_slots", means.shape[0], int(degree), means, campos, d_rows, _f32c(v_colors), d_grad_rows,
         ids, slot_of, v_means, _st())


# --------------------------------------------------------------------------- densify / prune / opacity reset
def densify_classify(grad_accum, denom, log_scale, opacity_logit, grad_threshold, dense_threshold, min_opacity,
                     world_size_threshold=None):
    N = log_scale.shape[0]
    cls = torch.empty(N, dtype=torch.uint8, device=log_scale.device)
    use_ws = world_size_threshold is not None
    call("clmgs_densify_classify", N, _f32c(grad_accum), _f32c(denom), _f32c(log_scale), _f32c(opacity_logit),
         float(grad_threshold), float(dense_threshold), float(min_opacity), float(world_size_threshold or 0.0),
         int(use_ws), cls, _st())
    return cls


def prune_mask_to_class(mask):
    m = mask.to(torch.uint8).contiguous()
    cls = torch.empty_like(m)
    call("clmgs_prune_mask_to_class", m.shape[0], m, cls, _st())
    return cls


def densify_plan(cls):
    """-> (src int32 [nA+nB+2nC] device, noise_row int32 [nC] device, (nA, nB, nS, nC) host ints).  Synchronises."""
    N, dev = cls.shape[0], cls.device
    totals = torch.zeros(4, dtype=torch.int64, device=dev)
    src = torch.empty(max(3 * N, 1), dtype=torch.int32, device=dev)  # worst case: every row cloned + split
    noise_row = torch.empty(max(N, 1), dtype=torch.int32, device=dev)
    nbytes = call("clmgs_densify_plan_scratch_bytes", N)
    scratch = _scratch(nbytes, dev)
    call("clmgs_densify_plan", N, cls, src, noise_row, totals, scratch, scratch.numel(), _st())
    nA, nB, nS, nC = (int(x) for x in totals.cpu().tolist())
    return src[:nA + nB + 2 * nC], noise_row[:nC], (nA, nB, nS, nC)


def densify_apply(counts, src, noise_row, noise, tensors_in, tensors_out):
    """tensors_in / tensors_out: 12 tensors {xyz, opacity, log_scale, quat, exp_avg x4, exp_avg_sq x4}."""
    import numpy as np
    nA, nB, nS, nC = counts
    assert len(tensors_in) == 12 and len(tensors_out) == 12
    for t in list(tensors_in) + list(tensors_out):
        assert t.dtype == torch.float32 and t.is_contiguous()
    pin = np.array([t.data_ptr() for t in tensors_in], np.uint64)
    pout = np.array([t.data_ptr() for t in tensors_out], np.uint64)
    call("clmgs_densify_apply", nA, nB, nC, nS, src, noise_row, noise, pin, pout, _st())


def host_rows_densify(param_buffer, exp_avg_buffer, exp_avg_sq_buffer, n_old, src_host, nA, n_tail, n_threads=0):
    import numpy as np
    src_host = np.ascontiguousarray(src_host, np.int32)
    cols, cap = param_buffer.shape[1], param_buffer.shape[0]
    assert exp_avg_buffer.shape[0] >= cap and exp_avg_sq_buffer.shape[0] >= cap
    call("clmgs_host_rows_densify", param_buffer, exp_avg_buffer, exp_avg_sq_buffer, cols, int(n_old), int(cap),
         src_host, int(nA), int(n_tail), int(n_threads))


def reset_opacity(opacity_logit, exp_avg, exp_avg_sq, max_opacity=0.01):
    call("clmgs_reset_opacity", opacity_logit.shape[0], opacity_logit, exp_avg, exp_avg_sq, float(max_opacity), _st())


def radix_sort_pairs_u32(keys, vals, end_bit=32):
    """Stable sort of (uint32-as-int32 tensor keys, int32 vals) on the low end_bit bits -> (keys_sorted, vals_sorted)."""
    import numpy as np
    n, dev = keys.shape[0], keys.device
    k0, v0 = keys.contiguous().clone(), vals.contiguous().clone()
    k1, v1 = torch.empty_like(k0), torch.empty_like(v0)
    sc = _scratch(call("clmgs_radix_sort_pairs_u32_scratch_bytes", n), dev)
    flag = np.zeros(1, np.int32)
    call("clmgs_radix_sort_pairs_u32", n, k0, k1, v0, v1, int(end_bit), sc, sc.numel(), flag, _st())
    return (k1, v1) if flag[0] else (k0, v0)


def cpu_adam_lazy_catch_up(param, exp_avg, exp_avg_sq, col_lr, rows, last_step, through_step, mark_step, b1, b2, eps,
                           n_threads=0):
    """rows: int32 numpy / CPU tensor of row ids, or None for all rows of `param`."""
    import numpy as np
    if rows is None:
        rows_arg, n = None, param.shape[0]
    else:
        rows_arg = rows if hasattr(rows, "data_ptr") else np.ascontiguousarray(rows, np.int32)
        n = rows_arg.shape[0]
    call("clmgs_cpu_adam_lazy_catch_up", param, exp_avg, exp_avg_sq, param.shape[1], np.ascontiguousarray(col_lr, np.float32),
         rows_arg, int(n), last_step, int(through_step), int(mark_step), float(b1), float(b2), float(eps), int(n_threads))
