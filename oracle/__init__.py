"""CPU oracle for the CLM-GS hot path -- TEST INFRASTRUCTURE ONLY.

numpy front-end over ``clmgs_oracle.c`` (plain C, gcc).  Only ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs may import this package; the product under
``clm-gs_b200/`` never does (tests/test_layout.py greps for it).

Each wrapper cites the reference file:line its C body restates.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "libclmgs_oracle.so")


def build(force=False):
    src = os.path.join(_HERE, "clmgs_oracle.c")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-s", "-B" if force else "-s"])
    return _SO


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = C.CDLL(_SO)
        _lib.orc_expf.restype = C.c_float
        _lib.orc_expf.argtypes = [C.c_float]
        _lib.orc_isect_count.restype = C.c_int64
        _lib.orc_loss_fwd_bwd.restype = C.c_float
        _lib.orc_tile_n_bits.restype = C.c_int
    return _lib


def _p(a):
    if a is None:
        return None
    assert a.flags["C_CONTIGUOUS"], "oracle wants contiguous arrays"
    return a.ctypes.data_as(C.c_void_p)


def _f32(a):
    return None if a is None else np.ascontiguousarray(a, dtype=np.float32)


f = C.c_float
i32 = C.c_int
i64 = C.c_int64


def activate(log_scale=None, raw_quat=None, opacity_logit=None):
    """exp / normalize / sigmoid (strategies/base_gaussian_model.py:42-50)."""
    ls, rq, ol = _f32(log_scale), _f32(raw_quat), _f32(opacity_logit)
    n = next(a for a in (ls, rq, ol) if a is not None).shape[0]
    s = np.empty_like(ls) if ls is not None else None
    q = np.empty_like(rq) if rq is not None else None
    o = np.empty_like(ol) if ol is not None else None
    lib().orc_activate(i64(n), _p(ls), _p(rq), _p(ol), _p(s), _p(q), _p(o))
    return s, q, o


def cull_mask(means, scales, quats, viewmats, Ks, width, height, eps2d=0.3, near=0.01,
              far=1e10, radius_clip=0.0):
    """[B,N] uint8 keep mask (CSRC/fully_fused_projection_packed_fwd.cu:52-131)."""
    means, scales, quats, viewmats, Ks = map(_f32, (means, scales, quats, viewmats, Ks))
    N, B = means.shape[0], viewmats.shape[0]
    mask = np.zeros((B, N), np.uint8)
    lib().orc_cull_mask(i64(N), i32(B), _p(means), _p(scales), _p(quats), _p(viewmats), _p(Ks),
                        i32(width), i32(height), f(eps2d), f(near), f(far), f(radius_clip), _p(mask))
    return mask


def projection_fwd(means, quats, scales, viewmats, Ks, width, height, eps2d=0.3, near=0.01,
                   far=1e10, radius_clip=0.0, calc_compensations=False):
    """Dense projection (CSRC/fully_fused_projection_fwd.cu:19-129). Culled rows are 0."""
    means, scales, quats, viewmats, Ks = map(_f32, (means, scales, quats, viewmats, Ks))
    N, Cn = means.shape[0], viewmats.shape[0]
    radii = np.zeros((Cn, N), np.int32)
    means2d = np.zeros((Cn, N, 2), np.float32)
    depths = np.zeros((Cn, N), np.float32)
    conics = np.zeros((Cn, N, 3), np.float32)
    comp = np.zeros((Cn, N), np.float32) if calc_compensations else None
    lib().orc_projection_fwd(i32(Cn), i64(N), _p(means), _p(quats), _p(scales), _p(viewmats), _p(Ks),
                             i32(width), i32(height), f(eps2d), f(near), f(far), f(radius_clip),
                             _p(radii), _p(means2d), _p(depths), _p(conics), _p(comp))
    return radii, means2d, depths, conics, comp


def projection_bwd(means, quats, scales, viewmats, Ks, width, height, radii, conics,
                   v_means2d, v_depths, v_conics):
    """VJP of the dense projection (CSRC/fully_fused_projection_bwd.cu:19-185)."""
    means, scales, quats, viewmats, Ks = map(_f32, (means, scales, quats, viewmats, Ks))
    conics, v_means2d, v_depths, v_conics = map(_f32, (conics, v_means2d, v_depths, v_conics))
    radii = np.ascontiguousarray(radii, np.int32)
    N, Cn = means.shape[0], viewmats.shape[0]
    v_means = np.zeros((N, 3), np.float32)
    v_quats = np.zeros((N, 4), np.float32)
    v_scales = np.zeros((N, 3), np.float32)
    lib().orc_projection_bwd(i32(Cn), i64(N), _p(means), _p(quats), _p(scales), _p(viewmats), _p(Ks),
                             i32(width), i32(height), _p(radii), _p(conics), _p(v_means2d),
                             _p(v_depths), _p(v_conics), _p(v_means), _p(v_quats), _p(v_scales))
    return v_means, v_quats, v_scales


def sh_fwd(degree, dirs, coeffs, masks=None):
    """CSRC/compute_sh_fwd.cu:10-29 + CSRC/spherical_harmonics.cuh:11-91."""
    dirs, coeffs = _f32(dirs), _f32(coeffs)
    K = coeffs.shape[-2]
    N = dirs.size // 3
    colors = np.zeros(dirs.shape, np.float32)
    m = None if masks is None else np.ascontiguousarray(masks, np.uint8)
    lib().orc_sh_fwd(i64(N), i32(K), i32(degree), _p(dirs), _p(coeffs), _p(m), _p(colors))
    return colors


def sh_bwd_inplace(degree, dirs, coeffs, v_coeffs, v_colors, masks=None):
    """CK/compute_sh_bwd.cu:276-306: v_coeffs += ..., returns v_dirs."""
    dirs, coeffs, v_colors = _f32(dirs), _f32(coeffs), _f32(v_colors)
    assert v_coeffs.dtype == np.float32 and v_coeffs.flags["C_CONTIGUOUS"]
    K = coeffs.shape[-2]
    N = dirs.size // 3
    v_dirs = np.zeros(dirs.shape, np.float32)
    m = None if masks is None else np.ascontiguousarray(masks, np.uint8)
    lib().orc_sh_bwd(i64(N), i32(K), i32(degree), _p(dirs), _p(coeffs), _p(m), _p(v_colors),
                     _p(v_coeffs), _p(v_dirs))
    return v_dirs


def isect_tiles(means2d, radii, depths, tile_size, tile_width, tile_height, sort=True):
    """CSRC/isect_tiles.cu:15-232 -> (tiles_per_gauss i32 [C,N], isect_ids i64, flatten_ids i32)."""
    means2d, depths = _f32(means2d), _f32(depths)
    radii = np.ascontiguousarray(radii, np.int32)
    Cn, N = radii.shape
    tpg = np.zeros((Cn, N), np.int32)
    n = lib().orc_isect_count(i32(Cn), i64(N), _p(means2d), _p(radii), i32(tile_size),
                              i32(tile_width), i32(tile_height), _p(tpg))
    ids = np.zeros(n, np.int64)
    flat = np.zeros(n, np.int32)
    lib().orc_isect_emit_sort(i32(Cn), i64(N), _p(means2d), _p(radii), _p(depths), i32(tile_size),
                              i32(tile_width), i32(tile_height), i32(1 if sort else 0), _p(ids), _p(flat))
    return tpg, ids, flat


def isect_offset_encode(isect_ids, n_cameras, tile_width, tile_height):
    """CSRC/isect_tiles.cu:234-301."""
    ids = np.ascontiguousarray(isect_ids, np.int64)
    off = np.zeros((n_cameras, tile_height, tile_width), np.int32)
    lib().orc_isect_offset_encode(i64(ids.shape[0]), _p(ids), i32(n_cameras), i32(tile_width),
                                  i32(tile_height), _p(off))
    return off


def rasterize_fwd(means2d, conics, colors, opacities, width, height, tile_size, isect_offsets,
                  flatten_ids, backgrounds=None):
    """CSRC/rasterize_to_pixels_fwd.cu:14-160 -> colors [C,H,W,3], alphas [C,H,W,1], last_ids, examined."""
    means2d, conics, colors, opacities = map(_f32, (means2d, conics, colors, opacities))
    bg = _f32(backgrounds)
    off = np.ascontiguousarray(isect_offsets, np.int32)
    flat = np.ascontiguousarray(flatten_ids, np.int32)
    Cn, th, tw = off.shape
    N = opacities.shape[-1]
    rc = np.zeros((Cn, height, width, 3), np.float32)
    ra = np.zeros((Cn, height, width, 1), np.float32)
    last = np.zeros((Cn, height, width), np.int32)
    ex = i64(0)
    lib().orc_rasterize_fwd(i32(Cn), i64(N), i64(flat.shape[0]), _p(means2d), _p(conics), _p(colors),
                            _p(opacities), _p(bg), i32(width), i32(height), i32(tile_size), i32(tw),
                            i32(th), _p(off), _p(flat), _p(rc), _p(ra), _p(last), C.byref(ex))
    return rc, ra, last, ex.value


def rasterize_bwd(means2d, conics, colors, opacities, width, height, tile_size, isect_offsets,
                  flatten_ids, render_alphas, last_ids, v_render_colors, v_render_alphas=None,
                  backgrounds=None, absgrad=False):
    """CSRC/rasterize_to_pixels_bwd.cu:14-249 -> v_means2d, v_conics, v_colors, v_opacities[, abs]."""
    means2d, conics, colors, opacities = map(_f32, (means2d, conics, colors, opacities))
    bg = _f32(backgrounds)
    off = np.ascontiguousarray(isect_offsets, np.int32)
    flat = np.ascontiguousarray(flatten_ids, np.int32)
    ra, vrc, vra = _f32(render_alphas), _f32(v_render_colors), _f32(v_render_alphas)
    last = np.ascontiguousarray(last_ids, np.int32)
    Cn, th, tw = off.shape
    N = opacities.shape[-1]
    vm = np.zeros((Cn, N, 2), np.float32)
    vc = np.zeros((Cn, N, 3), np.float32)
    vcol = np.zeros((Cn, N, 3), np.float32)
    vo = np.zeros((Cn, N), np.float32)
    vabs = np.zeros((Cn, N, 2), np.float32) if absgrad else None
    lib().orc_rasterize_bwd(i32(Cn), i64(N), i64(flat.shape[0]), _p(means2d), _p(conics), _p(colors),
                            _p(opacities), _p(bg), i32(width), i32(height), i32(tile_size), i32(tw),
                            i32(th), _p(off), _p(flat), _p(ra), _p(last), _p(vrc), _p(vra), _p(vabs),
                            _p(vm), _p(vc), _p(vcol), _p(vo))
    return vm, vc, vcol, vo, vabs


def loss_fwd_bwd(image_chw, gt_u8_chw, lambda_dssim=0.2, want_grad=True):
    """0.8*L1 + 0.2*(1-SSIM) and d/d image (strategies/base_engine.py:79-94, CK/ssim.cu:187-366)."""
    img = _f32(image_chw)
    gt = np.ascontiguousarray(gt_u8_chw, np.uint8)
    _, H, W = img.shape
    v = np.zeros_like(img) if want_grad else None
    ssim, l1 = f(0), f(0)
    loss = lib().orc_loss_fwd_bwd(i32(H), i32(W), _p(img), _p(gt), f(lambda_dssim), _p(v),
                                  C.byref(ssim), C.byref(l1))
    return float(loss), v, float(ssim.value), float(l1.value)


def selective_adam(param, grad, exp_avg, exp_avg_sq, mask, lr, b1, b2, eps):
    """CK/adam.cu:12-41 (in place; no bias correction)."""
    N = mask.shape[0]
    M = param.size // N
    m = np.ascontiguousarray(mask, np.uint8)
    lib().orc_selective_adam(_p(param), _p(_f32(grad)), _p(exp_avg), _p(exp_avg_sq), _p(m), f(lr),
                             f(b1), f(b2), f(eps), i64(N), i32(M))


def adam_dense(param, grad, exp_avg, exp_avg_sq, lr, b1, b2, eps, step):
    """torch.optim.Adam(fused=True) update with bias correction (optimizer.py:127-129)."""
    lib().orc_adam_dense(_p(param), _p(_f32(grad)), _p(exp_avg), _p(exp_avg_sq), f(lr), f(b1), f(b2),
                         f(eps), i64(step), i64(param.size))


def cpu_adam_rows(param, grad, exp_avg, exp_avg_sq, rows, col_lr, b1, b2, eps, step, scale,
                  zero_grad=True):
    """cpu-adam batched_sparse_step body (source missing; semantics from optimizer.py:130-142, ENG:320-327)."""
    rows = np.ascontiguousarray(rows, np.int32)
    col_lr = _f32(col_lr)
    lib().orc_cpu_adam_rows(_p(param), _p(grad), _p(exp_avg), _p(exp_avg_sq), _p(rows),
                            i64(rows.shape[0]), i32(param.shape[1]), _p(col_lr), f(b1), f(b2), f(eps),
                            i64(step), f(scale), i32(1 if zero_grad else 0))


def expf(x):
    x = np.ascontiguousarray(x, np.float32)
    out = np.empty_like(x)
    L = lib()
    for k, v in enumerate(x.ravel()):
        out.ravel()[k] = L.orc_expf(float(v))
    return out
