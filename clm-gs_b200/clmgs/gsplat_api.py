"""gsplat-shaped operators over libclmgs.so -- same names, arguments and return tuples as
``gsplat.cuda._wrapper`` (submodules/gsplat/gsplat/cuda/_wrapper.py) for the five ops CLM uses
(engine.py:7-13) plus the ``rasterization()`` composite (gsplat/rendering.py:17-402).

Differentiable through torch.autograd.Function exactly where the reference's are
(_wrapper.py:691-927,1115-1145), including the packed projection (ids from the cull kernel, backward
through the dense projection-backward kernel) with optional sparse gradients.  Any number of colour
channels (the 3-channel kernels run once per group of three: alpha compositing is linear in the
colours, so results and gradients are those of one wide pass), all five render modes (RGB, D, ED, RGB+D,
RGB+ED).  Restrictions, raised as errors (not silently ignored): tile_size 16, quats/scales (not
precomputed covars), "classic" rasterize mode.
"""
import math

import torch

from . import ops


class _Projection(torch.autograd.Function):
    @staticmethod
    def forward(ctx, means, quats, scales, viewmats, Ks, width, height, eps2d, near_plane, far_plane, radius_clip,
                calc_compensations):
        radii, means2d, depths, conics, comp = ops.projection_fwd(means, quats, scales, viewmats, Ks, width, height,
                                                                  eps2d, near_plane, far_plane, radius_clip,
                                                                  calc_compensations)
        ctx.save_for_backward(means, quats, scales, viewmats, Ks, radii, conics)
        ctx.dims = (width, height, eps2d)
        ctx.mark_non_differentiable(radii)
        return radii, means2d, depths, conics, comp

    @staticmethod
    def backward(ctx, v_radii, v_means2d, v_depths, v_conics, v_comp):
        means, quats, scales, viewmats, Ks, radii, conics = ctx.saved_tensors
        width, height, eps2d = ctx.dims
        if v_comp is not None and v_comp.abs().sum() != 0:
            raise NotImplementedError("gradient through compensations is not on the CLM path")
        z = lambda t, ref: torch.zeros_like(ref) if t is None else t.contiguous()
        v_means, v_quats, v_scales = ops.projection_bwd(means, quats, scales, viewmats, Ks, width, height, eps2d, radii,
                                                        conics, z(v_means2d, conics[..., :2]), z(v_depths, conics[..., 0]),
                                                        z(v_conics, conics))
        return v_means, v_quats, v_scales, None, None, None, None, None, None, None, None, None


class _ProjectionPacked(torch.autograd.Function):
    """_FullyFusedProjectionPacked (_wrapper.py:930-1112).  Forward = the one-pass cull with packed outputs (the
    PACKED radius rule, fully_fused_projection_packed_fwd.cu:117-119).  Backward: the packed upstream gradients are
    scattered into dense [C, N] tensors and go through the dense projection-backward kernel, which recomputes the
    forward from means / quats / scales and only needs `radii > 0` to know which pairs exist."""

    @staticmethod
    def forward(ctx, means, quats, scales, viewmats, Ks, width, height, eps2d, near_plane, far_plane, radius_clip,
                sparse_grad):
        out = ops.cull(means, scales, quats, viewmats, Ks, width, height, True, eps2d, near_plane, far_plane,
                       radius_clip, packed_outputs=True)
        cam, gid = out["camera_ids"], out["gaussian_ids"]
        ctx.save_for_backward(means, quats, scales, viewmats, Ks, cam, gid, out["radii"], out["conics"])
        ctx.cfg = (width, height, eps2d, sparse_grad)
        ctx.mark_non_differentiable(cam, gid, out["radii"])
        return cam, gid, out["radii"], out["means2d"], out["depths"], out["conics"]

    @staticmethod
    def backward(ctx, v_cam, v_gid, v_radii, v_means2d, v_depths, v_conics):
        means, quats, scales, viewmats, Ks, cam, gid, radii, conics = ctx.saved_tensors
        width, height, eps2d, sparse_grad = ctx.cfg
        C, N = viewmats.shape[0], means.shape[0]
        flat = cam * N + gid

        def dense(src, tail, dtype=torch.float32):
            t = torch.zeros((C * N,) + tail, dtype=dtype, device=means.device)
            if src is not None:
                t[flat] = src.to(dtype)
            return t.view((C, N) + tail)

        v_means, v_quats, v_scales = ops.projection_bwd(means, quats, scales, viewmats, Ks, width, height, eps2d,
                                                        dense(radii, (), torch.int32), dense(conics, (3,)),
                                                        dense(v_means2d, (2,)), dense(v_depths, ()), dense(v_conics, (3,)))
        if sparse_grad:  # _wrapper.py:1078-1101: COO gradients over the Gaussians that are visible somewhere
            rows = torch.unique(gid)
            sp = lambda g: torch.sparse_coo_tensor(rows[None], g[rows], size=g.shape, is_coalesced=True,
                                                   check_invariants=False)
            v_means, v_quats, v_scales = sp(v_means), sp(v_quats), sp(v_scales)
        return v_means, v_quats, v_scales, None, None, None, None, None, None, None, None, None


def fully_fused_projection(means, covars, quats, scales, viewmats, Ks, width, height, eps2d=0.3, near_plane=0.01,
                           far_plane=1e10, radius_clip=0.0, packed=False, sparse_grad=False,
                           calc_compensations=False):
    """_wrapper.py:141-219.  dense -> (radii [C,N] i32, means2d, depths, conics, compensations);
    packed -> (camera_ids i64, gaussian_ids i64, radii, means2d, depths, conics, compensations)."""
    if covars is not None:
        raise NotImplementedError("precomputed covars are not on the CLM path; pass quats and scales")
    if sparse_grad and not packed:
        raise ValueError("sparse_grad is only supported with packed=True (as in gsplat, _wrapper.py:190-191)")
    if packed:
        if calc_compensations:
            raise NotImplementedError("compensations (antialiased mode) are not on the CLM path")
        res = _ProjectionPacked.apply(means.contiguous(), quats.contiguous(), scales.contiguous(), viewmats.contiguous(),
                                      Ks.contiguous(), int(width), int(height), eps2d, near_plane, far_plane, radius_clip,
                                      bool(sparse_grad))
        return tuple(res) + (None,)
    return _Projection.apply(means.contiguous(), quats.contiguous(), scales.contiguous(), viewmats.contiguous(),
                             Ks.contiguous(), int(width), int(height), eps2d, near_plane, far_plane, radius_clip,
                             calc_compensations)


class _SphericalHarmonics(torch.autograd.Function):
    @staticmethod
    def forward(ctx, degree, dirs, coeffs, masks):
        colors = ops.sh_fwd(degree, dirs, coeffs, masks)
        ctx.save_for_backward(dirs, coeffs, masks)
        ctx.degree = degree
        return colors

    @staticmethod
    def backward(ctx, v_colors):
        dirs, coeffs, masks = ctx.saved_tensors
        v_coeffs, v_dirs = ops.sh_bwd(ctx.degree, dirs, coeffs, v_colors.contiguous(), None, masks,
                                      compute_v_dirs=ctx.needs_input_grad[1])
        return None, v_dirs, v_coeffs, None


def spherical_harmonics(degrees_to_use, dirs, coeffs, masks=None):
    """_wrapper.py:17-47."""
    assert (degrees_to_use + 1) ** 2 <= coeffs.shape[-2], coeffs.shape
    assert dirs.shape[:-1] == coeffs.shape[:-2] and dirs.shape[-1] == 3 and coeffs.shape[-1] == 3
    return _SphericalHarmonics.apply(int(degrees_to_use), dirs.contiguous(), coeffs.contiguous(), masks)


@torch.no_grad()
def isect_tiles(means2d, radii, depths, tile_size, tile_width, tile_height, sort=True, packed=False, n_cameras=None,
                camera_ids=None, gaussian_ids=None):
    """_wrapper.py:277-348."""
    if packed:
        return ops.isect_tiles(means2d, radii, depths, tile_size, tile_width, tile_height, sort, n_cameras, camera_ids)
    return ops.isect_tiles(means2d, radii, depths, tile_size, tile_width, tile_height, sort)


@torch.no_grad()
def isect_offset_encode(isect_ids, n_cameras, tile_width, tile_height):
    """_wrapper.py:351-368."""
    return ops.isect_offset_encode(isect_ids, n_cameras, tile_width, tile_height)


class _RasterizeToPixels(torch.autograd.Function):
    @staticmethod
    def forward(ctx, means2d, conics, colors, opacities, backgrounds, width, height, tile_size, isect_offsets,
                flatten_ids, absgrad):
        rc, ra, last = ops.rasterize_fwd(means2d, conics, colors, opacities, width, height, tile_size, isect_offsets,
                                         flatten_ids, backgrounds)
        ctx.save_for_backward(means2d, conics, colors, opacities, backgrounds, isect_offsets, flatten_ids, ra, last)
        ctx.cfg = (width, height, tile_size, absgrad)
        return rc, ra

    @staticmethod
    def backward(ctx, v_rc, v_ra):
        means2d, conics, colors, opacities, backgrounds, offsets, flat, ra, last = ctx.saved_tensors
        width, height, tile_size, absgrad = ctx.cfg
        v_means2d, v_conics, v_colors, v_opac, v_abs = ops.rasterize_bwd(
            means2d, conics, colors, opacities, width, height, tile_size, offsets, flat, ra, last, v_rc.contiguous(),
            None if v_ra is None else v_ra.contiguous(), backgrounds, absgrad)
        if absgrad:
            means2d.absgrad = v_abs
        v_bg = None
        if backgrounds is not None and ctx.needs_input_grad[4]:
            v_bg = (v_rc * (1.0 - ra).float()).sum(dim=(1, 2))
        return v_means2d, v_conics, v_colors, v_opac, v_bg, None, None, None, None, None, None


def rasterize_to_pixels(means2d, conics, colors, opacities, image_width, image_height, tile_size, isect_offsets,
                        flatten_ids, backgrounds=None, packed=False, absgrad=False):
    """_wrapper.py:371-497 -> (render_colors [C,H,W,3], render_alphas [C,H,W,1])."""
    D = colors.shape[-1]
    if backgrounds is not None:
        assert backgrounds.shape[-1] == D, (backgrounds.shape, D)
    means2d, conics, opacities = means2d.contiguous(), conics.contiguous(), opacities.contiguous()
    if D == 3:
        return _RasterizeToPixels.apply(means2d, conics, colors.contiguous(), opacities, backgrounds, int(image_width),
                                        int(image_height), int(tile_size), isect_offsets, flatten_ids, absgrad)
    # Any other channel count: groups of three through the 3-channel kernels (the reference pads to its next
    # compiled width and chunks above channel_chunk, rendering.py:329-358).  Compositing is linear in the colours, the
    # alpha / transmittance recursion does not depend on them, and autograd adds the groups' gradients of means2d /
    # conics / opacities -- the same sums one wide kernel forms.  The alphas (and their gradient) come from group 0.
    outs, alphas = [], None
    for c0 in range(0, D, 3):
        w = min(3, D - c0)
        col = colors[..., c0:c0 + w]
        bg = None if backgrounds is None else backgrounds[..., c0:c0 + w]
        if w < 3:
            col = torch.nn.functional.pad(col, (0, 3 - w))
            bg = None if bg is None else torch.nn.functional.pad(bg, (0, 3 - w))
        rc, ra = _RasterizeToPixels.apply(means2d, conics, col.contiguous(), opacities,
                                          None if bg is None else bg.contiguous(), int(image_width), int(image_height),
                                          int(tile_size), isect_offsets, flatten_ids, absgrad and c0 == 0)
        outs.append(rc[..., :w])
        if alphas is None:
            alphas = ra
    return torch.cat(outs, dim=-1), alphas


def rasterization(means, quats, scales, opacities, colors, viewmats, Ks, width, height, near_plane=0.01,
                  far_plane=1e10, radius_clip=0.0, eps2d=0.3, sh_degree=None, packed=True, tile_size=16,
                  backgrounds=None, render_mode="RGB", sparse_grad=False, absgrad=False, rasterize_mode="classic",
                  channel_chunk=32):
    """gsplat.rasterization() (gsplat/rendering.py:17-402): same arguments, same returns (render_colors [C,H,W,X],
    render_alphas [C,H,W,1], meta with the keys of rendering.py:384-401).  packed=True really is the packed data
    flow (meta["camera_ids"] / meta["gaussian_ids"], [nnz, ...] intermediates); render_mode RGB / D / ED / RGB+D /
    RGB+ED; colours [N,D] / [C,N,D] or SH coefficients [N,K,3] / [C,N,K,3] with sh_degree."""
    assert render_mode in ("RGB", "D", "ED", "RGB+D", "RGB+ED"), render_mode
    if rasterize_mode != "classic":
        raise NotImplementedError("rasterize_mode='antialiased' (compensations) is not on the CLM path")
    C, N = viewmats.shape[0], means.shape[0]
    assert means.shape == (N, 3) and quats.shape == (N, 4) and scales.shape == (N, 3) and opacities.shape == (N,)
    assert viewmats.shape == (C, 4, 4) and Ks.shape == (C, 3, 3)
    if sh_degree is None:
        assert (colors.dim() == 2 and colors.shape[0] == N) or (colors.dim() == 3 and colors.shape[:2] == (C, N)), colors.shape
    else:
        assert (colors.dim() == 3 and colors.shape[0] == N and colors.shape[2] == 3) or (
            colors.dim() == 4 and colors.shape[:2] == (C, N) and colors.shape[3] == 3), colors.shape
        assert (sh_degree + 1) ** 2 <= colors.shape[-2], colors.shape
    proj = fully_fused_projection(means, None, quats, scales, viewmats, Ks, width, height, eps2d, near_plane, far_plane,
                                  radius_clip, packed, sparse_grad, False)
    if packed:
        camera_ids, gaussian_ids, radii, means2d, depths, conics, _ = proj
        opac = opacities[gaussian_ids]                       # [nnz]
    else:
        radii, means2d, depths, conics, _ = proj
        opac = opacities.repeat(C, 1)                        # [C, N]
        camera_ids = gaussian_ids = None
    tile_width, tile_height = math.ceil(width / float(tile_size)), math.ceil(height / float(tile_size))
    tiles_per_gauss, isect_ids, flatten_ids = isect_tiles(means2d, radii, depths, tile_size, tile_width, tile_height,
                                                          packed=packed, n_cameras=C, camera_ids=camera_ids,
                                                          gaussian_ids=gaussian_ids)
    isect_offsets = isect_offset_encode(isect_ids, C, tile_width, tile_height)
    if sh_degree is None:
        if packed:
            cols = colors[gaussian_ids] if colors.dim() == 2 else colors[camera_ids, gaussian_ids]
        else:
            cols = colors.expand(C, -1, -1) if colors.dim() == 2 else colors
    else:
        camtoworlds = torch.inverse(viewmats)
        if packed:
            dirs = means[gaussian_ids, :] - camtoworlds[camera_ids, :3, 3]
            shs = colors[gaussian_ids] if colors.dim() == 3 else colors[camera_ids, gaussian_ids]
        else:
            dirs = means[None, :, :] - camtoworlds[:, None, :3, 3]
            shs = colors.expand(C, -1, -1, -1) if colors.dim() == 3 else colors
        cols = spherical_harmonics(sh_degree, dirs.contiguous(), shs.contiguous(), masks=radii > 0)
        cols = torch.clamp_min(cols + 0.5, 0.0)              # rendering.py:315
    if render_mode in ("RGB+D", "RGB+ED"):
        cols = torch.cat((cols, depths[..., None]), dim=-1)
        if backgrounds is not None:
            backgrounds = torch.cat([backgrounds, torch.zeros(C, 1, device=backgrounds.device)], dim=-1)
    elif render_mode in ("D", "ED"):
        cols = depths[..., None]
        if backgrounds is not None:
            backgrounds = torch.zeros(C, 1, device=backgrounds.device)
    render_colors, render_alphas = rasterize_to_pixels(means2d, conics, cols, opac, width, height, tile_size,
                                                       isect_offsets, flatten_ids, backgrounds, packed=packed,
                                                       absgrad=absgrad)
    if render_mode in ("ED", "RGB+ED"):   # accumulated depth -> expected depth (rendering.py:373-381)
        render_colors = torch.cat([render_colors[..., :-1], render_colors[..., -1:] / render_alphas.clamp(min=1e-10)],
                                  dim=-1)
    meta = dict(camera_ids=camera_ids, gaussian_ids=gaussian_ids, radii=radii, means2d=means2d, depths=depths,
                conics=conics, opacities=opac, tile_width=tile_width, tile_height=tile_height,
                tiles_per_gauss=tiles_per_gauss, isect_ids=isect_ids, flatten_ids=flatten_ids,
                isect_offsets=isect_offsets, width=width, height=height, tile_size=tile_size)
    return render_colors, render_alphas, meta
