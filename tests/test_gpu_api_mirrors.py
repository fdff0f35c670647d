"""The reference-shaped python operators (gsplat / clm_kernels names) run through autograd
and agree with the oracle: a reference-style forward+backward of one camera, written the way
strategies/clm_offload/engine.py:30-127 writes it."""
import math

import numpy as np
import pytest
import torch

import oracle
from oracle import train_step as TS
from clmgs import gsplat_api as G, clm_kernels_api as CK, scene as S, model

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def test_reference_style_pipeline_with_autograd():
    W, H, n = 112, 80, 15_000
    spec = S.SceneSpec("t", n, W, H, 1, "object", 60.0, 0)
    g = S.make_gaussians(spec)
    V, K, P = S.make_cameras(spec, 1, g["extent"])
    gt = S.make_gt_images(spec, 1)[0]
    st = TS.new_state(g)
    sc, qu, op = oracle.activate(g["log_scale"], g["raw_quat"], g["opacity_logit"][:, 0])
    ids = np.nonzero(oracle.cull_mask(g["xyz"], sc, qu, V, K, W, H)[0])[0]
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(DEV)
    xyz = t(g["xyz"][ids]).requires_grad_()
    ls = t(g["log_scale"][ids]).requires_grad_()
    rq = t(g["raw_quat"][ids]).requires_grad_()
    ol = t(g["opacity_logit"][ids]).requires_grad_()
    shs = t(g["sh"][ids])
    viewmat, Kt = t(V[0]), t(K[0])
    # --- engine.py:51-119, with the reference's own call shapes ---------------------------------
    radii, means2d, depths, conics, _ = G.fully_fused_projection(
        means=xyz, covars=None, quats=torch.nn.functional.normalize(rq, dim=-1), scales=torch.exp(ls),
        viewmats=viewmat.unsqueeze(0), Ks=Kt.unsqueeze(0), width=W, height=H, packed=False)
    means2d.retain_grad()
    dirs = xyz[None] - t(P[0])[None, None]
    with torch.no_grad():
        col0 = G.spherical_harmonics(3, dirs, shs.reshape(1, -1, 16, 3))
    col_det = col0.detach().requires_grad_()
    colors = torch.clamp_min(col_det + 0.5, 0.0)
    tw, th = math.ceil(W / 16.0), math.ceil(H / 16.0)
    _, isect_ids, flatten_ids = G.isect_tiles(means2d, radii, depths, 16, tw, th, packed=False)
    offsets = G.isect_offset_encode(isect_ids, 1, tw, th)
    img, _ = G.rasterize_to_pixels(means2d, conics, colors, torch.sigmoid(ol).squeeze(1).unsqueeze(0), W, H, 16,
                                   offsets, flatten_ids, backgrounds=None)
    img = img.squeeze(0).permute(2, 0, 1).contiguous()
    gt_f = torch.clamp(t(gt) / 255.0, 0.0, 1.0)
    ssim = CK.fused_ssim(img.unsqueeze(0), gt_f.unsqueeze(0))
    loss = 0.8 * torch.abs(img - gt_f).mean() + 0.2 * (1.0 - ssim)
    loss.backward()
    shs_grad = torch.zeros(len(ids), 48, device=DEV)
    v_dirs = CK.spherical_harmonics_bwd_inplace(3, dirs, shs.reshape(1, -1, 16, 3), shs_grad, col_det.grad)
    dirs.backward(v_dirs)
    # --- oracle --------------------------------------------------------------------------------------
    fw = TS.render_camera(st, ids, V[0], K[0], P[0], W, H, 3)
    ref_loss, v_chw, _, _ = oracle.loss_fwd_bwd(np.ascontiguousarray(fw["rc"][0].transpose(2, 0, 1)), gt)
    grads = {k: np.zeros_like(st[k]) for k in ("xyz", "log_scale", "raw_quat", "opacity_logit")}
    sh_g = np.zeros_like(st["sh"])
    TS.backward_camera(st, ids, fw, V[0], K[0], W, H, 3, np.ascontiguousarray(v_chw.transpose(1, 2, 0)), grads, sh_g)
    assert abs(loss.item() - ref_loss) < 2e-5
    rel = lambda a, b: np.abs(a - b).max() / max(np.abs(b).max(), 1e-20)
    assert rel(xyz.grad.cpu().numpy(), grads["xyz"][ids]) < 1e-3
    assert rel(ls.grad.cpu().numpy(), grads["log_scale"][ids]) < 1e-3
    assert rel(rq.grad.cpu().numpy(), grads["raw_quat"][ids]) < 1e-3
    assert rel(ol.grad.cpu().numpy(), grads["opacity_logit"][ids]) < 1e-3
    assert rel(shs_grad.cpu().numpy(), sh_g[ids]) < 1e-3


@pytest.mark.parametrize("sh_degree", [None, 3])
def test_rasterization_composite(sh_degree):
    W, H, n = 96, 64, 8_000
    spec = S.SceneSpec("t", n, W, H, 2, "object", 60.0, 0)
    g = S.make_gaussians(spec)
    V, K, P = S.make_cameras(spec, 2, g["extent"])
    sc, qu, op = oracle.activate(g["log_scale"], g["raw_quat"], g["opacity_logit"][:, 0])
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(DEV)
    colors = t(g["sh"].reshape(n, 16, 3)) if sh_degree is not None else t(np.abs(g["sh"][:, :3]))
    rc, ra, meta = G.rasterization(t(g["xyz"]), t(qu), t(sc), t(op), colors, t(V), t(K), W, H, sh_degree=sh_degree,
                                   packed=False)
    assert rc.shape == (2, H, W, 3) and ra.shape == (2, H, W, 1)
    assert set(["radii", "means2d", "depths", "conics", "isect_offsets", "flatten_ids", "tile_width"]) <= set(meta)
    # oracle image for camera 1
    radii, m2, dep, con, _ = oracle.projection_fwd(g["xyz"], qu, sc, V, K, W, H)
    tw, th = (W + 15) // 16, (H + 15) // 16
    _, ids, flat = oracle.isect_tiles(m2, radii, dep, 16, tw, th)
    off = oracle.isect_offset_encode(ids, 2, tw, th)
    if sh_degree is None:
        cols = np.broadcast_to(np.abs(g["sh"][:, :3])[None], (2, n, 3)).copy()
    else:
        cols = np.stack([np.maximum(oracle.sh_fwd(3, g["xyz"] - P[c][None], g["sh"].reshape(n, 16, 3)) + 0.5, 0)
                         for c in range(2)]).astype(np.float32)
    ref, ref_a, _, _ = oracle.rasterize_fwd(m2, con, cols, np.broadcast_to(op[None], (2, n)).copy(), W, H, 16, off, flat)
    assert (np.abs(rc.cpu().numpy() - ref).max(-1) > 1e-4).mean() <= 1e-4
    np.testing.assert_array_equal(meta["flatten_ids"].cpu().numpy(), flat)



def _composite_inputs(n=8_000, W=96, H=64, C=2):
    spec = S.SceneSpec("t", n, W, H, C, "object", 60.0, 0)
    g = S.make_gaussians(spec)
    V, K, P = S.make_cameras(spec, C, g["extent"])
    sc, qu, op = oracle.activate(g["log_scale"], g["raw_quat"], g["opacity_logit"][:, 0])
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(DEV)
    return dict(means=t(g["xyz"]), quats=t(qu), scales=t(sc), opac=t(op), shs=t(g["sh"].reshape(n, 16, 3)), V=t(V), K=t(K),
                W=W, H=H, n=n, C=C)


@pytest.mark.parametrize("sh_degree", [None, 3])
def test_rasterization_packed_equals_dense_with_gradients(sh_degree):
    """gsplat's own test strategy (tests/test_basic.py:257-404, tests/test_rasterization.py): packed and dense data
    flows give the same image and the same gradients; meta carries camera_ids / gaussian_ids in packed mode."""
    d = _composite_inputs()
    res = {}
    for packed in (False, True):
        leaves = [d[k].clone().requires_grad_() for k in ("means", "quats", "scales", "opac")]
        col = (d["shs"] if sh_degree is not None else d["shs"][:, 0, :].abs()).clone().requires_grad_()
        rc, ra, meta = G.rasterization(*leaves, col, d["V"], d["K"], d["W"], d["H"], sh_degree=sh_degree, packed=packed,
                                       backgrounds=torch.full((d["C"], 3), 0.3, device=DEV))
        torch.manual_seed(0)
        w = torch.randn_like(rc)
        ((rc * w).sum() + ra.sum() * 0.1).backward()
        res[packed] = (rc.detach(), ra.detach(), [x.grad for x in leaves + [col]], meta)
    (rc0, ra0, g0, m0), (rc1, ra1, g1, m1) = res[False], res[True]
    assert m0["camera_ids"] is None and m1["camera_ids"] is not None
    nnz = m1["gaussian_ids"].shape[0]
    assert m1["means2d"].shape == (nnz, 2) and m1["radii"].shape == (nnz,) and m1["opacities"].shape == (nnz,)
    # the packed radius rule keeps a superset of the dense rule's Gaussians; what it adds cannot reach a pixel
    assert nnz >= int((m0["radii"] > 0).sum())
    assert float((rc0 - rc1).abs().max()) < 1e-4 and float((ra0 - ra1).abs().max()) < 1e-4
    for a, b in zip(g0, g1):
        assert float((a - b).abs().max()) <= 1e-3 * float(a.abs().max()) + 1e-7


def test_rasterization_sparse_grad_and_render_modes():
    d = _composite_inputs()
    col = d["shs"][:, 0, :].abs().contiguous()
    args = (d["means"], d["quats"], d["scales"], d["opac"], col, d["V"], d["K"], d["W"], d["H"])
    rgb, a_rgb, _ = G.rasterization(*args, packed=False)
    dep, a_d, meta = G.rasterization(*args, packed=False, render_mode="D")
    rgbd, _, _ = G.rasterization(*args, packed=False, render_mode="RGB+D")
    rgbed, _, _ = G.rasterization(*args, packed=True, render_mode="RGB+ED")
    ed, _, _ = G.rasterization(*args, packed=False, render_mode="ED")
    assert dep.shape[-1] == 1 and rgbd.shape[-1] == 4 and rgbed.shape[-1] == 4 and ed.shape[-1] == 1
    assert torch.equal(rgbd[..., :3], rgb) and torch.equal(a_d, a_rgb)
    assert float((rgbd[..., 3:] - dep).abs().max()) < 1e-5
    torch.testing.assert_close(ed, dep / a_rgb.clamp(min=1e-10), rtol=1e-5, atol=1e-6)
    torch.testing.assert_close(rgbed[..., 3:], ed, rtol=1e-4, atol=1e-4)
    # depth channel against the oracle: the accumulated depth is the compositing of per-Gaussian depths
    m2, con, rad, dp = (meta[k].cpu().numpy() for k in ("means2d", "conics", "radii", "depths"))
    cols = np.repeat(dp[..., None], 3, -1).astype(np.float32)
    ref, _, _, _ = oracle.rasterize_fwd(m2, con, cols, meta["opacities"].cpu().numpy(), d["W"], d["H"], 16,
                                        meta["isect_offsets"].cpu().numpy(), meta["flatten_ids"].cpu().numpy())
    assert (np.abs(dep.cpu().numpy()[..., 0] - ref[..., 0]) > 1e-4 * max(1.0, float(dp[rad > 0].max()))).mean() <= 1e-4
    # 7 colour channels = three groups of three through the 3-channel kernels
    torch.manual_seed(1)
    c7 = torch.rand(d["n"], 7, device=DEV).requires_grad_()
    means = d["means"].clone().requires_grad_()
    out7, _, _ = G.rasterization(means, d["quats"], d["scales"], d["opac"], c7, d["V"], d["K"], d["W"], d["H"], packed=False)
    assert out7.shape[-1] == 7
    parts = [G.rasterization(d["means"], d["quats"], d["scales"], d["opac"],
                             torch.nn.functional.pad(c7.detach()[:, i:i + 3], (0, max(0, i + 3 - 7))), d["V"], d["K"], d["W"],
                             d["H"], packed=False)[0] for i in (0, 3, 6)]
    torch.testing.assert_close(out7.detach(), torch.cat([parts[0], parts[1], parts[2][..., :1]], -1), rtol=0, atol=0)
    out7.sum().backward()
    assert c7.grad.shape == c7.shape and float(c7.grad.abs().sum()) > 0 and float(means.grad.abs().sum()) > 0
    # sparse gradients (packed only): same values as the dense gradients, COO layout
    leaves = [d[k].clone().requires_grad_() for k in ("means", "quats", "scales")]
    rc, _, _ = G.rasterization(*leaves, d["opac"], col, d["V"], d["K"], d["W"], d["H"], packed=True, sparse_grad=True)
    rc.sum().backward()
    dense = [d[k].clone().requires_grad_() for k in ("means", "quats", "scales")]
    rc2, _, _ = G.rasterization(*dense, d["opac"], col, d["V"], d["K"], d["W"], d["H"], packed=True)
    rc2.sum().backward()
    for a, b in zip(leaves, dense):
        assert a.grad.is_sparse and not b.grad.is_sparse
        # two backward runs: the rasteriser's atomics sum in a different order
        assert float((a.grad.to_dense() - b.grad).abs().max()) <= 1e-3 * float(b.grad.abs().max())
    with pytest.raises(ValueError):
        G.rasterization(*args, packed=False, sparse_grad=True)
