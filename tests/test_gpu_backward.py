"""GPU parity: backward path, loss, host-row movement, bitmaps and optimisers of
libclmgs.so (through the C ABI) vs the CPU oracle / numpy restatements.
north_star tolerances: gradients 1e-3 relative, integer work bit-exact."""
import numpy as np
import pytest
import torch

import oracle
from clmgs import ops, scene as S

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _t(a):
    return torch.from_numpy(np.ascontiguousarray(a)).to(DEV)


def _relerr(got, ref):
    return np.abs(got - ref).max() / max(np.abs(ref).max(), 1e-20)


def _scene(n, W, H, ncam, index=0):
    spec = S.SceneSpec("t", n, W, H, ncam, "object", 60.0, index)
    g = S.make_gaussians(spec)
    V, K, P = S.make_cameras(spec, ncam, g["extent"])
    return spec, g, V, K, P


@pytest.mark.parametrize("ncam", [1, 3])
def test_projection_bwd(ncam):
    spec, g, V, K, _ = _scene(30_000, 320, 200, ncam)
    scales, quats, _ = oracle.activate(g["log_scale"], g["raw_quat"], None)
    radii, means2d, depths, conics, _ = oracle.projection_fwd(g["xyz"], quats, scales, V, K, 320, 200)
    rng = np.random.default_rng(0)
    vm2 = rng.standard_normal(means2d.shape).astype(np.float32)
    vd = rng.standard_normal(depths.shape).astype(np.float32)
    vc = rng.standard_normal(conics.shape).astype(np.float32)
    ref = oracle.projection_bwd(g["xyz"], quats, scales, V, K, 320, 200, radii, conics, vm2, vd, vc)
    got = ops.projection_bwd(_t(g["xyz"]), _t(quats), _t(scales), _t(V), _t(K), 320, 200, 0.3, _t(radii), _t(conics),
                             _t(vm2), _t(vd), _t(vc))
    for nm, r, t in zip(("v_means", "v_quats", "v_scales"), ref, got):
        a = t.cpu().numpy()
        # per-element: 1e-3 relative to the row's magnitude
        denom = np.abs(r).max(axis=1, keepdims=True) + 1e-6 * np.abs(r).max()
        assert (np.abs(a - r) / denom).max() < 1e-3, nm


@pytest.mark.parametrize("deg", [0, 1, 2, 3, 4])
def test_sh_bwd(deg):
    rng = np.random.default_rng(deg)
    N, K = 30_007, (25 if deg == 4 else 16)
    dirs = rng.standard_normal((N, 3)).astype(np.float32)
    coeffs = rng.standard_normal((N, K, 3)).astype(np.float32)
    v_colors = rng.standard_normal((N, 3)).astype(np.float32)
    base = rng.standard_normal((N, K, 3)).astype(np.float32)
    ref_vc = base.copy()
    ref_vd = oracle.sh_bwd_inplace(deg, dirs, coeffs, ref_vc, v_colors)
    vc = _t(base)
    _, vd = ops.sh_bwd(deg, _t(dirs), _t(coeffs), _t(v_colors), v_coeffs=vc)
    np.testing.assert_allclose(vc.cpu().numpy(), ref_vc, rtol=1e-4, atol=1e-5)
    np.testing.assert_allclose(vd.cpu().numpy(), ref_vd, rtol=1e-3, atol=1e-4)
    vc2, _ = ops.sh_bwd(deg, _t(dirs), _t(coeffs), _t(v_colors))  # overwrite semantics (gsplat)
    np.testing.assert_allclose(vc2.cpu().numpy(), ref_vc - base, rtol=1e-4, atol=1e-5)


@pytest.mark.parametrize("W,H,n,bg,absgrad", [(200, 120, 20_000, False, False), (123, 77, 10_000, True, True)])
def test_rasterize_bwd_vs_oracle(W, H, n, bg, absgrad):
    spec, g, V, K, P = _scene(n, W, H, 2)
    scales, quats, opac = oracle.activate(g["log_scale"], g["raw_quat"], g["opacity_logit"][:, 0])
    radii, means2d, depths, conics, _ = oracle.projection_fwd(g["xyz"], quats, scales, V, K, W, H)
    rng = np.random.default_rng(3)
    colors = rng.random((2, n, 3)).astype(np.float32)
    opacs = np.broadcast_to(opac[None], (2, n)).copy()
    bgs = rng.random((2, 3)).astype(np.float32) if bg else None
    tw, th = (W + 15) // 16, (H + 15) // 16
    _, ids, flat = oracle.isect_tiles(means2d, radii, depths, 16, tw, th)
    off = oracle.isect_offset_encode(ids, 2, tw, th)
    rc_r, ra_r, last_r, _ = oracle.rasterize_fwd(means2d, conics, colors, opacs, W, H, 16, off, flat, bgs)
    v_rc = rng.standard_normal(rc_r.shape).astype(np.float32)
    v_ra = rng.standard_normal(ra_r.shape).astype(np.float32)
    ref = oracle.rasterize_bwd(means2d, conics, colors, opacs, W, H, 16, off, flat, ra_r, last_r, v_rc, v_ra, bgs,
                               absgrad=absgrad)
    # feed the ORACLE's forward outputs so that this test isolates the backward kernel
    got = ops.rasterize_bwd(_t(means2d), _t(conics), _t(colors), _t(opacs), W, H, 16, _t(off), _t(flat), _t(ra_r),
                            _t(last_r), _t(v_rc), _t(v_ra), None if bgs is None else _t(bgs), absgrad=absgrad)
    names = ["v_means2d", "v_conics", "v_colors", "v_opacities", "v_means2d_abs"]
    for nm, r, t in zip(names, ref, got):
        if r is None:
            assert t is None
            continue
        a = t.cpu().numpy()
        err = _relerr(a, r)
        print(nm, "rel err", err)
        assert err < 1e-3, (nm, err)


@pytest.mark.parametrize("H,W", [(45, 70), (128, 96), (300, 413)])
def test_loss_fwd_bwd(H, W):
    rng = np.random.default_rng(H)
    img = rng.random((3, H, W)).astype(np.float32)
    gt = (rng.random((3, H, W)) * 255).astype(np.uint8)
    loss, v, ssim, l1 = oracle.loss_fwd_bwd(img, gt)
    for hwc in (False, True):
        im = _t(img.transpose(1, 2, 0)) if hwc else _t(img)
        out, vg = ops.loss_fwd_bwd(im, _t(gt), 0.2, hwc=hwc)
        o = out.cpu().numpy()
        assert abs(o[0] - loss) < 1e-5 and abs(o[1] - l1) < 1e-6 and abs(o[2] - ssim) < 1e-5
        vg = vg.cpu().numpy()
        if hwc:
            vg = vg.transpose(2, 0, 1)
        assert _relerr(vg, v) < 1e-3


def test_loss_matches_reference_golden(golden):
    g = golden("loss")
    out, vg = ops.loss_fwd_bwd(_t(g["image"]), _t(g["gt_u8"]), 0.2, hwc=False)
    assert abs(out[0].item() - float(g["loss"])) < 1e-5
    assert _relerr(vg.cpu().numpy(), g["v_image"]) < 1e-3


def test_fused_ssim_dropin():
    rng = np.random.default_rng(1)
    a = rng.random((2, 3, 50, 61)).astype(np.float32)
    b = rng.random((2, 3, 50, 61)).astype(np.float32)
    s, v = ops.fused_ssim_fwd_bwd(_t(a), _t(b))
    # oracle: lambda=1 -> loss = 1 - ssim, v = -d ssim
    tot, grads = 0.0, []
    for k in range(2):
        gt_u8 = None
        # oracle takes uint8 gt; emulate float gt through its internals by using img2*255 rounding-free path:
        # compare against torch conv2d reference instead (utils/loss_utils.py:_ssim restated)
    import torch.nn.functional as F
    w1 = torch.tensor([0.001028380123898387, 0.0075987582094967365, 0.036000773310661316, 0.10936068743467331,
                       0.21300552785396576, 0.26601171493530273, 0.21300552785396576, 0.10936068743467331,
                       0.036000773310661316, 0.0075987582094967365, 0.001028380123898387], dtype=torch.float64)
    win = (w1[:, None] * w1[None, :]).expand(3, 1, 11, 11)
    ta = torch.tensor(a, dtype=torch.float64, requires_grad=True)
    tb = torch.tensor(b, dtype=torch.float64)
    mu1 = F.conv2d(ta, win, padding=5, groups=3); mu2 = F.conv2d(tb, win, padding=5, groups=3)
    s11 = F.conv2d(ta * ta, win, padding=5, groups=3) - mu1 ** 2
    s22 = F.conv2d(tb * tb, win, padding=5, groups=3) - mu2 ** 2
    s12 = F.conv2d(ta * tb, win, padding=5, groups=3) - mu1 * mu2
    m = ((2 * mu1 * mu2 + 1e-4) * (2 * s12 + 9e-4)) / ((mu1 ** 2 + mu2 ** 2 + 1e-4) * (s11 + s22 + 9e-4))
    ref = m.mean()
    (gr,) = torch.autograd.grad(ref, [ta])
    assert abs(s.item() - ref.item()) < 1e-5
    assert _relerr(v.cpu().numpy(), gr.numpy()) < 1e-3


def test_activate_bwd_scatter():
    spec, g, V, K, _ = _scene(20_000, 64, 64, 1)
    rng = np.random.default_rng(0)
    ids = np.sort(rng.choice(20_000, 7_000, replace=False)).astype(np.int64)
    vm = rng.standard_normal((7000, 3)).astype(np.float32); vs = rng.standard_normal((7000, 3)).astype(np.float32)
    vq = rng.standard_normal((7000, 4)).astype(np.float32); vo = rng.standard_normal(7000).astype(np.float32)
    ls = torch.tensor(g["log_scale"][ids], dtype=torch.float64, requires_grad=True)
    rq = torch.tensor(g["raw_quat"][ids], dtype=torch.float64, requires_grad=True)
    ol = torch.tensor(g["opacity_logit"][ids, 0], dtype=torch.float64, requires_grad=True)
    L = (torch.exp(ls) * torch.tensor(vs)).sum() + (torch.nn.functional.normalize(rq, dim=-1) * torch.tensor(vq)).sum() \
        + (torch.sigmoid(ol) * torch.tensor(vo)).sum()
    gls, grq, gol = torch.autograd.grad(L, [ls, rq, ol])
    N = 20_000
    gx = torch.ones(N, 3, device=DEV); gl = torch.ones(N, 3, device=DEV)
    gq = torch.ones(N, 4, device=DEV); go = torch.ones(N, 1, device=DEV)
    ops.activate_bwd_scatter(_t(g["log_scale"]), _t(g["raw_quat"]), _t(g["opacity_logit"]), _t(ids), _t(vm), _t(vs),
                             _t(vq), _t(vo), gx, gl, gq, go)
    gx, gl, gq, go = (t.cpu().numpy() for t in (gx, gl, gq, go))
    np.testing.assert_allclose(gx[ids] - 1, vm, rtol=1e-6, atol=1e-6)
    np.testing.assert_allclose(gl[ids] - 1, gls.numpy(), rtol=1e-4, atol=1e-5)
    np.testing.assert_allclose(gq[ids] - 1, grq.numpy(), rtol=1e-4, atol=1e-5)
    np.testing.assert_allclose(go[ids, 0] - 1, gol.numpy(), rtol=1e-4, atol=1e-5)
    rest = np.ones(N, bool); rest[ids] = False
    assert (gx[rest] == 1).all() and (go[rest] == 1).all()


def test_rows_gather_scatter_pinned_host():
    rng = np.random.default_rng(0)
    N, n = 50_000, 12_345
    h = torch.from_numpy(rng.standard_normal((N, 48)).astype(np.float32)).pin_memory()
    ids = np.sort(rng.choice(N, n, replace=False)).astype(np.int64)
    d = torch.empty(n, 48, device=DEV)
    ops.rows_gather(d, h, _t(ids))
    torch.cuda.synchronize()
    assert d.cpu().numpy().tobytes() == h.numpy()[ids].tobytes()
    # scatter back: overwrite and accumulate
    hg = torch.zeros(N, 48).pin_memory()
    grad = _t(rng.standard_normal((n, 48)).astype(np.float32))
    ops.rows_scatter_grad(grad, hg, _t(ids), accum=False)
    ops.rows_scatter_grad(grad, hg, _t(ids), accum=True)
    torch.cuda.synchronize()
    ref = np.zeros((N, 48), np.float32); ref[ids] = 2 * grad.cpu().numpy()
    np.testing.assert_array_equal(hg.numpy(), ref)
    flags = torch.zeros(4, dtype=torch.int32).pin_memory()
    ops.set_signal(flags, 2, 1)
    torch.cuda.synchronize()
    assert flags.tolist() == [0, 0, 1, 0]


@pytest.mark.parametrize("accum,bsz", [(1, 4), (2, 4), (2, 8)])
def test_rows_retention_list_api_and_fused(accum, bsz):
    rng = np.random.default_rng(1)
    N = 40_000
    h = torch.from_numpy(rng.standard_normal((N, 48)).astype(np.float32)).pin_memory()
    hn = h.numpy()
    vis = rng.random((bsz, N)) < 0.4
    filters = [np.nonzero(vis[k])[0].astype(np.int64) for k in range(bsz)]
    # ordered bitmap: camera k <-> bit bsz-1-k
    bm = np.zeros(N, np.uint8)
    for k in range(bsz):
        bm[filters[k]] |= np.uint8(1 << (bsz - 1 - k))
    bitmap = _t(bm)
    hg = torch.zeros(N, 48).pin_memory()
    ref_hg = np.zeros((N, 48), np.float32)
    # camera 0 load
    cur = torch.empty(len(filters[0]), 48, device=DEV)
    ops.rows_gather(cur, h, _t(filters[0]))
    cur_grad = torch.zeros_like(cur)
    ref_grad = np.zeros((len(filters[0]), 48), np.float32)
    for k in range(bsz):
        ids_this = _t(filters[k])
        assert cur.cpu().numpy().tobytes() == hn[filters[k]].tobytes(), k
        # "backward": add a known per-camera gradient
        gk = rng.standard_normal(ref_grad.shape).astype(np.float32)
        cur_grad += _t(gk); ref_grad = ref_grad + gk
        if k < bsz - 1:
            ids_next = _t(filters[k + 1])
            nxt = torch.empty(len(filters[k + 1]), 48, device=DEV)
            src_map = ops.rows_prefetch_next(nxt, h, cur, ids_this, ids_next, bitmap, bsz - 1 - k)
            nxt_grad = torch.empty_like(nxt)
            ops.rows_grad_handoff(cur_grad, nxt_grad, hg, ids_this, len(filters[k + 1]), src_map, bitmap,
                                  bsz - 1 - k, bsz - 1 - (k + 1), accum=accum)
            # reference bookkeeping
            both = vis[k] & vis[k + 1]
            pos_this = np.cumsum(vis[k]) - 1; pos_next = np.cumsum(vis[k + 1]) - 1
            new_ref = np.zeros((len(filters[k + 1]), 48), np.float32)
            d_ids = np.nonzero(both)[0]
            new_ref[pos_next[d_ids]] = ref_grad[pos_this[d_ids]]
            g_ids = np.nonzero(vis[k] & ~vis[k + 1])[0]
            ref_hg[g_ids] += ref_grad[pos_this[g_ids]]
            sm = src_map.cpu().numpy()[: len(filters[k + 1])]
            exp_sm = np.where(vis[k][filters[k + 1]], pos_this[filters[k + 1]], -1)
            np.testing.assert_array_equal(sm, exp_sm)
            cur, cur_grad, ref_grad = nxt, nxt_grad, new_ref
            np.testing.assert_array_equal(cur_grad.cpu().numpy(), ref_grad)
        else:
            ops.rows_grad_handoff(cur_grad, None, hg, ids_this, 0, None, bitmap, bsz - 1 - k, 0, accum=accum)
            ref_hg[filters[k]] += ref_grad
    torch.cuda.synchronize()
    np.testing.assert_allclose(hg.numpy(), ref_hg, rtol=0, atol=0)
    # list-based (reference-shaped) API on one transition
    k = 1
    both = vis[k] & vis[k + 1]; only_next = ~vis[k] & vis[k + 1]
    pos_this = np.cumsum(vis[k]) - 1; pos_next = np.cumsum(vis[k + 1]) - 1
    host_ids = np.nonzero(only_next)[0].astype(np.int32); d_ids = np.nonzero(both)[0]
    prev = _t(hn[filters[k]])
    out = torch.empty(len(filters[k + 1]), 48, device=DEV)
    ops.rows_gather_retention(out, h, prev, _t(host_ids), _t(pos_this[d_ids].astype(np.int32)),
                              _t(pos_next[host_ids].astype(np.int32)), _t(pos_next[d_ids].astype(np.int32)))
    assert out.cpu().numpy().tobytes() == hn[filters[k + 1]].tobytes()


@pytest.mark.parametrize("bsz", [4, 8, 16, 32, 64])
def test_bitmap_ops(bsz):
    rng = np.random.default_rng(bsz)
    N = 100_003
    vis = rng.random((bsz, N)) < rng.uniform(0.05, 0.6, (bsz, 1))
    wb = ops.word_bytes_for(bsz)
    npdt = {1: np.uint8, 2: np.uint16, 4: np.uint32, 8: np.uint64}[wb]
    tdt = ops._WORD_DTYPE[wb]
    # cull-convention bitmap (bit c = camera c) built with scatter_to_bit
    bitmap = torch.zeros(N, dtype=tdt, device=DEV)
    for c in range(bsz):
        ops.scatter_to_bit(bitmap, _t(np.nonzero(vis[c])[0].astype(np.int64)), c)
    words = np.zeros(N, np.uint64)
    for c in range(bsz):
        words |= vis[c].astype(np.uint64) << np.uint64(c)
    np.testing.assert_array_equal(bitmap.cpu().numpy().view(npdt).astype(np.uint64), words)
    # distance matrix on a sample
    sample = np.sort(rng.choice(N, N // 32, replace=False)).astype(np.int64)
    dist = ops.bitmap_distance(bitmap, bsz, _t(sample)).cpu().numpy()
    sv = vis[:, sample].astype(np.int32)
    ref_dist = (sv[:, None, :] ^ sv[None, :, :]).sum(-1)
    np.testing.assert_array_equal(dist, ref_dist)
    # order -> permuted bitmap, ffs, cnt_d, grouping
    order = rng.permutation(bsz).astype(np.int32)
    ordered = ops.bitmap_permute(bitmap, bsz, order)
    ow = np.zeros(N, np.uint64)
    for i, c in enumerate(order):
        ow |= vis[c].astype(np.uint64) << np.uint64(bsz - 1 - i)
    np.testing.assert_array_equal(ordered.cpu().numpy().view(npdt).astype(np.uint64), ow)
    ffs = ops.extract_ffs(ordered)
    ref_ffs = np.zeros(N, np.uint8)
    for b in range(bsz - 1, -1, -1):
        ref_ffs[((ow >> np.uint64(b)) & np.uint64(1)) == 1] = b + 1
    np.testing.assert_array_equal(ffs.cpu().numpy(), ref_ffs)
    cnt_d = ops.compute_cnt_d(ordered, bsz).cpu().numpy()
    ref_cnt = np.array([(vis[order[k]] & vis[order[k + 1]]).sum() for k in range(bsz - 1)], np.int32)
    np.testing.assert_array_equal(cnt_d, ref_cnt)
    out_ids, counts = ops.group_by_ffs(ffs, bsz)
    ref_sorted = np.argsort(ref_ffs, kind="stable").astype(np.int32)
    np.testing.assert_array_equal(out_ids.cpu().numpy(), ref_sorted)
    np.testing.assert_array_equal(counts.cpu().numpy(), np.bincount(ref_ffs, minlength=bsz + 1))


def test_gpu_adam_kernels():
    rng = np.random.default_rng(0)
    N = 20_011
    for M in (1, 3, 4):
        p = rng.standard_normal((N, M)).astype(np.float32); g = rng.standard_normal((N, M)).astype(np.float32)
        m = rng.standard_normal((N, M)).astype(np.float32) * 0.1; v = rng.random((N, M)).astype(np.float32) * 0.01
        mask = rng.random(N) < 0.5
        pr, mr, vr = p.copy(), m.copy(), v.copy()
        oracle.selective_adam(pr, g * 0.25, mr, vr, mask, 1e-2, 0.9 ** 4, 0.999 ** 4, 1e-15 / 2)
        tp, tm, tv = _t(p), _t(m), _t(v)
        ops.selective_adam(tp, _t(g), tm, tv, _t(mask), 1e-2, 0.9 ** 4, 0.999 ** 4, 1e-15 / 2, grad_scale=0.25)
        np.testing.assert_allclose(tp.cpu().numpy(), pr, rtol=1e-5, atol=1e-6)
        np.testing.assert_allclose(tm.cpu().numpy(), mr, rtol=1e-5, atol=1e-7)
        np.testing.assert_allclose(tv.cpu().numpy(), vr, rtol=1e-5, atol=1e-9)
        assert (tp.cpu().numpy()[~mask] == p[~mask]).all()
        pr, mr, vr = p.copy(), m.copy(), v.copy()
        oracle.adam_dense(pr, g * 0.25, mr, vr, 1e-2, 0.9, 0.999, 1e-15, 3)
        tp, tm, tv = _t(p), _t(m), _t(v)
        ops.adam_dense(tp, _t(g), tm, tv, 1e-2, 0.9, 0.999, 1e-15, 3, grad_scale=0.25)
        np.testing.assert_allclose(tp.cpu().numpy(), pr, rtol=1e-5, atol=1e-6)


@pytest.mark.parametrize("deg", [0, 1, 2, 3])
def test_engine_fused_sh_colors_fwd_bwd(deg):
    """clmgs_sh_colors_fwd/_bwd (dirs = means - campos, clamp_min(c + 0.5, 0), += into retained grads,
    v_means += v_dirs) vs the oracle composed the way engine.py:67-78,709-716 composes it."""
    rng = np.random.default_rng(10 + deg)
    n = 50_003
    means = rng.standard_normal((n, 3)).astype(np.float32) * 2
    campos = np.array([0.3, -4.0, 0.5], np.float32)
    coeffs = (rng.standard_normal((n, 16, 3)) * 0.4).astype(np.float32)
    v_colors = rng.standard_normal((n, 3)).astype(np.float32)
    base_g = rng.standard_normal((n, 48)).astype(np.float32)
    base_m = rng.standard_normal((n, 3)).astype(np.float32)
    dirs = means - campos[None]
    raw = oracle.sh_fwd(deg, dirs, coeffs)
    col_ref = np.maximum(raw + np.float32(0.5), 0)
    col = ops.sh_colors_fwd(deg, _t(means), _t(campos), _t(coeffs.reshape(n, 48))).cpu().numpy()
    np.testing.assert_allclose(col, col_ref, rtol=1e-5, atol=1e-5)
    # keep away from the clamp edge, where a 1-ulp colour difference legitimately flips the mask
    safe = (np.abs(raw + 0.5) > 1e-4).all(1)
    vmask = v_colors * (raw + np.float32(0.5) >= 0)
    g_ref = base_g.reshape(n, 16, 3).copy()
    vd_ref = oracle.sh_bwd_inplace(deg, dirs, coeffs, g_ref, vmask)
    g, vm = _t(base_g), _t(base_m)
    ops.sh_colors_bwd(deg, _t(means), _t(campos), _t(coeffs.reshape(n, 48)), _t(v_colors), g, vm)
    np.testing.assert_allclose(g.cpu().numpy()[safe], g_ref.reshape(n, 48)[safe], rtol=1e-4, atol=1e-5)
    np.testing.assert_allclose(vm.cpu().numpy()[safe], (base_m + vd_ref)[safe], rtol=1e-3, atol=1e-4)
