"""Densify / prune: oracle restatement pinned against first principles, the host-side
in-place table compaction of libclmgs.so against numpy, and the schedule helper.  CPU only."""
import numpy as np
import pytest

import oracle as O
from oracle import densify as D
from oracle import train_step as TS
from clmgs import _lib, densification, scene as S


def _state(n=4000, seed=0):
    spec = S.SceneSpec("t", n, 64, 64, 4, "object", 60.0, 0)
    g = S.make_gaussians(spec)
    st = TS.new_state(g)
    rng = np.random.default_rng(seed)
    st["xyz_gradient_accum"] = (rng.random(n) * 1e-3).astype(np.float32)
    st["denom"] = rng.integers(0, 3, n).astype(np.float32)
    st["max_radii2D"] = np.zeros(n, np.float32)
    for k in ("xyz", "log_scale", "raw_quat", "opacity_logit", "sh"):
        st["m_" + k] = rng.standard_normal(st[k].shape).astype(np.float32)
        st["v_" + k] = rng.random(st[k].shape).astype(np.float32)
    return st, float(g["extent"]), rng


def test_det_log_is_within_2ulp_of_libm():
    x = np.concatenate([np.float32([1e-40, 1e-38, 1e-30, 1e-3, 0.5, 1.0, 2.0, 1e10, 3.4e38]),
                        np.exp(np.random.default_rng(0).uniform(-80, 80, 100000)).astype(np.float32)])
    got = D.det_log(x).astype(np.float64)
    ref = np.log(x.astype(np.float64))
    ulp = np.spacing(np.abs(ref).astype(np.float32)).astype(np.float64)
    assert np.all(np.abs(got - ref) <= 2.0 * ulp + 1e-45)
    assert np.isneginf(D.det_log(np.float32([0.0]))[0]) and np.isnan(D.det_log(np.float32([-1.0]))[0])


def _discrete_state(n=3000, seed=0):
    """Rows far from every threshold, so the expected classes follow from exact arithmetic."""
    st, extent, rng = _state(n, seed)
    st["log_scale"] = np.log(rng.choice(np.float32([0.01, 0.1, 1.0]), (n, 3))).astype(np.float32)
    st["opacity_logit"] = rng.choice(np.float32([-9.0, 2.0]), (n, 1)).astype(np.float32)
    st["xyz_gradient_accum"] = rng.choice(np.float32([0.0, 1.0]), n).astype(np.float32)
    st["denom"] = rng.choice(np.float32([0.0, 1.0, 2.0]), n).astype(np.float32)
    return st, extent, rng


def expected_plan(st, extent, pd, min_opacity, use_ws):
    """Independent statement of the final row order: kept originals, kept clones, kept children x2."""
    with np.errstate(all="ignore"):
        g = st["xyz_gradient_accum"].astype(np.float64) / st["denom"].astype(np.float64)
    g[np.isnan(g)] = 0
    smax = np.exp(st["log_scale"].astype(np.float64)).max(1)
    opac = 1 / (1 + np.exp(-st["opacity_logit"][:, 0].astype(np.float64)))
    clone = (g >= 0.0002) & (smax <= pd * extent)
    split = (g >= 0.0002) & (smax > pd * extent)
    drop = opac < min_opacity
    drop_self = drop | (use_ws & (smax > 0.1 * extent))
    drop_child = drop | (use_ws & (smax / 1.6 > 0.1 * extent))
    A = np.nonzero(~split & ~drop_self)[0]
    B = np.nonzero(clone & ~drop_self)[0]
    Cc = np.nonzero(split & ~drop_child)[0]
    rank = np.cumsum(split) - 1
    return A, B, Cc, rank[Cc], int(split.sum())


@pytest.mark.parametrize("use_ws", [False, True])
def test_oracle_densify_matches_first_principles(use_ws):
    st, extent, rng = _discrete_state()
    pd = 0.05  # 0.15: scales {0.01, 0.1} clone, 1.0 splits; world-size limit 0.3 drops 1.0 and its 0.625 children
    before = {k: np.copy(v) for k, v in st.items() if isinstance(v, np.ndarray)}
    A, B, Cc, rank, ns = expected_plan(before, extent, pd, 0.005, use_ws)
    assert D.count_split(st, 0.0002, extent, pd) == ns and ns > 100 and len(B) > 100
    noise = rng.standard_normal((2 * ns, 3)).astype(np.float32)
    nc, ns2 = D.densify_and_prune(st, 0.0002, 0.005, extent, 20 if use_ws else None, pd, noise)
    assert ns2 == ns
    src = np.concatenate([A, B, Cc, Cc])
    assert st["xyz"].shape[0] == len(src)
    nA, nB, nC = len(A), len(B), len(Cc)
    if use_ws:
        assert nC == 0  # every child of a scale-1.0 parent is still larger than 0.3
    for k in ("sh", "raw_quat", "opacity_logit"):
        np.testing.assert_array_equal(st[k], before[k][src])
    np.testing.assert_array_equal(st["xyz"][:nA + nB], before["xyz"][src[:nA + nB]])
    for k in ("xyz", "log_scale", "raw_quat", "opacity_logit", "sh"):
        np.testing.assert_array_equal(st["m_" + k][:nA], before["m_" + k][A])
        np.testing.assert_array_equal(st["v_" + k][:nA], before["v_" + k][A])
        assert not st["m_" + k][nA:].any() and not st["v_" + k][nA:].any()
    # children: scale / 1.6 and a sample inside the parent's frame
    if nC:
        kids = slice(nA + nB, None)
        np.testing.assert_allclose(np.exp(st["log_scale"][kids]), np.exp(before["log_scale"][src[kids]]) / 1.6, rtol=1e-5)
        z = np.concatenate([noise[rank], noise[ns + rank]])
        q = before["raw_quat"][src[kids]].astype(np.float64)
        q /= np.linalg.norm(q, axis=1, keepdims=True)
        w, x, y, zz = q.T
        R = np.stack([1 - 2 * (y * y + zz * zz), 2 * (x * y - w * zz), 2 * (x * zz + w * y),
                      2 * (x * y + w * zz), 1 - 2 * (x * x + zz * zz), 2 * (y * zz - w * x),
                      2 * (x * zz - w * y), 2 * (y * zz + w * x), 1 - 2 * (x * x + y * y)], 1).reshape(-1, 3, 3)
        exp_xyz = np.einsum("nij,nj->ni", R, z * np.exp(before["log_scale"][src[kids]].astype(np.float64))) + \
            before["xyz"][src[kids]]
        np.testing.assert_allclose(st["xyz"][kids], exp_xyz, rtol=1e-4, atol=1e-5)
    assert not st["denom"].any() and st["denom"].shape[0] == len(src)


def test_oracle_reset_opacity():
    st, _, _ = _state(500)
    D.reset_opacity(st)
    o = O.activate(None, None, st["opacity_logit"][:, 0])[2]
    assert o.max() <= 0.0100001 and not st["m_opacity_logit"].any()


@pytest.mark.parametrize("n_old,frac_keep,n_tail", [(50_000, 0.9, 7000), (300_000, 0.7, 20_000), (20_000, 1.0, 100), (30_000, 0.0, 5), (1, 1.0, 0)])
def test_host_rows_densify_matches_numpy(n_old, frac_keep, n_tail):
    rng = np.random.default_rng(1)
    cols, cap = 48, n_old + n_tail + 10
    tabs = [rng.standard_normal((cap, cols)).astype(np.float32) for _ in range(3)]
    ref = [t.copy() for t in tabs]
    keep = rng.random(n_old) < frac_keep
    if frac_keep == 1.0:
        keep[:] = True
    A = np.nonzero(keep)[0].astype(np.int32)
    tail = rng.integers(0, n_old, n_tail).astype(np.int32)
    src = np.concatenate([A, tail])
    L = _lib.lib()
    L.call("clmgs_host_rows_densify", tabs[0], tabs[1], tabs[2], cols, n_old, cap, src, len(A), n_tail, 4)
    n_new = len(A) + n_tail
    np.testing.assert_array_equal(tabs[0][:n_new], ref[0][src])
    for k in (1, 2):
        np.testing.assert_array_equal(tabs[k][:len(A)], ref[k][A])
        assert not tabs[k][len(A):n_new].any()


def test_host_rows_densify_rejects_bad_lists():
    L = _lib.lib()
    t = [np.zeros((10, 48), np.float32) for _ in range(3)]
    with pytest.raises(_lib.ClmgsError):
        L.call("clmgs_host_rows_densify", t[0], t[1], t[2], 48, 8, 10, np.int32([3, 2]), 2, 0, 1)  # not ascending
    with pytest.raises(_lib.ClmgsError):
        L.call("clmgs_host_rows_densify", t[0], t[1], t[2], 48, 8, 10, np.int32([0, 9]), 2, 0, 1)  # out of range
    with pytest.raises(_lib.ClmgsError):
        L.call("clmgs_host_rows_densify", t[0], t[1], t[2], 48, 8, 10, np.arange(8, dtype=np.int32), 8, 3, 1)  # capacity


def test_schedule_helper_matches_restatement():
    for it in range(0, 7000, 37):
        for bsz in (1, 4, 64):
            for interval in (100, 3000):
                assert densification.check_update_at_this_iter(it, bsz, interval, 0) == \
                    D.check_update_at_this_iter(it, bsz, interval, 0)
    assert densification.check_update_at_this_iter(596, 4, 100, 0) is False
    assert densification.check_update_at_this_iter(597, 4, 100, 0) is True
    assert densification.check_update_at_this_iter(600, 4, 100, 0) is True
