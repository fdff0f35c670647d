"""Densify / prune / opacity reset on the split GPU + pinned-host state vs the sequential oracle restatement
(SURVEY.md 8f row 1).  Row selection and final row order must be identical; copied values bit-equal."""
import numpy as np
import pytest
import torch

from oracle import densify as D
from oracle import train_step as TS
from clmgs import densification, engine, model, scene as S

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
_KEYS = (("xyz", "_xyz", "xyz"), ("log_scale", "_scaling", "scaling"), ("raw_quat", "_rotation", "rotation"),
         ("opacity_logit", "_opacity", "opacity"))


def _setup(n, bsz=4, W=112, H=80, capacity_factor=3, seed=0, shape="object"):
    spec = S.SceneSpec("t", n, W, H, bsz, shape, 60.0 if shape == "object" else 50.0, 0)
    g = S.make_gaussians(spec)
    V, K, P = S.make_cameras(spec, bsz, g["extent"])
    gts = S.make_gt_images(spec, bsz)
    args = model.TrainArgs(bsz=bsz, cpu_adam_threads=4, densification_stats=True, percent_dense=0.05)
    gm = model.GaussianModelCLMOffload(3, DEV).create_from_arrays(g["xyz"], g["log_scale"], g["raw_quat"],
                                                                 g["opacity_logit"], g["sh"],
                                                                 capacity=capacity_factor * n)
    gm.training_setup(args)
    st = TS.new_state(g)
    rng = np.random.default_rng(seed)
    # random statistics and moments on both sides
    st["xyz_gradient_accum"] = (rng.random(n) * 1e-3).astype(np.float32)
    st["denom"] = rng.integers(0, 3, n).astype(np.float32)
    st["max_radii2D"] = np.zeros(n, np.float32)
    gm.xyz_gradient_accum = torch.from_numpy(st["xyz_gradient_accum"]).to(DEV).reshape(n, 1)
    gm.denom = torch.from_numpy(st["denom"]).to(DEV).reshape(n, 1)
    for ok, _, gk in _KEYS:
        st["m_" + ok] = rng.standard_normal(st[ok].shape).astype(np.float32)
        st["v_" + ok] = rng.random(st[ok].shape).astype(np.float32)
        gm.optimizer.gpu_adam.state[gk]["exp_avg"].copy_(torch.from_numpy(st["m_" + ok]))
        gm.optimizer.gpu_adam.state[gk]["exp_avg_sq"].copy_(torch.from_numpy(st["v_" + ok]))
    st["m_sh"] = rng.standard_normal(st["sh"].shape).astype(np.float32)
    st["v_sh"] = rng.random(st["sh"].shape).astype(np.float32)
    gm.optimizer.cpu_adam.exp_avg.copy_(torch.from_numpy(st["m_sh"]))
    gm.optimizer.cpu_adam.exp_avg_sq.copy_(torch.from_numpy(st["v_sh"]))
    return spec, g, V, K, P, gts, args, gm, st, rng


def _compare(gm, st, n_exact_rows=None):
    n = st["xyz"].shape[0]
    assert gm.n_gaussians == n and gm._parameters.shape[0] == n
    for ok, attr, gk in _KEYS:
        got = getattr(gm, attr).cpu().numpy()
        if ok in ("xyz", "log_scale"):
            np.testing.assert_allclose(got, st[ok], rtol=2e-6, atol=1e-6, err_msg=ok)
            if n_exact_rows is not None:
                np.testing.assert_array_equal(got[:n_exact_rows], st[ok][:n_exact_rows])
        else:
            np.testing.assert_array_equal(got, st[ok], err_msg=ok)
        np.testing.assert_array_equal(gm.optimizer.gpu_adam.state[gk]["exp_avg"].cpu().numpy(), st["m_" + ok])
        np.testing.assert_array_equal(gm.optimizer.gpu_adam.state[gk]["exp_avg_sq"].cpu().numpy(), st["v_" + ok])
    np.testing.assert_array_equal(gm._parameters.numpy(), st["sh"])
    np.testing.assert_array_equal(gm.optimizer.cpu_adam.exp_avg.numpy(), st["m_sh"])
    np.testing.assert_array_equal(gm.optimizer.cpu_adam.exp_avg_sq.numpy(), st["v_sh"])
    assert gm.optimizer.cpu_adam.param.data_ptr() == gm._parameters.data_ptr()
    assert gm.denom.shape == (n, 1) and not gm.denom.any() and not gm.max_radii2D.any()


@pytest.mark.parametrize("n,max_screen_size", [(30_000, None), (30_000, 20), (1000, 20)])
def test_densify_and_prune_matches_oracle(n, max_screen_size):
    spec, g, V, K, P, gts, args, gm, st, rng = _setup(n)
    extent = float(g["extent"])
    ns = gm.count_split_parents(args.densify_grad_threshold, extent)
    assert ns == D.count_split(st, args.densify_grad_threshold, extent, args.percent_dense)
    noise = rng.standard_normal((2 * ns, 3)).astype(np.float32)
    info = gm.densify_and_prune(args.densify_grad_threshold, args.min_opacity, extent, max_screen_size,
                                noise=torch.from_numpy(noise))
    nc, ns2 = D.densify_and_prune(st, args.densify_grad_threshold, args.min_opacity, extent, max_screen_size,
                                  args.percent_dense, noise)
    assert info["n_split_parents"] == ns2 == ns and ns > 0 and (nc > 0 or n < 5000) and info["n_cloned"] <= nc
    assert info["n_after"] == st["xyz"].shape[0] != n
    _compare(gm, st, n_exact_rows=info["n_after"] - info["n_children"])


def test_prune_points_and_reset_opacity_match_oracle():
    spec, g, V, K, P, gts, args, gm, st, rng = _setup(20_000)
    mask = rng.random(20_000) < 0.3
    gm.prune_points(torch.from_numpy(mask))
    D.prune_points(st, mask)
    _compare(gm, st, n_exact_rows=st["xyz"].shape[0])
    gm.reset_opacity()
    D.reset_opacity(st)
    np.testing.assert_array_equal(gm._opacity.cpu().numpy(), st["opacity_logit"])
    assert not gm.optimizer.gpu_adam.state["opacity"]["exp_avg"].any()
    # nothing pruned / everything pruned
    gm.prune_points(torch.zeros(gm.n_gaussians, dtype=torch.bool))
    _compare_sizes = gm.n_gaussians
    assert _compare_sizes == st["xyz"].shape[0]
    gm.prune_points(torch.ones(gm.n_gaussians, dtype=torch.bool))
    assert gm.n_gaussians == 0 and gm._parameters.shape[0] == 0


def test_capacity_overflow_is_an_error():
    spec, g, V, K, P, gts, args, gm, st, rng = _setup(5000, capacity_factor=1)
    gm.xyz_gradient_accum.fill_(1.0)
    gm.denom.fill_(1.0)
    with pytest.raises(RuntimeError, match="pinned buffers"):
        gm.densify_and_prune(args.densify_grad_threshold, 0.0, float(g["extent"]), None)
    assert gm.n_gaussians == 5000  # untouched


def test_training_continues_after_densification():
    """Iterate -> densify (statistics from the iteration itself) -> iterate, against the oracle doing the same."""
    W, H, bsz = 112, 80, 4
    spec, g, V, K, P, gts, args, gm, st, rng = _setup(20_000, bsz, W, H)
    n = 20_000
    for k in ("xyz_gradient_accum", "denom", "max_radii2D"):
        st.pop(k)
    gm.xyz_gradient_accum.zero_()
    gm.denom.zero_()
    for ok, _, gk in _KEYS:
        st["m_" + ok][:] = 0
        st["v_" + ok][:] = 0
        gm.optimizer.gpu_adam.state[gk]["exp_avg"].zero_()
        gm.optimizer.gpu_adam.state[gk]["exp_avg_sq"].zero_()
    st["m_sh"][:] = 0
    st["v_sh"][:] = 0
    gm.optimizer.cpu_adam.exp_avg.zero_()
    gm.optimizer.cpu_adam.exp_avg_sq.zero_()
    cams = [model.Camera(V[i], K[i], gts[i], DEV, uid=i) for i in range(bsz)]
    comm, gen = torch.cuda.Stream(), torch.Generator().manual_seed(0)
    losses, order, _ = engine.clm_offload_train_one_batch(gm, None, cams, gm.parameters_grad_buffer, None, None, comm, gen)
    ref_losses, _, _ = TS.train_one_batch(st, V, K, P, gts, W, H, args, order)
    np.testing.assert_allclose([float(l) for l in losses], ref_losses, rtol=2e-5, atol=2e-6)
    # the selection depends on the accumulated gradient norms (1e-3 relative): use the product's statistics on
    # both sides so that the comparison is about the restructuring, not about rows on the threshold
    st["xyz_gradient_accum"] = gm.xyz_gradient_accum[:, 0].cpu().numpy()
    st["denom"] = gm.denom[:, 0].cpu().numpy()
    np.testing.assert_array_equal(st["max_radii2D"], gm.max_radii2D.cpu().numpy())
    for ok, attr, gk in _KEYS:  # same for the parameters: adopt the product's post-step values
        st[ok] = getattr(gm, attr).cpu().numpy().copy()
        st["m_" + ok] = gm.optimizer.gpu_adam.state[gk]["exp_avg"].cpu().numpy().copy()
        st["v_" + ok] = gm.optimizer.gpu_adam.state[gk]["exp_avg_sq"].cpu().numpy().copy()
    st["sh"] = gm._parameters.numpy().copy()
    st["m_sh"] = gm.optimizer.cpu_adam.exp_avg.numpy().copy()
    st["v_sh"] = gm.optimizer.cpu_adam.exp_avg_sq.numpy().copy()
    extent = float(g["extent"])
    thr = float(np.median(st["xyz_gradient_accum"][st["denom"] > 0] / st["denom"][st["denom"] > 0]))
    ns = gm.count_split_parents(thr, extent)
    noise = rng.standard_normal((2 * ns, 3)).astype(np.float32)
    info = gm.densify_and_prune(thr, args.min_opacity, extent, None, noise=torch.from_numpy(noise))
    D.densify_and_prune(st, thr, args.min_opacity, extent, None, args.percent_dense, noise)
    assert info["n_after"] > n and info["n_cloned"] > 0 and info["n_children"] > 0
    _compare(gm, st)
    assert not gm.parameters_grad_buffer.numpy().any()
    losses, order, _ = engine.clm_offload_train_one_batch(gm, None, cams, gm.parameters_grad_buffer, None, None, comm, gen)
    ref_losses, _, _ = TS.train_one_batch(st, V, K, P, gts, W, H, args, order)
    np.testing.assert_allclose([float(l) for l in losses], ref_losses, rtol=2e-5, atol=2e-6)
    err = np.abs(gm.optimizer.cpu_adam.exp_avg.numpy() - st["m_sh"]).max() / np.abs(st["m_sh"]).max()
    assert err < 1e-3, err


def test_schedule_driver():
    spec, g, V, K, P, gts, args, gm, st, rng = _setup(5000)

    class Scene:
        cameras_extent = float(g["extent"])

    args.densify_from_iter, args.densification_interval, args.opacity_reset_interval = 8, 12, 24
    ran = {}
    for it in range(1, 60, args.bsz):
        gm.xyz_gradient_accum.copy_(torch.rand(gm.n_gaussians, 1, device=DEV) * 2.2e-4)
        gm.denom.fill_(1.0)
        did = densification.gsplat_densification(it, Scene, gm)
        if did:
            ran[it] = sorted(did)
    assert ran == {9: ["densify_and_prune"], 21: ["densify_and_prune", "reset_opacity"], 33: ["densify_and_prune"],
                   45: ["densify_and_prune", "reset_opacity"], 57: ["densify_and_prune"]}, ran
