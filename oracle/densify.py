"""CPU restatement of the reference's densify / prune / opacity-reset step -- TEST INFRASTRUCTURE.

Follows the reference literally, step by step, with every intermediate concatenation
(strategies/base_gaussian_model.py:364-388 ``densify_and_prune``;
strategies/clm_offload/gaussian_model.py:801-831 ``densify_and_clone``, :741-799
``densify_and_split``, :578-611 ``prune_points`` / :522-576 ``_prune_unified_adam``, :613-700
``cat_tensors_to_unified_adam``, :701-739 ``densification_postfix``, :488-520 ``reset_opacity``),
so that the product's fused classify -> plan -> apply formulation is checked against the sequential
semantics rather than against itself.  exp / sigmoid / log are the deterministic ones shared with
the CUDA side (orc_expf / orc_sigmoidf / orc_logf, <= 2 ulp from torch's).

Parity unpinned: the reference has no test for this step and its classes need CUDA + pinned memory,
so it cannot be run here; the restatement is pinned against first principles in
tests/test_oracle_pin.py (counts, ordering, moment bookkeeping).

State: dict with xyz [N,3], log_scale [N,3], raw_quat [N,4], opacity_logit [N,1], sh [N,48] and
Adam moments m_* / v_* of each (oracle/train_step.py ``new_state``), plus the densification
statistics xyz_gradient_accum [N], denom [N], max_radii2D [N].
"""
import ctypes as C

import numpy as np

import oracle as O

_KEYS = ("xyz", "log_scale", "raw_quat", "opacity_logit", "sh")


def det_log(x):
    x = np.ascontiguousarray(x, np.float32)
    y = np.empty_like(x)
    O.lib().orc_logf_array(C.c_int64(x.size), x.ctypes.data_as(C.c_void_p), y.ctypes.data_as(C.c_void_p))
    return y


def _scaling(st):
    return O.activate(st["log_scale"], None, None)[0]


def _opacity(st):
    return O.activate(None, None, st["opacity_logit"].reshape(-1))[2]


def _cat(st, new):
    """cat_tensors_to_unified_adam + densification_postfix: append rows, zero moments, reset statistics."""
    for k in _KEYS:
        ext = np.asarray(new[k], np.float32).reshape((-1,) + st[k].shape[1:])
        st[k] = np.concatenate([st[k], ext])
        st["m_" + k] = np.concatenate([st["m_" + k], np.zeros_like(ext)])
        st["v_" + k] = np.concatenate([st["v_" + k], np.zeros_like(ext)])
    n = st["xyz"].shape[0]
    for k in ("xyz_gradient_accum", "denom", "max_radii2D"):
        st[k] = np.zeros(n, np.float32)


def prune_points(st, mask):
    """GM:578-611: keep ~mask everywhere (parameters, moments, statistics)."""
    keep = ~np.asarray(mask, bool)
    for k in _KEYS:
        for pre in ("", "m_", "v_"):
            st[pre + k] = st[pre + k][keep]
    for k in ("xyz_gradient_accum", "denom", "max_radii2D"):
        st[k] = st[k][keep]


def build_rotation(r):
    """utils/general_utils.py:311-334."""
    r = np.asarray(r, np.float32)
    norm = np.sqrt(r[:, 0] * r[:, 0] + r[:, 1] * r[:, 1] + r[:, 2] * r[:, 2] + r[:, 3] * r[:, 3])
    q = r / norm[:, None]
    R = np.zeros((q.shape[0], 3, 3), np.float32)
    w, x, y, z = q[:, 0], q[:, 1], q[:, 2], q[:, 3]
    R[:, 0, 0] = 1 - 2 * (y * y + z * z)
    R[:, 0, 1] = 2 * (x * y - w * z)
    R[:, 0, 2] = 2 * (x * z + w * y)
    R[:, 1, 0] = 2 * (x * y + w * z)
    R[:, 1, 1] = 1 - 2 * (x * x + z * z)
    R[:, 1, 2] = 2 * (y * z - w * x)
    R[:, 2, 0] = 2 * (x * z - w * y)
    R[:, 2, 1] = 2 * (y * z + w * x)
    R[:, 2, 2] = 1 - 2 * (x * x + y * y)
    return R


def densify_and_clone(st, grads, grad_threshold, scene_extent, percent_dense):
    """GM:801-831."""
    sel = np.abs(grads) >= np.float32(grad_threshold)
    sel &= _scaling(st).max(1) <= np.float32(percent_dense * scene_extent)
    _cat(st, {k: st[k][sel] for k in _KEYS})
    return int(sel.sum())


def densify_and_split(st, grads, grad_threshold, scene_extent, percent_dense, noise, n_copies=2):
    """GM:741-799.  ``noise`` [n_copies * n_selected, 3] standard normal stands in for torch.normal's draw
    (samples = noise * stds, the way torch.normal(mean=0, std=stds) scales its standard normals)."""
    n_init = st["xyz"].shape[0]
    padded = np.zeros(n_init, np.float32)
    padded[:grads.shape[0]] = grads
    scaling = _scaling(st)
    sel = padded >= np.float32(grad_threshold)
    sel &= scaling.max(1) > np.float32(percent_dense * scene_extent)
    ns = int(sel.sum())
    noise = np.asarray(noise, np.float32).reshape(-1, 3)
    assert noise.shape[0] == n_copies * ns, (noise.shape, ns)
    stds = np.tile(scaling[sel], (n_copies, 1))
    samples = noise * stds
    rots = np.tile(build_rotation(st["raw_quat"][sel]), (n_copies, 1, 1))
    new_xyz = np.einsum("nij,nj->ni", rots, samples).astype(np.float32) + np.tile(st["xyz"][sel], (n_copies, 1))
    new_scaling = det_log(np.tile(scaling[sel], (n_copies, 1)) / np.float32(0.8 * n_copies))
    new = dict(xyz=new_xyz, log_scale=new_scaling, raw_quat=np.tile(st["raw_quat"][sel], (n_copies, 1)),
               opacity_logit=np.tile(st["opacity_logit"][sel], (n_copies, 1)), sh=np.tile(st["sh"][sel], (n_copies, 1)))
    _cat(st, new)
    prune_points(st, np.concatenate([sel, np.zeros(n_copies * ns, bool)]))
    return ns


def count_split(st, grad_threshold, scene_extent, percent_dense):
    """How many parents densify_and_split will select (to size the noise draw): the cloned copies have zero
    padded gradient, so the selection is decided on the current rows alone."""
    with np.errstate(divide="ignore", invalid="ignore"):
        g = st["xyz_gradient_accum"] / st["denom"]
    g[np.isnan(g)] = 0.0
    sel = (g >= np.float32(grad_threshold)) & (_scaling(st).max(1) > np.float32(percent_dense * scene_extent))
    return int(sel.sum())


def densify_and_prune(st, max_grad, min_opacity, extent, max_screen_size, percent_dense, noise):
    """strategies/base_gaussian_model.py:364-388.  Returns (n_cloned, n_split)."""
    with np.errstate(divide="ignore", invalid="ignore"):
        grads = (st["xyz_gradient_accum"] / st["denom"]).astype(np.float32)
    grads[np.isnan(grads)] = 0.0
    nc = densify_and_clone(st, grads, max_grad, extent, percent_dense)
    ns = densify_and_split(st, grads, max_grad, extent, percent_dense, noise)
    prune_mask = _opacity(st) < np.float32(min_opacity)
    if max_screen_size:
        big_vs = st["max_radii2D"] > max_screen_size  # all False: the postfix zeroed it (base:374-381)
        big_ws = _scaling(st).max(1) > np.float32(0.1 * extent)
        prune_mask = prune_mask | big_vs | big_ws
    prune_points(st, prune_mask)
    return nc, ns


def reset_opacity(st):
    """GM:488-520 (inverse_sigmoid: utils/general_utils.py:145)."""
    o = np.minimum(_opacity(st), np.float32(0.01))
    st["opacity_logit"] = det_log(o / (np.float32(1.0) - o)).reshape(-1, 1)
    st["m_opacity_logit"] = np.zeros_like(st["opacity_logit"])
    st["v_opacity_logit"] = np.zeros_like(st["opacity_logit"])


def check_update_at_this_iter(iteration, bsz, update_interval, update_residual):
    """utils/general_utils.py:130-142."""
    lo = iteration % update_interval
    hi = lo + bsz
    if lo <= update_residual < hi:
        return True
    return lo <= update_residual + update_interval < hi
