"""CPU restatement of one ``clm_offload_train_one_batch`` (strategies/clm_offload/engine.py:338-925)
built only from the oracle's C functions -- TEST INFRASTRUCTURE and the timed CPU baseline.

The camera order is an INPUT (fast-tsp is a time-budgeted heuristic, so parity tests
fix the order); everything else follows the reference: cull -> per camera {gather + activations,
projection, SH colour + clamp, tile intersection + sort, rasterise, loss, backward chain,
gradient accumulation} -> Adam on the GPU-resident attributes (dense torch rule or selective) and
cpu-adam on the SH rows that were touched.

State is a dict of float32 numpy arrays: xyz, log_scale, raw_quat, opacity_logit [N,1], sh [N,48]
plus Adam moments m_*/v_* and step counters.
"""
import math

import numpy as np

import oracle as O


def new_state(g):
    st = dict(xyz=g["xyz"].copy(), log_scale=g["log_scale"].copy(), raw_quat=g["raw_quat"].copy(),
              opacity_logit=g["opacity_logit"].reshape(-1, 1).copy(), sh=np.array(g["sh"], np.float32, copy=True))
    for k in ("xyz", "log_scale", "raw_quat", "opacity_logit", "sh"):
        st["m_" + k] = np.zeros_like(st[k])
        st["v_" + k] = np.zeros_like(st[k])
    st["gpu_step"] = 0
    st["cpu_step"] = 0
    return st


def hyper(args):
    """GM:180-211 batch scaling: lr*sqrt(B), eps/sqrt(B), betas**B."""
    s = math.sqrt(args.bsz)
    return dict(
        lr=dict(xyz=args.position_lr_init * args.spatial_lr_scale * s, opacity_logit=args.opacity_lr * s,
                log_scale=args.scaling_lr * s, raw_quat=args.rotation_lr * s),
        col_lr=np.array([args.feature_lr * s] * 3 + [args.feature_lr / 20.0 * s] * 45, np.float32),
        b1=0.9 ** args.bsz, b2=0.999 ** args.bsz, eps=1e-15 / s)


def render_camera(st, ids, viewmat, K, campos, W, H, sh_degree, background=None):
    scales, quats, opac = O.activate(st["log_scale"][ids], st["raw_quat"][ids], st["opacity_logit"][ids, 0])
    means = np.ascontiguousarray(st["xyz"][ids])
    radii, means2d, depths, conics, _ = O.projection_fwd(means, quats, scales, viewmat[None], K[None], W, H)
    dirs = means - campos[None].astype(np.float32)
    coeffs = np.ascontiguousarray(st["sh"][ids]).reshape(-1, 16, 3)
    raw_colors = O.sh_fwd(sh_degree, dirs, coeffs)
    colors = np.maximum(raw_colors + np.float32(0.5), np.float32(0.0))
    tw, th = (W + 15) // 16, (H + 15) // 16
    _, isect_ids, flat = O.isect_tiles(means2d, radii, depths, 16, tw, th)
    off = O.isect_offset_encode(isect_ids, 1, tw, th)
    bg = None if background is None else np.asarray(background, np.float32).reshape(1, 3)
    rc, ra, last, examined = O.rasterize_fwd(means2d, conics, colors[None], opac[None], W, H, 16, off, flat, bg)
    return dict(means=means, scales=scales, quats=quats, opac=opac, radii=radii, means2d=means2d, depths=depths,
                conics=conics, dirs=dirs, coeffs=coeffs, raw_colors=raw_colors, colors=colors, off=off, flat=flat,
                rc=rc, ra=ra, last=last, bg=bg, examined=examined, n_isects=len(flat))


def backward_camera(st, ids, fw, viewmat, K, W, H, sh_degree, v_render_hwc, grads, sh_grad):
    vm2, vcon, vcol, vop, _ = O.rasterize_bwd(fw["means2d"], fw["conics"], fw["colors"][None], fw["opac"][None], W, H,
                                              16, fw["off"], fw["flat"], fw["ra"], fw["last"], v_render_hwc[None],
                                              None, fw["bg"])
    v_means, v_quats, v_scales = O.projection_bwd(fw["means"], fw["quats"], fw["scales"], viewmat[None], K[None], W,
                                                  H, fw["radii"], fw["conics"], vm2, np.zeros_like(fw["depths"]), vcon)
    v_colors = vcol[0] * (fw["raw_colors"] + np.float32(0.5) >= 0)  # clamp_min backward (engine.py:78)
    rows = np.zeros((len(ids), 16, 3), np.float32)
    v_dirs = O.sh_bwd_inplace(sh_degree, fw["dirs"], fw["coeffs"], rows, v_colors)
    sh_grad[ids] += rows.reshape(-1, 48)
    v_means = v_means + v_dirs
    # activation backward (strategies/base_gaussian_model.py:42-50)
    grads["xyz"][ids] += v_means
    grads["log_scale"][ids] += v_scales * fw["scales"]
    q = st["raw_quat"][ids]
    nn = np.maximum(np.linalg.norm(q, axis=1, keepdims=True), 1e-12)
    qn = q / nn
    grads["raw_quat"][ids] += (v_quats - (v_quats * qn).sum(1, keepdims=True) * qn) / nn
    o = fw["opac"]
    grads["opacity_logit"][ids, 0] += vop[0] * o * (1 - o)
    return vm2


def train_one_batch(st, viewmats, Ks, camposs, gts_u8, W, H, args, ordered_cams, background=None):
    """Returns (losses in processed order, filters in processed order, per-camera stats)."""
    N, bsz = st["xyz"].shape[0], len(ordered_cams)
    hp = hyper(args)
    scales_all, quats_all, _ = O.activate(st["log_scale"], st["raw_quat"], None)
    mask = O.cull_mask(st["xyz"], scales_all, quats_all, viewmats, Ks, W, H, radius_clip=args.radius_clip)
    filters = [np.nonzero(mask[c])[0].astype(np.int64) for c in ordered_cams]
    grads = {k: np.zeros_like(st[k]) for k in ("xyz", "log_scale", "raw_quat", "opacity_logit")}
    sh_grad = np.zeros_like(st["sh"])
    losses, stats = [], []
    for k, c in enumerate(ordered_cams):
        ids = filters[k]
        fw = render_camera(st, ids, viewmats[c], Ks[c], camposs[c], W, H, args.sh_degree, background)
        img_chw = np.ascontiguousarray(fw["rc"][0].transpose(2, 0, 1))
        loss, v_chw, _, _ = O.loss_fwd_bwd(img_chw, gts_u8[c], args.lambda_dssim)
        losses.append(loss)
        vm2 = backward_camera(st, ids, fw, viewmats[c], Ks[c], W, H, args.sh_degree,
                              np.ascontiguousarray(v_chw.transpose(1, 2, 0)), grads, sh_grad)
        if getattr(args, "densification_stats", False):  # gaussian_model.py:833-851
            for key in ("max_radii2D", "xyz_gradient_accum", "denom"):
                st.setdefault(key, np.zeros(N, np.float32))
            st["max_radii2D"][ids] = np.maximum(st["max_radii2D"][ids], fw["radii"][0].astype(np.float32))
            gg = vm2[0] * np.array([W * 0.5, H * 0.5], np.float32)
            st["xyz_gradient_accum"][ids] += np.sqrt((gg * gg).sum(1))
            st["denom"][ids] += 1
        stats.append(dict(n_visible=len(ids), n_isects=fw["n_isects"], examined=fw["examined"]))
    touched = mask[list(ordered_cams)].any(0)
    # GPU-resident attributes (engine.py:870-882)
    st["gpu_step"] += 1
    for key in ("xyz", "opacity_logit", "log_scale", "raw_quat"):
        g = grads[key] / np.float32(bsz)
        if args.sparse_adam:
            O.selective_adam(st[key], g, st["m_" + key], st["v_" + key], touched, hp["lr"][key], hp["b1"], hp["b2"],
                             hp["eps"])
        else:
            O.adam_dense(st[key], g, st["m_" + key], st["v_" + key], hp["lr"][key], hp["b1"], hp["b2"], hp["eps"],
                         st["gpu_step"])
    # offloaded SH rows (engine.py:301-335)
    st["cpu_step"] += 1
    rows = (np.nonzero(touched)[0] if args.sparse_adam else np.arange(N)).astype(np.int32)
    O.cpu_adam_rows(st["sh"], sh_grad, st["m_sh"], st["v_sh"], rows, hp["col_lr"], hp["b1"], hp["b2"], hp["eps"],
                    st["cpu_step"], 1.0 / bsz)
    return losses, filters, stats
