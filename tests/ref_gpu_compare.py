"""Timing + parity report: libclmgs.so vs the REFERENCE's own CUDA kernels on a B200 (the asserting version of the
parity half is tests/test_gpu_vs_reference_cuda.py; this script also TIMES every stage of both).

Needs oracle/_ref/ (built in the build container by `python oracle/build_ref.py` from the sources
under /root/reference, compiled for sm_100 with the reference's flags incl. --use_fast_math).
Run on the GPU box:   python tests/ref_gpu_compare.py [--n 6000000]

Reports (a) parity on identical inputs -- cull-mask / radii mismatch rates against a fast-math
build, exactness of the integer stages given identical inputs, image and gradient differences --
and (b) device time of each stage, reference kernel vs ours.  Output: gpurun_out/ref_compare.json.
"""
import argparse
import importlib.util
import json
import math
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "clm-gs_b200")]
import numpy as np
import torch

from clmgs import ops, scene as S


def load_ref(name, sub):
    path = os.path.join(ROOT, "oracle", "_ref", sub, name + ".so")
    if not os.path.exists(path):
        raise SystemExit("missing %s -- run oracle/build_ref.py in the build container" % path)
    spec = importlib.util.spec_from_file_location(name, path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def timeit(fn, reps=10, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(reps):
        fn()
    e.record()
    torch.cuda.synchronize()
    return s.elapsed_time(e) / reps


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=6_000_000)
    ap.add_argument("--bsz", type=int, default=4)
    a = ap.parse_args()
    G = load_ref("gsplat_ref_csrc", "gsplat")
    CK = load_ref("clm_kernels_ref_C", "clm_kernels")
    dev = "cuda:0"
    base = S.CONFIGS["c2_bicycle_6m"]
    spec = S.SceneSpec(base.name, a.n, base.width, base.height, a.bsz, base.shape, base.fovx_deg, base.index)
    W, H = spec.width, spec.height
    g = S.make_gaussians(spec)
    V, K, P = S.make_cameras(spec, a.bsz, g["extent"])
    t = lambda x: torch.from_numpy(np.ascontiguousarray(x)).to(dev)
    xyz, ls, rq, ol = t(g["xyz"]), t(g["log_scale"]), t(g["raw_quat"]), t(g["opacity_logit"])
    sh = t(g["sh"])
    Vt, Kt = t(V), t(K)
    scales, quats = torch.exp(ls), torch.nn.functional.normalize(rq, dim=-1)   # the reference's activations
    out = dict(n=a.n, bsz=a.bsz, width=W, height=H)

    # ---- S1 cull: reference packed projection (2 passes + cumsum + .item()) vs ours ------------------
    def ref_cull():
        r = G.fully_fused_projection_packed_fwd(xyz, None, quats, scales, Vt, Kt, W, H, 0.3, 0.01, 1e10, 0.0, False)
        return r[1], r[2]
    cam_r, gid_r = ref_cull()
    ours = ops.cull(xyz, scales, quats, Vt, Kt, W, H, activated=True, packed_outputs=True)
    mask_r = torch.zeros(a.bsz, a.n, dtype=torch.bool, device=dev); mask_r[cam_r, gid_r] = True
    mask_o = torch.zeros(a.bsz, a.n, dtype=torch.bool, device=dev); mask_o[ours["camera_ids"], ours["gaussian_ids"]] = True
    mism = int((mask_r != mask_o).sum())
    out["cull"] = dict(ref_nnz=int(mask_r.sum()), ours_nnz=int(mask_o.sum()), mask_mismatches=mism,
                       mismatch_rate=mism / float(a.bsz * a.n), ref_ms=timeit(ref_cull, 5),
                       ours_ms=timeit(lambda: ops.cull(xyz, ls, rq, Vt, Kt, W, H, activated=False), 5))

    # ---- one camera: filtered set -------------------------------------------------------------------------
    ids = ours["gaussian_ids"][: ours["counts"][0]]
    n = ids.shape[0]
    m, s_, q_, o_ = ops.gather_activate(xyz, ls, rq, ol, ids)
    v1, k1 = Vt[:1].contiguous(), Kt[:1].contiguous()
    ref_p = lambda: G.fully_fused_projection_fwd(m, None, q_, s_, v1, k1, W, H, 0.3, 0.01, 1e10, 0.0, False)
    our_p = lambda: ops.projection_fwd(m, q_, s_, v1, k1, W, H)
    rr, rm2, rd, rc, _ = ref_p()
    orad, om2, od, oc, _ = our_p()
    both = (rr > 0) & (orad > 0)
    out["projection_fwd"] = dict(n=n, radii_mismatch_rate=float((rr != orad).float().mean()),
                                 max_abs_means2d=float((rm2 - om2)[both].abs().max()),
                                 max_rel_conics=float(((rc - oc)[both].abs() / (rc[both].abs() + 1e-6)).max()),
                                 depth_bits_mismatch_rate=float((rd.view(torch.int32) != od.view(torch.int32))[both].float().mean()),
                                 ref_ms=timeit(ref_p), ours_ms=timeit(our_p))

    # ---- S4 on IDENTICAL inputs (ours): integer stage, must be exact -----------------------------------------
    tw, th = math.ceil(W / 16.0), math.ceil(H / 16.0)
    ref_i = lambda: G.isect_tiles(om2, orad, od, None, None, 1, 16, tw, th, True, True)
    our_i = lambda: ops.isect_tiles(om2, orad, od, 16, tw, th)
    rt, rids, rflat = ref_i()
    ot, oids, oflat = our_i()
    roff = G.isect_offset_encode(rids, 1, tw, th)
    ooff = ops.isect_offset_encode(oids, 1, tw, th)
    out["isect"] = dict(n_isects=int(oids.shape[0]), tiles_per_gauss_equal=bool(torch.equal(rt, ot)),
                        isect_ids_equal=bool(torch.equal(rids, oids)), flatten_ids_equal=bool(torch.equal(rflat, oflat)),
                        offsets_equal=bool(torch.equal(roff, ooff)), ref_ms=timeit(ref_i, 5), ours_ms=timeit(our_i, 5),
                        ref_offsets_ms=timeit(lambda: G.isect_offset_encode(rids, 1, tw, th)),
                        ours_offsets_ms=timeit(lambda: ops.isect_offset_encode(oids, 1, tw, th)))

    # ---- SH colour ---------------------------------------------------------------------------------------------
    campos = t(P[0])
    dirs = (m - campos[None]).contiguous()
    coeffs = torch.empty(n, 48, device=dev)
    ops.rows_gather(coeffs, sh, ids)
    c3 = coeffs.view(n, 16, 3)
    ref_sh = lambda: G.compute_sh_fwd(3, dirs, c3, None)
    our_sh = lambda: ops.sh_colors_fwd(3, m, campos, coeffs)
    col_r = torch.clamp_min(ref_sh() + 0.5, 0.0)
    col_o = our_sh()
    out["sh_fwd"] = dict(max_abs=float((col_r - col_o).abs().max()), ref_ms=timeit(ref_sh), ours_ms=timeit(our_sh),
                         note="ref time excludes its dirs subtraction and clamp (separate torch ops)")

    # ---- S5 rasterise on identical inputs ----------------------------------------------------------------------------
    colors = col_o[None].contiguous()
    opac = o_[None].contiguous()
    ref_rf = lambda: G.rasterize_to_pixels_fwd(om2, oc, colors, opac, None, W, H, 16, ooff, oflat)
    our_rf = lambda: ops.rasterize_fwd(om2, oc, colors, opac, W, H, 16, ooff, oflat)
    r_rc, r_ra, r_last = ref_rf()
    o_rc, o_ra, o_last = our_rf()
    out["raster_fwd"] = dict(max_abs_image=float((r_rc - o_rc).abs().max()),
                             frac_pixels_over_1e4=float(((r_rc - o_rc).abs().amax(-1) > 1e-4).float().mean()),
                             last_ids_mismatch_rate=float((r_last != o_last).float().mean()),
                             ref_ms=timeit(ref_rf), ours_ms=timeit(our_rf))
    v_rc = torch.randn_like(r_rc)
    v_ra = torch.zeros_like(r_ra)
    ref_rb = lambda: G.rasterize_to_pixels_bwd(om2, oc, colors, opac, None, W, H, 16, ooff, oflat, r_ra, r_last, v_rc, v_ra, False)
    our_rb = lambda: ops.rasterize_bwd(om2, oc, colors, opac, W, H, 16, ooff, oflat, r_ra, r_last, v_rc, None)
    _, rv_m, rv_c, rv_col, rv_o = ref_rb()
    ov_m, ov_c, ov_col, ov_o, _ = our_rb()
    rel = lambda x, y: float((x - y).abs().max() / y.abs().max().clamp_min(1e-20))
    out["raster_bwd"] = dict(rel_v_means2d=rel(ov_m, rv_m), rel_v_conics=rel(ov_c, rv_c), rel_v_colors=rel(ov_col, rv_col),
                             rel_v_opacities=rel(ov_o, rv_o), ref_ms=timeit(ref_rb, 5), ours_ms=timeit(our_rb, 5))

    # ---- projection bwd / SH bwd -----------------------------------------------------------------------------------------
    vd = torch.zeros_like(od)
    ref_pb = lambda: G.fully_fused_projection_bwd(m, None, q_, s_, v1, k1, W, H, 0.3, orad, oc, None, rv_m, vd, rv_c, None, False)
    our_pb = lambda: ops.projection_bwd(m, q_, s_, v1, k1, W, H, 0.3, orad, oc, rv_m, None, rv_c)
    r_vm, _, r_vq, r_vs, _ = ref_pb()
    o_vm, o_vq, o_vs = our_pb()
    out["projection_bwd"] = dict(rel_v_means=rel(o_vm, r_vm), rel_v_quats=rel(o_vq, r_vq), rel_v_scales=rel(o_vs, r_vs),
                                 ref_ms=timeit(ref_pb), ours_ms=timeit(our_pb))
    vcol = rv_col[0].contiguous()
    g_ref = torch.zeros(n, 16, 3, device=dev)
    g_our = torch.zeros(n, 48, device=dev)
    ref_sb = lambda: CK.compute_sh_bwd_inplace(16, 3, dirs, c3, g_ref, None, vcol, True)
    vm_acc = torch.zeros(n, 3, device=dev)
    our_sb = lambda: ops.sh_colors_bwd(3, m, campos, coeffs, vcol, g_our, vm_acc)
    # one call each for parity (both accumulate): the clamp mask is applied by autograd in the reference
    mask = (G.compute_sh_fwd(3, dirs, c3, None) + 0.5 >= 0).float()
    vcol_masked = (vcol * mask).contiguous()
    g_ref.zero_(); CK.compute_sh_bwd_inplace(16, 3, dirs, c3, g_ref, None, vcol_masked, True)
    g_our.zero_(); our_sb()
    out["sh_bwd"] = dict(rel_v_coeffs=rel(g_our.view(n, 16, 3), g_ref), ref_ms=timeit(ref_sb), ours_ms=timeit(our_sb))

    # ---- loss ---------------------------------------------------------------------------------------------------------------
    img = o_rc[0].permute(2, 0, 1).contiguous()
    gt_u8 = t(S.make_gt_images(spec, 1)[0])
    gt = torch.clamp(gt_u8 / 255.0, 0.0, 1.0)

    def ref_loss():
        mp, d1, d2, d3 = CK.fusedssim(1e-4, 9e-4, img[None], gt[None], True)
        dl = torch.full_like(img[None], 1.0 / img.numel())
        return mp, CK.fusedssim_backward(1e-4, 9e-4, img[None], gt[None], dl, d1, d2, d3)
    mp, gr = ref_loss()
    lv, vg = ops.loss_fwd_bwd(o_rc[0], gt_u8, 0.2, hwc=True)
    ref_total = 0.8 * (img - gt).abs().mean() + 0.2 * (1 - mp.mean())
    s_our, g_our = ops.fused_ssim_fwd_bwd(img[None].contiguous(), gt[None].contiguous())
    out["loss"] = dict(abs_loss_diff=float((ref_total - lv[0]).abs()), abs_ssim_diff=float((mp.mean() - s_our).abs()),
                       rel_ssim_grad=rel(g_our, gr),
                       ref_ms=timeit(lambda: ref_loss()), ours_ms=timeit(lambda: ops.loss_fwd_bwd(o_rc[0], gt_u8, 0.2, hwc=True)),
                       note="ref time = fusedssim + fusedssim_backward only (its L1 is a separate torch.compile kernel); "
                            "ours = SSIM + L1 forward and backward")

    # ---- SH row gather from pinned host: reference 4-byte kernel (grid 32) vs ours ------------------------------------------------
    h = torch.from_numpy(g["sh"]).pin_memory()
    d_out = torch.empty(n, 48, device=dev)
    ref_g = lambda: CK.send_shs2gpu_stream(d_out, h, ids, 32, 256)
    our_g = lambda: ops.rows_gather(d_out, h, ids)
    out["rows_gather_host"] = dict(rows=n, ref_ms=timeit(ref_g, 3, 1), ours_ms=timeit(our_g, 3, 1),
                                   gb=n * 192 / 1e9)
    for k, v in out.items():
        if isinstance(v, dict) and "ref_ms" in v:
            v["speedup"] = round(v["ref_ms"] / v["ours_ms"], 2)
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    json.dump(out, open(os.path.join(ROOT, "gpurun_out", "ref_compare.json"), "w"), indent=1)
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
