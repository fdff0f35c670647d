"""CPU check of the cull kernel's rejection pass (csrc/proj.cu: make_precam / surely_culled): a float32 numpy
restatement of it must never reject a (camera, Gaussian) pair that the oracle's full test keeps.

    python tests/cull_pretest_check.py [n_gaussians=200000]

Scenes: the three bench scene shapes plus adversarial sets (Gaussians next to the camera, on the frustum edges,
very large / very anisotropic).  Prints the fraction of pairs the pass rejects next to the fraction the full test
rejects.  Test infrastructure (uses oracle/)."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "clm-gs_b200")]
import oracle  # noqa: E402
from clmgs import scene as S  # noqa: E402

f32 = np.float32


def pretest(means, scales, viewmats, Ks, W, H, eps2d=0.3, near=0.01, far=1e10):
    """[B,N] bool: True = surely culled."""
    means = means.astype(f32); scales = scales.astype(f32)
    smax2 = (np.abs(scales).max(1) ** 2).astype(f32)
    l1 = np.abs(means).sum(1, dtype=f32)
    out = np.zeros((viewmats.shape[0], means.shape[0]), bool)
    for c in range(viewmats.shape[0]):
        vm = viewmats[c].astype(f32); K = Ks[c].astype(f32)
        fx, fy, cx, cy = K[0, 0], K[1, 1], K[0, 2], K[1, 2]
        lim_x = f32(1.3) * (f32(0.5) * f32(W) / fx); lim_y = f32(1.3) * (f32(0.5) * f32(H) / fy)
        R, t = vm[:3, :3], vm[:3, 3]
        rs = (R * R).sum(dtype=f32); rmax = np.abs(R).max(); tmax = np.abs(t).max()
        xrow = fx * R[0] + cx * R[2]; x0 = fx * t[0] + cx * t[2]
        yrow = fy * R[1] + cy * R[2]; y0 = fy * t[1] + cy * t[2]
        g = f32(1.9073486e-6) * (max(abs(fx), abs(fy)) + max(abs(cx), abs(cy)))
        kx = f32(1.001) * (rs / f32(3)) * (fx * fx * (1 + lim_x * lim_x) + fy * fy * (1 + lim_y * lim_y))
        z = ((R[2, 0] * means[:, 0] + R[2, 1] * means[:, 1]) + R[2, 2] * means[:, 2]) + t[2]
        with np.errstate(all="ignore"):
            rz = (f32(1) / z).astype(f32)
            u = (kx * smax2) * (rz * rz) + f32(eps2d + 0.32)
            Rb = f32(3.06) * np.sqrt(u) + f32(1.07)
            mx = (xrow[0] * means[:, 0] + xrow[1] * means[:, 1] + xrow[2] * means[:, 2] + x0) * rz
            my = (yrow[0] * means[:, 0] + yrow[1] * means[:, 1] + yrow[2] * means[:, 2] + y0) * rz
            Rr = (l1 * (g * rmax) + g * tmax) * rz + Rb
            Rx = f32(1e-5) * np.abs(mx) + Rr; Ry = f32(1e-5) * np.abs(my) + Rr
            screen = (mx + Rx <= 0) | (mx - Rx >= W) | (my + Ry <= 0) | (my - Ry >= H)
        out[c] = (z < near) | (z > far) | screen
    return out


def check(name, means, scales, quats, V, K, W, H):
    keep = oracle.cull_mask(means, scales, quats, V, K, W, H).astype(bool)
    rej = pretest(means, scales, V, K, W, H)
    bad = rej & keep
    print("%-28s pairs %9d  full test rejects %.4f  pass 1 rejects %.4f  (%.1f %% of the rejectable)  WRONG %d" % (
        name, keep.size, 1 - keep.mean(), rej.mean(), 100 * rej.sum() / max(1, (~keep).sum()), bad.sum()))
    return int(bad.sum())


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 200_000
    wrong = 0
    rng = np.random.default_rng(0)
    for cfg, B in (("c4_bigcity_102m", 64), ("c2_bicycle_6m", 16), ("c3_rubble4k_28m", 16)):
        base = S.CONFIGS[cfg]
        spec = S.SceneSpec(base.name, base.n_gaussians, base.width, base.height, B, base.shape, base.fovx_deg, base.index)
        g = S.make_gaussians(spec, n=n)
        V, K, _ = S.make_cameras(spec, B, g["extent"])
        sc, q, _ = oracle.activate(g["log_scale"], g["raw_quat"], g["opacity_logit"])
        wrong += check(cfg, g["xyz"], sc, q, V, K, spec.width, spec.height)
        # adversarial variants of the same scene
        cam_pos = np.stack([-(V[c][:3, :3].T @ V[c][:3, 3]) for c in range(B)])
        m2 = g["xyz"].copy()
        pick = rng.integers(0, B, n)
        # (a) next to a camera: within 0.005 .. 2 scene units of its centre, any direction
        m2a = cam_pos[pick] + rng.normal(size=(n, 3)).astype(f32) * rng.uniform(0.005, 2.0, (n, 1)).astype(f32)
        wrong += check(cfg + " near-camera", m2a.astype(f32), sc, q, V, K, spec.width, spec.height)
        # (b) huge and very anisotropic Gaussians
        sc_b = (sc * rng.choice([1, 30, 300], (n, 1)) * np.array([1, 1, 0.01], f32)).astype(f32)
        wrong += check(cfg + " huge/anisotropic", g["xyz"], sc_b, q, V, K, spec.width, spec.height)
        # (c) on the frustum edges of camera `pick`: direction at the image border, random depth
        Kp, Vp = K[pick], V[pick]
        u = np.where(rng.random(n) < 0.5, 0.0, spec.width) + rng.normal(0, 3, n)
        v = rng.uniform(0, spec.height, n)
        zc = rng.uniform(0.02, 50, n)
        pc = np.stack([(u - Kp[:, 0, 2]) / Kp[:, 0, 0] * zc, (v - Kp[:, 1, 2]) / Kp[:, 1, 1] * zc, zc], 1)
        m2c = np.einsum("nij,nj->ni", np.transpose(Vp[:, :3, :3], (0, 2, 1)), pc - Vp[:, :3, 3])
        wrong += check(cfg + " frustum-edge", m2c.astype(f32), sc, q, V, K, spec.width, spec.height)
    print("TOTAL WRONG", wrong)
    return 1 if wrong else 0


if __name__ == "__main__":
    sys.exit(main())
