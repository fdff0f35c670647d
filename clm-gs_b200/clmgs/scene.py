"""Deterministic synthetic scenes and cameras.

Replaces the reference's COLMAP / MatrixCity readers (scene/, OUT OF SCOPE) for
tests and benchmarks; only the conventions are kept: ``viewmat`` is the
row-major world-to-camera matrix (= ``camera.world_view_transform.T``,
strategies/clm_offload/engine.py:44) and ``K`` comes from the field of view
exactly as scene/cameras.py:110-126 builds it.

Everything is numpy on the host; callers upload what they need.
"""
from dataclasses import dataclass
import math
import os

import numpy as np


@dataclass
class SceneSpec:
    name: str
    n_gaussians: int
    width: int
    height: int
    batch: int
    shape: str  # "object" | "aerial"
    fovx_deg: float
    index: int  # config index -> seeds


# BASELINE.json configs[0..4]
CONFIGS = {
    "c1_plumbing_100k": SceneSpec("c1_plumbing_100k", 100_000, 256, 256, 8, "object", 60.0, 0),
    "c2_bicycle_6m": SceneSpec("c2_bicycle_6m", 6_000_000, 1237, 822, 4, "object", 60.0, 1),
    "c3_rubble4k_28m": SceneSpec("c3_rubble4k_28m", 28_000_000, 3840, 2160, 4, "aerial", 50.0, 2),
    "c4_bigcity_102m": SceneSpec("c4_bigcity_102m", 102_000_000, 1920, 1080, 64, "aerial", 50.0, 3),
    "c5_stress_500m": SceneSpec("c5_stress_500m", 500_000_000, 1920, 1080, 64, "aerial", 50.0, 4),
}


def make_gaussians(spec: SceneSpec, n=None, chunk=1_000_000, sh_out=None, sh_rows=None, want_sh=True, workers=0):
    """Returns dict of float32 arrays: xyz[N,3], log_scale[N,3], raw_quat[N,4],
    opacity_logit[N,1], sh[rows,48].

    Generation is chunked with one RNG per (scene seed, chunk index), so any row range can be
    produced independently: ``sh_rows=(lo, hi)`` makes only that shard of the SH table (multi-GPU
    ranks), ``n`` makes only the first n Gaussians (a uniform subsample -- ids are i.i.d.), and
    ``sh_out`` may be a preallocated (pinned) array that is filled in place."""
    N = int(n if n is not None else spec.n_gaussians)   # how many to generate (a prefix of the scene)
    NF = int(spec.n_gaussians)                          # the scene's size: fixes spacing / extent
    seed = 1234 + spec.index
    if spec.shape == "object":
        spacing = (4.0 / 3.0 * math.pi * 27.0 / NF) ** (1.0 / 3.0)
        L = 3.0
    else:
        L = 0.5 * math.sqrt(NF / 1e4)
        spacing = (2 * L * 2 * L * 0.1 * L / NF) ** (1.0 / 3.0)
    mu = math.log(0.5 * spacing)
    xyz = np.empty((N, 3), np.float32)
    log_scale = np.empty((N, 3), np.float32)
    raw_quat = np.empty((N, 4), np.float32)
    opacity = np.empty((N, 1), np.float32)
    lo, hi = (0, N) if sh_rows is None else (int(sh_rows[0]), int(sh_rows[1]))
    sh = None
    if want_sh:
        sh = sh_out if sh_out is not None else np.empty((hi - lo, 48), np.float32)
        assert sh.shape[0] == hi - lo
    def gen_chunk(ci):
        s0 = ci * chunk
        e0 = min(N, s0 + chunk)
        m = e0 - s0
        # one stream per attribute, so that a shorter chunk (n < chunk) is a prefix of the full one
        rng = np.random.default_rng([seed, ci, 0])
        r_s, r_q, r_o = (np.random.default_rng([seed, ci, k]) for k in (1, 2, 3))
        if spec.shape == "object":
            # N(0, I) truncated to |x| < 3: draw extra and keep the first m accepted
            p = rng.standard_normal((int(m * 1.05) + 64, 3), dtype=np.float32)
            p = p[(p * p).sum(1) < 9.0]
            while len(p) < m:
                q = rng.standard_normal((m, 3), dtype=np.float32)
                p = np.concatenate([p, q[(q * q).sum(1) < 9.0]])
            xyz[s0:e0] = p[:m]
        else:
            u = rng.random((m, 3), dtype=np.float32)
            xyz[s0:e0, 0] = (2 * u[:, 0] - 1) * L
            xyz[s0:e0, 1] = (2 * u[:, 1] - 1) * L
            xyz[s0:e0, 2] = u[:, 2] * 0.1 * L
        log_scale[s0:e0] = mu + 0.4 * r_s.standard_normal((m, 3), dtype=np.float32)
        q = r_q.standard_normal((m, 4), dtype=np.float32)
        raw_quat[s0:e0] = q / np.linalg.norm(q, axis=1, keepdims=True)
        opacity[s0:e0] = 1.5 * r_o.standard_normal((m, 1), dtype=np.float32)
        a_, b_ = max(s0, lo), min(e0, hi)
        if want_sh and a_ < b_:
            dc = 0.5 * np.random.default_rng([seed, ci, 7]).standard_normal((m, 3), dtype=np.float32)
            rest = 0.1 * np.random.default_rng([seed, ci, 8]).standard_normal((m, 45), dtype=np.float32)
            sh[a_ - lo:b_ - lo, :3] = dc[a_ - s0:b_ - s0]
            sh[a_ - lo:b_ - lo, 3:] = rest[a_ - s0:b_ - s0]

    n_chunks = (N + chunk - 1) // chunk
    workers = min(n_chunks, workers or (os.cpu_count() or 1))
    if workers <= 1:
        for ci in range(n_chunks):
            gen_chunk(ci)
    else:
        # chunks are independent (own RNG streams, disjoint output rows) and numpy's generators drop the GIL
        import concurrent.futures as cf
        with cf.ThreadPoolExecutor(max_workers=workers) as ex:
            list(ex.map(gen_chunk, range(n_chunks)))
    return dict(xyz=xyz, log_scale=log_scale, raw_quat=raw_quat, opacity_logit=opacity, sh=sh,
                extent=L, spacing=spacing)


def look_at(eye, target, up=(0.0, 0.0, 1.0)):
    """Row-major 4x4 world->camera, OpenCV axes (x right, y down, z forward)."""
    eye = np.asarray(eye, np.float64)
    fwd = np.asarray(target, np.float64) - eye
    fwd /= np.linalg.norm(fwd)
    right = np.cross(fwd, np.asarray(up, np.float64))
    right /= np.linalg.norm(right)
    down = np.cross(fwd, right)
    R = np.stack([right, down, fwd])  # rows
    V = np.eye(4)
    V[:3, :3] = R
    V[:3, 3] = -R @ eye
    return V.astype(np.float32)


def intrinsics(width, height, fovx_deg):
    """scene/cameras.py:110-126 with FoVy chosen for square pixels."""
    tanx = math.tan(math.radians(fovx_deg) * 0.5)
    fx = width / (2 * tanx)
    fy = fx
    return np.array([[fx, 0, width / 2.0], [0, fy, height / 2.0], [0, 0, 1]], np.float32)


def make_cameras(spec: SceneSpec, n_cameras, extent, seed_offset=0):
    """Returns viewmats[n,4,4], Ks[n,3,3], campos[n,3] float32."""
    rng = np.random.default_rng(4321 + spec.index + 1000 * seed_offset)
    V, Kk, P = [], [], []
    K = intrinsics(spec.width, spec.height, spec.fovx_deg)
    for _ in range(n_cameras):
        if spec.shape == "object":
            ang = rng.uniform(0, 2 * math.pi)
            eye = np.array([4 * math.cos(ang), 4 * math.sin(ang), rng.uniform(-0.5, 1.5)])
            tgt = np.zeros(3)
        else:
            L = extent
            eye = np.array([rng.uniform(-0.8 * L, 0.8 * L), rng.uniform(-0.8 * L, 0.8 * L),
                            0.25 * L * rng.uniform(0.8, 1.2)])
            yaw = rng.uniform(0, 2 * math.pi)
            pitch = math.radians(45 + rng.uniform(-10, 10))
            d = np.array([math.cos(yaw) * math.cos(pitch), math.sin(yaw) * math.cos(pitch), -math.sin(pitch)])
            tgt = eye + d
        V.append(look_at(eye, tgt))
        Kk.append(K)
        P.append(eye.astype(np.float32))
    return np.stack(V), np.stack(Kk), np.stack(P)


def make_gt_images(spec: SceneSpec, n_cameras, seed=999):
    """uint8 [n,3,H,W] low-pass noise: bilinear upsample of 1/16-res uniform noise."""
    rng = np.random.default_rng(seed)
    H, W = spec.height, spec.width
    h, w = max(2, H // 16), max(2, W // 16)
    out = np.empty((n_cameras, 3, H, W), np.uint8)
    ys = np.linspace(0, h - 1, H)
    xs = np.linspace(0, w - 1, W)
    y0 = np.floor(ys).astype(int).clip(0, h - 2)
    x0 = np.floor(xs).astype(int).clip(0, w - 2)
    fy = (ys - y0)[:, None].astype(np.float32)
    fx = (xs - x0)[None, :].astype(np.float32)
    for i in range(n_cameras):
        lo = rng.random((3, h, w), dtype=np.float32)
        a = lo[:, y0][:, :, x0]
        b = lo[:, y0][:, :, x0 + 1]
        c = lo[:, y0 + 1][:, :, x0]
        d = lo[:, y0 + 1][:, :, x0 + 1]
        img = (a * (1 - fx) + b * fx) * (1 - fy) + (c * (1 - fx) + d * fx) * fy
        out[i] = (img * 255.0 + 0.5).astype(np.uint8)
    return out


def morton_order(xyz, bits=10, device=None):
    """Permutation that sorts points by the Morton (Z-order) code of their position, `bits` per axis, each axis scaled
    to its own range.  Training is invariant to the order of the Gaussians; the order decides how scattered the SH rows
    of one camera are in the pinned tables (random 192-byte reads against runs of neighbouring rows).  Computed with
    torch on `device` when given (a 100 M-point sort takes seconds there), else with numpy."""
    def spread(v, lib):   # 10 bits -> every third bit of 30
        v = (v | (v << 16)) & 0x030000FF
        v = (v | (v << 8)) & 0x0300F00F
        v = (v | (v << 4)) & 0x030C30C3
        v = (v | (v << 2)) & 0x09249249
        return v
    assert 1 <= bits <= 10
    if device is not None:
        import torch
        p = torch.from_numpy(np.ascontiguousarray(xyz, np.float32)).to(device)
        lo, hi = p.min(0).values, p.max(0).values
        q = ((p - lo) / (hi - lo).clamp_min(1e-30) * float((1 << bits) - 1)).round_().to(torch.int64)
        del p
        code = spread(q[:, 0], torch) | (spread(q[:, 1], torch) << 1) | (spread(q[:, 2], torch) << 2)
        del q
        return torch.argsort(code, stable=True).cpu().numpy()
    p = np.asarray(xyz, np.float32)
    lo, hi = p.min(0), p.max(0)
    q = np.rint((p - lo) / np.maximum(hi - lo, 1e-30) * float((1 << bits) - 1)).astype(np.uint32)
    code = spread(q[:, 0], np) | (spread(q[:, 1], np) << np.uint32(1)) | (spread(q[:, 2], np) << np.uint32(2))
    return np.argsort(code, kind="stable")
