"""Golden fixture for a C4-SHAPED whole training iteration (BASELINE.json configs[3] shape: aerial scene, 1920x1080,
batch 64 -> 64-bit visibility words, sparse Adam, no densification; 1 M Gaussians so that the CPU oracle finishes in
minutes).  Written by the oracle (oracle/train_step.py, itself pinned by tests/test_oracle_pin.py); the GPU test
tests/test_gpu_engine.py::test_c4_shaped_iteration_matches_golden runs the same seeded inputs through the CUDA
engine and compares.  Takes ~7 minutes on 8 cores, which is why its output is committed instead of recomputed
in the test:

    python tests/golden/make_c4_iteration.py        # writes tests/golden/c4_iteration.npz

Stored: per-camera losses / visible counts / tile-intersection counts (camera order = input order; the engine's TSP
order only changes fp32 summation order), and the post-step state (parameters + both Adam moments of every attribute)
of a sample of Gaussians: every 355th touched one and every 1000th untouched one.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path[:0] = [ROOT, os.path.join(ROOT, "clm-gs_b200")]

N, W, H, BSZ, N_GT = 1_000_000, 1920, 1080, 64, 4


def inputs():
    from clmgs import model, scene as S
    spec = S.SceneSpec("c4_shaped_1m", N, W, H, BSZ, "aerial", 50.0, 3)
    g = S.make_gaussians(spec)
    V, K, P = S.make_cameras(spec, BSZ, g["extent"])
    gts = S.make_gt_images(spec, N_GT)
    args = model.TrainArgs(bsz=BSZ, sparse_adam=True)
    return spec, g, V, K, P, gts, args


def main():
    from oracle import train_step as TS
    spec, g, V, K, P, gts, args = inputs()
    st = TS.new_state(g)
    gts_all = gts[np.arange(BSZ) % N_GT]
    losses, filters, stats = TS.train_one_batch(st, V, K, P, gts_all, W, H, args, list(range(BSZ)))
    touched = np.zeros(N, bool)
    for f in filters:
        touched[f] = True
    t_ids, u_ids = np.nonzero(touched)[0], np.nonzero(~touched)[0]
    sample = np.sort(np.concatenate([t_ids[::355], u_ids[::1000]])).astype(np.int64)
    out = dict(losses=np.asarray(losses, np.float32), n_visible=np.array([s["n_visible"] for s in stats], np.int64),
               n_isects=np.array([s["n_isects"] for s in stats], np.int64),
               examined=np.array([s["examined"] for s in stats], np.int64),
               n_touched=np.int64(touched.sum()), sample=sample, sample_touched=touched[sample])
    for k in ("xyz", "log_scale", "raw_quat", "opacity_logit", "sh"):
        for pre in ("", "m_", "v_"):
            out[pre + k] = st[pre + k][sample]
    np.savez_compressed(os.path.join(HERE, "c4_iteration.npz"), **out)
    print("wrote c4_iteration.npz:", {k: getattr(v, "shape", v) for k, v in out.items()})


if __name__ == "__main__":
    main()
