"""Time the densify / prune step at a BASELINE.json scene size on one B200.

    python tools/densify_bench.py [--config c2_bicycle_6m] [--reps 3]

Prints one JSON line: GPU part (classify + plan + apply; CUDA events), host part (in-place compaction of the
three pinned tables; wall clock) and the bytes each moves."""
import argparse
import json
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "clm-gs_b200"))
from clmgs import model, ops, scene as S  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", default="c2_bicycle_6m")
    ap.add_argument("--reps", type=int, default=3)
    a = ap.parse_args()
    spec = S.CONFIGS[a.config]
    N = spec.n_gaussians
    g = S.make_gaussians(spec)
    cap = int(N * 1.5)
    out = []
    for rep in range(a.reps):
        gm = model.GaussianModelCLMOffload(3, "cuda:0").create_from_arrays(g["xyz"], g["log_scale"], g["raw_quat"],
                                                                          g["opacity_logit"], g["sh"], capacity=cap)
        args = model.TrainArgs(bsz=spec.batch)
        gm.training_setup(args)
        gen = torch.Generator(device="cuda:0").manual_seed(rep)
        gm.xyz_gradient_accum.copy_(torch.rand(N, 1, device="cuda:0", generator=gen) * 2.5e-4)  # ~20 % above threshold
        gm.denom.fill_(1.0)
        extent = float(g["extent"])
        scale_med = float(torch.exp(gm._scaling).max(1).values.median())
        gm.percent_dense = scale_med / extent  # half clone, half split
        torch.cuda.synchronize()
        e0, e1, e2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
        t0 = time.perf_counter()
        e0.record()
        cls = ops.densify_classify(gm.xyz_gradient_accum, gm.denom, gm._scaling, gm._opacity, 0.0002,
                                   gm.percent_dense * extent, 0.005, None)
        src, noise_row, (nA, nB, nS, nC) = ops.densify_plan(cls)
        e1.record()
        torch.cuda.synchronize()
        t_plan = time.perf_counter() - t0
        t1 = time.perf_counter()
        info = gm._restructure(cls, generator=gen)
        torch.cuda.synchronize()
        t_total = time.perf_counter() - t1
        n_new = info["n_after"]
        out.append(dict(n_before=N, n_after=n_new, cloned=info["n_cloned"], split=info["n_split_parents"],
                        pruned=info["n_pruned"], classify_plan_ms=round(e0.elapsed_time(e1), 3),
                        restructure_wall_ms=round(t_total * 1e3, 1), plan_wall_ms=round(t_plan * 1e3, 1)))
        del gm
    best = min(out, key=lambda o: o["restructure_wall_ms"])
    best["host_table_bytes_moved"] = int(3 * 192 * best["n_after"] * 2)
    best["config"] = a.config
    print(json.dumps(best))


if __name__ == "__main__":
    main()
