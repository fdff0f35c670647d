"""CPU-side product code (no GPU needed): the C ABI loads and exports every
declared symbol, host Adam and the TSP heuristic work, layout rules hold."""
import os
import re
import subprocess
import sys
import threading
import time

import numpy as np
import pytest

import oracle

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_abi_exports_every_declared_symbol():
    from clmgs import _lib
    L = _lib.lib()
    assert len(L.protos) >= 38
    assert L.missing_symbols() == []
    assert L.call("clmgs_abi_version") == 1
    out = subprocess.check_output(["nm", "-D", "--defined-only", _lib.LIB_PATH], text=True)
    exported = set(re.findall(r"\b(clmgs_\w+)\b", out))
    assert set(L.protos) <= exported


def test_abi_reports_errors_without_a_gpu():
    from clmgs import _lib
    import pytest
    with pytest.raises(_lib.ClmgsError, match="degree"):
        _lib.call("clmgs_sh_fwd", 10, 16, 7, None, None, None, None, None)
    with pytest.raises(_lib.ClmgsError, match="tile_size"):
        _lib.call("clmgs_rasterize_fwd", 1, 1, 0, None, None, None, None, None, 16, 16, 8, 1, 1, None, None, None,
                  None, None, None)


def test_product_never_touches_the_oracle():
    bad = []
    for dp, _, fs in os.walk(os.path.join(ROOT, "clm-gs_b200")):
        if "build" in dp.split(os.sep):
            continue
        for f in fs:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h")):
                s = open(os.path.join(dp, f), errors="ignore").read()
                if re.search(r"import oracle|from oracle|clmgs_oracle|oracle/", s):
                    bad.append(os.path.join(dp, f))
    assert not bad, bad


def test_cpu_adam_matches_oracle_and_waits_for_signals():
    from clmgs import ops
    rng = np.random.default_rng(0)
    N, bsz = 5000, 4
    p = rng.standard_normal((N, 48)).astype(np.float32)
    g = rng.standard_normal((N, 48)).astype(np.float32)
    m = np.zeros_like(p); v = np.zeros_like(p)
    col_lr = np.array([2.5e-3] * 3 + [2.5e-3 / 20] * 45, np.float32)
    key = rng.integers(0, bsz + 1, N)
    rows = np.argsort(key, kind="stable").astype(np.int32)
    offs = np.concatenate([[0], np.cumsum(np.bincount(key, minlength=bsz + 1))]).astype(np.int64)
    for sparse in (0, 1):
        pr, gr, mr, vr = p.copy(), g.copy(), m.copy(), v.copy()
        sel = rows[offs[1]:] if sparse else rows
        oracle.cpu_adam_rows(pr, gr, mr, vr, sel, col_lr, 0.9 ** 4, 0.999 ** 4, 1e-15 / 2, 2, 0.25)
        pp, gg, mm, vv = p.copy(), g.copy(), m.copy(), v.copy()
        signal = np.zeros(bsz, np.int32)

        def fire():
            for k in range(bsz):
                time.sleep(0.02)
                signal[k] = 1

        th = threading.Thread(target=fire); th.start()
        t0 = time.time()
        first = 1 if sparse else 0   # sparse_adam: the untouched group is simply not passed
        seg_b, seg_e = offs[first:-1], offs[first + 1:]
        seg_sig = np.arange(first, bsz + 1) - 1   # group m waits on flag m-1; group 0 (-1) does not wait
        prog = np.zeros(1, np.int64)
        ops.cpu_adam_batched_sparse_step(pp, gg, mm, vv, col_lr, rows, seg_b, seg_e, seg_sig, signal, 0.9 ** 4,
                                         0.999 ** 4, 1e-15 / 2, 2, 0.25, n_threads=3, progress=prog)
        assert prog[0] == len(seg_b)
        dt = time.time() - t0
        th.join()
        assert dt > 0.06, "did not wait for the per-camera signals"
        np.testing.assert_allclose(pp, pr, rtol=1e-5, atol=1e-7)
        np.testing.assert_allclose(mm, mr, rtol=1e-5, atol=1e-8)
        np.testing.assert_allclose(vv, vr, rtol=1e-5, atol=1e-10)
        assert not gg[sel].any()
        if sparse:
            untouched = rows[: offs[1]]
            assert (pp[untouched] == p[untouched]).all() and (gg[untouched] == g[untouched]).all()


def test_tsp_tour_is_a_permutation_and_not_worse_than_identity():
    from clmgs import ops
    rng = np.random.default_rng(0)
    for n in (4, 8, 16, 64):
        pts = rng.random((n, 2))
        d = (np.linalg.norm(pts[:, None] - pts[None], axis=-1) * 1000).astype(np.int32)
        tour = ops.tsp_tour(d)
        assert sorted(tour) == list(range(n)) and tour[0] == 0
        cost = lambda t: sum(d[t[i], t[(i + 1) % n]] for i in range(n))
        assert cost(tour) <= cost(list(range(n)))
    assert ops.tsp_tour(np.zeros((1, 1), np.int32)) == [0]


def test_cpu_adam_zero_grad_segments_match_reading_zero_rows():
    from clmgs import ops
    rng = np.random.default_rng(3)
    N = 2000
    p = rng.standard_normal((N, 48)).astype(np.float32)
    m = rng.standard_normal((N, 48)).astype(np.float32) * 0.1
    v = rng.random((N, 48)).astype(np.float32) * 0.01
    col_lr = np.full(48, 1e-3, np.float32)
    rows = np.arange(N, dtype=np.int32)
    sig = np.zeros(1, np.int32)
    outs = []
    for zg in (None, [1]):
        pp, mm, vv, gg = p.copy(), m.copy(), v.copy(), np.zeros((N, 48), np.float32)
        ops.cpu_adam_batched_sparse_step(pp, gg, mm, vv, col_lr, rows, [0], [N], [-1], sig, 0.9, 0.999, 1e-15, 5, 0.25,
                                         n_threads=2, seg_zero_grad=zg)
        outs.append((pp, mm, vv))
    for a, b in zip(outs[0], outs[1]):
        np.testing.assert_array_equal(a, b)


def test_new_entry_points_validate_arguments_without_a_gpu():
    """Argument checks come before any CUDA call, so they can be exercised on a machine without a GPU."""
    from clmgs import _lib
    import pytest
    L = _lib.lib()
    with pytest.raises(_lib.ClmgsError, match="out of range"):
        L.call("clmgs_densify_classify", -1, None, None, None, None, 0.0, 0.0, 0.0, 0.0, 0, None, None)
    with pytest.raises(_lib.ClmgsError, match="null pointer"):
        L.call("clmgs_densify_classify", 10, None, None, None, None, 0.0, 0.0, 0.0, 0.0, 0, None, None)
    with pytest.raises(_lib.ClmgsError, match="totals"):
        L.call("clmgs_densify_plan", 10, None, None, None, None, None, 0, None)
    with pytest.raises(_lib.ClmgsError, match="bad counts"):
        L.call("clmgs_densify_apply", 1, 0, 2, 1, None, None, None, None, None, None)  # more kept parents than parents
    with pytest.raises(_lib.ClmgsError, match="max_opacity"):
        L.call("clmgs_reset_opacity", 4, np.zeros(4, np.float32), np.zeros(4, np.float32), np.zeros(4, np.float32),
               1.5, None)
    with pytest.raises(_lib.ClmgsError, match="null pointer"):
        L.call("clmgs_radix_sort_pairs_u32", 8, None, None, None, None, 32, None, 0, None, None)
    with pytest.raises(_lib.ClmgsError, match="null output"):
        L.call("clmgs_isect_sorted_offsets", 1, 8, None, None, None, None, None, 16, 2, 2, 4, None, None, None, 0, None)
    # sizes that need no device work succeed without touching the GPU
    assert L.call("clmgs_densify_classify", 0, None, None, None, None, 0.0, 0.0, 0.0, 0.0, 0, None, None) == 0
    assert L.call("clmgs_radix_sort_pairs_u32", 0, None, None, None, None, 32, None, 0, None, None) == 0
    assert L.call("clmgs_densify_plan_scratch_bytes", 1_000_000) > 0
    assert L.call("clmgs_radix_sort_pairs_u32_scratch_bytes", 46_000_000) >= (46_000_000 // 4096) * 256 * 4


def test_lazy_dense_adam_is_bit_identical_to_dense():
    """Postponing the zero-gradient updates of untouched rows and replaying them later must reproduce dense Adam
    exactly: same parameters and moments, bit for bit, after every materialisation."""
    import torch
    from clmgs import model
    rng = np.random.default_rng(5)
    n, steps = 3000, 14
    p0 = rng.standard_normal((n, 48)).astype(np.float32)

    def make():
        param = torch.from_numpy(p0.copy())
        return model.CpuAdam(param, [3, 45], [2.5e-3, 1.25e-4], [0.9 ** 4, 0.999 ** 4], 5e-16, n_threads=3, capacity=n + 10)

    dense, lazy = make(), make()
    gd, gl = torch.zeros(n, 48), torch.zeros(n, 48)
    sig = np.ones(1, np.int32)
    for it in range(steps):
        touched = np.nonzero(rng.random(n) < (0.05 if it % 5 else 0.6))[0].astype(np.int32)
        untouched = np.setdiff1d(np.arange(n, dtype=np.int32), touched).astype(np.int32)
        g = rng.standard_normal((len(touched), 48)).astype(np.float32)
        gd[touched] = torch.from_numpy(g)
        gl[touched] = torch.from_numpy(g)
        # dense: untouched rows (zero gradient, not read) + touched rows
        rows = np.concatenate([untouched, touched]).astype(np.int32)
        dense.batched_sparse_step(gd, rows, [0, len(untouched)], [len(untouched), n], [-1, -1], sig, 0.25, None, [1, 0])
        # lazy: replay what the touched rows skipped, then this step for them only
        if not lazy.lazy_pending:
            lazy.last_step[:n] = lazy.step_count
        lazy.catch_up(touched, include_this_step=True)
        lazy.lazy_pending = True
        lazy.batched_sparse_step(gl, touched, [0], [len(touched)], [-1], sig, 0.25, None, [0])
        assert not gl.numpy().any() and not gd.numpy().any()
        if it in (3, steps - 1):  # evaluation / densification / saving points
            lazy.catch_up(None)
            assert not lazy.lazy_pending and (lazy.last_step[:n] == lazy.step_count).all()
            np.testing.assert_array_equal(lazy.param.numpy(), dense.param.numpy())
            np.testing.assert_array_equal(lazy.exp_avg.numpy(), dense.exp_avg.numpy())
            np.testing.assert_array_equal(lazy.exp_avg_sq.numpy(), dense.exp_avg_sq.numpy())
    assert np.abs(dense.param.numpy() - p0).max() > 1e-3  # the optimiser did move things


def test_lazy_catch_up_rejects_bad_arguments():
    from clmgs import _lib
    import pytest
    L = _lib.lib()
    t = [np.zeros((4, 48), np.float32) for _ in range(3)]
    lr, ls = np.zeros(48, np.float32), np.zeros(4, np.int32)
    with pytest.raises(_lib.ClmgsError, match="bad arguments"):
        L.call("clmgs_cpu_adam_lazy_catch_up", t[0], t[1], t[2], 48, lr, None, 4, ls, 3, 2, 0.9, 0.999, 1e-8, 1)
    ls[1] = -5
    with pytest.raises(_lib.ClmgsError, match="negative"):
        L.call("clmgs_cpu_adam_lazy_catch_up", t[0], t[1], t[2], 48, lr, None, 4, ls, 3, 3, 0.9, 0.999, 1e-8, 1)


def test_cpu_adam_avx2_and_avx512_rows_are_bit_identical_and_stats_are_reported():
    """The host Adam picks AVX-512 rows where the CPU has them and AVX2 rows elsewhere; the two must give the same
    bits (they are built without FMA contraction).  Each ISA runs in its own process (the choice is made once)."""
    import subprocess
    import sys
    code = r'''
import sys, hashlib, numpy as np
sys.path[:0] = [%r, %r]
from clmgs import ops, _lib
rng = np.random.default_rng(11)
N = 30011
p = rng.standard_normal((N, 48)).astype(np.float32); g = rng.standard_normal((N, 48)).astype(np.float32)
m = (rng.standard_normal((N, 48)) * 0.1).astype(np.float32); v = (rng.random((N, 48)) * 0.01).astype(np.float32)
col_lr = np.array([2.5e-3] * 3 + [2.5e-3 / 20] * 45, np.float32)
rows = rng.permutation(N).astype(np.int32)
sig = np.ones(2, np.int32)
for step in (1, 2, 3):   # persistent pool: several calls, different thread counts
    ops.cpu_adam_batched_sparse_step(p, g, m, v, col_lr, rows, [0, N // 3], [N // 3, N], [0, 1], sig, 0.9, 0.999, 1e-15,
                                     step, 0.25, n_threads=1 + step, seg_zero_grad=[1, 0])
    g[:] = rng.standard_normal((N, 48)).astype(np.float32)
st = np.zeros(4)
_lib.call("clmgs_cpu_adam_last_stats", st)
assert st[2] == N and st[3] == N // 3 and st[0] > 0 and st[1] >= 0, st
print(hashlib.sha256(p.tobytes() + m.tobytes() + v.tobytes()).hexdigest())
''' % (ROOT, os.path.join(ROOT, "clm-gs_b200"))
    digests = {}
    for isa in ("avx2", "auto"):
        env = dict(os.environ, CLMGS_CPU_ADAM_ISA=isa)
        r = subprocess.run([sys.executable, "-c", code], env=env, capture_output=True, text=True, timeout=300)
        assert r.returncode == 0, r.stderr[-2000:]
        digests[isa] = r.stdout.strip().splitlines()[-1]
    assert digests["avx2"] == digests["auto"], digests


def test_cpu_adam_packed_gradients_equal_table_gradients():
    """clmgs_cpu_adam_batched_sparse_step_packed: the gradient of list entry r comes from row r - packed_base of a
    packed staging buffer; same update as reading (and re-zeroing) the [N, 48] table, table untouched."""
    from clmgs import ops
    rng = np.random.default_rng(5)
    N = 6000
    p = rng.standard_normal((N, 48)).astype(np.float32)
    m = (rng.standard_normal((N, 48)) * 0.1).astype(np.float32)
    v = (rng.random((N, 48)) * 0.01).astype(np.float32)
    col_lr = np.array([2.5e-3] * 3 + [2.5e-3 / 20] * 45, np.float32)
    rows = rng.permutation(N).astype(np.int32)
    n_unt = 1500                      # list = [untouched | group A | group B]
    edges = [n_unt, 3500, N]
    g_table = np.zeros((N, 48), np.float32)
    g_packed = rng.standard_normal((N - n_unt, 48)).astype(np.float32)
    g_table[rows[n_unt:]] = g_packed
    sig = np.ones(2, np.int32)
    seg_b, seg_e, seg_s, seg_z = [0, edges[0], edges[1]], [edges[0], edges[1], edges[2]], [-1, 0, 1], [1, 0, 0]
    a = [x.copy() for x in (p, m, v)]
    ops.cpu_adam_batched_sparse_step(a[0], g_table, a[1], a[2], col_lr, rows, seg_b, seg_e, seg_s, sig, 0.9, 0.999, 1e-15,
                                     3, 0.25, n_threads=3, seg_zero_grad=seg_z)
    assert not g_table.any()
    b = [x.copy() for x in (p, m, v)]
    untouched_table = np.full((N, 48), 7.0, np.float32)   # must not be read or written
    keep = g_packed.copy()
    ops.cpu_adam_batched_sparse_step(b[0], untouched_table, b[1], b[2], col_lr, rows, seg_b, seg_e, seg_s, sig, 0.9, 0.999,
                                     1e-15, 3, 0.25, n_threads=3, seg_zero_grad=seg_z, grad_packed=g_packed,
                                     packed_base=n_unt)
    for x, y in zip(a, b):
        np.testing.assert_array_equal(x, y)
    assert (untouched_table == 7.0).all() and np.array_equal(g_packed, keep)
    with pytest.raises(Exception):    # a gradient-reading segment before packed_base is refused
        ops.cpu_adam_batched_sparse_step(b[0], untouched_table, b[1], b[2], col_lr, rows, [0], [10], [-1], sig, 0.9, 0.999,
                                         1e-15, 4, 0.25, n_threads=1, seg_zero_grad=[0], grad_packed=g_packed,
                                         packed_base=n_unt)


def test_morton_order_is_a_locality_improving_permutation():
    """scene.morton_order (bench.py --spatial-order): a permutation; numpy and torch paths agree; neighbours in the
    order are neighbours in space."""
    from clmgs import scene as S
    rng = np.random.default_rng(3)
    p = (rng.random((50_000, 3)) * np.array([200.0, 200.0, 10.0])).astype(np.float32)
    perm = S.morton_order(p)
    assert np.array_equal(np.sort(perm), np.arange(len(p)))
    assert np.array_equal(perm, S.morton_order(p, device="cpu"))
    step = lambda a: np.linalg.norm(np.diff(a, axis=0), axis=1).mean()
    assert step(p[perm]) < 0.1 * step(p)
