"""clmgs_isect_count_static: the tile count for hosts that cannot hand the library any scratch
memory (no device allocator of their own, every byte of HBM already assigned, or a loop that must
not hit a synchronising cudaMalloc/cudaFree).  It keeps its temporary in a fixed array inside
libclmgs.so instead.

Checked here, through the C ABI: two calls in a row (the second smaller, so the array still holds
the first call's values past the part it uses) give exactly what clmgs_isect_count gives with
caller scratch -- integer outputs, so the tolerance is zero, as in test_gpu_forward.test_isect_exact
-- and an oversized call is refused.  Collected last (file name) so a failure here cannot hide
another test's result under `-x`.
"""
import numpy as np
import pytest
import torch

from clmgs import _lib

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _inputs(n, seed, W=640, H=400):
    rng = np.random.default_rng(seed)
    means2d = (rng.random((n, 2)) * [W + 80, H + 80] - 40).astype(np.float32)
    radii = rng.integers(0, 40, n).astype(np.int32)  # zeros = culled Gaussians
    return torch.from_numpy(means2d).to(DEV), torch.from_numpy(radii).to(DEV), (W + 15) // 16, (H + 15) // 16


def _with_caller_scratch(means2d, radii, tw, th):
    n = radii.shape[0]
    tpg = torch.empty(n, dtype=torch.int32, device=DEV)
    cum = torch.empty(n, dtype=torch.int64, device=DEV)
    sb = _lib.call("clmgs_isect_count_scratch_bytes", n)
    scratch = torch.empty(max(sb, 16), dtype=torch.uint8, device=DEV)
    _lib.call("clmgs_isect_count", n, means2d, radii, 16, tw, th, tpg, cum, scratch, scratch.numel(),
              torch.cuda.current_stream().cuda_stream)
    return tpg, cum


def _static(means2d, radii, tw, th):
    n = radii.shape[0]
    tpg = torch.full((n,), -1, dtype=torch.int32, device=DEV)   # caller still owns inputs and outputs
    cum = torch.full((n,), -1, dtype=torch.int64, device=DEV)
    _lib.call("clmgs_isect_count_static", n, means2d, radii, 16, tw, th, tpg, cum,
              torch.cuda.current_stream().cuda_stream)
    return tpg, cum


def test_static_scratch_matches_caller_scratch_twice_in_a_row():
    for n, seed in ((300_017, 1), (41_003, 2)):   # big first, then small: stale entries stay behind
        means2d, radii, tw, th = _inputs(n, seed)
        ref_tpg, ref_cum = _with_caller_scratch(means2d, radii, tw, th)
        tpg, cum = _static(means2d, radii, tw, th)
        torch.cuda.synchronize()
        assert torch.equal(tpg, ref_tpg), n
        assert torch.equal(cum, ref_cum), n
        assert int(cum[-1]) == int(ref_tpg.sum()) and int(cum[-1]) > 0


def test_static_scratch_refuses_more_than_its_limit():
    limit = _lib.call("clmgs_isect_count_static_max")
    assert limit == 16 * 1024 * 1024
    with pytest.raises(_lib.ClmgsError, match="exceed the built-in scratch"):
        _lib.call("clmgs_isect_count_static", limit + 1, None, None, 16, 4, 4, None, None, None)
