"""The multi-GPU engine (clmgs/engine_dist.py) against the single-GPU engine on the same global batch, launched the
way bench.py --gpus N is: torchrun, one process per GPU, NCCL.  With one GPU the same script runs as a world of 1 --
every collective degenerates but the whole code path (node-shared SH table, plan / merge, chunked owner flush with
flags, reduce-scatter -> sharded Adam -> all-gather) executes; with >= 2 GPUs it runs as a world of 2 as well."""
import os
import socket
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


def _run(world, extra):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(world), "--master-addr",
           "127.0.0.1", "--master-port", str(_free_port()), os.path.join(ROOT, "tests", "dist_parity_check.py")] + extra
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, (r.stdout[-3000:], r.stderr[-3000:])
    assert "DIST PARITY OK" in r.stdout, r.stdout[-3000:]


@pytest.mark.parametrize("extra", [[], ["--sparse"]], ids=["dense_adam", "sparse_adam"])
def test_dist_engine_world_1(extra):
    _run(1, extra)


@pytest.mark.parametrize("extra", [[], ["--sparse"], ["--nccl"], ["--sparse", "--nccl"]],
                         ids=["dense_adam_peer", "sparse_adam_peer", "dense_adam_nccl", "sparse_adam_nccl"])
def test_dist_engine_world_2(extra):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs (run with gpurun --gpus 2); the world-of-1 test above covers the code path")
    _run(2, extra)
