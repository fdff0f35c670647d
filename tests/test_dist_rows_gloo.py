"""World-size-2 gloo test (CPU) of the sharded SH-row exchange used by the multi-GPU engine:
plan -> fetch rows from owners -> return gradients, summed at the owner."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


def _worker(rank, world, port, N, needs, out_dir):
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    sys.path[:0] = [root, os.path.join(root, "clm-gs_b200")]
    from clmgs.dist_rows import RowExchange, ShardMap
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    shard = ShardMap(N, world)
    lo, hi = shard.range_of(rank)
    table = (torch.arange(N, dtype=torch.float32)[:, None] * 100 + torch.arange(48, dtype=torch.float32)[None])
    my_rows = table[lo:hi].clone()
    my_grad = torch.zeros(hi - lo, 48)
    ex = RowExchange(shard, rank)
    need = torch.from_numpy(needs[rank])
    plan = ex.plan(need)
    rows = ex.fetch(plan, lambda idx: my_rows[idx])
    assert torch.equal(rows, table[need]), "fetched rows differ from the global table"
    grads = rows * 0 + (rank + 1)  # rank r contributes r+1 to every row it touched
    ex.return_grads(plan, grads, lambda idx, g: my_grad.index_add_(0, idx, g))
    np.save(os.path.join(out_dir, "grad_%d.npy" % rank), my_grad.numpy())
    dist.barrier()
    dist.destroy_process_group()


def test_row_exchange_two_ranks(tmp_path):
    N, world = 1001, 2
    rng = np.random.default_rng(0)
    needs = [np.sort(rng.choice(N, 400, replace=False)).astype(np.int64),
             np.sort(rng.choice(N, 250, replace=False)).astype(np.int64)]
    mp.spawn(_worker, args=(world, _free_port(), N, needs, str(tmp_path)), nprocs=world, join=True)
    ref = np.zeros((N, 48), np.float32)
    for r in range(world):
        ref[needs[r]] += r + 1
    got = np.concatenate([np.load(tmp_path / ("grad_%d.npy" % r)) for r in range(world)])
    np.testing.assert_array_equal(got, ref)


def test_row_exchange_empty_request(tmp_path):
    N, world = 64, 2
    needs = [np.arange(0, 64, 3, dtype=np.int64), np.zeros(0, np.int64)]  # rank 1 needs nothing
    mp.spawn(_worker, args=(world, _free_port(), N, needs, str(tmp_path)), nprocs=world, join=True)
    got = np.concatenate([np.load(tmp_path / ("grad_%d.npy" % r)) for r in range(world)])
    ref = np.zeros((N, 48), np.float32); ref[needs[0]] = 1
    np.testing.assert_array_equal(got, ref)
