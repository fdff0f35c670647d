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


def test_shard_map_equal_shards():
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    sys.path[:0] = [root, os.path.join(root, "clm-gs_b200")]
    from clmgs.dist_rows import ShardMap
    for N, G in ((1001, 2), (102_000_000, 8), (7, 8), (64, 1)):
        sm = ShardMap(N, G)
        assert sm.bounds[0] == 0 and sm.bounds[-1] == N and sm.shard * G >= N
        assert all(0 <= sm.bounds[g + 1] - sm.bounds[g] <= sm.shard for g in range(G))
        assert all(sm.bounds[g] == min(N, g * sm.shard) for g in range(G + 1))   # rank g's slice of a padded tensor


def _merge_worker(rank, world, port, N, needs, out_dir):
    """plan -> owner-side merge (no sort) -> gradient return accumulated through dst_index, as engine_dist does it;
    plus the node-shared host table: every rank writes its rows, every rank reads all rows."""
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    sys.path[:0] = [root, os.path.join(root, "clm-gs_b200")]
    from clmgs.dist_rows import RowExchange, ShardMap, merge_requests
    from clmgs.shm import SharedHostTable
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    shard = ShardMap(N, world)
    lo, hi = shard.range_of(rank)
    table = SharedHostTable(N, 48)
    table.array[lo:hi] = (np.arange(lo, hi, dtype=np.float32)[:, None] * 100 + np.arange(48, dtype=np.float32)[None])
    dist.barrier()
    full = np.arange(N, dtype=np.float32)[:, None] * 100 + np.arange(48, dtype=np.float32)[None]
    assert np.array_equal(table.array, full), "a rank does not see the other ranks' rows of the shared table"
    ex = RowExchange(shard, rank)
    need = torch.from_numpy(needs[rank])
    plan = ex.plan(need)
    uniq, dst_index, untouched, mask = merge_requests(plan.recv_local, hi - lo, want_untouched=True)
    assert torch.equal(uniq, torch.unique(plan.recv_local)) and bool((uniq[1:] > uniq[:-1]).all())
    assert torch.equal(uniq[dst_index], plan.recv_local)
    assert torch.equal(torch.sort(torch.cat([uniq, untouched]))[0], torch.arange(hi - lo))
    grads = torch.full((need.shape[0], 48), float(rank + 1))
    recv = torch.empty(plan.recv_local.shape[0], 48)
    dist.all_to_all_single(recv, grads, output_split_sizes=plan.recv_counts, input_split_sizes=plan.send_counts)
    acc = torch.zeros(uniq.shape[0], 48).index_add_(0, dst_index, recv)
    my_grad = torch.zeros(hi - lo, 48)
    my_grad[uniq] = acc
    np.save(os.path.join(out_dir, "grad_%d.npy" % rank), my_grad.numpy())
    dist.barrier()
    table.close()
    dist.destroy_process_group()


def test_owner_merge_and_shared_table_two_ranks(tmp_path):
    N, world = 1003, 2
    rng = np.random.default_rng(1)
    needs = [np.sort(rng.choice(N, 500, replace=False)).astype(np.int64),
             np.sort(rng.choice(N, 300, replace=False)).astype(np.int64)]
    mp.spawn(_merge_worker, args=(world, _free_port(), N, needs, str(tmp_path)), nprocs=world, join=True)
    ref = np.zeros((N, 48), np.float32)
    for r in range(world):
        ref[needs[r]] += r + 1
    got = np.concatenate([np.load(tmp_path / ("grad_%d.npy" % r)) for r in range(world)])
    np.testing.assert_array_equal(got, ref)


def _chunk_worker(rank, world, port, N, K, needs, tags, out_dir):
    """plan_chunked end to end on CPU: K exchanges, owner accumulates through dst_index, a row is 'final' after the
    exchange given by final_counts; afterwards every touched row holds the sum of all ranks' contributions."""
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    sys.path[:0] = [root, os.path.join(root, "clm-gs_b200")]
    from clmgs.dist_rows import RowExchange, ShardMap, plan_chunked
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    shard = ShardMap(N, world)
    lo, hi = shard.range_of(rank)
    ex = RowExchange(shard, rank)
    need, tag = torch.from_numpy(needs[rank]), torch.from_numpy(tags[rank])
    plan = plan_chunked(ex, need, tag, K, hi - lo, want_untouched=True)
    # slot order = (chunk, id); chunk slices owner-major
    assert torch.equal(torch.sort(plan.send_order)[0], need)
    off = 0
    for c in range(K):
        n_c = sum(plan.send_counts[c])
        ids_c = plan.send_order[off:off + n_c]
        assert bool((ids_c[1:] > ids_c[:-1]).all()) and torch.equal(ids_c, need[tag == c])
        off += n_c
    acc = torch.zeros(plan.acc_order.shape[0], 48)
    final_seen = torch.zeros(plan.acc_order.shape[0], dtype=torch.bool)
    grads = torch.full((need.shape[0], 48), float(rank + 1))      # rows in slot order; value identifies the sender
    off, f_off = 0, 0
    for c in range(K):
        n_c = sum(plan.send_counts[c])
        recv = torch.empty(sum(plan.recv_counts[c]), 48)
        dist.all_to_all_single(recv, grads[off:off + n_c].contiguous(), output_split_sizes=plan.recv_counts[c],
                               input_split_sizes=plan.send_counts[c])
        acc.index_add_(0, plan.dst_index[c], recv)
        # peer-push mode: the requester already knows each row's slot in its owner's accumulator
        slots = torch.empty(sum(plan.recv_counts[c]), dtype=torch.int64)
        dist.all_to_all_single(slots, plan.remote_slot[off:off + n_c].contiguous(), output_split_sizes=plan.recv_counts[c],
                               input_split_sizes=plan.send_counts[c])
        assert torch.equal(slots, plan.dst_index[c])
        # nothing arrives for a row after its final chunk
        assert not bool(final_seen[plan.dst_index[c]].any()), "a contribution arrived after the row was declared final"
        final_seen[f_off:f_off + plan.final_counts[c]] = True
        off += n_c
        f_off += plan.final_counts[c]
    assert bool(final_seen.all()) and f_off == plan.acc_order.shape[0]
    my_grad = torch.zeros(hi - lo, 48)
    my_grad[plan.acc_order] = acc
    assert torch.equal(torch.sort(torch.cat([plan.acc_order, plan.untouched]))[0], torch.arange(hi - lo))
    assert torch.equal(plan.touched_mask.nonzero().flatten(), torch.sort(plan.acc_order)[0])
    np.save(os.path.join(out_dir, "grad_%d.npy" % rank), my_grad.numpy())
    dist.barrier()
    dist.destroy_process_group()


def test_chunked_gradient_return_two_ranks(tmp_path):
    N, world, K = 1003, 2, 4
    rng = np.random.default_rng(2)
    needs = [np.sort(rng.choice(N, 600, replace=False)).astype(np.int64),
             np.sort(rng.choice(N, 350, replace=False)).astype(np.int64)]
    tags = [rng.integers(0, K, len(n)).astype(np.int64) for n in needs]
    tags[1][:] = np.where(np.arange(len(tags[1])) % 3 == 0, K - 1, tags[1])   # uneven chunks
    mp.spawn(_chunk_worker, args=(world, _free_port(), N, K, needs, tags, str(tmp_path)), nprocs=world, join=True)
    ref = np.zeros((N, 48), np.float32)
    for r in range(world):
        ref[needs[r]] += r + 1
    got = np.concatenate([np.load(tmp_path / ("grad_%d.npy" % r)) for r in range(world)])
    np.testing.assert_array_equal(got, ref)
