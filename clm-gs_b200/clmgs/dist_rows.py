"""Gaussian-sharded SH rows across ranks (SURVEY.md 8e -- new design, the reference is single-GPU).

Rank g owns the contiguous id range [bounds[g], bounds[g+1]) of the offloaded table: its pinned
host rows, gradient table and host-Adam moments.  A rank that renders a camera needs rows owned by
others; per batch it asks every owner once (ids), owners gather from their pinned shard and the rows
travel over NCCL all-to-all-v (NVLink); gradients of touched rows go back the same way and are summed
at the owner (several ranks may touch one row).

Device-agnostic on purpose: the only collectives are torch.distributed.all_to_all_single, and row
gather / accumulate are callbacks (CUDA kernels in the engine, plain indexing in the gloo CPU tests).
"""
from dataclasses import dataclass
from typing import Callable, List

import numpy as np
import torch
import torch.distributed as dist


class ShardMap:
    """Equal shards of `shard` = ceil(N / world) ids (the last one may be shorter): the same partition serves the
    host tables and the NCCL reduce-scatter / all-gather of the GPU-resident attributes, which need equal counts
    (tensors are padded to world x shard rows)."""

    def __init__(self, n_total: int, world: int):
        self.n_total, self.world = int(n_total), int(world)
        self.shard = -(-self.n_total // self.world)
        self.bounds = [min(self.n_total, g * self.shard) for g in range(self.world + 1)]

    def range_of(self, rank):
        return self.bounds[rank], self.bounds[rank + 1]

    def n_local(self, rank):
        return self.bounds[rank + 1] - self.bounds[rank]


@dataclass
class Plan:
    need_ids: torch.Tensor      # ascending global ids this rank needs (its cameras' union)
    send_counts: List[int]      # how many of need_ids each owner holds (owner-major == ascending)
    recv_counts: List[int]      # how many ids each rank asks me for
    recv_local: torch.Tensor    # the requested rows as LOCAL row indices of my shard, requester-major


class RowExchange:
    def __init__(self, shard: ShardMap, rank: int, group=None):
        self.shard, self.rank, self.group = shard, rank, group
        self.world = shard.world

    def plan(self, need_ids: torch.Tensor) -> Plan:
        dev = need_ids.device
        inner = torch.tensor(self.shard.bounds[1:-1], dtype=need_ids.dtype, device=dev)
        cuts = torch.searchsorted(need_ids, inner).tolist() if self.world > 1 else []
        edges = [0] + cuts + [need_ids.shape[0]]
        send_counts = [edges[i + 1] - edges[i] for i in range(self.world)]
        cnt_in = torch.tensor(send_counts, dtype=torch.int64, device=dev)
        cnt_out = torch.empty(self.world, dtype=torch.int64, device=dev)
        dist.all_to_all_single(cnt_out, cnt_in, group=self.group)
        recv_counts = cnt_out.tolist()
        recv_ids = torch.empty(sum(recv_counts), dtype=need_ids.dtype, device=dev)
        dist.all_to_all_single(recv_ids, need_ids.contiguous(), output_split_sizes=recv_counts,
                               input_split_sizes=send_counts, group=self.group)
        lo, hi = self.shard.range_of(self.rank)
        recv_local = recv_ids - lo
        return Plan(need_ids, send_counts, recv_counts, recv_local)

    def fetch(self, plan: Plan, gather_local: Callable[[torch.Tensor], torch.Tensor], cols=48) -> torch.Tensor:
        """rows[i] = table row of need_ids[i].  gather_local(local_idx) -> [n, cols] from my shard."""
        send_rows = gather_local(plan.recv_local)
        rows = torch.empty(plan.need_ids.shape[0], cols, dtype=send_rows.dtype, device=send_rows.device)
        dist.all_to_all_single(rows, send_rows.contiguous(), output_split_sizes=plan.send_counts,
                               input_split_sizes=plan.recv_counts, group=self.group)
        return rows

    def return_grads(self, plan: Plan, grad_rows: torch.Tensor,
                     accumulate_local: Callable[[torch.Tensor, torch.Tensor], None]):
        """grad_rows aligned with need_ids; every owner gets accumulate_local(local_idx, rows) with the
        contributions of all ranks (local_idx may repeat across requesters: the callback must sum)."""
        recv = torch.empty(plan.recv_local.shape[0], grad_rows.shape[1], dtype=grad_rows.dtype, device=grad_rows.device)
        dist.all_to_all_single(recv, grad_rows.contiguous(), output_split_sizes=plan.recv_counts,
                               input_split_sizes=plan.send_counts, group=self.group)
        accumulate_local(plan.recv_local, recv)


def merge_requests(recv_local: torch.Tensor, n_local: int, want_untouched: bool = False):
    """Owner side of the sparse reduce-scatter: several ranks return gradients for the same row.  From the
    requester-major list of requested LOCAL rows (ascending per requester, repeats across requesters) build, without
    sorting (a flag table over the shard + its compaction -- the lists are already ordered):
      uniq       ascending distinct local rows (= the rows this rank's host Adam updates)
      dst_index  position of every received row in `uniq` (where its gradient is accumulated)
      untouched  ascending local rows nobody asked for (dense Adam still decays their moments), or None.
    One host sync (the size of uniq).  Device-agnostic: torch indexing only."""
    dev = recv_local.device
    mask = torch.zeros(n_local, dtype=torch.bool, device=dev)
    mask[recv_local] = True
    uniq = torch.nonzero(mask).flatten()
    pos = torch.empty(n_local, dtype=torch.int64, device=dev)
    pos[uniq] = torch.arange(uniq.shape[0], dtype=torch.int64, device=dev)
    dst_index = pos[recv_local]
    untouched = torch.nonzero(~mask).flatten() if want_untouched else None
    return uniq, dst_index, untouched, mask


@dataclass
class ChunkPlan:
    """Gradient return in K chunks (see plan_chunked)."""
    K: int
    send_order: torch.Tensor          # this rank's needed ids in slot order: (last chunk, id) -- chunk slices are owner-major
    send_counts: List[List[int]]      # [K][G]: rows of chunk c owned by rank o
    recv_counts: List[List[int]]      # [K][G]: rows of chunk c arriving from rank s
    acc_order: torch.Tensor           # owner: touched local rows in accumulator order: (final chunk, local row)
    final_counts: List[int]           # [K]: rows whose last contribution arrives with chunk c
    dst_index: List[torch.Tensor]     # [K]: accumulator slot of every row received in exchange c (source-major)
    untouched: torch.Tensor           # owner: local rows nobody touches (or None)
    touched_mask: torch.Tensor        # owner: bool [n_local]
    rows_from_peers: int
    remote_slot: torch.Tensor = None  # aligned with send_order: the row's slot in ITS OWNER's accumulator (peer push)


def plan_chunked(exchange: "RowExchange", need_ids: torch.Tensor, last_chunk: torch.Tensor, K: int, n_local: int,
                 want_untouched: bool = False) -> ChunkPlan:
    """Sparse reduce-scatter of touched-row gradients, pipelined in K chunks of the local camera sequence.

    A rank's gradient for a row is final once the last of ITS cameras that sees the row is done (chunk
    last_chunk[i] for need_ids[i]); it is sent with that chunk's exchange.  An owner may get contributions for one row
    from several ranks in different exchanges, so the row's sum is final -- and can go to the owner's host Adam --
    after the latest of them.  All lists involved are ascending, so the whole plan is flag tables and compactions
    (no sort).  need_ids: ascending int64 global ids; last_chunk: int64 in [0, K) aligned with need_ids."""
    G, dev = exchange.world, need_ids.device
    shard = exchange.shard
    lo, _ = shard.range_of(exchange.rank)
    inner = torch.tensor(shard.bounds[1:-1], dtype=torch.int64, device=dev)
    cuts = torch.searchsorted(need_ids, inner).tolist() if G > 1 else []
    tot_edges = [0] + cuts + [need_ids.shape[0]]
    send_tot = [tot_edges[i + 1] - tot_edges[i] for i in range(G)]
    cnt_out = torch.empty(G, dtype=torch.int64, device=dev)
    dist.all_to_all_single(cnt_out, torch.tensor(send_tot, dtype=torch.int64, device=dev), group=exchange.group)
    recv_tot = cnt_out.tolist()
    # ids and chunk tags travel as one word (owner-major = ascending order)
    recv_packed = torch.empty(sum(recv_tot), dtype=torch.int64, device=dev)
    dist.all_to_all_single(recv_packed, (need_ids * K + last_chunk).contiguous(), output_split_sizes=recv_tot,
                           input_split_sizes=send_tot, group=exchange.group)
    recv_local = torch.div(recv_packed, K, rounding_mode="floor") - lo
    recv_tag = recv_packed % K
    # requester: slots in (chunk, id) order; per chunk the rows are ascending = owner-major
    order_parts, send_counts = [], []
    for c in range(K):
        ids_c = need_ids[last_chunk == c]
        order_parts.append(ids_c)
        cc = torch.searchsorted(ids_c, inner).tolist() if G > 1 else []
        e = [0] + cc + [ids_c.shape[0]]
        send_counts.append([e[i + 1] - e[i] for i in range(G)])
    send_order = torch.cat(order_parts) if order_parts else need_ids
    # owner: final chunk of every touched local row = latest chunk in which anybody sends it
    fc = torch.full((n_local,), -1, dtype=torch.int64, device=dev)
    if recv_local.numel():
        fc.scatter_reduce_(0, recv_local, recv_tag, reduce="amax", include_self=True)
    acc_parts, final_counts = [], []
    for c in range(K):
        rows_c = torch.nonzero(fc == c).flatten()
        acc_parts.append(rows_c)
        final_counts.append(int(rows_c.shape[0]))
    acc_order = torch.cat(acc_parts) if acc_parts else torch.zeros(0, dtype=torch.int64, device=dev)
    pos = torch.empty(n_local, dtype=torch.int64, device=dev)
    pos[acc_order] = torch.arange(acc_order.shape[0], dtype=torch.int64, device=dev)
    seg_edges = torch.tensor(np.cumsum([0] + recv_tot), dtype=torch.int64, device=dev)
    dst_index, recv_counts = [], []
    for c in range(K):
        sel = torch.nonzero(recv_tag == c).flatten()          # keeps source-major, ascending order
        dst_index.append(pos[recv_local[sel]])
        e = torch.searchsorted(sel, seg_edges).tolist()
        recv_counts.append([e[i + 1] - e[i] for i in range(G)])
    touched = fc >= 0
    untouched = torch.nonzero(~touched).flatten() if want_untouched else None
    # every requester also learns where its rows live in their owners' accumulators, so that the gradient return can
    # be a direct accumulate into peer memory instead of send + receive + accumulate
    slot_at_owner = torch.empty(need_ids.shape[0], dtype=torch.int64, device=dev)
    dist.all_to_all_single(slot_at_owner, pos[recv_local].contiguous(), output_split_sizes=send_tot,
                           input_split_sizes=recv_tot, group=exchange.group)
    remote_slot = torch.cat([slot_at_owner[last_chunk == c] for c in range(K)]) if K else slot_at_owner
    return ChunkPlan(K, send_order, send_counts, recv_counts, acc_order, final_counts, dst_index, untouched, touched,
                     int(need_ids.shape[0] - send_tot[exchange.rank]), remote_slot)
