"""Row-sharded serving over several GPUs of one box: one process per GPU, torch.distributed for the
plumbing (NCCL over NVLink on GPUs; gloo on CPU for the host-logic tests).

The reference's serving layer is not distributed (src/site/markdown/docs/admin.md:253-254); its
intra-query parallelism is one task per LSH partition on a thread pool
(ALSServingModel.java:295-333) followed by `sorted(DESC).limit(howMany)` (:334).  Here the same
shape is spread over GPUs: Y is split by item rows, every rank scans its shard (oryx_topn_device)
and ONE exchange step -- an all-gather of howMany (score, row) pairs plus a count per rank -- is
followed by the merge (oryx_merge_topn_device) on every rank.  YtY partials are summed with an
all-reduce, which is exactly the reference's sum over partitions (ALSServingModel.java:362-368).

This module holds only host logic (who owns which row, buffer layout of the exchange, global row
ids); the compute is injected (`local_topn`, `merge`) so the N>1 path can be exercised on CPU with
gloo, where tests inject the oracle.
"""
from __future__ import annotations

from dataclasses import dataclass

import numpy as np


@dataclass(frozen=True)
class ShardPlan:
    """Contiguous block sharding: global row g lives on rank g // rows_per_rank as local row g % rows_per_rank."""
    world: int
    rows_per_rank: int

    @staticmethod
    def for_rows(n_rows: int, world: int) -> "ShardPlan":
        return ShardPlan(world, (n_rows + world - 1) // world)

    def owner(self, global_row: int) -> int:
        return global_row // self.rows_per_rank

    def local(self, global_row: int) -> int:
        return global_row % self.rows_per_rank

    def to_global(self, rank: int, local_row):
        return rank * self.rows_per_rank + local_row

    def local_range(self, rank: int, n_rows: int):
        lo = rank * self.rows_per_rank
        return lo, max(lo, min(n_rows, lo + self.rows_per_rank))

    def route(self, global_rows: np.ndarray):
        """Split an update batch by owner: {rank: (positions in the batch, local rows)}."""
        g = np.asarray(global_rows, dtype=np.int64)
        owners = g // self.rows_per_rank
        out = {}
        for r in np.unique(owners):
            pos = np.nonzero(owners == r)[0]
            out[int(r)] = (pos, (g[pos] % self.rows_per_rank).astype(np.int32))
        return out

    def split_exclusions(self, rank: int, excluded_global) -> np.ndarray:
        """The part of a known-items exclusion list that lives on `rank`, as local rows."""
        g = np.asarray(excluded_global, dtype=np.int64)
        if g.size == 0:
            return np.zeros(0, dtype=np.int32)
        mine = g[(g // self.rows_per_rank) == rank]
        return (mine % self.rows_per_rank).astype(np.int32)


def exchange_and_merge(dist, torch, plan: ShardPlan, how_many: int, local_scores, local_rows, local_n, merge,
                       buffers=None):
    """The one exchange step of a sharded query.

    local_scores [how_many] f64, local_rows [how_many] i32, local_n [1] i32 are this rank's result
    (tensors on the rank's device).  All ranks receive all lists (all-gather: N x howMany x 12 B, a
    latency-bound message) and call `merge(g_scores, g_rows, g_n, plan)` -> (rows int64, scores f64).
    """
    world = plan.world
    # flat receive buffers (rank-major), viewed as [world, how_many] for the merge; pass `buffers`
    # (the tuple this function would allocate) to reuse them across queries
    if buffers is None:
        buffers = (torch.empty(world * how_many, dtype=local_scores.dtype, device=local_scores.device),
                   torch.empty(world * how_many, dtype=local_rows.dtype, device=local_rows.device),
                   torch.empty(world, dtype=local_n.dtype, device=local_n.device))
    g_scores, g_rows, g_n = buffers
    dist.all_gather_into_tensor(g_scores, local_scores)
    dist.all_gather_into_tensor(g_rows, local_rows)
    dist.all_gather_into_tensor(g_n, local_n)
    return merge(g_scores.view(world, how_many), g_rows.view(world, how_many), g_n, plan)


def exchange_packed(dist, torch, plan: ShardPlan, local_record, gathered=None):
    """Single-collective form of the exchange: every rank contributes one packed record (uint8 tensor, see
    oryx_merge_topn_packed) and receives all of them, rank-major."""
    if gathered is None:
        gathered = torch.empty(plan.world * local_record.numel(), dtype=local_record.dtype,
                               device=local_record.device)
    dist.all_gather_into_tensor(gathered, local_record)
    return gathered


def allreduce_gram(dist, torch, gram_partial):
    """YtY = sum of per-shard Grams (ALSServingModel.java:362-368), fp64, in place."""
    dist.all_reduce(gram_partial, op=dist.ReduceOp.SUM)
    return gram_partial
