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


class PipelinedShardedTopN:
    """Back-to-back sharded queries without a bubble between them.

    The scan of query i+1 does not depend on the exchange of query i, so the all-gather and the merge
    run on a side stream (on the SM the model leaves free, ``reserved_sms=1``) while the main stream
    already scans the next query.  Two packed result records alternate; ``ready`` events order
    scan -> exchange, ``consumed`` events keep scan i+2 from overwriting a record the all-gather of
    query i still reads.  (Measured alternative, rejected: three records with the exchange trailing by
    one query so that no event sits between a finishing kernel and the next scan -- 0.857 ms/step
    against 0.758 ms/step for this arrangement at N=2.)
    ``drain()`` makes the main stream wait for everything submitted so far.
    """

    NB = 2

    def __init__(self, ob, model, dist, torch, plan: ShardPlan, how_many: int, device):
        self.ob, self.model, self.dist, self.torch, self.plan, self.hm = ob, model, dist, torch, plan, how_many
        self.rec_bytes = ob.packed_record_bytes(how_many)
        self.side = torch.cuda.Stream(device=device)
        z = lambda n, dt: torch.zeros(n, dtype=dt, device=device)
        nb = self.NB
        self.d_rec = [z(self.rec_bytes, torch.uint8) for _ in range(nb)]
        self.g_rec = [z(plan.world * self.rec_bytes, torch.uint8) for _ in range(nb)]
        self.m_rows = [z(how_many, torch.int64) for _ in range(nb)]
        self.m_scores = [z(how_many, torch.float64) for _ in range(nb)]
        self.m_n = [z(1, torch.int32) for _ in range(nb)]
        self.ready = [torch.cuda.Event() for _ in range(nb)]
        self.consumed = [torch.cuda.Event() for _ in range(nb)]
        self.i = 0

    def submit(self, kind: int, d_query_ptr: int):
        torch, hm = self.torch, self.hm
        b = self.i % self.NB
        main = torch.cuda.current_stream()
        if self.i >= self.NB:
            main.wait_event(self.consumed[b])
        rec = self.d_rec[b]
        self.model.topn_device(kind, d_query_ptr, hm, rec[hm * 8:].data_ptr(), rec.data_ptr(),
                               rec[hm * 12:].data_ptr(), main.cuda_stream)
        self.ready[b].record(main)
        with torch.cuda.stream(self.side):
            self.side.wait_event(self.ready[b])
            self.dist.all_gather_into_tensor(self.g_rec[b], self.d_rec[b])
            self.consumed[b].record(self.side)
            self.ob.merge_topn_packed(self.g_rec[b].data_ptr(), self.plan.world, self.rec_bytes, hm,
                                      self.plan.rows_per_rank, self.m_rows[b].data_ptr(),
                                      self.m_scores[b].data_ptr(), self.m_n[b].data_ptr(), self.side.cuda_stream)
        self.i += 1

    def drain(self):
        self.torch.cuda.current_stream().wait_stream(self.side)

    def merged(self, i: int):
        b = i % self.NB
        return self.m_rows[b], self.m_scores[b], self.m_n[b]


class P2PShardedTopN:
    """Sharded queries whose exchange step is the library's peer-memory kernel (oryx_topn_sharded_device):
    the finishing kernel's record is stored straight into every peer's inbox over NVLink and merged
    there -- no NCCL call on the query path.  ``dist`` is used once, to hand the 64-byte inbox handles
    to the other processes."""

    def __init__(self, ob, model, dist, torch, plan: ShardPlan, how_many: int, device):
        self.ob, self.model, self.torch, self.plan, self.hm = ob, model, torch, plan, how_many
        index = device.index if hasattr(device, "index") else int(device)
        self.x = ob.Exchange(index, plan.world, dist.get_rank(), how_many)
        handles = [None] * plan.world
        dist.all_gather_object(handles, self.x.handle())
        failure = None
        try:
            self.x.connect(handles)
        except Exception as e:  # noqa: BLE001 -- every rank must reach the vote below
            failure = e
        # nobody publishes before every inbox is mapped, and either all ranks use this path or none does
        vote = torch.tensor([0 if failure else 1], dtype=torch.int32, device=device)
        dist.all_reduce(vote, op=dist.ReduceOp.MIN)
        if int(vote.item()) == 0:
            self.x.close()
            raise RuntimeError(f"peer mapping of the exchange inboxes failed on at least one rank ({failure})")
        z = lambda n, dt: torch.zeros(n, dtype=dt, device=device)
        self.m_rows, self.m_scores, self.m_n = z(how_many, torch.int64), z(how_many, torch.float64), z(1, torch.int32)

    def submit(self, kind: int, d_query_ptr: int, stream=None):
        st = self.torch.cuda.current_stream() if stream is None else stream
        self.x.topn(self.model, kind, d_query_ptr, self.hm, self.plan.rows_per_rank, self.m_rows.data_ptr(),
                    self.m_scores.data_ptr(), self.m_n.data_ptr(), st.cuda_stream)

    def drain(self):
        pass  # everything is ordered on the submitting stream

    def merged(self, i: int = 0):
        return self.m_rows, self.m_scores, self.m_n

    def check(self):
        """Synchronises and raises if the last exchange timed out waiting for a peer (count == -2)."""
        if int(self.m_n.item()) == -2:
            raise RuntimeError("a peer did not deliver its record within ORYX_EXCHANGE_TIMEOUT_MS")

    def close(self):
        self.x.close()


def allreduce_gram(dist, torch, gram_partial):
    """YtY = sum of per-shard Grams (ALSServingModel.java:362-368), fp64, in place."""
    dist.all_reduce(gram_partial, op=dist.ReduceOp.SUM)
    return gram_partial


def connect_peer_stores(ob, dist, torch, model, rank: int, world: int, peers=None):
    """Every rank (one process per GPU) describes its X and Y stores to the others and maps theirs
    (oryx_peer_stores_*): returns a ``PeerStores`` for ``Solver.fold_in_rows_peers``.  Collective: all ranks call
    it together -- also again, with ``peers`` = the previous object, after any rank's store was reallocated."""
    peers = peers if peers is not None else ob.PeerStores(model, world, rank)
    mine = torch.frombuffer(bytearray(peers.handle()), dtype=torch.uint8)
    cuda = dist.get_backend() == "nccl"
    if cuda:
        mine = mine.to(torch.device("cuda", torch.cuda.current_device()))
    gathered = [torch.empty_like(mine) for _ in range(world)]
    dist.all_gather(gathered, mine)
    peers.connect([bytes(g.cpu().numpy().tobytes()) for g in gathered])
    dist.barrier()  # nobody gathers before every rank has mapped what it needs
    return peers


def merge_batch_topn(shard_results, row_stride: int, how_many: int):
    """Merge of per-shard batched answers (the host half of a sharded oryx_topn_batch).

    ``shard_results[r] = (rows [nq x how_many] int32 local rows, scores [nq x how_many] float64,
    counts [nq] int32)`` as returned by ``Model.topn_batch`` on shard r.  Returns global rows (int64),
    scores and counts with the ordering of ``merge_topn_kernel``: score descending, global row ascending.
    The lists are tiny (world x how_many per query); the device does the scoring, not this.
    """
    import numpy as np
    nq = shard_results[0][0].shape[0]
    out_rows = np.zeros((nq, how_many), dtype=np.int64)
    out_scores = np.zeros((nq, how_many), dtype=np.float64)
    out_cnt = np.zeros(nq, dtype=np.int32)
    for q in range(nq):
        rows, scores = [], []
        for r, (sr, ss, sc) in enumerate(shard_results):
            k = int(sc[q])
            if k < 0:
                raise ValueError(f"shard {r} handed query {q} back unanswered")
            rows.append(sr[q, :k].astype(np.int64) + r * row_stride)
            scores.append(ss[q, :k])
        rows, scores = np.concatenate(rows), np.concatenate(scores)
        order = np.lexsort((rows, -scores))[:how_many]
        out_cnt[q] = order.size
        out_rows[q, :order.size] = rows[order]
        out_scores[q, :order.size] = scores[order]
    return out_rows, out_scores, out_cnt


def merge_batch_topn_device(ob, torch, shard_results, row_stride: int, how_many: int, device, gather=None):
    """Device half of a sharded oryx_topn_batch: every shard's (rows, scores, counts) for the page are stacked in
    device memory -- by ``gather`` (e.g. three NCCL all-gathers between processes) or, inside one process, by plain
    copies -- and merged by ONE launch of oryx_merge_topn_batch_device (one CTA per query).  Returns device tensors
    (rows int64 [nq x how_many], scores float64, counts int32)."""
    nq = shard_results[0][0].shape[0]
    world = len(shard_results)
    d_rows = torch.stack([torch.as_tensor(r[0], dtype=torch.int32) for r in shard_results]).to(device).contiguous()
    d_scores = torch.stack([torch.as_tensor(r[1], dtype=torch.float64) for r in shard_results]).to(device).contiguous()
    d_n = torch.stack([torch.as_tensor(r[2], dtype=torch.int32) for r in shard_results]).to(device).contiguous()
    if gather is not None:
        d_rows, d_scores, d_n = gather(d_rows, d_scores, d_n)
        world = d_n.shape[0]
    out_rows = torch.zeros((nq, how_many), dtype=torch.int64, device=device)
    out_scores = torch.zeros((nq, how_many), dtype=torch.float64, device=device)
    out_n = torch.zeros(nq, dtype=torch.int32, device=device)
    ob.merge_topn_batch_device(d_scores.data_ptr(), d_rows.data_ptr(), d_n.data_ptr(), world, nq, how_many, row_stride,
                               out_rows.data_ptr(), out_scores.data_ptr(), out_n.data_ptr(),
                               torch.cuda.current_stream(device).cuda_stream)
    return out_rows, out_scores, out_n
