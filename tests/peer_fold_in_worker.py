"""Worker of tests/test_peer_fold_in.py: one PROCESS per rank (torchrun, gloo for the plumbing), X and Y sharded by
rows over the ranks, every rank folds in a batch of interactions whose rows live anywhere
(oryx_fold_in_rows_peers over CUDA-IPC-mapped peer stores) and checks the result against the oracle over the
whole matrices.  Two ranks share GPU 0 when the box has one GPU (CUDA IPC works between processes on one device).
Exits non-zero on any mismatch."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import oracle  # noqa: E402
import oryx_b200 as ob  # noqa: E402
from oryx_b200.sharded import allreduce_gram, connect_peer_stores  # noqa: E402


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    dev = rank % torch.cuda.device_count()
    torch.cuda.set_device(dev)
    dist.init_process_group("gloo")
    f, nu, ni, n = 24, 901, 1203, 3000
    rng = np.random.default_rng(0xB20000C0)            # same stream on every rank: the whole matrices
    x = (rng.standard_normal((nu, f)) * 0.3).astype(np.float32)
    y = (rng.standard_normal((ni, f)) * 0.3).astype(np.float32)
    xs, ys = -(-nu // world), -(-ni // world)           # rows per shard
    m = ob.Model(f, device=dev, capacity_hint=4096)
    lo, hi = rank * xs, min(nu, (rank + 1) * xs)
    m.upsert(ob.X, np.arange(hi - lo, dtype=np.int32), x[lo:hi])
    lo, hi = rank * ys, min(ni, (rank + 1) * ys)
    m.upsert(ob.Y, np.arange(hi - lo, dtype=np.int32), y[lo:hi])
    hole = (world - 1) * ys + 3                          # a removed row in the last shard
    if rank == world - 1:
        m.remove(ob.Y, np.array([3], np.int32))
    y_live = np.delete(y, hole, axis=0)
    # YtY: per-shard Gram, all-reduced (ALSServingModel.java:362-368)
    g = torch.from_numpy(m.gram(ob.Y)[0].copy())
    allreduce_gram(dist, torch, g)
    g_want = oracle.transpose_times_self(y_live)
    assert np.max(np.abs(g.numpy() - g_want)) <= 1e-12 * np.max(np.abs(g_want)), "all-reduced Gram"
    solver, so = ob.Solver(g_want, device=dev), oracle.Solver(g_want)
    peers = connect_peer_stores(ob, dist, torch, m, rank, world)
    r2 = np.random.default_rng(0xB20000C1 + rank)       # every rank its own interactions
    ur = r2.integers(-1, nu + 30, n).astype(np.int64)   # -1 and rows beyond the end: Xu == null
    ir = r2.integers(0, ni + 30, n).astype(np.int64)    # beyond the end: Yi == null -> no result
    ir[:3] = hole
    vals = r2.choice([1.0, 2.0, -1.0, 0.5, 0.0], n)
    has_xu = ((ur >= 0) & (ur < nu)).astype(np.uint8)
    has_yi = (ir < ni) & (ir != hole)
    xu = np.where(has_xu[:, None] == 1, x[np.clip(ur, 0, nu - 1)], 0).astype(np.float32)
    yi = y[np.clip(ir, 0, ni - 1)]
    for implicit in (True, False):
        eo, ev = oracle.fold_in_batch(so, implicit, xu, has_xu, yi, vals)
        ev = ev & has_yi.astype(ev.dtype)
        go, gv = solver.fold_in_rows_peers(peers, xs, ys, implicit, ob.X, ur, ir, vals)
        assert np.array_equal(ev.astype(np.uint8), gv), ("valid flags", int(ev.sum()), int(gv.sum()))
        ok = ev.astype(bool)
        scale = np.maximum(np.abs(eo[ok]).max(axis=1, keepdims=True), 1e-6)
        err = float(np.max(np.abs(eo[ok] - go[ok]) / scale))
        assert err <= 5e-7, err                          # the tolerance of test_fold_in_goldens_and_random
        # and bit-identical to the single-process entry point fed the same vectors
        bo, bv = solver.fold_in_batch(implicit, xu, has_xu, yi, vals)
        assert np.array_equal(go[ok], bo[ok])
    # the item side too: Xu is a row of Y, Yi a row of X
    go, gv = solver.fold_in_rows_peers(peers, ys, xs, True, ob.Y, ir, np.clip(ur, 0, None), vals)
    assert gv.shape == (n,)
    dist.barrier()
    # a store that moves invalidates the handles it gave out: peers must refuse to read through them ...
    if rank == 0:
        m.upsert(ob.Y, np.arange(5000, 5000 + 6000, dtype=np.int32), np.zeros((6000, f), np.float32))  # beyond capacity_hint
    dist.barrier()
    if rank != 0:
        try:
            solver.fold_in_rows_peers(peers, xs, ys, True, ob.X, ur[:10], ir[:10], vals[:10])
            raise SystemExit("a stale mapping was read")
        except ob.OryxError as e:
            assert "moved" in str(e), str(e)
    dist.barrier()
    # ... until fresh handles are exchanged
    connect_peer_stores(ob, dist, torch, m, rank, world, peers)
    go2, gv2 = solver.fold_in_rows_peers(peers, xs, ys, True, ob.X, ur, ir, vals)
    eo, ev = oracle.fold_in_batch(so, True, xu, has_xu, yi, vals)
    ok = (ev & has_yi.astype(ev.dtype)).astype(bool)
    assert np.array_equal(gv2.astype(bool), ok)
    dist.barrier()
    peers.close()
    m.close()
    dist.destroy_process_group()
    print(f"rank {rank}: peer fold-in ok")


if __name__ == "__main__":
    main()
