"""The `extras` of bench.py's JSON line: BASELINE.json configs[2..4] measured in the driver's own run.

  N = 1   c3_batch   4096 queries x 10 M x 128 f, top-100, tcgen05 batch path (configs[2])
          c4         YtY over 10 M x 256 f + computeUpdatedXu for 1 M interactions (configs[3])
  N > 1   parity_vs_oracle_sharded   ~1 M rows per rank, LSH on, sharded query == oracle over the concatenation
          gram_allreduce             per-shard oryx_gram + NCCL fp64 all-reduce == one model's Gram
          c5_strong                  100 M x 256 f split over the ranks, LSH 0.3 / 32 cores, compacted (configs[4])

Every leg checks its results against the CPU oracle (or, where the oracle cannot finish in seconds, against
the single-GPU path that is oracle-checked elsewhere) and reports the check next to the number.  The oracle
is only ever the checker here, never inside a timed region.
"""
from __future__ import annotations

import json
import os
import statistics
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
FP64_NOMINAL_TFLOPS = 37.0   # B200 binary64 peak: nominal 37; measured here 36.6 (DFMA) / 37.1 (DMMA m8n8k4), tools/fp64_probe.cu


def peaks() -> dict:
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    out = {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0, "source": "fallback (B200_PROFILING.md)"}
    if os.path.exists(p):
        try:
            d = json.load(open(p))
            out.update({k: float(d[k]) for k in ("hbm_gbs", "bf16_tflops", "bf16_tflops_sustained") if k in d})
            out["source"] = "measured (MEASURED_PEAKS.json)"
        except (ValueError, OSError):
            pass
    return out


def gen_rows(seed: int, shard: int, rows: int, f: int) -> np.ndarray:
    """N(0,1) fp32 rows of shard `shard`: the same bits on whichever rank generates them."""
    out = np.empty((rows, f), dtype=np.float32)
    pieces = 8
    bounds = np.linspace(0, rows, pieces + 1).astype(np.int64)
    from concurrent.futures import ThreadPoolExecutor

    def fill(i):
        np.random.default_rng([seed, shard, i]).standard_normal(out=out[bounds[i]:bounds[i + 1]], dtype=np.float32)

    with ThreadPoolExecutor(pieces) as ex:
        list(ex.map(fill, range(pieces)))
    return out


# ------------------------------------------------------------------------------------------ N = 1

def c3_batch(ob, model, y_host, n_items: int, f: int, threads: int) -> dict:
    """configs[2]: 4096 queries x n_items x f, top-100, on the model bench.py already loaded."""
    nq, hm = 4096, 100
    pk = peaks()
    q = np.random.default_rng(0xB2000301).standard_normal((nq, f), dtype=np.float32)
    model.topn_batch(ob.DOT, q[:256], hm)          # warm: bf16 shadow of Y, scratch, kernel attributes
    model.topn_batch(ob.DOT, q, hm)                # grow the scratch to the full page once
    runs0 = model.debug_batch_timings()["runs"]
    walls, phases = [], []
    for _ in range(3):
        t0 = time.perf_counter()
        rows, scores, cnt = model.topn_batch(ob.DOT, q, hm)
        walls.append(time.perf_counter() - t0)
        phases.append(model.debug_batch_timings())
    on_tensor_cores = model.debug_batch_timings()["runs"] == runs0 + 3
    best = min(range(3), key=lambda i: phases[i]["score_ms"])
    ph = phases[best]
    device_ms = ph["convert_ms"] + ph["score_ms"] + ph["threshold_ms"] + ph["refine_ms"]
    flops = 2.0 * nq * n_items * f                 # algorithmic, not the padded count
    parity = None
    if y_host is not None:
        import oracle
        parity = True
        for i in (0, 1, 511, 1024, 2047, 3000, 4000, 4095):
            er, es = oracle.topn_mt(oracle.DOT, q[i], y_host, hm, threads)
            k = int(cnt[i])
            parity = parity and k == len(er) and np.array_equal(rows[i, :k], er) and np.array_equal(scores[i, :k], es)
    handed_back = int((cnt < 0).sum())
    return {
        "workload": f"{nq} queries x {n_items} items x {f} f, top-{hm} (BASELINE.json configs[2]), host buffers in and out",
        "answered_by_tcgen05_path": bool(on_tensor_cores), "queries_handed_back_to_the_scan": handed_back,
        "phase_ms": {k: ph[k] for k in ("convert_ms", "score_ms", "threshold_ms", "refine_ms")},
        "device_ms_per_batch": device_ms, "wall_ms_per_batch_host_buffers": 1000.0 * min(walls),
        "queries_per_sec_device": nq / (device_ms / 1000.0), "queries_per_sec_e2e": nq / min(walls),
        "candidate_subblocks_per_query": {"mean": ph["cand_mean"], "max": ph["cand_max"]},
        "roofline_score_kernel": {"bound": "tensor", "achieved": flops / (ph["score_ms"] / 1000.0) / 1e12,
                                  "peak": pk["bf16_tflops"], "unit": "TFLOP/s",
                                  "frac": flops / (ph["score_ms"] / 1000.0) / 1e12 / pk["bf16_tflops"],
                                  "peak_kind": "bf16 burst, " + pk["source"]},
        "roofline_whole_batch": {"bound": "tensor", "achieved": flops / (device_ms / 1000.0) / 1e12,
                                 "peak": pk["bf16_tflops_sustained"], "unit": "TFLOP/s",
                                 "frac": flops / (device_ms / 1000.0) / 1e12 / pk["bf16_tflops_sustained"],
                                 "peak_kind": "bf16 sustained, " + pk["source"]},
        "algorithmic_flops_per_batch": flops,
        "parity_vs_oracle": parity, "parity_sample": "8 of the 4096 queries against oracle.topn_mt on the full matrix: rows and binary64 scores bit-equal",
    }


def c4_gram_foldin(ob, torch, dev, threads: int, n_items: int = 10_000_000, f: int = 256, n_updates: int = 1_000_000) -> dict:
    """configs[3]: YtY over n_items x 256 f and computeUpdatedXu for 1 M interactions."""
    import oracle
    pk = peaks()
    sub = 1_000_000                                   # rows checked against the oracle
    gen = torch.Generator(device=dev)
    gen.manual_seed(0xB2000401)
    n_users = 1_000_000
    model = ob.Model(f, implicit=True, device=dev.index, capacity_hint=n_items)
    y_sub = np.empty((min(sub, n_items), f), dtype=np.float32)
    for off in range(0, n_items, 1_000_000):
        c = min(1_000_000, n_items - off)
        block = torch.randn((c, f), generator=gen, device=dev, dtype=torch.float32) * 0.3
        torch.cuda.synchronize()
        model.upsert_device(ob.Y, off, block.data_ptr(), c)
        if off < sub:
            y_sub[off:off + c] = block[: len(y_sub) - off].cpu().numpy()
        del block
    x_sub_rows = 4096
    x_block = torch.randn((n_users, f), generator=gen, device=dev, dtype=torch.float32) * 0.3
    torch.cuda.synchronize()
    model.upsert_device(ob.X, 0, x_block.data_ptr(), n_users)
    x_sub = x_block[:x_sub_rows].cpu().numpy()
    del x_block
    # ---- Gram
    g, used = model.gram(ob.Y)                        # warm
    ms = []
    for _ in range(5):
        g, used = model.gram(ob.Y)
        ms.append(model.debug_gram_ms())
    gram_ms = statistics.median(ms)
    alg_bytes = n_items * f * 4
    # parity: a model holding the first `sub` rows (same kernel path: it is above the large-store threshold)
    with ob.Model(f, implicit=True, device=dev.index, capacity_hint=len(y_sub)) as small:
        small.upsert(ob.Y, np.arange(len(y_sub), dtype=np.int32), y_sub)
        g_sub, _ = small.gram(ob.Y)
    t0 = time.perf_counter()
    e_sub = oracle.transpose_times_self_mt(y_sub, threads)
    oracle_s = time.perf_counter() - t0
    rel_err = float(np.max(np.abs(g_sub - e_sub)) / np.max(np.abs(e_sub)))
    # full size: the Gram is additive over row ranges, so the subset's exact Gram bounds the full one's diagonal share
    # ---- fold-in: new user vectors for n_updates (user, item, strength) interactions (ALSSpeedModelManager.java:182-204)
    solver = ob.Solver(g, device=dev.index)
    rng = np.random.default_rng(0xB2000402)
    ur = rng.integers(0, n_users + n_users // 50, n_updates).astype(np.int32)     # ~2 % unknown users (Xu == null)
    ir = rng.integers(0, n_items, n_updates).astype(np.int32)
    vals = rng.choice([1.0, 1.0, 1.0, 2.0, 0.5, -1.0], n_updates)
    ur[:x_sub_rows] = np.arange(x_sub_rows)                                      # rows the parity check has host copies of
    ir[:x_sub_rows] = rng.integers(0, len(y_sub), x_sub_rows)
    solver.fold_in_rows(model, True, ob.X, ur[:200_000], ir[:200_000], vals[:200_000])   # warm: scratch, pinned staging
    t0 = time.perf_counter()
    out, valid = solver.fold_in_rows(model, True, ob.X, ur, ir, vals)
    fold_wall = time.perf_counter() - t0
    ft = solver.debug_fold_in_timings()
    # the same call into a page-locked result array (what a JNI direct buffer registered once, or pinned memory of any
    # host runtime, gives): the rows are copied straight into it
    pinned_wall = pinned_same = None
    try:
        pin = torch.empty((n_updates, f), dtype=torch.float32, pin_memory=True)
        out_p, valid_p = solver.fold_in_rows(model, True, ob.X, ur, ir, vals, out=pin.numpy())
        t0 = time.perf_counter()
        out_p, valid_p = solver.fold_in_rows(model, True, ob.X, ur, ir, vals, out=pin.numpy())
        pinned_wall = time.perf_counter() - t0
        vm = valid == 1
        pinned_same = bool(np.array_equal(valid_p, valid) and np.array_equal(out_p[vm], out[vm]))
        del pin, out_p
    except Exception as e:  # noqa: BLE001
        pinned_same = repr(e)
    e_solver = oracle.Solver(g)
    npar = 512
    eo, ev = oracle.fold_in_batch(e_solver, True, x_sub[:npar], np.ones(npar, np.uint8), y_sub[ir[:npar]], vals[:npar])
    fold_ok = bool(np.array_equal(valid[:npar], ev))
    sel = ev == 1
    fold_rel = float(np.max(np.abs(out[:npar][sel] - eo[sel])) / np.max(np.abs(eo[sel]))) if sel.any() else 0.0
    fold_bytes = n_updates * 4 * f * 4               # read Xu, Yi, write 2 vectors' worth per interaction
    fold_flops = 2.0 * n_updates * f * f             # one binary64 F x F product per interaction (this call: one side)
    k_ms = ft["kernel_ms"]
    solver.close()
    model.close()
    return {
        "workload": f"YtY over {n_items} x {f} f + computeUpdatedXu for {n_updates} interactions (BASELINE.json configs[3])",
        "gram": {
            "device_ms": gram_ms, "runs_ms": ms, "rows_used": int(used),
            "roofline": {"bound": "hbm", "achieved": alg_bytes / (gram_ms / 1000.0) / 1e9, "peak": pk["hbm_gbs"], "unit": "GB/s",
                         "frac": alg_bytes / (gram_ms / 1000.0) / 1e9 / pk["hbm_gbs"], "algorithmic_bytes": alg_bytes,
                         "peak_kind": pk["source"]},
            "flops_rate_tflops": 2.0 * n_items * f * f / (gram_ms / 1000.0) / 1e12,
            "max_rel_err_vs_oracle": rel_err, "parity_rows": len(y_sub), "oracle_seconds": round(oracle_s, 1),
            "parity_note": "same kernel on the first rows of the same matrix vs oracle.transpose_times_self_mt "
                           "(exact float-multiply / double-add per partition, partials summed); error relative to the largest entry",
            "symmetric": bool(np.array_equal(g, g.T)),
        },
        "fold_in": {
            "fold_in_device_ms": k_ms, "chunks": ft["chunks"], "wall_ms_host_buffers_out": 1000.0 * fold_wall,
            "wall_ms_page_locked_out": 1000.0 * pinned_wall if pinned_wall is not None else None,
            "page_locked_result_identical": pinned_same,
            "interactions_per_sec_device": n_updates / (k_ms / 1000.0) if k_ms > 0 else None,
            "roofline_hbm": {"achieved": fold_bytes / (k_ms / 1000.0) / 1e9 if k_ms > 0 else None, "peak": pk["hbm_gbs"], "unit": "GB/s",
                             "frac": fold_bytes / (k_ms / 1000.0) / 1e9 / pk["hbm_gbs"] if k_ms > 0 else None,
                             "algorithmic_bytes": fold_bytes},
            "roofline_fp64": {"achieved": fold_flops / (k_ms / 1000.0) / 1e12 if k_ms > 0 else None, "peak": FP64_NOMINAL_TFLOPS,
                              "unit": "TFLOP/s", "frac": fold_flops / (k_ms / 1000.0) / 1e12 / FP64_NOMINAL_TFLOPS if k_ms > 0 else None,
                              "peak_kind": "binary64 rate: nominal 37, measured 36.6 DFMA / 37.1 DMMA on this pool (profiles/r02_fp64_probe.txt)"},
            "valid_flags_match_oracle": fold_ok, "max_rel_err_vs_oracle": fold_rel, "parity_rows": npar,
            "valid_fraction": float(valid.mean()),
        },
    }


# ------------------------------------------------------------------------------------------ N > 1

def sharded_parity_and_gram(ob, torch, dist, dev, rank: int, world: int, threads: int) -> dict:
    """~1 M rows per rank, LSH on (sample-rate 0.3 / 32 cores): the sharded query (peer-memory exchange, NCCL
    fallback) against oracle.topn over the concatenation, before and after compaction; then per-shard Grams
    all-reduced over NCCL against one model holding every row."""
    from oryx_b200.sharded import P2PShardedTopN, PipelinedShardedTopN, ShardPlan, allreduce_gram
    import oracle
    f, hm = 128, 10
    rows_per_rank = 1_000_000 if world <= 4 else 500_000
    H, bits = ob.lsh_plan(0.3, 32)
    hv = np.random.default_rng(0xB2000501).standard_normal((H, f), dtype=np.float32)
    plan = ShardPlan(world, rows_per_rank)
    y_local = gen_rows(0xB2000502, rank, rows_per_rank, f)
    model = ob.Model(f, hash_vectors=hv, max_bits_differing=bits, device=dev.index, capacity_hint=rows_per_rank, reserved_sms=1)
    model.upsert(ob.Y, np.arange(rows_per_rank, dtype=np.int32), y_local)
    try:
        pipe = P2PShardedTopN(ob, model, dist, torch, plan, hm, dev)
        exchange = "peer-memory kernel"
    except RuntimeError:
        pipe = PipelinedShardedTopN(ob, model, dist, torch, plan, hm, dev)
        exchange = "NCCL all-gather + merge"
    nq = 8
    q_host = np.random.default_rng(0xB2000503).standard_normal((nq, f), dtype=np.float32)
    d_q = torch.from_numpy(q_host).to(dev)

    def run_queries():
        got = []
        for i in range(nq):
            pipe.submit(ob.DOT if i % 2 == 0 else ob.COSINE, d_q[i].data_ptr())
            pipe.drain()
            torch.cuda.synchronize()
            mr, ms_, mn = pipe.merged(i)
            k = int(mn.item())
            got.append((mr[:k].cpu().numpy().copy(), ms_[:k].cpu().numpy().copy()))
        return got

    got_before = run_queries()
    model.compact(ob.Y)
    got_after = run_queries()
    ok = None
    if rank == 0:
        y_all = np.concatenate([gen_rows(0xB2000502, r, rows_per_rank, f) for r in range(world)])
        part = oracle.lsh_index_rows(hv, y_all)
        ok = True
        for i in range(nq):
            kind = oracle.DOT if i % 2 == 0 else oracle.COSINE
            cand = oracle.lsh_candidates(H, bits, oracle.lsh_index(hv, q_host[i]))
            er, es = oracle.topn(kind, q_host[i], y_all, hm, part=part, cand=cand)
            for got in (got_before, got_after):
                gr, gs = got[i]
                ok = ok and len(gr) == len(er) and np.array_equal(gr, er) and np.array_equal(gs, es)
    # ---- Gram: per-shard partial, NCCL all-reduce (ALSServingModel.java:362-368), vs one model with all rows
    g_local, _ = model.gram(ob.Y)
    d_g = torch.from_numpy(g_local).to(dev)
    torch.cuda.synchronize()
    dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    allreduce_gram(dist, torch, d_g.clone())          # warm the communicator for this size
    torch.cuda.synchronize()
    e0.record()
    allreduce_gram(dist, torch, d_g)
    e1.record()
    torch.cuda.synchronize()
    allreduce_ms = e0.elapsed_time(e1)
    gram_ms = model.debug_gram_ms()
    # the same sum as ONE kernel over peer memory (oryx_gram_allreduce_device): no NCCL call; bits equal on every rank
    peer = None
    if hasattr(pipe, "x"):
        try:
            d_p = torch.from_numpy(g_local).to(dev)
            d_st = torch.zeros(1, dtype=torch.int32, device=dev)
            st_ptr = torch.cuda.current_stream(dev).cuda_stream
            pipe.x.gram_allreduce_device(d_p.clone().data_ptr(), f, d_st.data_ptr(), st_ptr)   # warm-up
            torch.cuda.synchronize()
            dist.barrier()
            e0.record()
            pipe.x.gram_allreduce_device(d_p.data_ptr(), f, d_st.data_ptr(), st_ptr)
            e1.record()
            torch.cuda.synchronize()
            t = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            same = torch.tensor([d_p.cpu().numpy().tobytes() == d_g.cpu().numpy().tobytes()], dtype=torch.int32, device=dev)
            ref = d_p.clone()
            dist.broadcast(ref, 0)
            same_everywhere = torch.tensor([int(torch.equal(ref, d_p))], dtype=torch.int32, device=dev)
            dist.all_reduce(same_everywhere, op=dist.ReduceOp.MIN)
            peer = {"peer_memory_kernel_ms_max_over_ranks": float(t.item()), "status": int(d_st.item()),
                    "max_rel_diff_vs_nccl_sum": float((d_p - d_g).abs().max().item() / d_g.abs().max().item()),
                    "same_bits_as_nccl_sum": bool(same.item()),
                    "note": "ranks are added in rank order; NCCL's own order differs for N > 2, so its sum may differ in the last bit",
                    "bits_equal_on_every_rank": bool(same_everywhere.item())}
        except Exception as e:  # noqa: BLE001
            peer = {"error": repr(e)}
    gram_rel = gram_oracle_rel = None
    if rank == 0:
        with ob.Model(f, device=dev.index, capacity_hint=len(y_all)) as whole:
            for off in range(0, len(y_all), 1_000_000):
                whole.upsert(ob.Y, np.arange(off, min(off + 1_000_000, len(y_all)), dtype=np.int32), y_all[off:off + 1_000_000])
            g_whole, _ = whole.gram(ob.Y)
        g_sum = d_g.cpu().numpy()
        gram_rel = float(np.max(np.abs(g_sum - g_whole)) / np.max(np.abs(g_whole)))
        if len(y_all) <= 2_000_000:
            e = oracle.transpose_times_self_mt(y_all, threads)
            gram_oracle_rel = float(np.max(np.abs(g_sum - e)) / np.max(np.abs(e)))
    # ---- fold-in with X and Y sharded over the ranks (separate processes): every rank gathers the rows its
    # interactions name from their owners over NVLink (oryx_peer_stores_* / oryx_fold_in_rows_peers); the solver
    # is built from the all-reduced Gram (ALSSpeedModelManager.java:173-204 with ALSServingModel.java:362-368)
    fold = None
    try:
        from oryx_b200.sharded import connect_peer_stores
        x_per_rank, n_int = 200_000, 200_000
        model.upsert(ob.X, np.arange(x_per_rank, dtype=np.int32), gen_rows(0xB2000504, rank, x_per_rank, f))
        g_all = d_g.cpu().numpy()
        solver = ob.Solver(g_all, device=dev.index)
        peers = connect_peer_stores(ob, dist, torch, model, rank, world)
        r3 = np.random.default_rng(0xB2000505 + rank)
        ur = r3.integers(0, x_per_rank * world, n_int).astype(np.int64)
        ir = r3.integers(0, rows_per_rank * world, n_int).astype(np.int64)
        vals = r3.choice([1.0, 2.0, 0.5, -1.0], n_int)
        solver.fold_in_rows_peers(peers, x_per_rank, rows_per_rank, True, ob.X, ur, ir, vals)  # warm-up: sizes the pinned staging
        dist.barrier()
        t0 = time.perf_counter()
        out, valid = solver.fold_in_rows_peers(peers, x_per_rank, rows_per_rank, True, ob.X, ur, ir, vals)
        wall_ms = (time.perf_counter() - t0) * 1e3
        tm = solver.debug_fold_in_timings()
        t = torch.tensor([wall_ms, tm["kernel_ms"]], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        remote = float(np.mean(ur // x_per_rank != rank)), float(np.mean(ir // rows_per_rank != rank))
        fold = {"interactions_per_rank": n_int, "wall_ms_max_over_ranks": float(t[0].item()),
                "fold_in_kernel_ms_max_over_ranks": float(t[1].item()),
                "fraction_of_rows_on_other_ranks": {"xu": remote[0], "yi": remote[1]},
                "interactions_per_sec_all_ranks": n_int * world / (float(t[0].item()) / 1e3)}
        if rank == 0:
            ns = 512
            shards = {r: gen_rows(0xB2000504, r, x_per_rank, f) for r in set((ur[:ns] // x_per_rank).tolist())}
            x_rows = np.stack([shards[int(u // x_per_rank)][int(u % x_per_rank)] for u in ur[:ns]])
            y_rows = y_all[ir[:ns]]
            so = oracle.Solver(g_all)
            eo, ev = oracle.fold_in_batch(so, True, x_rows, np.ones(ns, np.uint8), y_rows, vals[:ns])
            okv = ev.astype(bool)
            scale = np.maximum(np.abs(eo[okv]).max(axis=1, keepdims=True), 1e-6)
            fold["valid_flags_match_oracle"] = bool(np.array_equal(ev, valid[:ns]))
            fold["max_rel_err_vs_oracle"] = float(np.max(np.abs(eo[okv] - out[:ns][okv]) / scale)) if okv.any() else 0.0
            fold["parity_rows"] = ns
        dist.barrier()
        peers.close()
    except Exception as e:  # noqa: BLE001 -- reported, never silently dropped
        fold = {"error": repr(e)}
    if hasattr(pipe, "close"):
        pipe.close()
    model.close()
    return {"parity_vs_oracle_sharded": ok, "fold_in_peers": fold,
            "parity_note": f"{rows_per_rank} rows x {f} f per rank, LSH sample-rate 0.3/32 cores ({H} hashes, {bits} bits), {nq} queries "
                           f"(dot and cosine), before and after oryx_items_compact, merged global rows and binary64 scores "
                           f"bit-equal to oracle.topn over the concatenation; exchange: {exchange}",
            "gram_allreduce": {"per_shard_gram_ms": gram_ms, "nccl_allreduce_ms": allreduce_ms, "peer_memory": peer, "bytes": f * f * 8,
                               "max_rel_diff_vs_single_model_gram": gram_rel, "max_rel_err_vs_oracle": gram_oracle_rel,
                               "rows_total": rows_per_rank * world}}


def c5_strong(ob, torch, dist, dev, rank: int, world: int, total: int = 100_000_000, f: int = 256, steps: int = 100) -> dict:
    """configs[4]: `total` items x 256 f split by rows over the ranks, LSH candidate filtering on, per-GPU top-10,
    peer-memory exchange; strong scaling (the total is fixed)."""
    from oryx_b200.sharded import P2PShardedTopN, PipelinedShardedTopN, ShardPlan
    plan = ShardPlan.for_rows(total, world)
    lo, hi = plan.local_range(rank, total)
    n = hi - lo
    H, bits = ob.lsh_plan(0.3, 32)
    hv = np.random.default_rng(4).standard_normal((H, f), dtype=np.float32)
    model = ob.Model(f, hash_vectors=hv, max_bits_differing=bits, device=dev.index, capacity_hint=n, reserved_sms=1)
    gen = torch.Generator(device=dev)
    gen.manual_seed(100 + rank)
    t0 = time.perf_counter()
    for off in range(0, n, 1_000_000):
        c = min(1_000_000, n - off)
        blk = torch.randn((c, f), generator=gen, device=dev)
        torch.cuda.synchronize()
        model.upsert_device(ob.Y, off, blk.data_ptr(), c)
        del blk
    t_load = time.perf_counter() - t0
    t0 = time.perf_counter()
    model.compact(ob.Y)
    t_compact = time.perf_counter() - t0
    hm, nq = 10, 32
    d_q = torch.from_numpy(np.random.default_rng(9).standard_normal((nq, f), dtype=np.float32)).to(dev)
    try:
        pipe = P2PShardedTopN(ob, model, dist, torch, plan, hm, dev)
        exchange = "peer-memory kernel"
    except RuntimeError:
        pipe = PipelinedShardedTopN(ob, model, dist, torch, plan, hm, dev)
        exchange = "NCCL all-gather + merge"
    stream = torch.cuda.current_stream(dev)
    for i in range(5):
        pipe.submit(ob.DOT, d_q[i % nq].data_ptr())
    pipe.drain()
    torch.cuda.synchronize()
    dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for i in range(steps):
        pipe.submit(ob.DOT, d_q[i % nq].data_ptr())
    pipe.drain()
    e1.record(stream)
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item()) / steps
    cand_frac = sum(1 for p in range(1 << H) if bin(p).count("1") <= bits) / float(1 << H)
    # ---- the same sharded, LSH-filtered model answering a PAGE of queries: every rank runs the tensor-core batch
    # path on its shard (candidate filtering in the score epilogue and the refine step), the per-shard lists are
    # all-gathered and merged by one launch (oryx_merge_topn_batch_device)
    pages = None
    try:
        from oryx_b200.sharded import merge_batch_topn_device
        n_page = 1024
        q_page = np.random.default_rng(10).standard_normal((n_page, f), dtype=np.float32)
        q_page[:nq] = d_q.cpu().numpy()           # the first queries are the ones the single-query path answers

        def gather(d_rows, d_scores, d_n):
            outs = []
            for t_ in (d_rows, d_scores, d_n):
                parts = [torch.empty_like(t_[0]) for _ in range(world)]
                dist.all_gather(parts, t_[0].contiguous())
                outs.append(torch.stack(parts))
            return outs

        def one_page():
            res = model.topn_batch(ob.DOT, q_page, hm)
            return merge_batch_topn_device(ob, torch, [res], plan.rows_per_rank, hm, dev, gather=gather)

        one_page()                                 # builds the bf16 shadow of the shard, sizes the scratch
        torch.cuda.synchronize()
        dist.barrier()
        reps, walls = 3, []
        for _ in range(reps):
            t0 = time.perf_counter()
            m_rows, m_scores, m_n = one_page()
            torch.cuda.synchronize()
            walls.append((time.perf_counter() - t0) * 1e3)
        wall = sum(walls) / reps / 1e3
        tw = torch.tensor([wall], dtype=torch.float64, device=dev)
        dist.all_reduce(tw, op=dist.ReduceOp.MAX)
        bt = model.debug_batch_timings()
        # the page must give the single-query path's answers (both are exact)
        same = True
        for i in range(min(nq, 8)):
            pipe.submit(ob.DOT, d_q[i].data_ptr())
            pipe.drain()
            torch.cuda.synchronize()
            sr, ss, sn = pipe.merged(i)
            k = int(sn.item())
            same = same and int(m_n[i].item()) == k and torch.equal(m_rows[i, :k], sr[:k]) and torch.equal(m_scores[i, :k], ss[:k])
        pages = {"queries_per_page": n_page, "wall_ms_per_page_max_over_ranks": float(tw.item()) * 1e3,
                 "queries_per_sec": n_page / float(tw.item()),
                 "device_phase_ms_last_page_this_rank": {k: bt[k] for k in ("convert_ms", "score_ms", "threshold_ms", "refine_ms")},
                 "candidate_subblocks_mean": bt["cand_mean"], "wall_ms_each_page_this_rank": walls,
                 "answered_by_tcgen05_path": bool(bt["runs"] >= 1),
                 "equal_to_single_query_sharded_path": bool(same),
                 "note": "host buffers in and out per rank; the tensor cores score every tile, LSH filtering is applied to "
                         "the maxima and in the refine step"}
    except Exception as e:  # noqa: BLE001
        pages = {"error": repr(e)}
    if hasattr(pipe, "close"):
        pipe.close()
    model.close()
    return {"workload": f"{total} items x {f} f over {world} GPUs, LSH sample-rate 0.3/32 cores ({H} hashes, {bits} bits: "
                        f"{cand_frac:.3f} of the partitions are candidates), top-{hm}, compacted (BASELINE.json configs[4])",
            "exchange": exchange, "items_per_gpu": n, "ms_per_query": ms, "queries_per_sec": 1000.0 / ms,
            "nominal_item_scores_per_sec": total * 1000.0 / ms,
            "GBps_per_gpu_on_candidate_rows": cand_frac * n * f * 4 / (ms * 1e-3) / 1e9,
            "pages": pages,
            "load_s_per_gpu": round(t_load, 1), "compact_s": round(t_compact, 1), "steps": steps,
            "timing": "CUDA events on the submitting stream, max over ranks"}
