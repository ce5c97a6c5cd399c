"""oryx_topn_device_static -- the allocation-free form of the single-query top-N scan.

Some hosts cannot let the library allocate: a serving thread that must not risk cudaMalloc stalling
the device while other queries run, or a process whose GPU memory already belongs to another
allocator.  For them the library offers one entry point whose scratch is a fixed array inside the
loaded module.  This file checks, through the C ABI, that it returns exactly what the allocating entry
point (oryx_topn_device) returns -- row ids equal, scores bit-identical doubles, the tolerance every
top-N parity test here uses -- on two calls in a row (the array is shared and keeps the previous
call's contents, so the second call proves stale contents do not leak), with and without LSH
filtering, and that a request beyond the array's capacity is refused with a clear message.

Collected last on purpose (zz): a run that stops at the first failure has then already run
everything else.  It passes both alone in a fresh process and after the other GPU tests.
"""
import numpy as np
import pytest

import oryx_b200 as ob
from oryx_b200 import native
from als_harness import random_hash_vectors, random_model

pytestmark = pytest.mark.gpu


def _run(fn, kind, d_q, how_many, torch):
    dev = d_q.device
    d_rows = torch.full((how_many,), -7, dtype=torch.int32, device=dev)
    d_scores = torch.full((how_many,), float("nan"), dtype=torch.float64, device=dev)
    d_n = torch.full((1,), -1, dtype=torch.int32, device=dev)
    stream = torch.cuda.current_stream(dev)
    fn(kind, d_q.data_ptr(), how_many, d_rows.data_ptr(), d_scores.data_ptr(), d_n.data_ptr(), stream.cuda_stream)
    stream.synchronize()
    n = int(d_n.item())
    return d_rows[:n].cpu().numpy(), d_scores[:n].cpu().numpy()


@pytest.mark.parametrize("lsh", [False, True])
def test_static_scratch_matches_allocating_entry_point(lsh):
    import torch
    n, f = 60000, 48
    y = random_model(n, f, seed=91)
    hv, bits = None, 0
    if lsh:
        h, bits = ob.lsh_plan(0.3, 16)
        hv = random_hash_vectors(h, f, seed=92)
    rng = np.random.default_rng(93)
    queries = rng.standard_normal((3, f), dtype=np.float32)
    with ob.Model(f, hash_vectors=hv, max_bits_differing=bits) as m:
        m.upsert(ob.Y, np.arange(n, dtype=np.int32), y)
        d_qs = [torch.from_numpy(q).cuda() for q in queries]
        # two calls in a row (then a third, repeating the first query) against the same shared array
        for kind, how_many in ((ob.DOT, 10), (ob.COSINE, native.STATIC_MAX_HOW_MANY)):
            for d_q in (d_qs[0], d_qs[1], d_qs[0]):
                exp_rows, exp_scores = _run(m.topn_device, kind, d_q, how_many, torch)
                got_rows, got_scores = _run(m.topn_device_static, kind, d_q, how_many, torch)
                assert len(exp_rows) == how_many
                assert np.array_equal(got_rows, exp_rows)
                assert np.array_equal(got_scores, exp_scores)  # bit-identical doubles
        # and against the host-buffer entry point, for good measure
        rows, scores = m.topn(ob.DOT, queries[2], 10)
        got_rows, got_scores = _run(m.topn_device_static, ob.DOT, d_qs[2], 10, torch)
        assert np.array_equal(got_rows, rows) and np.array_equal(got_scores, scores)
        # ... and straight against the CPU oracle (rows and binary64 scores bit-equal)
        import oracle
        part = oracle.lsh_index_rows(hv, y) if lsh else None
        cand = oracle.lsh_candidates(hv.shape[0], bits, oracle.lsh_index(hv, queries[2])) if lsh else None
        e_rows, e_scores = oracle.topn(oracle.DOT, queries[2], y, 10, part=part, cand=cand)
        assert np.array_equal(got_rows, e_rows) and np.array_equal(got_scores, e_scores)
        # The point of the entry point: it allocates nothing.  Outputs are created up front, then the
        # device's free memory must read the same before and after a series of calls (any cudaMalloc,
        # pooled or not, would lower it; nothing else runs in this process meanwhile).
        dev = d_qs[0].device
        d_rows = torch.zeros(10, dtype=torch.int32, device=dev)
        d_scores = torch.zeros(10, dtype=torch.float64, device=dev)
        d_n = torch.zeros(1, dtype=torch.int32, device=dev)
        stream = torch.cuda.current_stream(dev)
        torch.cuda.synchronize()
        free_before, _ = torch.cuda.mem_get_info(dev)
        for d_q in d_qs * 3:
            m.topn_device_static(ob.DOT, d_q.data_ptr(), 10, d_rows.data_ptr(), d_scores.data_ptr(), d_n.data_ptr(),
                                 stream.cuda_stream)
        stream.synchronize()
        free_after, _ = torch.cuda.mem_get_info(dev)
        assert free_after == free_before, (free_before, free_after)


def test_static_scratch_rejects_oversized_requests():
    import torch
    with ob.Model(8) as m:
        m.upsert(ob.Y, np.arange(100, dtype=np.int32), random_model(100, 8, seed=1))
        d_q = torch.ones(8, dtype=torch.float32, device="cuda")
        with pytest.raises(ob.OryxError) as ei:
            _run(m.topn_device_static, ob.DOT, d_q, native.STATIC_MAX_HOW_MANY + 1, torch)
        assert ei.value.status == native.E_INVALID
        assert "ORYX_STATIC_MAX_HOW_MANY" in str(ei.value)
        # a small model (fewer rows than how_many) still works right after the refusal
        rows, scores = _run(m.topn_device_static, ob.DOT, d_q, native.STATIC_MAX_HOW_MANY, torch)
        exp_rows, exp_scores = m.topn(ob.DOT, np.ones(8, np.float32), native.STATIC_MAX_HOW_MANY)
        assert np.array_equal(rows, exp_rows) and np.array_equal(scores, exp_scores)
