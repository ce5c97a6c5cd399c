"""C3 measurement: 4096 queries x 10M items x 128 f, top-100, through oryx_topn_batch (host buffers)."""
import ctypes as C
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import oryx_b200 as ob
from oryx_b200 import native

n = int(sys.argv[1]) if len(sys.argv) > 1 else 10_000_000
nq = int(sys.argv[2]) if len(sys.argv) > 2 else 4096
how_many = int(sys.argv[3]) if len(sys.argv) > 3 else 100
F = int(sys.argv[4]) if len(sys.argv) > 4 else 128
KIND = ob.COSINE if (len(sys.argv) > 5 and sys.argv[5] == "cosine") else ob.DOT   # /similarity scores by cosine
dev = torch.device("cuda", 0)
m = ob.Model(F, capacity_hint=n)
g = torch.Generator(device=dev); g.manual_seed(0x0B2000001)
for off in range(0, n, 1_000_000):
    c = min(1_000_000, n - off)
    blk = torch.randn((c, F), generator=g, device=dev); torch.cuda.synchronize()
    m.upsert_device(ob.Y, off, blk.data_ptr(), c)
    del blk
q = np.random.default_rng(3).standard_normal((nq, F), dtype=np.float32)
L = native.lib()
L.oryx_debug_batch_timings.restype = C.c_int32
L.oryx_debug_batch_timings.argtypes = [C.c_void_p, C.POINTER(C.c_double)]
res = []
if nq < 32:  # small batches ride on a bf16 shadow that a larger batch (or an earlier call) has built
    m.topn_batch(KIND, np.random.default_rng(4).standard_normal((32, F), dtype=np.float32), how_many)
for rep in range(3):
    t0 = time.perf_counter()
    rows, scores, cnt = m.topn_batch(KIND, q, how_many)
    wall = time.perf_counter() - t0
    t = (C.c_double * 9)()
    L.oryx_debug_batch_timings(m._h, t)
    conv, score, thr, ref, flops, cand_mean, cand_max = [float(x) for x in t][:7]
    res.append(dict(wall_ms=wall * 1e3, convert_ms=conv, score_ms=score, threshold_ms=thr, refine_ms=ref,
                    score_tflops=flops / (score * 1e-3) / 1e12, handed_back=int((cnt < 0).sum()),
                    candidate_subblocks_mean=cand_mean, candidate_subblocks_max=cand_max))
    print(json.dumps(res[-1]))
# spot-check against the single-query path
ok = True
for i in sorted({0, min(1, nq - 1), nq // 2, nq - 1}):
    r1, s1 = m.topn(KIND, q[i], how_many)
    ok = ok and np.array_equal(r1, rows[i, :cnt[i]]) and np.array_equal(s1, scores[i, :cnt[i]])
best = min(res, key=lambda r: r["score_ms"])
peak = 1624.3
try:
    peak = json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")))["bf16_tflops"]
except Exception:
    pass
print(json.dumps({"config": f"{nq} queries x {n} items x {F} f, top-{how_many}, {'cosine (/similarity)' if KIND == ob.COSINE else 'dot (/recommendToMany)'}", "matches_single_query_path": bool(ok),
                  "score_kernel_frac_of_bf16_peak": best["score_tflops"] / peak, "bf16_peak_tflops": peak,
                  "queries_per_sec_device": nq / ((best["convert_ms"] + best["score_ms"] + best["threshold_ms"] + best["refine_ms"]) * 1e-3),
                  "queries_per_sec_wall": nq / (min(r["wall_ms"] for r in res) * 1e-3)}))
