"""Where does a small LSH-filtered shard lose time?  One GPU, 12.5 M x 256 f (the per-GPU share of C5 at N = 8):
pipelined oryx_topn_device calls, LSH on (compacted) vs an unfiltered store holding the same number of candidate
bytes, vs the device timestamps of single calls.

  python tools/bench_c5_tail.py [rows] [features]
"""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import oryx_b200 as ob

n = int(sys.argv[1]) if len(sys.argv) > 1 else 12_500_000
F = int(sys.argv[2]) if len(sys.argv) > 2 else 256
dev = torch.device("cuda", 0)
HM, NQ, STEPS = 10, 32, 200
d_q = torch.from_numpy(np.random.default_rng(9).standard_normal((NQ, F), dtype=np.float32)).to(dev)
stream = torch.cuda.current_stream(dev)
out = {}


def load(model, rows):
    g = torch.Generator(device=dev); g.manual_seed(100)
    for off in range(0, rows, 1_000_000):
        c = min(1_000_000, rows - off)
        blk = torch.randn((c, F), generator=g, device=dev); torch.cuda.synchronize()
        model.upsert_device(ob.Y, off, blk.data_ptr(), c)
        del blk


def pipelined(model, label):
    d_rows = torch.zeros(HM, dtype=torch.int32, device=dev); d_sc = torch.zeros(HM, dtype=torch.float64, device=dev)
    d_n = torch.zeros(1, dtype=torch.int32, device=dev)
    def step(i):
        model.topn_device(ob.DOT, d_q[i % NQ].data_ptr(), HM, d_rows.data_ptr(), d_sc.data_ptr(), d_n.data_ptr(), stream.cuda_stream)
    for i in range(10): step(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for i in range(STEPS): step(i)
    e1.record(stream); torch.cuda.synchronize()
    out[label + "_ms_per_query_pipelined"] = e0.elapsed_time(e1) / STEPS
    # one at a time (latency)
    ts = []
    for i in range(20):
        torch.cuda.synchronize()
        e0.record(stream); step(i); e1.record(stream); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    out[label + "_ms_single_median"] = float(np.median(ts))


H, bits = ob.lsh_plan(0.3, 32)
hv = np.random.default_rng(4).standard_normal((H, F), dtype=np.float32)
for reserved in (0, 1):
    m = ob.Model(F, hash_vectors=hv, max_bits_differing=bits, device=0, capacity_hint=n, reserved_sms=reserved)
    load(m, n); m.compact(ob.Y)
    pipelined(m, f"lsh_reserved{reserved}")
    m.close()
cand = int(n * 37 / 256)
m = ob.Model(F, device=0, capacity_hint=cand)
load(m, cand)
pipelined(m, "unfiltered_same_bytes")
m.close()
out["candidate_rows"] = cand
out["candidate_GB"] = cand * F * 4 / 1e9
for k in list(out):
    if k.endswith("_pipelined"):
        out[k.replace("_ms_per_query_pipelined", "_GBps")] = out["candidate_GB"] / out[k] * 1e3
print(json.dumps(out, indent=1))

# device timestamps of single calls (ns relative to the earliest scan CTA's start)
import ctypes as C
from oryx_b200 import native
L = native.lib()
L.oryx_debug_scan_timestamps.restype = C.c_int32
L.oryx_debug_scan_timestamps.argtypes = [C.c_void_p, C.c_int32, C.POINTER(C.c_ulonglong)]
m = ob.Model(F, hash_vectors=hv, max_bits_differing=bits, device=0, capacity_hint=n)
load(m, n); m.compact(ob.Y)
d_rows = torch.zeros(HM, dtype=torch.int32, device=dev); d_sc = torch.zeros(HM, dtype=torch.float64, device=dev)
d_n = torch.zeros(1, dtype=torch.int32, device=dev)
for i in range(5):
    m.topn_device(ob.DOT, d_q[i].data_ptr(), HM, d_rows.data_ptr(), d_sc.data_ptr(), d_n.data_ptr(), stream.cuda_stream)
torch.cuda.synchronize()
stamps = []
for i in range(5):
    L.oryx_debug_scan_timestamps(m._h, 1, None)
    m.topn_device(ob.DOT, d_q[i].data_ptr(), HM, d_rows.data_ptr(), d_sc.data_ptr(), d_n.data_ptr(), stream.cuda_stream)
    ts = (C.c_ulonglong * 16)()
    L.oryx_debug_scan_timestamps(m._h, 0, ts)
    t = [int(x) for x in ts]
    stamps.append({"scan_first_end_us": (t[1] - t[0]) / 1e3, "scan_last_end_us": (t[2] - t[0]) / 1e3,
                   "finish_start_us": (t[3] - t[0]) / 1e3, "select_done_us": (t[4] - t[0]) / 1e3,
                   "rescore_done_us": (t[5] - t[0]) / 1e3, "finish_end_us": (t[6] - t[0]) / 1e3,
                   "select_phases_us_after_finish_start": [(t[i] - t[3]) / 1e3 for i in (8, 9, 10, 4)]})
print(json.dumps(stamps, indent=1))
m.close()
