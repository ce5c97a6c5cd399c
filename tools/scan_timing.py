"""Development aid: where does one scan_topn launch spend its time? (globaltimer stamps)"""
import ctypes as C
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import oryx_b200 as ob
from oryx_b200 import native

n = int(sys.argv[1]) if len(sys.argv) > 1 else 10_000_000
F = int(sys.argv[2]) if len(sys.argv) > 2 else 128
lsh = len(sys.argv) > 3 and sys.argv[3] == "lsh"
dev = torch.device("cuda", 0)
if lsh:
    H, bits = ob.lsh_plan(0.3, 32)
    hv = np.random.default_rng(4).standard_normal((H, F), dtype=np.float32)
    m = ob.Model(F, hash_vectors=hv, max_bits_differing=bits, capacity_hint=n)
else:
    m = ob.Model(F, capacity_hint=n)
g = torch.Generator(device=dev); g.manual_seed(1)
for off in range(0, n, 1_000_000):
    c = min(1_000_000, n - off)
    blk = torch.randn((c, F), generator=g, device=dev); torch.cuda.synchronize()
    m.upsert_device(ob.Y, off, blk.data_ptr(), c)
if lsh:
    m.compact(ob.Y)
q = np.random.default_rng(0).standard_normal((8, F), dtype=np.float32)
L = native.lib()
L.oryx_debug_scan_timestamps.restype = C.c_int32
L.oryx_debug_scan_timestamps.argtypes = [C.c_void_p, C.c_int32, C.POINTER(C.c_uint64)]
for i in range(3): m.topn(ob.DOT, q[i], 10)
for i in range(5):
    L.oryx_debug_scan_timestamps(m._h, 1, None)
    m.topn(ob.DOT, q[i], 10)
    out = (C.c_uint64 * 8)()
    L.oryx_debug_scan_timestamps(m._h, 0, out)
    t = [int(x) for x in out]
    t0 = t[0]
    print("scan_end first/last: %.1f / %.1f us | merge start %.1f | select +%.1f | rescore +%.1f | rank+write +%.1f | total %.1f us (event %.1f)" % (
        (t[1]-t0)/1e3, (t[2]-t0)/1e3, (t[3]-t0)/1e3, (t[4]-t[3])/1e3, (t[5]-t[4])/1e3, (t[6]-t[5])/1e3, (t[6]-t0)/1e3, m.stats()["last_query_ms"]*1e3))
