"""LSH hashing throughput: rows x features hashed by H hash vectors (upsert_device path, rows resident).
  python tools/bench_lsh.py [rows] [features] [sample_rate] [cores]"""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import oryx_b200 as ob
n = int(sys.argv[1]) if len(sys.argv) > 1 else 10_000_000
f = int(sys.argv[2]) if len(sys.argv) > 2 else 128
rate = float(sys.argv[3]) if len(sys.argv) > 3 else 1.0
cores = int(sys.argv[4]) if len(sys.argv) > 4 else 32
H, bits = ob.lsh_plan(rate, cores)
hv = np.random.default_rng(4).standard_normal((H, f), dtype=np.float32)
dev = torch.device("cuda", 0)
m = ob.Model(f, hash_vectors=hv, max_bits_differing=bits, device=0, capacity_hint=n)
gen = torch.Generator(device=dev); gen.manual_seed(1)
blk = torch.randn((1_000_000, f), generator=gen, device=dev)
torch.cuda.synchronize()
m.upsert_device(ob.Y, 0, blk.data_ptr(), 1_000_000)          # warm
t0 = time.perf_counter()
for off in range(0, n, 1_000_000):
    m.upsert_device(ob.Y, off, blk.data_ptr(), min(1_000_000, n - off))
dt = time.perf_counter() - t0
print(json.dumps({"rows": n, "features": f, "hashes": H, "upsert_device_ms_total": dt * 1e3,
                  "note": "wall time of hash + scatter + bookkeeping per 1M-row chunk; see the ncu launch list for the hash kernel alone"}))
m.close()
