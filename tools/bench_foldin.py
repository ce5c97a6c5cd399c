"""Fold-in kernel timing: 1M interactions x F features against device-resident X / Y (oryx_fold_in_rows).
  python tools/bench_foldin.py [features] [interactions]"""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import oryx_b200 as ob
f = int(sys.argv[1]) if len(sys.argv) > 1 else 256
n = int(sys.argv[2]) if len(sys.argv) > 2 else 1_000_000
dev = torch.device("cuda", 0)
ni, nu = 2_000_000, 1_000_000
m = ob.Model(f, device=0, capacity_hint=ni)
gen = torch.Generator(device=dev); gen.manual_seed(1)
for off in range(0, ni, 1_000_000):
    blk = torch.randn((1_000_000, f), generator=gen, device=dev) * 0.3
    torch.cuda.synchronize(); m.upsert_device(ob.Y, off, blk.data_ptr(), 1_000_000); del blk
blk = torch.randn((nu, f), generator=gen, device=dev) * 0.3
torch.cuda.synchronize(); m.upsert_device(ob.X, 0, blk.data_ptr(), nu); del blk
g, _ = m.gram(ob.Y)
s = ob.Solver(g)
rng = np.random.default_rng(2)
ur = rng.integers(0, nu, n).astype(np.int32); ir = rng.integers(0, ni, n).astype(np.int32)
vals = rng.choice([1.0, 2.0, 0.5], n)
s.fold_in_rows(m, True, ob.X, ur[:200000], ir[:200000], vals[:200000])
res = []
for _ in range(3):
    t0 = time.perf_counter(); out, valid = s.fold_in_rows(m, True, ob.X, ur, ir, vals); wall = time.perf_counter() - t0
    t = s.debug_fold_in_timings(); t["wall_ms_python"] = wall * 1e3; res.append(t)
pin = torch.empty((n, f), dtype=torch.float32, pin_memory=True)
pres = []
for _ in range(3):
    t0 = time.perf_counter(); out_p, valid_p = s.fold_in_rows(m, True, ob.X, ur, ir, vals, out=pin.numpy()); wall = time.perf_counter() - t0
    t = s.debug_fold_in_timings(); t["wall_ms_python"] = wall * 1e3; pres.append(t)
ok = valid == 1
same = bool(np.array_equal(valid_p, valid) and np.array_equal(out_p[ok], out[ok]))
print(json.dumps({"features": f, "interactions": n, "runs": res, "runs_page_locked_result": pres, "page_locked_result_identical": same}))
s.close(); m.close()
