"""Gram timing: items x features fp32 on one GPU, tensor-core kernel vs the floating-point large-store kernel.
  python tools/bench_gram.py [items] [features] [reps]      (ORYX_GRAM_TC_DIAG_CTAS=n tunes the role split)"""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import oryx_b200 as ob

n = int(sys.argv[1]) if len(sys.argv) > 1 else 10_000_000
f = int(sys.argv[2]) if len(sys.argv) > 2 else 256
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 5
dev = torch.device("cuda", 0)
m = ob.Model(f, device=0, capacity_hint=n)
gen = torch.Generator(device=dev); gen.manual_seed(1)
for off in range(0, n, 1_000_000):
    c = min(1_000_000, n - off)
    blk = torch.randn((c, f), generator=gen, device=dev) * 0.3
    torch.cuda.synchronize()
    m.upsert_device(ob.Y, off, blk.data_ptr(), c)
    del blk
g, used = m.gram(ob.Y)
ms = []
for _ in range(reps):
    g, used = m.gram(ob.Y)
    ms.append(m.debug_gram())
print(json.dumps({"items": n, "features": f, "runs": ms, "best_ms": min(x[0] for x in ms),
                  "GBps": n * f * 4 / (min(x[0] for x in ms) * 1e-3) / 1e9, "symmetric": bool(np.array_equal(g, g.T)),
                  "diag_mean": float(np.mean(np.diag(g))) / n}))
m.close()
