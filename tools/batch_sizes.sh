#!/bin/bash
# Latency / throughput of the tensor-core batch path at small batch sizes (10 M x 128 f, top-10).
for nq in 1 2 8 32 128 512 1024 4096; do
  python tools/bench_batch.py 10000000 $nq 10 128 2>&1 | tail -2 | head -1 > /tmp/bb.json
  python - "$nq" <<'PY'
import json, sys
d = json.load(open("/tmp/bb.json"))
dev = d["convert_ms"] + d["score_ms"] + d["threshold_ms"] + d["refine_ms"]
print(json.dumps({"nq": int(sys.argv[1]), "wall_ms": round(d["wall_ms"], 3), "device_ms": round(dev, 3),
                  "score_ms": round(d["score_ms"], 3), "threshold_ms": round(d["threshold_ms"], 3),
                  "refine_ms": round(d["refine_ms"], 3), "candidates_mean": d["candidate_subblocks_mean"],
                  "handed_back": d["handed_back"], "qps_wall": round(int(sys.argv[1]) / d["wall_ms"] * 1e3)}))
PY
done
