"""C5 measurement: items x 256 f sharded by rows over the GPUs of one box (torchrun, one rank per GPU),
LSH candidate filtering ON (sample-rate 0.3 on 32 cores -> 8 hashes, 2 bits differing = 37 of 256
partitions), per-GPU top-10, one packed all-gather over NCCL/NVLink, merge kernel.

  torchrun --nproc-per-node N tools/bench_c5.py [total_items] [features] [steps]
"""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import torch.distributed as dist
import oryx_b200 as ob
from oryx_b200.sharded import P2PShardedTopN, ShardPlan, exchange_packed

EXCHANGE = os.environ.get("ORYX_BENCH_EXCHANGE", "p2p")  # "p2p": peer-memory exchange kernel; "nccl": all-gather + merge

total = int(sys.argv[1]) if len(sys.argv) > 1 else 100_000_000
F = int(sys.argv[2]) if len(sys.argv) > 2 else 256
steps = int(sys.argv[3]) if len(sys.argv) > 3 else 100
rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1")); lr = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(lr); dev = torch.device("cuda", lr)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)
plan = ShardPlan.for_rows(total, world)
lo, hi = plan.local_range(rank, total)
n = hi - lo
H, bits = ob.lsh_plan(0.3, 32)
hv = np.random.default_rng(4).standard_normal((H, F), dtype=np.float32)
m = ob.Model(F, hash_vectors=hv, max_bits_differing=bits, device=lr, capacity_hint=n, reserved_sms=1 if world > 1 else 0)
g = torch.Generator(device=dev); g.manual_seed(100 + rank)
t0 = time.perf_counter()
for off in range(0, n, 1_000_000):
    c = min(1_000_000, n - off)
    blk = torch.randn((c, F), generator=g, device=dev); torch.cuda.synchronize()
    m.upsert_device(ob.Y, off, blk.data_ptr(), c)
    del blk
t_load = time.perf_counter() - t0
HM = 10
rec_bytes = ob.packed_record_bytes(HM)
d_rec = torch.zeros(rec_bytes, dtype=torch.uint8, device=dev)
g_rec = torch.zeros(world * rec_bytes, dtype=torch.uint8, device=dev)
m_rows = torch.zeros(HM, dtype=torch.int64, device=dev); m_scores = torch.zeros(HM, dtype=torch.float64, device=dev); m_n = torch.zeros(1, dtype=torch.int32, device=dev)
NQ = 32
d_q = torch.from_numpy(np.random.default_rng(9).standard_normal((NQ, F), dtype=np.float32)).to(dev)
stream = torch.cuda.current_stream(dev)
pipe = P2PShardedTopN(ob, m, dist, torch, plan, HM, dev) if (world > 1 and EXCHANGE == "p2p") else None

def step(i):
    if pipe is not None:
        pipe.submit(ob.DOT, d_q[i % NQ].data_ptr())
        return
    m.topn_device(ob.DOT, d_q[i % NQ].data_ptr(), HM, d_rec[HM * 8:].data_ptr(), d_rec.data_ptr(), d_rec[HM * 12:].data_ptr(), stream.cuda_stream)
    if world > 1:
        exchange_packed(dist, torch, plan, d_rec, g_rec)
        ob.merge_topn_packed(g_rec.data_ptr(), world, rec_bytes, HM, plan.rows_per_rank, m_rows.data_ptr(), m_scores.data_ptr(), m_n.data_ptr(), stream.cuda_stream)

def run(label):
    for i in range(5): step(i)
    torch.cuda.synchronize()
    if world > 1: dist.barrier(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for i in range(steps): step(i)
    e1.record(stream); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    if world > 1:
        t = torch.tensor([ms], dtype=torch.float64, device=dev); dist.all_reduce(t, op=dist.ReduceOp.MAX); ms = float(t.item())
    return ms / steps

ms_uncompacted = run("uncompacted")
free_b, _ = torch.cuda.mem_get_info(dev)
in_place = free_b < n * F * 4 * 1.05   # no room for a second copy of the shard: oryx_items_compact permutes in place
can_compact = True
t0 = time.perf_counter(); m.compact(ob.Y); t_compact = time.perf_counter() - t0
ms_compacted = run("compacted")
if rank == 0:
    frac = 37 / 256
    print(json.dumps({
        "config": f"{total} items x {F} f over {world} GPU(s), LSH sample-rate 0.3/32 cores ({H} hashes, {bits} bits -> 37 of 256 partitions), top-{HM}",
        "exchange": ("peer-memory kernel" if pipe is not None else "NCCL all-gather + merge") if world > 1 else None,
        "compacted": can_compact, "compaction_in_place": bool(in_place), "items_per_gpu": n, "load_s_per_gpu": round(t_load, 1), "compact_s": round(t_compact, 1),
        "ms_per_query_uncompacted": ms_uncompacted, "ms_per_query_compacted": ms_compacted,
        "qps": 1000.0 / ms_compacted, "nominal_item_scores_per_sec": total * 1000.0 / ms_compacted,
        "actual_item_scores_per_sec": frac * total * 1000.0 / ms_compacted,
        "effective_GBps_per_gpu_on_candidates": frac * n * F * 4 / (ms_compacted * 1e-3) / 1e9}))
m.close()
if world > 1: dist.destroy_process_group()
