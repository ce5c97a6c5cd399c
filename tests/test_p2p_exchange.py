"""Row-sharded queries over peer memory (oryx_exchange_* / oryx_topn_sharded_device): the publish /
wait / merge kernel against the oracle's top-N over the concatenated shards.  World 1 runs on any GPU
box; the 2-GPU cases need two visible devices (gpurun --gpus 2) and are skipped otherwise.  The
inter-process (CUDA IPC) connection is exercised by `bench.py --gpus N`."""
import numpy as np
import pytest
import torch

import oracle
import oryx_b200 as ob
from als_harness import random_model

pytestmark = pytest.mark.gpu


def _run(world, n_per, f, hm, steps, kind=ob.DOT, seed=0xB2000050):
    ys = [random_model(n_per, f, seed + r) for r in range(world)]
    y_all = np.concatenate(ys)
    models, xs, outs, streams, dqs = [], [], [], [], []
    qs = random_model(steps, f, seed + 100)
    for r in range(world):
        with torch.cuda.device(r):
            m = ob.Model(f, device=r)
            m.upsert(ob.Y, np.arange(n_per, dtype=np.int32), ys[r])
            models.append(m)
            xs.append(ob.Exchange(r, world, r, max(hm, 16)))
            dev = torch.device("cuda", r)
            outs.append([(torch.zeros(hm, dtype=torch.int64, device=dev), torch.zeros(hm, dtype=torch.float64, device=dev),
                          torch.zeros(1, dtype=torch.int32, device=dev)) for _ in range(steps)])
            streams.append(torch.cuda.Stream(device=dev))
            dqs.append(torch.from_numpy(qs).to(dev))
    ob.Exchange.connect_local(xs)
    # every rank enqueues all its queries back to back (no host synchronisation in between): the
    # exchange kernels of different ranks meet on the device
    for i in range(steps):
        for r in range(world):
            rows, scores, n = outs[r][i]
            xs[r].topn(models[r], kind, dqs[r][i].data_ptr(), hm, n_per, rows.data_ptr(), scores.data_ptr(),
                       n.data_ptr(), streams[r].cuda_stream)
    for r in range(world):
        streams[r].synchronize()
    for i in range(steps):
        want_rows, want_scores = oracle.topn(kind, qs[i], y_all, hm)
        for r in range(world):
            rows, scores, n = outs[r][i]
            k = int(n.item())
            assert k == len(want_rows), (i, r, k)
            assert rows[:k].cpu().numpy().tolist() == want_rows.tolist(), (i, r)
            assert scores[:k].cpu().numpy().tobytes() == want_scores.tobytes(), (i, r)
    for x in xs:
        x.close()
    for m in models:
        m.close()


def test_exchange_world_1_is_the_local_topn():
    _run(1, 30000, 32, 10, steps=6)


def test_exchange_world_1_short_shard_and_cosine():
    _run(1, 7, 16, 10, steps=3, kind=ob.COSINE)  # fewer rows than how_many


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
@pytest.mark.parametrize("hm", [10, 100])
def test_exchange_two_gpus_matches_oracle(hm):
    _run(2, 50000, 64, hm, steps=40)


@pytest.mark.skipif(torch.cuda.device_count() < 4, reason="needs four GPUs")
def test_exchange_four_gpus_matches_oracle():
    _run(4, 20000, 48, 10, steps=40)
