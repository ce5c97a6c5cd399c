"""A peer that never arrives must cost a bounded wait and an error, not a hang (exchange.cu: every wait on peer
memory carries ORYX_EXCHANGE_TIMEOUT_MS).  Two exchanges are connected on GPU 0 and only rank 0 ever issues its
call, so rank 1's record / Gram slices never land.  Runs in a child process: the timeout is read once per process."""
import os
import subprocess
import sys
import textwrap

import pytest

pytestmark = pytest.mark.gpu

CHILD = textwrap.dedent("""
    import sys, time
    import numpy as np, torch
    sys.path.insert(0, %r)
    import oryx_b200 as ob
    f, n, hm = 16, 5000, 10
    rng = np.random.default_rng(1)
    dev = torch.device("cuda", 0)
    m = ob.Model(f, device=0)
    m.upsert(ob.Y, np.arange(n, dtype=np.int32), rng.standard_normal((n, f), dtype=np.float32))
    xs = [ob.Exchange(0, 2, r, 16) for r in range(2)]
    ob.Exchange.connect_local(xs)
    q = torch.from_numpy(rng.standard_normal(f, dtype=np.float32)).to(dev)
    rows = torch.zeros(hm, dtype=torch.int64, device=dev); scores = torch.zeros(hm, dtype=torch.float64, device=dev)
    cnt = torch.zeros(1, dtype=torch.int32, device=dev)
    t0 = time.perf_counter()
    xs[0].topn(m, ob.DOT, q.data_ptr(), hm, n, rows.data_ptr(), scores.data_ptr(), cnt.data_ptr(),
               torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    assert int(cnt.item()) == -2, int(cnt.item())          # "a peer never arrived"
    assert 0.2 <= dt < 5.0, dt
    # the Gram all-reduce: status -2 on the device form, an error from the host form
    g = torch.ones((f, f), dtype=torch.float64, device=dev); st = torch.zeros(1, dtype=torch.int32, device=dev)
    xs[0].gram_allreduce_device(g.data_ptr(), f, st.data_ptr(), torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    assert int(st.item()) == -2
    try:
        xs[0].gram_sharded(m, ob.Y)
        raise SystemExit("no error from gram_sharded")
    except ob.OryxError as e:
        assert "did not arrive" in str(e), str(e)
    print("timeouts ok")
""")


def test_missing_peer_times_out_instead_of_hanging():
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    env = dict(os.environ, ORYX_EXCHANGE_TIMEOUT_MS="300")
    r = subprocess.run([sys.executable, "-c", CHILD % root], capture_output=True, text=True, timeout=300, env=env)
    assert r.returncode == 0 and "timeouts ok" in r.stdout, r.stdout[-2000:] + r.stderr[-2000:]
