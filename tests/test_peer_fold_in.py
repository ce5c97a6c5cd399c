"""Cross-PROCESS sharded fold-in (oryx_peer_stores_* + oryx_fold_in_rows_peers): two processes, each
holding a row shard of X and Y on its GPU, fold in interactions whose rows live on either -- checked against the
oracle over the whole matrices inside tests/peer_fold_in_worker.py.  Runs on a one-GPU box too (both processes on
GPU 0: CUDA IPC maps the other process's stores just the same)."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("world", [2, 3])
def test_fold_in_across_processes_matches_oracle(world):
    here = os.path.dirname(os.path.abspath(__file__))
    port = 29640 + world
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
           "--master-addr", "127.0.0.1", "--master-port", str(port), os.path.join(here, "peer_fold_in_worker.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert r.stdout.count("peer fold-in ok") == world
