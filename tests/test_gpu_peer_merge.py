"""GPU suite (-m gpu): the library-owned multi-process OR-allreduce (rpt_bf_peer_allreduce and friends, rpt_gpu.h).
Two processes with their own CUDA contexts — sharing one GPU when the box has one — merge row-range shards over CUDA
IPC and compare the merged words with a single build (tests/multi/peer_allreduce_check.py), once with
cuStreamWaitValue32 waits and once with the polling-kernel fallback."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("poll", [False, True])
def test_peer_allreduce_between_processes(rpt, poll):
    env = dict(os.environ)
    env.pop("RPT_SYNC_POLL", None)
    if poll:
        env["RPT_SYNC_POLL"] = "1"
        env["PEER_CHECK_ROWS"] = "300000"  # two contexts time-slice one GPU: a polling wait costs a time slice
    port = 29541 + (1 if poll else 0)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
           "127.0.0.1", "--master-port", str(port), os.path.join(ROOT, "tests", "multi", "peer_allreduce_check.py")]
    proc = subprocess.run(cmd, capture_output=True, text=True, timeout=600, env=env, cwd=ROOT)
    assert proc.returncode == 0, proc.stdout[-2000:] + proc.stderr[-4000:]
    assert "peer allreduce ok" in proc.stdout
