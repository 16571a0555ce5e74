"""The Distributed* operators at world 2-4 on ONE GPU: ranks are host threads of one process (cylon::net::ThreadConfig), each
with its own context; counts travel through host memory, payloads as device copies, and the peer-store shuffle writes into
the other ranks' buffers directly. Everything above the transport is the code the NCCL ranks run; every rank's result is held
to the reference's per-rank golden files (world 2 and 4) and to the oracle's per-rank results (tests/dist_checks.py).
peer = "0" sends every table through the general path (partition-ordered copy + all-to-all of column buffers)."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


# world 4 runs the golden-file checks only (reference per-rank CSVs + python/test/test_dist_rl.py counts): four Python rank
# threads share one interpreter lock, and the synthetic checks are already held at world 2 (peer stores) and 3 (general path)
@pytest.mark.parametrize("world,peer,scope", [(2, "1", "all"), (3, "0", "all"), (4, "1", "golden")])
def test_distributed_operators_with_thread_ranks(world, peer, scope):
    p = subprocess.run([sys.executable, os.path.join(ROOT, "tests", "thread_worker.py"), str(world), peer, scope], capture_output=True, text=True,
                       timeout=900, cwd=ROOT)
    assert p.returncode == 0, p.stdout[-3000:] + "\n" + p.stderr[-6000:]
    assert "THREAD_RANKS_OK world=%d peer=%s" % (world, peer) in p.stdout
