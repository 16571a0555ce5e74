"""Multi-GPU parity: launches tests/dist_worker.py under torch.distributed.run, one rank per GPU, NCCL transport.
Needs >= 2 GPUs on the box (gpurun --gpus N); skipped otherwise."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _ngpus():
    from cylon_b200 import capi
    return capi.device_count()


@pytest.mark.parametrize("world", [2, 4, 8])
def test_distributed_operators(world):
    if _ngpus() < world:
        pytest.skip("needs %d GPUs" % world)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(world), "--master-addr", "127.0.0.1",
           "--master-port", str(29500 + world), os.path.join(ROOT, "tests", "dist_worker.py")]
    p = subprocess.run(cmd, capture_output=True, text=True, timeout=900, cwd=ROOT)
    assert p.returncode == 0, p.stdout[-3000:] + "\n" + p.stderr[-6000:]
    assert "DIST_OK world=%d" % world in p.stdout
