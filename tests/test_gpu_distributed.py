"""Multi-GPU parity: launches tests/dist_worker.py under torch.distributed.run, one rank per GPU, NCCL transport.
Needs >= 2 GPUs on the box (gpurun --gpus N); skipped otherwise. peer = "0" forces the NCCL send/recv shuffle for every
table (the default sends fixed-width, null-free tables through NVLink peer stores instead). The worker's output is kept
under gpurun_out/ so that a run on the GPU box leaves a record (profiles/README.md points at the committed copies)."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _ngpus():
    from cylon_b200 import capi
    return capi.device_count()


@pytest.mark.parametrize("world,peer", [(2, "1"), (2, "0"), (4, "1"), (4, "0"), (8, "1")])
def test_distributed_operators(world, peer):
    if _ngpus() < world:
        pytest.skip("needs %d GPUs" % world)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(world), "--master-addr", "127.0.0.1",
           "--master-port", str(29500 + world), os.path.join(ROOT, "tests", "dist_worker.py")]
    env = dict(os.environ, CYLON_TEST_SHUFFLE_PEER=peer)
    p = subprocess.run(cmd, capture_output=True, text=True, timeout=900, cwd=ROOT, env=env)
    try:
        os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
        with open(os.path.join(ROOT, "gpurun_out", "dist_test_n%d_peer%s.log" % (world, peer)), "w") as f:
            f.write("returncode %d\n--- stdout\n%s\n--- stderr (tail)\n%s\n" % (p.returncode, p.stdout[-4000:], p.stderr[-4000:]))
    except OSError:
        pass
    assert p.returncode == 0, p.stdout[-3000:] + "\n" + p.stderr[-6000:]
    assert "DIST_OK world=%d peer=%s" % (world, peer) in p.stdout


@pytest.mark.parametrize("world", [2, 4])
def test_distributed_sort_over_nccl(world):
    """DistributedSort / MapToSortPartitions with one process per GPU (the same checks tests/test_gpu_sort.py runs with thread ranks)"""
    if _ngpus() < world:
        pytest.skip("needs %d GPUs" % world)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(world), "--master-addr", "127.0.0.1",
           "--master-port", str(29600 + world), os.path.join(ROOT, "tests", "dist_worker.py")]
    p = subprocess.run(cmd, capture_output=True, text=True, timeout=900, cwd=ROOT, env=dict(os.environ, CYLON_TEST_SCOPE="sort"))
    assert p.returncode == 0, p.stdout[-3000:] + "\n" + p.stderr[-6000:]
    assert "DIST_OK world=%d" % world in p.stdout

