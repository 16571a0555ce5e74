"""cylon::Sort, MapToSortPartitions and DistributedSort on the GPU against the oracle (tests/sort_checks.py): one rank here, ranks as
host threads of one process for the distributed forms (world 2 over peer stores is not used by the sort -- its shuffle takes the
general path -- so the transport is the thread communicator's all-to-all; world 3 covers unequal and empty ranks)."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_sort_and_range_partition_one_rank():
    from cylon_b200 import api
    from tests.sort_checks import run_local_checks
    ctx = api.CylonContext(device=0)
    assert run_local_checks(api, ctx) > 200
    ctx.finalize()


@pytest.mark.parametrize("world", [2, 3])
def test_distributed_sort_with_thread_ranks(world):
    p = subprocess.run([sys.executable, os.path.join(ROOT, "tests", "thread_worker.py"), str(world), "1", "sort"], capture_output=True, text=True,
                       timeout=600, cwd=ROOT)
    assert p.returncode == 0, p.stdout[-3000:] + "\n" + p.stderr[-6000:]
    assert "THREAD_RANKS_OK world=%d peer=1 checks=" % world in p.stdout
