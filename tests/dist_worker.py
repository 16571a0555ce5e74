"""One rank of the multi-GPU parity run (launched by tests/test_gpu_distributed.py through torch.distributed.run).
Each rank runs the Distributed* operators over NCCL on its own GPU and compares ITS result with the oracle's result
for that rank (the oracle simulates all ranks) and, where the reference ships per-rank golden files, with those."""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from cylon_b200 import api  # noqa: E402
from oracle import pyoracle as orc  # noqa: E402
from tests.dist_checks import run_checks  # noqa: E402

F64 = [np.int64, np.float64]


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    idt = torch.zeros(128, dtype=torch.uint8, device="cuda")
    if rank == 0:
        idt = torch.frombuffer(bytearray(api.nccl_unique_id()), dtype=torch.uint8).cuda()
    dist.broadcast(idt, 0)
    ctx = api.CylonContext(device=local, distributed=True, rank=rank, world_size=world, nccl_id=idt.cpu().numpy().tobytes())
    assert ctx.get_rank() == rank and ctx.get_world_size() == world
    peer = os.environ.get("CYLON_TEST_SHUFFLE_PEER", "1")
    ctx.add_config("cylon.shuffle.peer", peer)  # "0": every shuffle takes the NCCL send/recv path, also for fixed-width tables
    if os.environ.get("CYLON_TEST_SCOPE") == "sort":  # Sort / MapToSortPartitions / DistributedSort only (tests/sort_checks.py)
        from tests.dist_checks import _Recompute
        from tests.sort_checks import run_dist_checks
        checks = run_dist_checks(api, ctx, rank, world, _Recompute())
    else:
        checks = run_checks(api, ctx, rank, world)

    ctx.barrier()
    tot = torch.tensor([checks], device="cuda")
    dist.all_reduce(tot)
    if rank == 0:
        print("DIST_OK world=%d peer=%s checks=%d" % (world, peer, int(tot.item())))
    ctx.finalize()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
