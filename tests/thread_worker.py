"""Ranks as THREADS of this process on one GPU (cylon::net::ThreadConfig): the same checks as tests/dist_worker.py runs over
NCCL, so that the distributed operators' host logic, count exchange, packing and the peer-store scatter are exercised at
world > 1 on a single-GPU box. usage: thread_worker.py WORLD [PEER] [golden|sort]
("golden": only the checks against the reference's per-rank golden files and known answers; "sort": tests/sort_checks.py instead)"""
import os
import sys
import threading
import traceback

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from cylon_b200 import api  # noqa: E402
from tests.dist_checks import SharedExpectations, run_checks  # noqa: E402


def main():
    world = int(sys.argv[1])
    peer = sys.argv[2] if len(sys.argv) > 2 else "1"
    golden_only = len(sys.argv) > 3 and sys.argv[3] == "golden"
    sort_only = len(sys.argv) > 3 and sys.argv[3] == "sort"
    results, errors = [0] * world, []
    shared = SharedExpectations()

    def rank_main(rank):
        try:
            ctx = api.CylonContext(device=0, thread_group="ranks-%d" % os.getpid(), rank=rank, world_size=world)
            ctx.add_config("cylon.shuffle.peer", peer)
            assert ctx.get_rank() == rank and ctx.get_world_size() == world
            if sort_only:
                from tests.sort_checks import run_dist_checks
                results[rank] = run_dist_checks(api, ctx, rank, world, shared)
            else:
                results[rank] = run_checks(api, ctx, rank, world, shared, golden_only)
            ctx.barrier()
        except BaseException:
            errors.append("rank %d:\n%s" % (rank, traceback.format_exc()))
            os.write(2, errors[-1].encode())
            os._exit(3)  # the other ranks are waiting in a collective: end the process rather than hang

    threads = [threading.Thread(target=rank_main, args=(r,)) for r in range(world)]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    print("THREAD_RANKS_OK world=%d peer=%s checks=%d" % (world, peer, sum(results)))


if __name__ == "__main__":
    main()
