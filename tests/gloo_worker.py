"""One CPU rank of the world-2 gloo test of the shuffle's host logic (launched by tests/test_shuffle_plan_gloo.py).

What runs here is the PRODUCT's host arithmetic -- cyl_shuffle_plan from libcylon_b200.so, which needs no GPU -- wired the
way shuffle.cpp wires it: count exchange (all_gather over gloo instead of NCCL), plan, point-to-point payload moves into
the planned offsets. The partition ids and the expected per-rank result come from the oracle (checker only)."""
import ctypes as C
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from oracle import pyoracle as orc  # noqa: E402
from tests.helpers import synth_kv  # noqa: E402


def main():
    dist.init_process_group("gloo")
    rank, world = dist.get_rank(), dist.get_world_size()
    lib = C.CDLL(os.path.join(ROOT, "cylon_b200", "lib", "libcylon_b200.so"))

    # every rank owns a seeded table (key, payload); the oracle gives partition ids and the stable split order
    tables = []
    for r in range(world):
        k, v = synth_kv(5000 + 700 * r, 100000, 7, r)
        tables.append([orc.col(k), orc.col(v)])
    mine = tables[rank]
    pid, hist = orc.hash_partition([mine[0]], world)
    perm, part_off = orc.split_indices(pid, world)
    send_k, send_v = mine[0].data[perm], mine[1].data[perm]  # partition-ordered send buffers

    # count exchange (shuffle.cpp: one AllGather of the count vector)
    counts = torch.zeros(world, world, dtype=torch.int64)
    gathered = [torch.zeros(world, dtype=torch.int64) for _ in range(world)]
    dist.all_gather(gathered, torch.from_numpy(hist.astype(np.int64)))
    for r in range(world):
        counts[r] = gathered[r]
    cm = np.ascontiguousarray(counts.numpy())

    so, sc, ro, rc = (np.zeros(world, dtype=np.int64) for _ in range(4))
    total = C.c_int64()
    ret = lib.cyl_shuffle_plan(cm.ctypes.data_as(C.c_void_p), C.c_int32(world), C.c_int32(rank), so.ctypes.data_as(C.c_void_p),
                               sc.ctypes.data_as(C.c_void_p), ro.ctypes.data_as(C.c_void_p), rc.ctypes.data_as(C.c_void_p), C.byref(total))
    assert ret == 0
    assert np.array_equal(so, part_off[:world]) and np.array_equal(sc, hist.astype(np.int64))
    assert total.value == int(cm[:, rank].sum())

    # sender/receiver agreement: what I expect from p is what p plans to send me
    plans = [torch.zeros(world, dtype=torch.int64) for _ in range(world)]
    dist.all_gather(plans, torch.from_numpy(sc.copy()))
    for p in range(world):
        assert int(plans[p][rank]) == int(rc[p])

    # payload moves into the planned offsets (own slice copied locally, peers via isend/irecv)
    out_k = np.zeros(total.value, dtype=np.int64)
    out_v = np.zeros(total.value, dtype=np.float64)
    reqs, keep = [], []
    for p in range(world):
        s0, s1 = int(so[p]), int(so[p] + sc[p])
        r0, r1 = int(ro[p]), int(ro[p] + rc[p])
        if p == rank:
            out_k[r0:r1], out_v[r0:r1] = send_k[s0:s1], send_v[s0:s1]
            continue
        for src, dst in ((send_k, out_k), (send_v, out_v)):
            if s1 > s0:
                t = torch.from_numpy(np.ascontiguousarray(src[s0:s1]))
                keep.append(t)
                reqs.append(dist.isend(t, p))
            if r1 > r0:
                t = torch.from_numpy(dst[r0:r1])
                keep.append(t)
                reqs.append(dist.irecv(t, p))
    for q in reqs:
        q.wait()

    # the reference's result for this rank (oracle simulation of all ranks), chunk order by source rank
    exp = orc.shuffle(tables, [0])[rank]
    assert np.array_equal(out_k, exp[0].data) and np.array_equal(out_v, exp[1].data)
    # invalid inputs are refused
    assert lib.cyl_shuffle_plan(cm.ctypes.data_as(C.c_void_p), C.c_int32(world), C.c_int32(world), so.ctypes.data_as(C.c_void_p),
                                sc.ctypes.data_as(C.c_void_p), ro.ctypes.data_as(C.c_void_p), rc.ctypes.data_as(C.c_void_p), None) == 4
    dist.barrier()
    if rank == 0:
        print("GLOO_OK world=%d rows_rank0=%d" % (world, total.value))
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
