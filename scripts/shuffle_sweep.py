#!/usr/bin/env python
"""Shuffle-only sweep (BASELINE.json config 5b): cylon::Shuffle(t, {0}) on (int64 key, double payload) tables of
1 MB .. 4 GB per GPU. Launch with torchrun, one rank per GPU. Reports, per size, the max-over-ranks time (CUDA events on
the operator's stream), algbw = table bytes per GPU / time and busbw = bytes that actually cross NVLink per GPU / time
= (P-1)/P * table bytes / time, against the 900 GB/s per direction nominal NVLink figure."""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

import bench  # noqa: E402
from cylon_b200 import api  # noqa: E402


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
    ctx.set_stream(torch.cuda.current_stream().cuda_stream)
    max_mb = int(os.environ.get("SWEEP_MAX_MB", "4096"))
    sizes_mb = [s for s in (1, 4, 16, 64, 256, 1024, 4096) if s <= max_mb]
    results = []
    for mb in sizes_mb:
        rows = mb * (1 << 20) // 16
        k, v = bench.synth_table(rows, rows * world, 3, rank)
        t = api.Table.from_numpy(ctx, [k, v], names=["k", "v"])
        iters = 10 if mb <= 256 else 5
        for _ in range(3):
            t.shuffle([0]).free()
        torch.cuda.synchronize()
        dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        total_rows = 0
        for _ in range(iters):
            out = t.shuffle([0])
            total_rows = out.row_count
            out.free()
        e1.record()
        torch.cuda.synchronize()
        ms = torch.tensor([e0.elapsed_time(e1) / iters], dtype=torch.float64, device="cuda")
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        rows_all = torch.tensor([total_rows], dtype=torch.int64, device="cuda")
        dist.all_reduce(rows_all)
        assert int(rows_all.item()) == rows * world  # conservation (cpp/test/table_op_test.cpp:38-47)
        nbytes = rows * 16
        sec = float(ms.item()) / 1e3
        results.append({"mb_per_gpu": mb, "ms": round(float(ms.item()), 4), "algbw_GBps": round(nbytes / sec / 1e9, 1),
                        "busbw_GBps": round(nbytes * (world - 1) / world / sec / 1e9, 1),
                        "busbw_frac_of_900": round(nbytes * (world - 1) / world / sec / 1e9 / 900.0, 3)})
        t.free()
    if rank == 0:
        print(json.dumps({"op": "Shuffle(t,{0}) int64 key + double payload", "n_gpus": world, "sweep": results}))
    ctx.finalize()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
