#!/usr/bin/env python
"""bench.py -- DistributedJoin throughput (input rows/s) on N B200s, with roofline and CPU baseline.

Contract (see DESIGN.md "Measurement"):
  python bench.py --gpus N --steps K --warmup W            (N > 1: launched under torchrun, one rank per GPU)
  python bench.py --impl reference --gpus N --steps K --warmup W

A "step" is one cylon::DistributedJoin(inner, HASH) of two (int64 key, double payload) tables with `--rows` rows per
side per GPU (BASELINE.json configs[1]: 100M), keys uniform in [0, rows * N) (the reference's in-memory generator,
cpp/src/examples/ra_test_inmem_datagen_example.cpp:104-120). `value` = (N_L + N_R) * N / max-over-ranks step time
with the inputs already in HBM (CUDA events on the operator's stream); `e2e` = the same through host buffers
(pinned H2D of both tables and D2H of the whole result inside the timed region).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "DistributedJoin input rows/sec"
UNIT = "rows/s"


def splitmix64(x):
    with np.errstate(over="ignore"):
        x = x + np.uint64(0x9E3779B97F4A7C15)
        z = (x ^ (x >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        return z ^ (z >> np.uint64(31))


def synth_table(rows, key_range, seed, rank, out_keys=None, out_vals=None, chunk=1 << 24):
    """(int64 key uniform in [0, key_range), double payload in [DBL_MIN, 1)); deterministic in (seed, rank, row)."""
    keys = out_keys if out_keys is not None else np.empty(rows, dtype=np.int64)
    vals = out_vals if out_vals is not None else np.empty(rows, dtype=np.float64)
    # one independent counter-based stream per (seed, rank, column): random 64-bit start, then consecutive counters
    kbase = splitmix64(np.array([seed * 4096 + rank * 2 + 0], dtype=np.uint64))[0]
    vbase = splitmix64(np.array([seed * 4096 + rank * 2 + 1], dtype=np.uint64))[0]
    tiny = np.finfo(np.float64).tiny
    for s in range(0, rows, chunk):
        e = min(rows, s + chunk)
        with np.errstate(over="ignore"):
            i = np.arange(s, e, dtype=np.uint64)
            ki, vi = i + kbase, i + vbase
        keys[s:e] = (splitmix64(ki) % np.uint64(key_range)).astype(np.int64)
        v = (splitmix64(vi) >> np.uint64(11)).astype(np.float64) * (1.0 / 9007199254740992.0)
        np.maximum(v, tiny, out=v)
        vals[s:e] = v
    return keys, vals


class ClockSampler(threading.Thread):
    """samples nvidia-smi clocks / throttle reasons of one GPU while the timed region runs"""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        super().__init__(daemon=True)
        self.gpu = gpu_index
        self.samples = []
        self.stop_flag = threading.Event()

    def run(self):
        while not self.stop_flag.is_set():
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.gpu), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits"],
                                     capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.samples.append([f.strip() for f in out.splitlines()[0].split(",")])
            except Exception:
                pass
            self.stop_flag.wait(0.2)

    def summary(self):
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for s in self.samples:
            try:
                sm.append(float(s[1]))
                mx.append(float(s[2]))
            except (ValueError, IndexError):
                continue
            for name, flag in zip(names, s[4:8]):
                if flag.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def load_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


# algorithmic bytes per launch of each profiled kernel, per DESIGN.md "Kernels": inputs read once + outputs written once;
# hash-table / index / scratch traffic is NOT credited. n = rows per side, m = output rows.
def algorithmic_bytes(kernel, n, m):
    table = {
        "hash_join_build": n * 8,            # build keys in
        "hash_join_probe_count": n * 8,      # probe keys in
        "hash_join_probe_write": m * 16,     # (l, r) index pairs out
        "hash_join_scan": n * 4 + n * 8,     # counts in, offsets out
        "hash_join_init_slots": 0,
        "gather_fixed": m * 8 + m * 8,       # per launch: one 8-byte column, values in + out
        "fused_join_hist": n * 8,            # per launch (one table): keys in
        "fused_join_offsets": 0,
        "fused_join_scatter": n * 16 * 2,    # per launch (one table): (key, payload) columns in + packed records out
        "fused_join_scatter_coarse": n * 16 * 2,  # per launch (one table): columns in + records ordered by coarse digit out
        "fused_join_scatter_fine": n * 16 * 2,    # per launch (one table): records in + records ordered by bin out
        "fused_join_probe": 2 * n * 16 + m * 32,  # both record streams in + the 4 result columns out (= SURVEY 8d local join)
        "partition_scatter_peer": n * 16 * 2,  # per launch (one table): rows in + rows out (remote share over NVLink)
        "partition_tile_hist": n * 8,
        "partition_scatter": n * 16 * 2,     # whole (key, payload) table in + out
        "partition_scan": 0,
    }
    return table.get(kernel, 0)


def cpu_baseline(lk, lv, rk, rv, rows_sample, threads=1):
    """the reference's CPU join timed on this box: oracle/_ref (the reference's own JoinTables) when built, else the
    oracle port. Returns the cpu_baseline object."""
    from oracle import pyoracle as orc
    from oracle import pyref
    # keys of the sample keep the full workload's match rate (~1 match per probe row): fold them into [0, sample)
    lk = lk[:rows_sample] % rows_sample
    rk = rk[:rows_sample] % rows_sample
    lv, rv = lv[:rows_sample], rv[:rows_sample]
    t0 = time.perf_counter()
    if pyref.available():
        out = pyref.join_tables([orc.col(lk), orc.col(lv)], [orc.col(rk), orc.col(rv)], [0], [0], orc.INNER, orc.HASH)
        m = out[0].length
        kind = "reference"
    else:
        m, _cs = orc.join_i64_f64(lk, lv, rk, rv, orc.INNER, orc.HASH)
        kind = "port"
    dt = time.perf_counter() - t0
    return {"value": 2 * rows_sample / dt, "unit": UNIT, "cores": threads, "kind": kind,
            "sample": "inner hash join of %d x %d rows (keys folded into [0, %d)), %d output rows, %.2f s; "
                      "cylon::join::JoinTables, single-threaded as one MPI rank of the reference runs"
                      % (rows_sample, rows_sample, rows_sample, m, dt)}


def _ref_worker(a):
    rows, rank = a
    from oracle import pyoracle as orc
    from oracle import pyref
    lk, lv = synth_table(rows, rows, 1, rank)
    rk, rv = synth_table(rows, rows, 2, rank)
    t0 = time.perf_counter()
    if pyref.available():
        out = pyref.join_tables([orc.col(lk), orc.col(lv)], [orc.col(rk), orc.col(rv)], [0], [0], orc.INNER, orc.HASH)
        m = out[0].length
    else:
        m, _ = orc.join_i64_f64(lk, lv, rk, rv, orc.INNER, orc.HASH)
    return time.perf_counter() - t0, m


def run_reference(args):
    """--impl reference: the reference's CPU path on all host cores -- one single-threaded rank per core (how the
    reference scales: MPI ranks), each joining its own already-partitioned shard pair (the shuffle is free)."""
    if int(os.environ.get("RANK", "0")) != 0:
        return 0
    import multiprocessing as mp
    from oracle import pyoracle as orc
    from oracle import pyref
    cores = os.cpu_count() or 1
    rows = args.ref_rows
    if pyref.dist_available():
        # The reference's own cylon::DistributedJoin (table.cpp:482-502: hash partition, split, Arrow all-to-all, local
        # hash join) with one rank per host core. Ranks are threads of one process and the all-to-all goes through an
        # in-memory channel instead of MPI (oracle/ref_shim), which can only flatter the reference.
        lefts, rights = [], []
        for r in range(cores):
            lk, lv = synth_table(rows, rows * cores, 1, r)
            rk, rv = synth_table(rows, rows * cores, 2, r)
            lefts.append([orc.col(lk), orc.col(lv)])
            rights.append([orc.col(rk), orc.col(rv)])
        times, out_rows = [], 0
        for step in range(args.warmup + args.steps):
            per_rank, seconds = pyref.dist_op(pyref.DIST_JOIN, lefts, rights, p1=[0], p2=[0], join_type=orc.INNER, algorithm=orc.HASH,
                                              rows_only=True)
            out_rows = sum(per_rank)
            if step >= args.warmup:
                times.append(seconds)
        ms = 1e3 * float(np.mean(times))
        value = 2.0 * rows * cores / (ms / 1e3)
        sample = ("cylon::DistributedJoin of the reference, %d ranks (threads, in-memory all-to-all) x (%d x %d rows), %d output rows "
                  "per step, whole operator timed" % (cores, rows, rows, out_rows))
        line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": "int64", "data": "synthetic",
                "config": {"workload": "DistributedJoin inner hash join, int64 key + double payload, uniform keys in [0, rows*ranks); CPU "
                                       "sample of %d rows per side per rank, %d ranks (one per host core)" % (rows, cores)},
                "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "reference", "sample": sample},
                "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
        print(json.dumps(line))
        return 0
    times = []
    with mp.get_context("fork").Pool(cores) as pool:
        for step in range(args.warmup + args.steps):
            res = pool.map(_ref_worker, [(rows, r) for r in range(cores)])
            if step >= args.warmup:
                times.append(max(r[0] for r in res))
    ms = 1e3 * float(np.mean(times))
    value = 2.0 * rows * cores / (ms / 1e3)
    kind = "reference" if pyref.available() else "port"
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "int64", "data": "synthetic",
            "config": {"workload": "inner hash join, int64 key + double payload, uniform keys; CPU sample of %d rows per side per rank, "
                                   "%d single-threaded ranks (one per host core), shards pre-partitioned" % (rows, cores)},
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": kind,
                             "sample": "%d ranks x (%d x %d rows) per step, timed join only (max over ranks)" % (cores, rows, rows)},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line))
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--rows", type=int, default=100_000_000, help="rows per side per GPU (BASELINE configs[1]: 100M)")
    ap.add_argument("--cpu-rows", type=int, default=4_000_000, help="rows per side of the CPU baseline sample")
    ap.add_argument("--ref-rows", type=int, default=2_000_000, help="--impl reference: rows per side per rank")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--e2e-lanes", type=int, default=3, help="N=1 only: steps kept in flight by the e2e arm (one host thread + stream each)")
    ap.add_argument("--no-cpu", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)

    import ctypes as C

    import torch

    from cylon_b200 import api, capi

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
        idt = torch.zeros(128, dtype=torch.uint8, device="cuda")
        if rank == 0:
            idt = torch.frombuffer(bytearray(api.nccl_unique_id()), dtype=torch.uint8).cuda()
        dist.broadcast(idt, 0)
        ctx = api.CylonContext(device=local_rank, distributed=True, rank=rank, world_size=world, nccl_id=idt.cpu().numpy().tobytes())
    else:
        ctx = api.CylonContext(device=local_rank)
    ctx.set_stream(torch.cuda.current_stream().cuda_stream)

    rows = args.rows
    key_range = rows * world
    # pinned host copies of both tables (the e2e arm copies from these every step)
    lk, lv = capi.pinned_array(rows, np.int64), capi.pinned_array(rows, np.float64)
    rk, rv = capi.pinned_array(rows, np.int64), capi.pinned_array(rows, np.float64)
    synth_table(rows, key_range, 1, rank, lk, lv)
    synth_table(rows, key_range, 2, rank, rk, rv)

    def barrier():
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    # ---------------------------------------------------------------- device-resident arm (`value`)
    left = api.Table.from_numpy(ctx, [lk, lv], names=["k", "v"])
    right = api.Table.from_numpy(ctx, [rk, rv], names=["k", "v"])
    state = {"rows": 0}

    def step():
        out = left.distributed_join(right, "inner", "hash", [0], [0], "l_", "r_")
        state["rows"] = out.row_count
        out.free()

    sampler = ClockSampler(local_rank)  # samples through warm-up, the timed steps and the e2e steps (all under load)
    sampler.start()
    for _ in range(args.warmup):
        step()
    barrier()
    launches0 = capi.launch_count()
    capi.profile_enable(True)
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    for _ in range(args.steps):
        step()
    ev1.record()
    barrier()
    launches = capi.launch_count() - launches0
    capi.profile_enable(False)
    prof = capi.profile_report()
    t = torch.tensor([ev0.elapsed_time(ev1)], dtype=torch.float64, device="cuda")
    if dist is not None:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_per_step = float(t.item()) / args.steps
    value = 2.0 * rows * world / (ms_per_step / 1e3)
    out_rows = state["rows"]

    # ---------------------------------------------------------------- end-to-end arm (`e2e`): host buffers in and out
    e2e = None
    if not args.no_e2e:
        cap = int(out_rows * 1.05) + 1024
        host_out = [capi.pinned_array(cap, np.int64), capi.pinned_array(cap, np.float64), capi.pinned_array(cap, np.int64),
                    capi.pinned_array(cap, np.float64)]

        def e2e_step(cx=ctx, outs=host_out):
            l = api.Table.from_numpy(cx, [lk, lv], names=["k", "v"])
            r = api.Table.from_numpy(cx, [rk, rv], names=["k", "v"])
            out = l.distributed_join(r, "inner", "hash", [0], [0], "l_", "r_")
            mm = out.row_count
            assert mm <= cap
            for c in range(4):
                capi.check(capi.lib().cylon_table_column_to_host(out._h, C.c_int32(c), C.c_void_p(outs[c].ctypes.data), None, None))
            out.free()
            l.free()
            r.free()
            return mm

        # One GPU: two steps in flight. Two host threads, each with its own CylonContext on its own stream and its own
        # pinned result buffers, take alternate steps, so one step's result download (PCIe up) overlaps the next step's
        # input upload (PCIe down). Every step still uploads all of its inputs and downloads its whole result inside
        # the timed region. With torchrun (N > 1) steps stay strictly one after another on the single NCCL communicator.
        lanes = [(ctx, host_out)]
        lane_streams = []
        if world == 1 and args.steps >= 2:
            import threading
            for _ in range(1, max(1, min(args.e2e_lanes, args.steps))):
                cx = api.CylonContext(device=local_rank)
                lane_streams.append(torch.cuda.Stream())
                cx.set_stream(lane_streams[-1].cuda_stream)
                lanes.append((cx, [capi.pinned_array(cap, a.dtype) for a in host_out]))
        for cx, outs in lanes:
            e2e_step(cx, outs)
        barrier()
        t0 = time.perf_counter()
        if len(lanes) == 1:
            ev0.record()
            for _ in range(args.steps):
                mm = e2e_step()
            ev1.record()
            barrier()
            wall_ms = max(ev0.elapsed_time(ev1), 1e3 * (time.perf_counter() - t0))
        else:
            got = [0] * len(lanes)

            errors = []

            def lane_main(li):
                try:
                    torch.cuda.set_device(local_rank)  # the CUDA current device is per host thread
                    for sidx in range(li, args.steps, len(lanes)):
                        got[li] = e2e_step(*lanes[li])
                except BaseException as exc:  # a failed lane must fail the run, not vanish with its thread
                    errors.append(exc)

            threads = [threading.Thread(target=lane_main, args=(li,)) for li in range(len(lanes))]
            for th in threads:
                th.start()
            for th in threads:
                th.join()
            if errors:
                raise errors[0]
            torch.cuda.synchronize()
            wall_ms = 1e3 * (time.perf_counter() - t0)  # host clock around all K steps: spans both streams
            mm = got[0]
            assert all(g == mm for g in got)
            for cx, _ in lanes[1:]:
                cx.finalize()
        te = torch.tensor([wall_ms], dtype=torch.float64, device="cuda")
        if dist is not None:
            dist.all_reduce(te, op=dist.ReduceOp.MAX)
        e2e_ms = float(te.item()) / args.steps
        e2e = {"value": 2.0 * rows * world / (e2e_ms / 1e3), "unit": UNIT, "h2d_bytes_per_step": int(2 * rows * 16),
               "d2h_bytes_per_step": int(mm * 32), "ms_per_step": e2e_ms, "steps_in_flight": len(lanes),
               "timed_with": "host clock around all steps, device synchronised on both sides"}

    sampler.stop_flag.set()
    if rank != 0:
        ctx.finalize()
        return 0

    # ---------------------------------------------------------------- roofline of the dominant kernel
    peak, peak_src = load_peaks()
    m = out_rows
    dominant, dom_ms, dom_n = None, 0.0, 1
    kernels = {}
    for name, (n_l, ms) in sorted(prof.items()):
        if ms > dom_ms:
            dominant, dom_ms, dom_n = name, ms, n_l
        per_launch_ms = ms / max(1, n_l)
        ab = algorithmic_bytes(name, rows, m)
        kernels[name] = {"launches_per_step": n_l / args.steps, "ms_per_launch": round(per_launch_ms, 4),
                         "share_of_step": round(ms / args.steps / ms_per_step, 4),
                         "algorithmic_GBps": round(ab / 1e9 / (per_launch_ms / 1e3), 1) if per_launch_ms > 0 else None}
    roofline = None
    if dominant is not None:
        per_launch_ms = dom_ms / max(1, dom_n)
        ab = algorithmic_bytes(dominant, rows, m)
        achieved = ab / 1e9 / (per_launch_ms / 1e3)
        # dram__bytes_read+write per launch from the committed `ncu --set full` capture (only valid for the size it was taken at)
        traffic = None
        scope_to_kernel = {"fused_join_scatter": "void fused8_scatter_kernel<2>", "fused_join_probe": "void fused8_join_kernel<2, 2>",
                           "fused_join_scatter_coarse": "void fused8_radix_pass_kernel<2, 0>", "fused_join_scatter_fine": "void fused8_radix_pass_kernel<2, 1>",
                           "fused_join_hist": "fused8_block_hist_kernel"}
        try:
            if rows == 100_000_000:
                with open(os.path.join(ROOT, "profiles", "r01_traffic_100M.json")) as f:
                    traffic = json.load(f).get(scope_to_kernel.get(dominant, ""))
        except Exception:
            traffic = None
        roofline = {"kernel": dominant, "bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                    "traffic": traffic, "peak_source": peak_src, "algorithmic_bytes_per_launch": ab, "ms_per_launch": per_launch_ms}
    # whole local join against its algorithmic bytes ((N_L+N_R)*w + M*(w_L+w_R))
    step_bytes = 2 * rows * 16 + m * 32
    step_roofline = {"algorithmic_bytes": step_bytes, "achieved_GBps": step_bytes / 1e9 / (ms_per_step / 1e3),
                     "frac_of_peak": step_bytes / 1e9 / (ms_per_step / 1e3) / peak}

    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "int64",
            "data": "synthetic",
            "config": {"workload": "DistributedJoin inner hash join, %d rows per side per GPU, int64 key + double payload, uniform keys in [0, rows*N)"
                                   % rows, "rows_per_side_per_gpu": rows, "output_rows_rank0": m, "parallelism": "hash-partitioned, %d GPU(s)" % world,
                       "l2": "inputs (%.1f GB per step) far exceed the 126 MB L2; no flush needed" % (2 * rows * 16 / 1e9)},
            "gpu_launches": launches, "clocks": sampler.summary(), "roofline": roofline, "step_roofline": step_roofline,
            "kernels": kernels}
    if e2e is not None:
        line["e2e"] = e2e
    if not args.no_cpu and world == 1:
        line["cpu_baseline"] = cpu_baseline(np.asarray(lk), np.asarray(lv), np.asarray(rk), np.asarray(rv), min(args.cpu_rows, rows))
    print(json.dumps(line), flush=True)
    ctx.finalize()
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
