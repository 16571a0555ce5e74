#!/usr/bin/env python
"""bench.py -- throughput of the distributed relational hot path on N B200s, with roofline, parity and CPU baseline.

Contract (see DESIGN.md "Measurement"):
  python bench.py --gpus N --steps K --warmup W            (N > 1: launched under torchrun, one rank per GPU)
  python bench.py --impl reference --gpus N --steps K --warmup W
  python bench.py --config {join,sortjoin,groupby,setops,shuffle,sort} ...     (default join = BASELINE.json configs[1])

A "step" is one Distributed* operator over one batch of synthetic input (workloads.py):
  join      cylon::DistributedJoin(inner, HASH), 100M rows per side per GPU, (int64 key, double payload), uniform keys
  sortjoin  cylon::DistributedJoin(left|right|fullouter, SORT) on an (int64, int32) key, Zipf(1.1) left side, 50M rows per side per GPU
  groupby   cylon::DistributedHashGroupBy(sum, count, min, max), 125M rows per GPU, 1M distinct int64 groups
  setops    cylon::DistributedUnion|Intersect|Subtract / DistributedUnique on (int64, string, double) tables, 25M rows per GPU per table
  shuffle   cylon::Shuffle of an (int64, double) table of --shuffle-bytes per GPU (N = 1: the hash partition + split it starts with)
  sort      cylon::DistributedSort on the int64 key of an (int64, double) table, 100M rows per GPU (built in round 2, not yet measured)
`value` = input rows of all ranks / max-over-ranks step time with the inputs already in HBM (CUDA events on the operator's
stream); `e2e` = the same through host buffers (pinned H2D of the inputs and D2H of the whole result inside the timed
region). After the timed steps every rank's result is checked (`parity`): the full-size row count against a closed form of
the seeded generator, and a smaller run of the same operator, rank by rank, against the reference compiled in
oracle/_ref (else the oracle). A mismatch makes the run exit non-zero.
"""
import argparse
import faulthandler
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

# Streams share the device's hardware work queues (8 by default). The e2e arm keeps several steps in flight, each with its
# own stream, second stream and NCCL communicator; two streams aliased onto one queue would order one communicator's
# collective behind another's. Must be set before the CUDA context exists.
os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import workloads  # noqa: E402

UNIT = "rows/s"
NVLINK_PEAK_GBS = 900.0  # per direction per GPU (B200_PROFILING.md: nominal; measured peer copy 770)


# ---- the join benchmark's generator (kept byte for byte from round 1 so the numbers stay comparable) -------------------
def synth_table(rows, key_range, seed, rank, out_keys=None, out_vals=None, chunk=1 << 24):
    """(int64 key uniform in [0, key_range), double payload in [DBL_MIN, 1)); deterministic in (seed, rank, row)."""
    splitmix64 = workloads.splitmix64
    keys = out_keys if out_keys is not None else np.empty(rows, dtype=np.int64)
    vals = out_vals if out_vals is not None else np.empty(rows, dtype=np.float64)
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


def pinned_like(arr):
    from cylon_b200 import capi
    out = capi.pinned_array(len(arr), arr.dtype)
    out[:] = arr
    return out


def pinned_columns(cols):
    """numpy columns / (offsets, bytes) tuples -> the same in page-locked memory"""
    return [(pinned_like(c[0]), pinned_like(c[1])) if isinstance(c, tuple) else pinned_like(c) for c in cols]


def column_bytes(cols):
    return int(sum((c[0].nbytes + c[1].nbytes) if isinstance(c, tuple) else c.nbytes for c in cols))


def to_oracle(cols, names=None):
    from oracle import pyoracle as orc
    out = []
    for i, c in enumerate(cols):
        name = names[i] if names else "c%d" % i
        if isinstance(c, tuple):
            out.append(orc.HCol(orc.STRING, np.asarray(c[1]), np.asarray(c[0]), name=name))
        else:
            out.append(orc.col(np.asarray(c), name=name))
    return out


class Env:
    """process-wide plumbing: rank, device, torch.distributed group, library context"""

    def __init__(self):
        import torch

        from cylon_b200 import api
        self.torch = torch
        self.api = api
        self.rank = int(os.environ.get("RANK", "0"))
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.local_rank = int(os.environ.get("LOCAL_RANK", "0"))
        torch.cuda.set_device(self.local_rank)
        self.dist = None
        if self.world > 1:
            import torch.distributed as dist
            self.dist = dist
            dist.init_process_group("nccl", device_id=torch.device("cuda", self.local_rank))
        self.ctx = self.new_context()
        self.ctx.set_stream(torch.cuda.current_stream().cuda_stream)

    def new_context(self):
        torch, api = self.torch, self.api
        if self.world > 1:
            idt = torch.zeros(128, dtype=torch.uint8, device="cuda")
            if self.rank == 0:
                idt = torch.frombuffer(bytearray(api.nccl_unique_id()), dtype=torch.uint8).cuda()
            self.dist.broadcast(idt, 0)
            ctx = api.CylonContext(device=self.local_rank, distributed=True, rank=self.rank, world_size=self.world,
                                   nccl_id=idt.cpu().numpy().tobytes())
        else:
            ctx = api.CylonContext(device=self.local_rank)
        for kv in filter(None, os.environ.get("CYLON_CONFIG", "").split(";")):  # A/B runs: CYLON_CONFIG="cylon.join.overlap=0"
            k, v = kv.split("=", 1)
            ctx.add_config(k, v)
        return ctx

    def barrier(self):
        self.torch.cuda.synchronize()
        if self.dist is not None:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def allreduce(self, values, op="sum"):
        """list of python ints/floats -> reduced list (int64 / float64 on the device, NCCL)"""
        torch = self.torch
        is_float = any(isinstance(v, float) for v in values)
        t = torch.tensor(values, dtype=torch.float64 if is_float else torch.int64, device="cuda")
        if self.dist is not None:
            ops = {"sum": self.dist.ReduceOp.SUM, "max": self.dist.ReduceOp.MAX, "min": self.dist.ReduceOp.MIN}
            self.dist.all_reduce(t, op=ops[op])
        return t.tolist()

    def gather_objects(self, obj):
        """-> list with every rank's object on rank 0 (None elsewhere)"""
        if self.dist is None:
            return [obj]
        out = [None] * self.world if self.rank == 0 else None
        self.dist.gather_object(obj, out, dst=0)
        return out

    def global_bincount(self, values, domain):
        """int64 tensor[domain]: occurrences over ALL ranks (device; used by the closed-form row counts)"""
        torch = self.torch
        v = torch.from_numpy(np.ascontiguousarray(values)).cuda()
        h = torch.bincount(v, minlength=int(domain))
        del v
        if self.dist is not None:
            self.dist.all_reduce(h)
        return h


def split_checksum(cs):
    return [int(cs >> 32), int(cs & 0xFFFFFFFF)]


# ======================================================================================================== workloads
class Workload:
    name = ""
    metric = ""
    dtype = "int64"
    default_rows = 0

    def __init__(self, args, env):
        self.args, self.env = args, env
        self.rows = args.rows or self.default_rows
        self.rank, self.world = env.rank, env.world
        self.state = {}

    def host_inputs(self):           # -> list of tables, each a list of pinned columns
        raise NotImplementedError

    def names(self, t):
        return None

    def run(self, tables):           # one operator call over device tables -> result table
        raise NotImplementedError

    def input_rows(self):            # input rows of this rank per step
        raise NotImplementedError

    def step_algorithmic_bytes(self, out_rows, out_bytes):
        raise NotImplementedError

    def workload_text(self):
        raise NotImplementedError

    def kernel_bytes(self, kernel, out_rows):
        return 0

    def parity(self, out_rows_full):
        return None

    def cpu_baseline(self):
        return None

    def upload(self, ctx, host_tables):
        return [self.env.api.Table.from_numpy(ctx, t, names=self.names(i)) for i, t in enumerate(host_tables)]

    def result_bytes(self, out):
        total = 0
        for i in range(out.column_count):
            d, nbytes = out.column_desc(i)
            total += nbytes + (4 * (d.length + 1) if d.type in (13, 14) else 0) + ((d.length + 7) // 8 if d.validity else 0)
        return total


class JoinWorkload(Workload):
    name = "join"
    metric = "DistributedJoin input rows/sec"
    default_rows = 100_000_000

    def host_inputs(self):
        from cylon_b200 import capi
        rows, key_range = self.rows, self.rows * self.world
        lk, lv = capi.pinned_array(rows, np.int64), capi.pinned_array(rows, np.float64)
        rk, rv = capi.pinned_array(rows, np.int64), capi.pinned_array(rows, np.float64)
        synth_table(rows, key_range, 1, self.rank, lk, lv)
        synth_table(rows, key_range, 2, self.rank, rk, rv)
        return [[lk, lv], [rk, rv]]

    def names(self, t):
        return ["k", "v"]

    def run(self, tables):
        return tables[0].distributed_join(tables[1], "inner", "hash", [0], [0], "l_", "r_")

    def input_rows(self):
        return 2 * self.rows

    def step_algorithmic_bytes(self, out_rows, out_bytes):
        return 2 * self.rows * 16 + out_rows * 32  # (N_L + N_R) * w + M * (w_L + w_R)

    def workload_text(self):
        return ("DistributedJoin inner hash join, %d rows per side per GPU, int64 key + double payload, uniform keys in [0, rows*N)" % self.rows)

    def kernel_bytes(self, kernel, m):
        n = self.rows
        table = {
            "hash_join_build": n * 8, "hash_join_probe_count": n * 8, "hash_join_probe_write": m * 16, "hash_join_scan": n * 12,
            "gather_fixed": m * 16,
            "fused_join_hist": n * 8,                     # per launch (one table): keys in
            "fused_join_scatter": n * 32,                 # per launch (one table): columns in + packed records out
            "fused_join_scatter_coarse": n * 32,          # per launch: columns in + records ordered by coarse digit out
            "fused_join_scatter_fine": n * 32,            # per launch: records in + records ordered by bin out
            "fused_join_probe": 2 * n * 16 + m * 32,      # both record streams in + the 4 result columns out (= SURVEY 8d local join)
            "partition_scatter_peer": n * 32,             # per launch (one table): rows in + rows out (remote share over NVLink)
            "partition_tile_hist": n * 8, "partition_scatter": n * 32,
        }
        return table.get(kernel, 0)

    def parity(self, out_rows_full):
        env, api = self.env, self.env.api
        from oracle import pyoracle as orc
        from oracle import pyref
        # (1) full size: M = sum over keys of cntL[k] * cntR[k], from the generator's own keys
        key_range = self.rows * self.world
        lk, rk = self.state["host"][0][0], self.state["host"][1][0]
        hl = env.global_bincount(lk, key_range)
        hr = env.global_bincount(rk, key_range)
        expected = int((hl * hr).sum().item())
        del hl, hr
        env.torch.cuda.empty_cache()
        got = env.allreduce([int(out_rows_full)])[0]
        # (2) the same operator on a smaller pair of tables (large enough for the two-pass staged partition), rank by rank
        # against the reference's own DistributedJoin run on `world` ranks
        pr = min(self.rows, self.args.parity_rows or 2_200_000)
        sl = synth_table(pr, pr * self.world, 11, self.rank)
        sr = synth_table(pr, pr * self.world, 12, self.rank)
        L = api.Table.from_numpy(env.ctx, list(sl), names=["k", "v"])
        R = api.Table.from_numpy(env.ctx, list(sr), names=["k", "v"])
        out = L.distributed_join(R, "inner", "hash", [0], [0], "l_", "r_")
        allr = env.gather_objects([out.row_count] + split_checksum(out.checksum()))
        ok, checker = True, None
        if self.rank == 0:
            lefts = [to_oracle(synth_table(pr, pr * self.world, 11, r), ["k", "v"]) for r in range(self.world)]
            rights = [to_oracle(synth_table(pr, pr * self.world, 12, r), ["k", "v"]) for r in range(self.world)]
            if pyref.dist_available():
                exp = pyref.dist_op(pyref.DIST_JOIN, lefts, rights, p1=[0], p2=[0], join_type=orc.INNER, algorithm=orc.HASH)
                checker = "reference (oracle/_ref cylon::DistributedJoin, %d ranks)" % self.world
            else:
                exp = orc.distributed_join(lefts, rights, [0], [0], orc.INNER, orc.HASH, "l_", "r_")
                checker = "oracle (oracle/oracle.c)"
            for r in range(self.world):
                ok = ok and list(allr[r]) == [exp[r][0].length] + split_checksum(orc.table_checksum(exp[r]))
        ok = bool(env.allreduce([1 if ok else 0], "min")[0])
        return {"rows_ok": got == expected, "checksum_ok": ok, "rows": got, "expected_rows": expected,
                "sample": "%d rows per side per rank through the same operator, per-rank row count + order-independent checksum" % pr,
                "checker": checker}

    def cpu_baseline(self):
        from oracle import pyoracle as orc
        from oracle import pyref
        n = min(self.args.cpu_rows, self.rows)
        (lk, lv), (rk, rv) = self.state["host"]
        lk, rk = np.asarray(lk[:n]) % n, np.asarray(rk[:n]) % n  # keep ~1 match per probe row
        lv, rv = np.asarray(lv[:n]), np.asarray(rv[:n])
        t0 = time.perf_counter()
        if pyref.available():
            out = pyref.join_tables([orc.col(lk), orc.col(lv)], [orc.col(rk), orc.col(rv)], [0], [0], orc.INNER, orc.HASH)
            m, kind = out[0].length, "reference"
        else:
            m, _cs = orc.join_i64_f64(lk, lv, rk, rv, orc.INNER, orc.HASH)
            kind = "port"
        dt = time.perf_counter() - t0
        return {"value": 2 * n / dt, "unit": UNIT, "cores": 1, "kind": kind,
                "sample": "inner hash join of %d x %d rows (keys folded into [0, %d)), %d output rows, %.2f s; cylon::join::JoinTables, "
                          "single-threaded as one MPI rank of the reference runs" % (n, n, n, m, dt)}


JT_NAMES = {"left": 1, "right": 2, "fullouter": 3, "inner": 0}


class SortJoinWorkload(Workload):
    name = "sortjoin"
    metric = "DistributedJoin input rows/sec"
    default_rows = 50_000_000

    def host_inputs(self):
        l, r = workloads.sortjoin_tables(self.rows, self.world, self.rank)
        return [pinned_columns(l), pinned_columns(r)]

    def names(self, t):
        return ["k0", "k1", "v"]

    def run(self, tables):
        return tables[0].distributed_join(tables[1], self.args.join_type, "sort", [0, 1], [0, 1], "l_", "r_")

    def input_rows(self):
        return 2 * self.rows

    def step_algorithmic_bytes(self, out_rows, out_bytes):
        return 2 * self.rows * 20 + out_rows * 40

    def workload_text(self):
        return ("DistributedJoin %s SortJoin on an (int64, int32) key + double payload, %d rows per side per GPU, left side Zipf(1.1)-skewed "
                "over the right side's distinct tuples" % (self.args.join_type, self.rows))

    def kernel_bytes(self, kernel, m):
        n = self.rows
        return {"sort_join_merge": 2 * n * 20 + m * 40}.get(kernel, 0)

    def _sample(self, pr, r):
        l, rr = workloads.sortjoin_tables(pr, self.world, r, seed=13)
        keep = np.arange(pr) % 7 != 1  # some left rows lose their partner, some right rows stay unmatched
        return l, [c[keep] for c in rr]

    def parity(self, out_rows_full):
        env, api = self.env, self.env.api
        from oracle import pyoracle as orc
        from oracle import pyref
        jt = JT_NAMES[self.args.join_type]
        domain = self.rows * self.world
        expected = self.rows * self.world  # every left row has exactly one partner
        if jt in (orc.RIGHT, orc.FULL_OUTER):
            h = env.global_bincount(self.state["host"][0][0], domain)
            expected += domain - int((h > 0).sum().item())
            del h
            env.torch.cuda.empty_cache()
        got = env.allreduce([int(out_rows_full)])[0]
        pr = min(self.rows, self.args.parity_rows or 300_000)
        sl, sr = self._sample(pr, self.rank)
        L = api.Table.from_numpy(env.ctx, sl, names=self.names(0))
        R = api.Table.from_numpy(env.ctx, sr, names=self.names(1))
        out = L.distributed_join(R, self.args.join_type, "sort", [0, 1], [0, 1], "l_", "r_")
        allr = env.gather_objects([out.row_count] + split_checksum(out.checksum()))
        ok, checker = True, None
        if self.rank == 0:
            pairs = [self._sample(pr, r) for r in range(self.world)]
            lefts = [to_oracle(p[0], self.names(0)) for p in pairs]
            rights = [to_oracle(p[1], self.names(1)) for p in pairs]
            if pyref.dist_available():
                exp = pyref.dist_op(pyref.DIST_JOIN, lefts, rights, p1=[0, 1], p2=[0, 1], join_type=jt, algorithm=orc.SORT)
                checker = "reference (oracle/_ref cylon::DistributedJoin SORT, %d ranks)" % self.world
            else:
                exp = orc.distributed_join(lefts, rights, [0, 1], [0, 1], jt, orc.SORT, "l_", "r_")
                checker = "oracle (oracle/oracle.c)"
            for r in range(self.world):
                ok = ok and list(allr[r]) == [exp[r][0].length] + split_checksum(orc.table_checksum(exp[r]))
        ok = bool(env.allreduce([1 if ok else 0], "min")[0])
        return {"rows_ok": got == expected, "checksum_ok": ok, "rows": got, "expected_rows": expected,
                "sample": "%d rows per side per rank (1/7 of the right rows removed), per-rank row count + checksum" % pr, "checker": checker}

    def cpu_baseline(self):
        from oracle import pyoracle as orc
        from oracle import pyref
        if not pyref.dist_available():
            return None
        n = min(self.args.cpu_rows, self.rows)
        l, r = workloads.sortjoin_tables(n, 1, 0, seed=13)
        rows, dt = pyref.dist_op(pyref.DIST_JOIN, [to_oracle(l)], [to_oracle(r)], p1=[0, 1], p2=[0, 1], join_type=JT_NAMES[self.args.join_type],
                                 algorithm=orc.SORT, rows_only=True)
        return {"value": 2 * n / dt, "unit": UNIT, "cores": 1, "kind": "reference",
                "sample": "%s SortJoin of %d x %d rows, %d output rows, %.2f s, one rank" % (self.args.join_type, n, n, rows[0], dt)}


class GroupByWorkload(Workload):
    name = "groupby"
    metric = "DistributedHashGroupBy input rows/sec"
    dtype = "f64"
    default_rows = 125_000_000
    OPS = ["sum", "count", "min", "max"]

    def host_inputs(self):
        from cylon_b200 import capi
        k, v = capi.pinned_array(self.rows, np.int64), capi.pinned_array(self.rows, np.float64)
        workloads.groupby_table(self.rows, self.rank, self.args.groups, out_keys=k, out_vals=v)
        return [[k, v]]

    def names(self, t):
        return ["k", "v"]

    def run(self, tables):
        return tables[0].groupby([0], [1] * len(self.OPS), self.OPS, distributed=True)

    def input_rows(self):
        return self.rows

    def step_algorithmic_bytes(self, out_rows, out_bytes):
        return self.rows * 16 + out_rows * 40  # N * w in + G * 40 B out

    def workload_text(self):
        return ("DistributedHashGroupBy sum/count/min/max of a double over %d rows per GPU, %d distinct int64 groups" % (self.rows, self.args.groups))

    def kernel_bytes(self, kernel, g):
        n = self.rows
        return {"groupby_fused_aggregate": n * 16 + g * 40, "groupby_insert": n * 8, "groupby_aggregate": n * 16,
                "fused_join_hist": n * 8, "fused_join_scatter_coarse": n * 32, "fused_join_scatter_fine": n * 32, "fused_join_scatter": n * 32}.get(kernel, 0)

    @staticmethod
    def _check_rank(got, exp):
        """one rank's result (dict of numpy columns) against the expected table: keys/count/min/max exact, sum within 1e-12"""
        if sorted(got.keys()) != sorted(c.name for c in exp):
            return False
        kname = exp[0].name
        og, oe = np.argsort(got[kname], kind="stable"), np.argsort(exp[0].data, kind="stable")
        if len(og) != len(oe) or not np.array_equal(got[kname][og], exp[0].data[oe]):
            return False
        for c in exp[1:]:
            a, b = got[c.name][og], c.data[oe]
            if c.name.startswith("sum"):
                if not np.allclose(a, b, rtol=1e-12, atol=0):  # BASELINE.json: floating-point aggregates within 1e-12 relative
                    return False
            elif not np.array_equal(a, b):
                return False
        return True

    def parity(self, out_rows_full):
        env, api = self.env, self.env.api
        from oracle import pyoracle as orc
        from oracle import pyref
        # each group lives on exactly one rank after the shuffle: total output rows = distinct keys over all ranks
        h = env.global_bincount(np.asarray(self.state["host"][0][0]) // 7919, self.args.groups)
        expected = int((h > 0).sum().item())
        del h
        env.torch.cuda.empty_cache()
        got = env.allreduce([int(out_rows_full)])[0]
        pr = min(self.rows, self.args.parity_rows or 1_000_000)
        groups = max(1, min(self.args.groups, pr // 50))
        mine = workloads.groupby_table(pr, self.rank, groups, seed=14)
        # the checker library names its input columns c0, c1: same names here, so the output names compare too
        out = api.Table.from_numpy(env.ctx, mine, names=["c0", "c1"]).groupby([0], [1] * 4, self.OPS, distributed=True)
        res = {c["name"]: c["data"] for c in out.to_host()}
        allr = env.gather_objects(res)
        ok, checker = True, None
        if self.rank == 0:
            tabs = [to_oracle(workloads.groupby_table(pr, r, groups, seed=14), ["c0", "c1"]) for r in range(self.world)]
            codes = [orc.SUM, orc.COUNT, orc.MIN, orc.MAX]
            if pyref.dist_available():
                exp = pyref.dist_op(pyref.DIST_GROUPBY, tabs, p1=[0], p2=[x for o in codes for x in (1, o)])
                checker = "reference (oracle/_ref cylon::DistributedHashGroupBy, %d ranks)" % self.world
            else:
                exp = orc.distributed_hash_groupby(tabs, [0], [1] * 4, codes)
                checker = "oracle (oracle/oracle.c)"
            for r in range(self.world):
                ok = ok and self._check_rank(allr[r], exp[r])
        ok = bool(env.allreduce([1 if ok else 0], "min")[0])
        return {"rows_ok": got == expected, "checksum_ok": ok, "rows": got, "expected_rows": expected,
                "sample": "%d rows per rank over %d groups: per-rank keys, counts, min, max exact, sums within 1e-12 relative, column names" % (pr, groups),
                "checker": checker}

    def cpu_baseline(self):
        from oracle import pyoracle as orc
        from oracle import pyref
        if not pyref.dist_available():
            return None
        n = min(self.args.cpu_rows * 2, self.rows)
        groups = max(1, min(self.args.groups, n // 125))
        t = to_oracle(workloads.groupby_table(n, 0, groups, seed=14), ["k", "v"])
        codes = [orc.SUM, orc.COUNT, orc.MIN, orc.MAX]
        rows, dt = pyref.dist_op(pyref.DIST_GROUPBY, [t], p1=[0], p2=[x for o in codes for x in (1, o)], rows_only=True)
        return {"value": n / dt, "unit": UNIT, "cores": 1, "kind": "reference",
                "sample": "HashGroupBy sum/count/min/max of %d rows over %d groups, %.2f s, one rank" % (n, rows[0], dt)}


SETOP_NAMES = {"union": "DistributedUnion", "intersect": "DistributedIntersect", "subtract": "DistributedSubtract", "unique": "DistributedUnique"}


class SetOpsWorkload(Workload):
    name = "setops"
    dtype = "u8"
    default_rows = 25_000_000

    @property
    def metric(self):
        return SETOP_NAMES[self.args.setop] + " input rows/sec"

    def _tables(self, rows, rank, seed=5):
        n = 1 if self.args.setop == "unique" else 2
        rng = 50_000_000 if rows >= self.default_rows else max(1000, rows * self.world * 2)
        return [workloads.setop_table(rows, rank, w, seed=seed, int_range=rng) for w in range(n)]

    def host_inputs(self):
        return [pinned_columns(t) for t in self._tables(self.rows, self.rank)]

    def names(self, t):
        return ["a", "w", "f"]

    def run(self, tables):
        if self.args.setop == "unique":
            return tables[0].distributed_unique([0, 1])
        return getattr(tables[0], "distributed_" + self.args.setop)(tables[1])

    def input_rows(self):
        return self.rows * (1 if self.args.setop == "unique" else 2)

    def step_algorithmic_bytes(self, out_rows, out_bytes):
        return sum(column_bytes(t) for t in self.state["host"]) + out_bytes  # every input byte once + every output byte once

    def workload_text(self):
        return ("%s on (int64, string U{4..24} from a 1M-word dictionary, double) tables, %d rows per GPU per table, second table half overlapping"
                % (SETOP_NAMES[self.args.setop], self.rows))

    def parity(self, out_rows_full):
        env, api = self.env, self.env.api
        from oracle import pyoracle as orc
        op = self.args.setop
        host = self.state["host"]
        # closed form: rows are functions of their id, so distinct rows = distinct ids
        domain = int(max(int(np.asarray(t[0]).max()) for t in host)) + 1
        domain = int(env.allreduce([domain], "max")[0])
        hl = env.global_bincount(host[0][0], domain) > 0
        if op == "unique":
            expected = int(hl.sum().item())
        else:
            hr = env.global_bincount(host[1][0], domain) > 0
            expected = int({"union": hl | hr, "intersect": hl & hr, "subtract": hl & ~hr}[op].sum().item())
            del hr
        del hl
        env.torch.cuda.empty_cache()
        got = env.allreduce([int(out_rows_full)])[0]
        pr = min(self.rows, self.args.parity_rows or 200_000)
        mine = self._tables(pr, self.rank, seed=15)
        tabs = [api.Table.from_numpy(env.ctx, t, names=self.names(0)) for t in mine]
        out = self.run(tabs)
        allr = env.gather_objects([out.row_count] + split_checksum(out.checksum()))
        ok = True
        if self.rank == 0:
            per_rank = [self._tables(pr, r, seed=15) for r in range(self.world)]
            lefts = [to_oracle(p[0], self.names(0)) for p in per_rank]
            if op == "unique":
                exp = orc.distributed_unique(lefts, [0, 1])
            else:
                rights = [to_oracle(p[1], self.names(0)) for p in per_rank]
                exp = orc.distributed_set_op(lefts, rights, {"union": orc.UNION, "intersect": orc.INTERSECT, "subtract": orc.SUBTRACT}[op])
            for r in range(self.world):
                ok = ok and list(allr[r]) == [exp[r][0].length] + split_checksum(orc.table_checksum(exp[r]))
        ok = bool(env.allreduce([1 if ok else 0], "min")[0])
        return {"rows_ok": got == expected, "checksum_ok": ok, "rows": got, "expected_rows": expected,
                "sample": "%d rows per rank per table, per-rank row count + checksum" % pr,
                "checker": "oracle (oracle/oracle.c; the reference's cross-table string comparator is broken, SURVEY.md E3: parity unpinned)"}

    def cpu_baseline(self):
        from oracle import pyoracle as orc
        n = min(self.args.cpu_rows, self.rows)
        tabs = [to_oracle(t, self.names(0)) for t in self._tables(n, 0, seed=15)]
        t0 = time.perf_counter()
        if self.args.setop == "unique":
            out = orc.unique(tabs[0], [0, 1])
        else:
            out = orc.set_op(tabs[0], tabs[1], {"union": orc.UNION, "intersect": orc.INTERSECT, "subtract": orc.SUBTRACT}[self.args.setop])
        dt = time.perf_counter() - t0
        return {"value": n * len(tabs) / dt, "unit": UNIT, "cores": 1, "kind": "port",
                "sample": "%s of %d-row tables, %d output rows, %.2f s (oracle port: the reference's string comparator is unsafe to run)"
                          % (self.args.setop, n, out[0].length, dt)}


class ShuffleWorkload(Workload):
    name = "shuffle"
    metric = "Shuffle input rows/sec"

    def __init__(self, args, env):
        super().__init__(args, env)
        self.rows = args.rows or max(1, args.shuffle_bytes // 16)

    def host_inputs(self):
        from cylon_b200 import capi
        k, v = capi.pinned_array(self.rows, np.int64), capi.pinned_array(self.rows, np.float64)
        synth_table(self.rows, self.rows * self.world, 7, self.rank, k, v)
        return [[k, v]]

    def names(self, t):
        return ["k", "v"]

    def run(self, tables):
        if self.world == 1:  # Shuffle is the identity on one rank (table.cpp:910-921): time its first half, partition + split
            parts = tables[0].hash_partition([0], self.args.partitions)
            self.state["parts"] = parts[1:]
            return parts[0]
        return tables[0].shuffle([0])

    def input_rows(self):
        return self.rows

    def step_algorithmic_bytes(self, out_rows, out_bytes):
        return 2 * self.rows * 16

    def workload_text(self):
        if self.world == 1:
            return "HashPartition of a %d-row (int64, double) table into %d partitions (the local half of Shuffle)" % (self.rows, self.args.partitions)
        return "Shuffle of a %d-row (int64, double) table per GPU (%.1f MB per GPU)" % (self.rows, self.rows * 16 / 1e6)

    def kernel_bytes(self, kernel, m):
        return {"partition_scatter_peer": self.rows * 32, "partition_tile_hist": self.rows * 8, "partition_scatter": self.rows * 32}.get(kernel, 0)

    def parity(self, out_rows_full):
        env, api = self.env, self.env.api
        t = api.Table.from_numpy(env.ctx, self.state["host"][0], names=["k", "v"])
        cs_in = t.checksum()
        if self.world == 1:
            P = self.args.partitions
            parts = t.hash_partition([0], P)
            rows = sum(p.row_count for p in parts)
            cs_out = sum(p.checksum() for p in parts) & 0xFFFFFFFFFFFFFFFF
            placed = all(bool(np.all(np.asarray(p.to_host()[0]["data"]).astype(np.uint64).astype(np.uint32) % np.uint32(P) == np.uint32(i)))
                         for i, p in enumerate(parts) if p.row_count)
            return {"rows_ok": rows == self.rows, "checksum_ok": cs_out == cs_in and placed, "rows": rows, "expected_rows": self.rows,
                    "sample": "full size: rows and order-independent checksum conserved, every row in partition (uint32)key mod P",
                    "checker": "closed form (arrow_partition_kernels.cpp:84-87)"}
        out = t.shuffle([0])
        keys = np.asarray(out.to_host()[0]["data"]) if out.row_count else np.zeros(0, np.int64)
        placed = bool(np.all(keys.astype(np.uint64).astype(np.uint32) % np.uint32(self.world) == np.uint32(self.rank))) if len(keys) else True
        tot = env.allreduce([out.row_count] + split_checksum(out.checksum()) + split_checksum(cs_in) + [1 if placed else 0])
        cs_o = ((tot[1] << 32) + tot[2]) & 0xFFFFFFFFFFFFFFFF  # sums of 32-bit halves, compared modulo 2^64
        cs_i = ((tot[3] << 32) + tot[4]) & 0xFFFFFFFFFFFFFFFF
        return {"rows_ok": tot[0] == self.rows * self.world, "checksum_ok": cs_o == cs_i and tot[5] == self.world, "rows": tot[0],
                "expected_rows": self.rows * self.world,
                "sample": "full size: global rows and checksum conserved, every received row satisfies (uint32)key mod world == rank",
                "checker": "closed form (arrow_partition_kernels.cpp:84-87)"}

    def cpu_baseline(self):
        from oracle import pyref
        if not pyref.dist_available():
            return None
        n = min(self.args.cpu_rows * 2, self.rows)
        world = max(2, self.world)
        tabs = [to_oracle(synth_table(n, n * world, 7, r), ["k", "v"]) for r in range(world)]
        rows, dt = pyref.dist_op(pyref.DIST_SHUFFLE, tabs, p1=[0], rows_only=True)
        return {"value": n * world / dt, "unit": UNIT, "cores": world, "kind": "reference",
                "sample": "cylon::Shuffle of %d rows per rank on %d ranks (threads, in-memory all-to-all), %.2f s" % (n, world, dt)}


class SortWorkload(Workload):
    """cylon::DistributedSort (table.cpp:338-382): range partition on the key -> shuffle -> local sort."""
    name = "sort"
    metric = "DistributedSort input rows/sec"
    default_rows = 100_000_000

    def host_inputs(self):
        from cylon_b200 import capi
        k, v = capi.pinned_array(self.rows, np.int64), capi.pinned_array(self.rows, np.float64)
        synth_table(self.rows, self.rows * self.world, 9, self.rank, k, v)
        return [[k, v]]

    def names(self, t):
        return ["k", "v"]

    def run(self, tables):
        return tables[0].distributed_sort([0], True)

    def input_rows(self):
        return self.rows

    def step_algorithmic_bytes(self, out_rows, out_bytes):
        return self.rows * 16 + out_rows * 16  # every row in once, every row out once

    def workload_text(self):
        return ("DistributedSort ascending on the int64 key of a %d-row (int64, double) table per GPU, uniform keys in [0, rows*N), "
                "default SortOptions (1 %% sample, 16 bins per rank)" % self.rows)

    def kernel_bytes(self, kernel, m):
        # sort_indices: the whole argsort (key in, row order out; its radix passes move 12 B per row each way on top);
        # gather_fixed: per column, row order in + value in + value out
        return {"sort_indices": self.rows * 16, "gather_fixed": m * 24, "partition_scatter": self.rows * 32}.get(kernel, 0)

    def parity(self, out_rows_full):
        env, api = self.env, self.env.api
        from oracle import pyoracle as orc
        from oracle import pyref
        # (1) full size: rows conserved, every rank's keys ascending, and rank r's largest key <= rank r+1's smallest
        t = api.Table.from_numpy(env.ctx, self.state["host"][0], names=["k", "v"])
        cs_in = t.checksum()
        out = t.distributed_sort([0], True)
        keys = np.asarray(out.to_host()[0]["data"]) if out.row_count else np.zeros(0, np.int64)
        ascending = bool(np.all(keys[:-1] <= keys[1:])) if len(keys) > 1 else True
        bounds = env.gather_objects([int(keys[0]), int(keys[-1])] if len(keys) else None)
        if self.rank == 0:
            held = [b for b in bounds if b is not None]
            ascending = ascending and all(a[1] <= b[0] for a, b in zip(held, held[1:]))
        tot = env.allreduce([out.row_count] + split_checksum(out.checksum()) + split_checksum(cs_in) + [1 if ascending else 0])
        cs_o = ((tot[1] << 32) + tot[2]) & 0xFFFFFFFFFFFFFFFF
        cs_i = ((tot[3] << 32) + tot[4]) & 0xFFFFFFFFFFFFFFFF
        rows_ok = tot[0] == self.rows * self.world and cs_o == cs_i and tot[5] == self.world
        del out, t, keys
        # (2) a smaller table per rank with the whole column as the sample (then the reference does not sample at random either):
        # per-rank row count and checksum against the reference's own DistributedSort on `world` ranks
        pr = min(self.rows, self.args.parity_rows or 1_000_000)
        nb = 16 * self.world
        sk, sv = synth_table(pr, pr * self.world, 19, self.rank)
        S = api.Table.from_numpy(env.ctx, [sk, sv], names=["k", "v"])
        so = S.distributed_sort([0], True, nb, pr)
        sorted_here = bool(np.all(np.diff(np.asarray(so.to_host()[0]["data"])) >= 0)) if so.row_count > 1 else True
        allr = env.gather_objects([so.row_count, sorted_here] + split_checksum(so.checksum()))
        ok, checker = True, None
        if self.rank == 0:
            tabs = [to_oracle(synth_table(pr, pr * self.world, 19, r), ["k", "v"]) for r in range(self.world)]
            if pyref.dist_available():
                exp = pyref.dist_op(pyref.DIST_SORT, tabs, p1=[0], p2=[1], join_type=nb, algorithm=-1)
                checker = "reference (oracle/_ref cylon::DistributedSort, %d ranks)" % self.world
            else:
                exp = orc.distributed_sort(tabs, [0], [True], pr, nb)
                checker = "oracle (oracle/pyoracle.py)"
            for r in range(self.world):
                ok = ok and list(allr[r]) == [exp[r][0].length, True] + split_checksum(orc.table_checksum(exp[r]))
        ok = bool(env.allreduce([1 if ok else 0], "min")[0])
        return {"rows_ok": rows_ok, "checksum_ok": ok, "rows": tot[0], "expected_rows": self.rows * self.world,
                "sample": "full size: rows and checksum conserved, keys ascending within and across ranks; %d rows per rank with the whole "
                          "column as the sample: per-rank row count + checksum" % pr,
                "checker": checker}

    def cpu_baseline(self):
        from oracle import pyref
        if not pyref.dist_available():
            return None
        n = min(self.args.cpu_rows * 2, self.rows)
        k, v = self.state["host"][0]
        tab = to_oracle((np.asarray(k[:n]), np.asarray(v[:n])), ["k", "v"])
        _rows, dt = pyref.dist_op(pyref.DIST_SORT, [tab], p1=[0], p2=[1], rows_only=True)
        return {"value": n / dt, "unit": UNIT, "cores": 1, "kind": "reference",
                "sample": "cylon::DistributedSort (= Sort on one rank: std::sort of row indices + arrow Take) of %d rows, %.2f s" % (n, dt)}


WORKLOADS = {w.name: w for w in (JoinWorkload, SortJoinWorkload, GroupByWorkload, SetOpsWorkload, ShuffleWorkload, SortWorkload)}


# ======================================================================================================== reference arm
def run_reference(args):
    """--impl reference: the reference's own Distributed* operator on all host cores -- one rank per core (how the reference
    scales: MPI ranks; here threads of one process with an in-memory channel, oracle/ref_shim), whole operator timed."""
    if int(os.environ.get("RANK", "0")) != 0:
        return 0
    from oracle import pyoracle as orc
    from oracle import pyref
    cores = os.cpu_count() or 1
    rows = args.ref_rows
    cfg = args.config
    have_ref = pyref.dist_available()
    if cfg == "join":
        lefts = [to_oracle(synth_table(rows, rows * cores, 1, r), ["k", "v"]) for r in range(cores)]
        rights = [to_oracle(synth_table(rows, rows * cores, 2, r), ["k", "v"]) for r in range(cores)]
        call = lambda: pyref.dist_op(pyref.DIST_JOIN, lefts, rights, p1=[0], p2=[0], join_type=orc.INNER, algorithm=orc.HASH, rows_only=True)  # noqa: E731
        fallback = lambda: orc.distributed_join(lefts, rights, [0], [0], orc.INNER, orc.HASH)  # noqa: E731
        per_rank_rows, metric = 2 * rows, JoinWorkload.metric
        text = "DistributedJoin inner hash join, int64 key + double payload, uniform keys in [0, rows*ranks)"
    elif cfg == "sortjoin":
        jt = JT_NAMES[args.join_type]
        pairs = [workloads.sortjoin_tables(rows, cores, r) for r in range(cores)]
        lefts, rights = [to_oracle(p[0]) for p in pairs], [to_oracle(p[1]) for p in pairs]
        call = lambda: pyref.dist_op(pyref.DIST_JOIN, lefts, rights, p1=[0, 1], p2=[0, 1], join_type=jt, algorithm=orc.SORT, rows_only=True)  # noqa: E731
        fallback = lambda: orc.distributed_join(lefts, rights, [0, 1], [0, 1], jt, orc.SORT)  # noqa: E731
        per_rank_rows, metric = 2 * rows, SortJoinWorkload.metric
        text = "DistributedJoin %s SortJoin, (int64, int32) key, Zipf(1.1) left side" % args.join_type
    elif cfg == "groupby":
        tabs = [to_oracle(workloads.groupby_table(rows, r, args.groups), ["k", "v"]) for r in range(cores)]
        codes = [orc.SUM, orc.COUNT, orc.MIN, orc.MAX]
        call = lambda: pyref.dist_op(pyref.DIST_GROUPBY, tabs, p1=[0], p2=[x for o in codes for x in (1, o)], rows_only=True)  # noqa: E731
        fallback = lambda: orc.distributed_hash_groupby(tabs, [0], [1] * 4, codes)  # noqa: E731
        per_rank_rows, metric = rows, GroupByWorkload.metric
        text = "DistributedHashGroupBy sum/count/min/max, %d groups" % args.groups
    elif cfg == "shuffle":
        tabs = [to_oracle(synth_table(rows, rows * cores, 7, r), ["k", "v"]) for r in range(cores)]
        call = lambda: pyref.dist_op(pyref.DIST_SHUFFLE, tabs, p1=[0], rows_only=True)  # noqa: E731
        fallback = lambda: orc.shuffle(tabs, [0])  # noqa: E731
        per_rank_rows, metric = rows, ShuffleWorkload.metric
        text = "Shuffle of an (int64, double) table"
    elif cfg == "sort":  # whole column as the sample: the reference's own choice of "no sampling" (arrow_partition_kernels.cpp:375-381)
        tabs = [to_oracle(synth_table(rows, rows * cores, 9, r), ["k", "v"]) for r in range(cores)]
        call = lambda: pyref.dist_op(pyref.DIST_SORT, tabs, p1=[0], p2=[1], join_type=0, algorithm=-1, rows_only=True)  # noqa: E731
        fallback = lambda: orc.distributed_sort(tabs, [0], [True], rows, 0)  # noqa: E731
        per_rank_rows, metric = rows, SortWorkload.metric
        text = "DistributedSort ascending on the int64 key of an (int64, double) table"
    else:  # setops: the reference's cross-table string comparator reads out of bounds -- the oracle port is timed
        have_ref = False
        op = args.setop
        per = [[workloads.setop_table(rows, r, w) for w in range(1 if op == "unique" else 2)] for r in range(cores)]
        lefts = [to_oracle(p[0]) for p in per]
        if op == "unique":
            fallback = lambda: orc.distributed_unique(lefts, [0, 1])  # noqa: E731
        else:
            rights = [to_oracle(p[1]) for p in per]
            code = {"union": orc.UNION, "intersect": orc.INTERSECT, "subtract": orc.SUBTRACT}[op]
            fallback = lambda: orc.distributed_set_op(lefts, rights, code)  # noqa: E731
        call = None
        per_rank_rows = rows * (1 if op == "unique" else 2)
        metric = SETOP_NAMES[op] + " input rows/sec"
        text = "%s on (int64, string, double) tables" % SETOP_NAMES[op]
    kind = "reference" if have_ref else "port"
    times, out_rows = [], 0
    for step in range(args.warmup + args.steps):
        if have_ref:
            per_rank, seconds = call()
            out_rows = sum(per_rank)
        else:
            t0 = time.perf_counter()
            res = fallback()
            seconds = time.perf_counter() - t0
            out_rows = sum(t[0].length for t in res)
        if step >= args.warmup:
            times.append(seconds)
    ms = 1e3 * float(np.mean(times))
    used = cores if have_ref else 1
    value = float(per_rank_rows) * cores / (ms / 1e3)
    sample = ("%s of the %s, %d ranks (%s) x %d input rows per rank, %d output rows per step, whole operator timed"
              % (text, "reference (oracle/_ref)" if have_ref else "oracle port", cores,
                 "threads, in-memory all-to-all" if have_ref else "simulated in one thread", per_rank_rows, out_rows))
    line = {"impl": "reference", "metric": metric, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": WORKLOADS[cfg].dtype, "data": "synthetic",
            "config": {"workload": "%s; CPU sample of %d rows per table per rank, %d ranks (one per host core)" % (text, rows, cores)},
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": used, "kind": kind, "sample": sample},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line))
    return 0


# ======================================================================================================== GPU arm
def download_result(out, host_cols):
    """device -> host copy of every buffer of the result table into preallocated pinned buffers (data, offsets)"""
    import ctypes as C

    from cylon_b200 import capi
    total = 0
    for c in range(out.column_count):
        d, nbytes = out.column_desc(c)
        data, offs = host_cols[c]
        assert nbytes <= data.nbytes, "result column %d: %d bytes, host buffer has %d" % (c, nbytes, data.nbytes)
        capi.check(capi.lib().cylon_table_column_to_host(out._h, C.c_int32(c), C.c_void_p(data.ctypes.data),
                                                         C.c_void_p(offs.ctypes.data) if offs is not None else None, None))
        total += nbytes + (4 * (d.length + 1) if offs is not None else 0)
    return total


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", default="join", choices=sorted(WORKLOADS))
    ap.add_argument("--rows", type=int, default=0, help="rows per table per GPU (default: the BASELINE.json size of the config)")
    ap.add_argument("--join-type", default="left", choices=["left", "right", "fullouter", "inner"], help="sortjoin")
    ap.add_argument("--setop", default="union", choices=["union", "intersect", "subtract", "unique"], help="setops")
    ap.add_argument("--groups", type=int, default=1_000_000, help="groupby: distinct keys")
    ap.add_argument("--shuffle-bytes", type=int, default=1 << 30, help="shuffle: table bytes per GPU")
    ap.add_argument("--partitions", type=int, default=8, help="shuffle at N=1: number of partitions")
    ap.add_argument("--cpu-rows", type=int, default=4_000_000, help="rows per side of the CPU baseline sample")
    ap.add_argument("--ref-rows", type=int, default=2_000_000, help="--impl reference: rows per table per rank")
    ap.add_argument("--parity-rows", type=int, default=0, help="rows per rank of the parity sample (0 = per-config default)")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--e2e-lanes", type=int, default=3, help="steps kept in flight by the e2e arm (one host thread + stream each)")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-parity", action="store_true")
    args = ap.parse_args()
    # a stalled run (a rank that died, a collective left waiting) prints every thread's stack and exits instead of hanging
    faulthandler.dump_traceback_later(int(os.environ.get("BENCH_WATCHDOG_S", "1500")), exit=True)
    if args.impl == "reference":
        return run_reference(args)

    from cylon_b200 import capi

    for kv in filter(None, os.environ.get("CYLON_TUNE", "").split(",")):  # e.g. CYLON_TUNE=1=0,2=1 (cyl_tune_set: A/B timing of equivalent kernels)
        k, v = kv.split("=")
        capi.check(capi.lib().cyl_tune_set(int(k), int(v)))
    env = Env()
    torch, rank, world = env.torch, env.rank, env.world
    wl = WORKLOADS[args.config](args, env)
    host = wl.host_inputs()
    wl.state["host"] = host

    def free_parts():
        for p in wl.state.pop("parts", None) or []:
            p.free()

    # ---------------------------------------------------------------- device-resident arm (`value`)
    tables = wl.upload(env.ctx, host)
    info = {"rows": 0, "bytes": 0}

    def step():
        out = wl.run(tables)
        info["rows"] = out.row_count
        if not info["bytes"]:
            info["bytes"] = wl.result_bytes(out)
        out.free()
        free_parts()

    sampler = ClockSampler(env.local_rank)  # samples through warm-up, the timed steps and the e2e steps (all under load)
    sampler.start()
    for _ in range(args.warmup):
        step()
    env.barrier()
    launches0 = capi.launch_count()
    capi.profile_enable(True)
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    for _ in range(args.steps):
        step()
    ev1.record()
    env.barrier()
    launches = capi.launch_count() - launches0
    capi.profile_enable(False)
    prof = capi.profile_report()
    ms_per_step = env.allreduce([float(ev0.elapsed_time(ev1))], "max")[0] / args.steps
    value = float(wl.input_rows()) * world / (ms_per_step / 1e3)
    out_rows, out_bytes = info["rows"], info["bytes"]

    # ---------------------------------------------------------------- end-to-end arm (`e2e`): host buffers in and out
    e2e = None
    if not args.no_e2e:
        probe = wl.run(tables)  # result shape -> pinned buffers with slack (row counts differ a little between ranks' steps)
        shapes = []
        for c in range(probe.column_count):
            d, nbytes = probe.column_desc(c)
            shapes.append((int(nbytes * 1.05) + 4096, int(d.length * 1.05) + 1024 if d.type in (13, 14) else 0))
        probe.free()
        free_parts()

        def new_out_bufs():
            return [(capi.pinned_array(nb, np.uint8), capi.pinned_array(no + 1, np.int32) if no else None) for nb, no in shapes]

        d2h_bytes = [0]

        turn = {"next": 0, "failed": False}
        turn_cv = threading.Condition()

        def e2e_step(cx, outs, step_no=None):
            tabs = wl.upload(cx, host)
            # The operators of the steps in flight run one at a time, in step order, on every rank: each lane has its own
            # communicator, and two ranks that entered different lanes' collectives first would wait for each other for
            # good. Uploads and downloads of the neighbouring steps are what overlaps with an operator.
            if step_no is not None:
                with turn_cv:
                    turn_cv.wait_for(lambda: turn["next"] == step_no or turn["failed"])
                    if turn["failed"]:
                        raise RuntimeError("another lane failed")
            try:
                out = wl.run(tabs)
            except BaseException:
                with turn_cv:
                    turn["failed"] = True
                    turn_cv.notify_all()
                raise
            if step_no is not None:
                capi.sync(cx._stream)  # this lane's last collective has run before the next lane issues its first
                with turn_cv:
                    turn["next"] = step_no + 1
                    turn_cv.notify_all()
            mm = out.row_count
            d2h_bytes[0] = download_result(out, outs)
            out.free()
            free_parts()
            for t in tabs:
                t.free()
            return mm

        # Several steps in flight: host threads, each with its own CylonContext (at N > 1 its own NCCL communicator) on its
        # own stream and its own pinned result buffers, take alternate steps, so one step's result download (PCIe up)
        # overlaps the next step's input upload (PCIe down) and the operator of a third. Every step still uploads all of
        # its inputs and downloads its whole result inside the timed region. Lane i of every rank runs the same sequence
        # of steps on communicator i.
        nlanes = max(1, min(args.e2e_lanes, args.steps))
        if args.config == "shuffle" and world == 1:
            nlanes = 1  # wl.state["parts"] is per process
        lanes = [(env.ctx, new_out_bufs())]
        lane_streams = []
        for _ in range(1, nlanes):
            cx = env.new_context()
            lane_streams.append(torch.cuda.Stream())
            cx.set_stream(lane_streams[-1].cuda_stream)
            lanes.append((cx, new_out_bufs()))
        for cx, outs in lanes:
            e2e_step(cx, outs)
        env.barrier()
        t0 = time.perf_counter()
        errors = []

        def lane_main(li):
            try:
                torch.cuda.set_device(env.local_rank)  # the CUDA current device is per host thread
                for _s in range(li, args.steps, len(lanes)):
                    e2e_step(*lanes[li], step_no=_s)
            except BaseException as exc:  # a failed lane must fail the run, not vanish with its thread
                errors.append(exc)
                with turn_cv:
                    turn["failed"] = True
                    turn_cv.notify_all()

        if len(lanes) == 1:
            lane_main(0)
        else:
            threads = [threading.Thread(target=lane_main, args=(li,)) for li in range(len(lanes))]
            for th in threads:
                th.start()
            for th in threads:
                th.join()
        if errors:
            raise errors[0]
        torch.cuda.synchronize()
        wall_ms = 1e3 * (time.perf_counter() - t0)  # host clock around all K steps: spans every lane's stream
        env.barrier()
        for cx, _ in lanes[1:]:
            cx.finalize()
        e2e_ms = env.allreduce([float(wall_ms)], "max")[0] / args.steps
        h2d = sum(column_bytes(t) for t in host)
        e2e = {"value": float(wl.input_rows()) * world / (e2e_ms / 1e3), "unit": UNIT, "h2d_bytes_per_step": int(h2d),
               "d2h_bytes_per_step": int(d2h_bytes[0]), "ms_per_step": e2e_ms, "steps_in_flight": len(lanes),
               "pcie_GBps_per_rank": round((h2d + d2h_bytes[0]) / 1e9 / (e2e_ms / 1e3), 2),
               "timed_with": "host clock around all steps, device synchronised on both sides, max over ranks"}
    sampler.stop_flag.set()

    # ---------------------------------------------------------------- parity (after every timed region)
    parity = None
    if not args.no_parity:
        for t in tables:
            t.free()
        tables = []
        parity = wl.parity(out_rows)
    parity_bad = parity is not None and not (parity["rows_ok"] and parity["checksum_ok"])

    if rank != 0:
        env.ctx.finalize()
        if world > 1:
            env.dist.destroy_process_group()
        return 3 if parity_bad else 0

    # ---------------------------------------------------------------- roofline of the dominant kernel (rank 0's launches)
    peak, peak_src = load_peaks()
    m = out_rows
    dominant, dom_ms, dom_n = None, 0.0, 1
    kernels = {}
    for name, (n_l, ms) in sorted(prof.items()):
        if ms > dom_ms:
            dominant, dom_ms, dom_n = name, ms, n_l
        per_launch_ms = ms / max(1, n_l)
        ab = wl.kernel_bytes(name, m)
        kernels[name] = {"launches_per_step": n_l / args.steps, "ms_per_launch": round(per_launch_ms, 4),
                         "share_of_step": round(ms / args.steps / ms_per_step, 4),
                         "algorithmic_GBps": round(ab / 1e9 / (per_launch_ms / 1e3), 1) if per_launch_ms > 0 and ab else None}
    roofline = None
    if dominant is not None:
        per_launch_ms = dom_ms / max(1, dom_n)
        ab = wl.kernel_bytes(dominant, m)
        if dominant == "partition_scatter_peer" and world > 1:
            # NVLink-bound: bytes that leave this GPU per launch = (P-1)/P of the table, against 900 GB/s per direction
            out_b = wl.rows * 16 * (world - 1) / world
            ach = out_b / 1e9 / (per_launch_ms / 1e3)
            roofline = {"kernel": dominant, "bound": "nvlink", "achieved": ach, "peak": NVLINK_PEAK_GBS, "unit": "GB/s", "frac": ach / NVLINK_PEAK_GBS,
                        "traffic": None, "peak_source": "nominal NVLink 5 per direction per GPU (measured peer copy: 770 GB/s, B200_PROFILING.md)",
                        "algorithmic_bytes_per_launch": out_b, "ms_per_launch": per_launch_ms}
        else:
            achieved = ab / 1e9 / (per_launch_ms / 1e3)
            traffic = None
            try:  # dram__bytes_read+write per launch from the committed `ncu --set full` capture (valid for the size it was taken at)
                with open(os.path.join(ROOT, "profiles", "r02_traffic.json")) as f:
                    traffic = json.load(f).get("%s@%d" % (dominant, wl.rows))
            except Exception:
                traffic = None
            roofline = {"kernel": dominant, "bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                        "traffic": traffic, "peak_source": peak_src, "algorithmic_bytes_per_launch": ab, "ms_per_launch": per_launch_ms}
    step_bytes = wl.step_algorithmic_bytes(m, out_bytes)
    step_roofline = {"algorithmic_bytes": step_bytes, "achieved_GBps": step_bytes / 1e9 / (ms_per_step / 1e3),
                     "frac_of_peak": step_bytes / 1e9 / (ms_per_step / 1e3) / peak,
                     "note": "whole operator on rank 0's shard against the HBM peak (SURVEY.md 8d bytes); at N > 1 the step also holds the NVLink exchange"}

    line = {"metric": wl.metric, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": wl.dtype,
            "data": "synthetic",
            "config": {"workload": wl.workload_text(), "rows_per_table_per_gpu": wl.rows, "output_rows_rank0": m,
                       "parallelism": "hash-partitioned, %d GPU(s)" % world,
                       "l2": "inputs (%.2f GB per step per GPU) exceed the 126 MB L2; no flush needed" % (sum(column_bytes(t) for t in host) / 1e9)},
            "gpu_launches": launches, "clocks": sampler.summary(), "roofline": roofline, "step_roofline": step_roofline,
            "kernels": kernels}
    if e2e is not None:
        line["e2e"] = e2e
    if parity is not None:
        line["parity"] = parity
    if not args.no_cpu and world == 1:
        cb = wl.cpu_baseline()
        if cb is not None:
            line["cpu_baseline"] = cb
    print(json.dumps(line), flush=True)
    env.ctx.finalize()
    if world > 1:
        env.dist.destroy_process_group()
    if parity_bad:
        print("PARITY MISMATCH: %r" % (parity,), file=sys.stderr)
        return 3
    return 0


if __name__ == "__main__":
    sys.exit(main())
