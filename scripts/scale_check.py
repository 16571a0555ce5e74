#!/usr/bin/env python
"""Full-size single-GPU runs of the other BASELINE.json configs (3: outer SortJoin on a skewed 2-column key, 4: hash
GroupBy sum/count/min/max, 5a: set operations + unique with a string column), checked through size-independent
properties and timed with wall clock around the public API. Prints one JSON line per operator.
Usage: python scripts/scale_check.py [--scale 1.0]"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from cylon_b200 import api, capi  # noqa: E402


def timed(label, fn, **extra):
    capi.sync()
    t0 = time.perf_counter()
    out = fn()
    capi.sync()
    dt = time.perf_counter() - t0
    rec = {"op": label, "ms": round(dt * 1e3, 2)}
    rec.update(extra)
    return out, rec


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--scale", type=float, default=1.0)
    ap.add_argument("--only", default="", help="comma list of configs to run: 3,4,5")
    args = ap.parse_args()
    ctx = api.CylonContext(device=0)
    rng = np.random.default_rng(1)

    only = set(args.only.split(",")) if args.only else {"3", "4", "5"}
    if "3" in only:
        config3(ctx, rng, args)
    if "4" in only:
        config4(ctx, rng, args)
    if "5" in only:
        config5(ctx, rng, args)
    ctx.finalize()


def config3(ctx, rng, args):
    # ---------------- config 3: left/right/full-outer SortJoin, key (int64, int32), Zipf(1.1) left side
    n = int(50_000_000 * args.scale)
    rk0 = np.arange(n, dtype=np.int64)
    rng.shuffle(rk0)
    z = rng.zipf(1.1, n)
    lk0 = ((z - 1) % n).astype(np.int64)                      # rank-1 key ~ 12 % of the rows
    left = api.Table.from_numpy(ctx, [lk0, (lk0 % 65536).astype(np.int32), rng.random(n)], names=["k0", "k1", "v"])
    right = api.Table.from_numpy(ctx, [rk0, (rk0 % 65536).astype(np.int32), rng.random(n)], names=["k0", "k1", "w"])
    distinct_left = len(np.unique(lk0))
    for jt in ("left", "right", "outer"):
        for rep in range(2):
            out, rec = timed("sort_join_" + jt, lambda: left.join(right, jt, "sort", [0, 1], [0, 1]), rows_per_side=n)
        m = out.row_count
        # every left row matches exactly one right row (right keys are distinct and cover the left domain)
        expect = {"left": n, "right": n + (n - distinct_left), "outer": n + (n - distinct_left)}[jt]
        rec.update(out_rows=m, expected_rows=expect, ok=(m == expect), input_rows_per_s=round(2 * n / (rec["ms"] / 1e3)))
        print(json.dumps(rec), flush=True)
        out.free()
    left.free()
    right.free()



def config4(ctx, rng, args):
    # ---------------- config 4: hash GroupBy sum/count/min/max, 125M rows, 1M groups
    n = int(125_000_000 * args.scale)
    groups = 1_000_000
    g = rng.integers(0, groups, n).astype(np.int64) * 7919
    v = np.maximum(rng.random(n), np.finfo(np.float64).tiny)
    t = api.Table.from_numpy(ctx, [g, v], names=["g", "v"])
    for rep in range(2):
        capi.profile_enable(rep == 1)
        out, rec = timed("hash_groupby_sum_count_min_max", lambda: t.groupby(0, [1, 1, 1, 1], ["sum", "count", "min", "max"]), rows=n)
    capi.profile_enable(False)
    rec["kernels_ms"] = {k: round(v[1], 3) for k, v in capi.profile_report().items()}
    host = out.to_host()
    ok = (out.row_count == len(np.unique(g)) and int(host[2]["data"].sum()) == n
          and abs(float(host[1]["data"].sum()) - float(v.sum())) <= 1e-9 * float(v.sum())
          and float(host[3]["data"].min()) == float(v.min()) and float(host[4]["data"].max()) == float(v.max())
          and out.column_names == ["g", "sum_v", "count_v", "min_v", "max_v"])
    rec.update(groups=out.row_count, ok=bool(ok), rows_per_s=round(n / (rec["ms"] / 1e3)))
    print(json.dumps(rec), flush=True)
    out.free()
    t.free()



def config5(ctx, rng, args):
    # ---------------- config 5a: union / intersect / subtract / unique, 25M rows, (int64, string, double) rows
    n = int(25_000_000 * args.scale)
    words = np.array([("w%07d" % i)[: 4 + i % 21].encode() for i in range(1_000_000)], dtype=object)

    def table(keys):
        w = words[keys % 1_000_000]
        lens = np.fromiter((len(x) for x in w), dtype=np.int32, count=len(w))
        offs = np.zeros(len(w) + 1, dtype=np.int32)
        np.cumsum(lens, out=offs[1:])
        data = np.frombuffer(b"".join(w.tolist()), dtype=np.uint8)
        return api.Table.from_numpy(ctx, [keys, (offs, data), keys.astype(np.float64) * 0.5], names=["k", "s", "f"])

    ka = rng.integers(0, 50_000_000, n).astype(np.int64)
    kb = np.concatenate([ka[: n // 2], rng.integers(50_000_000, 100_000_000, n - n // 2).astype(np.int64)])  # 50 % overlap
    A, B = table(ka), table(kb)
    da, db = len(np.unique(ka)), len(np.unique(kb))
    dboth = len(np.intersect1d(ka, kb))
    for name, expect in (("union", da + db - dboth), ("intersect", dboth), ("subtract", da - dboth)):
        for rep in range(2):
            out, rec = timed("set_" + name, lambda: getattr(A, name)(B), rows_per_table=n)
        rec.update(out_rows=out.row_count, expected_rows=expect, ok=(out.row_count == expect), rows_per_s=round(2 * n / (rec["ms"] / 1e3)))
        print(json.dumps(rec), flush=True)
        out.free()
    for rep in range(2):
        out, rec = timed("unique_cols_0_1", lambda: A.unique([0, 1]), rows=n)
    rec.update(out_rows=out.row_count, expected_rows=da, ok=(out.row_count == da), rows_per_s=round(n / (rec["ms"] / 1e3)))
    print(json.dumps(rec), flush=True)


if __name__ == "__main__":
    main()
