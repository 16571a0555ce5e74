"""Table-level parity on one GPU: the C++ host API (through the C facade) against the oracle and the reference's
golden files. Per-rank goldens for world 2/4 are checked on ONE GPU by running the partition kernels per simulated
rank, routing partitions on the host and running the local operator -- everything but the NCCL transport itself,
which tests/test_gpu_distributed.py covers on real multi-GPU boxes."""
import json
import os

import numpy as np
import pytest

from oracle import pyoracle as orc
from tests.helpers import GOLDEN, assert_same_rows, from_orc, read_csv, to_orc

pytestmark = pytest.mark.gpu
KNOWN = json.load(open(os.path.join(GOLDEN, "known_answers.json")))
F64 = [np.int64, np.float64]


@pytest.fixture(scope="module")
def api():
    from cylon_b200 import api as a
    return a


@pytest.fixture(scope="module")
def ctx(api):
    c = api.CylonContext(device=0)
    yield c
    c.finalize()


def emulate_shuffle(api, ctx, tables, key_cols):
    """what Shuffle does at world W, minus NCCL: HashPartition on the GPU per rank, route partition p to rank p."""
    world = len(tables)
    parts = [from_orc(api, ctx, t).hash_partition(key_cols, world) for t in tables]
    return [orc.concat_tables([to_orc(parts[src][dst]) for src in range(world)]) for dst in range(world)]


@pytest.mark.parametrize("world", [1, 2, 4])
@pytest.mark.parametrize("algo", ["sort", "hash"])
def test_join_inner_golden(api, ctx, world, algo):
    lefts = [read_csv("csv1_%d.csv" % r, F64) for r in range(world)]
    rights = [read_csv("csv2_%d.csv" % r, F64) for r in range(world)]
    if world > 1:
        lefts, rights = emulate_shuffle(api, ctx, lefts, [0]), emulate_shuffle(api, ctx, rights, [0])
    for r in range(world):
        out = from_orc(api, ctx, lefts[r]).join(from_orc(api, ctx, rights[r]), "inner", algo, [0], [0])
        exp = read_csv("join_inner_%d_%d.csv" % (world, r), [np.int64, np.float64, np.int64, np.float64])
        assert_same_rows(to_orc(out), exp, "join_inner world %d rank %d" % (world, r))


@pytest.mark.parametrize("world", [1, 2, 4])
@pytest.mark.parametrize("algo", ["sort", "hash"])
def test_multi_join_golden(api, ctx, world, algo):
    lefts = [read_csv("multi_join1.csv") for _ in range(world)]
    rights = [read_csv("multi_join2.csv") for _ in range(world)]
    if world > 1:
        lefts, rights = emulate_shuffle(api, ctx, lefts, [0, 1]), emulate_shuffle(api, ctx, rights, [0, 1])
    for r in range(world):
        out = from_orc(api, ctx, lefts[r]).join(from_orc(api, ctx, rights[r]), "inner", algo, [0, 1], [0, 1], "l_", "r_")
        exp = read_csv("multi_join_%d_%d.csv" % (world, r), [np.int64] * 6)
        assert_same_rows(to_orc(out), exp, "multi_join world %d rank %d" % (world, r))
        assert out.column_names == ["l_a", "l_b", "l_c", "r_a", "r_b", "r_d"]


@pytest.mark.parametrize("world", [1, 2, 4])
@pytest.mark.parametrize("name", ["union", "subtract", "intersect"])
def test_set_op_golden(api, ctx, world, name):
    lefts = [read_csv("csv1_%d.csv" % r, F64) for r in range(world)]
    rights = [read_csv("csv2_%d.csv" % r, F64) for r in range(world)]
    if world > 1:
        lefts, rights = emulate_shuffle(api, ctx, lefts, [0, 1]), emulate_shuffle(api, ctx, rights, [0, 1])
    for r in range(world):
        out = getattr(from_orc(api, ctx, lefts[r]), name)(from_orc(api, ctx, rights[r]))
        exp = read_csv("%s_%d_%d.csv" % (name, world, r), F64)
        assert_same_rows(to_orc(out), exp, "%s world %d rank %d" % (name, world, r))


def test_unique_known_answer(api, ctx):
    t = from_orc(api, ctx, read_csv("duplicate_data_0.csv"))
    out = to_orc(t.unique([0]))
    assert sorted(out[3].values()) == KNOWN["unique_col3"]


def test_dist_rl_counts_single_gpu_emulation(api, ctx):
    k = KNOWN["test_dist_rl"]
    tabs = [read_csv("user_usage_tm_%d.csv" % (r + 1)) for r in range(4)]
    by_key = emulate_shuffle(api, ctx, tabs, [0])
    rows = [from_orc(api, ctx, t).join(from_orc(api, ctx, t), "inner", "hash", [0], [0]).row_count for t in by_key]
    assert rows == k["join_rows"]
    by_row = emulate_shuffle(api, ctx, tabs, [0, 1, 2, 3])
    assert [from_orc(api, ctx, t).union(from_orc(api, ctx, t)).row_count for t in by_row] == k["union_rows"]
    assert [from_orc(api, ctx, t).intersect(from_orc(api, ctx, t)).row_count for t in by_row] == k["intersect_rows"]
    assert [from_orc(api, ctx, t).subtract(from_orc(api, ctx, t)).row_count for t in by_row] == k["subtract_rows"]


def _rand_table(rng, n, dom, strings=True, nulls=False):
    k = rng.integers(0, dom, n).astype(np.int64)
    cols = [orc.col(k, name="k"), orc.col(rng.random(n), name="v"), orc.col(rng.integers(-5, 5, n).astype(np.int32), name="i")]
    if strings:
        cols.append(orc.col([b"s%d" % (x % 17) for x in rng.integers(0, 1000, n)], name="s"))
    if nulls:
        for c in cols[1:]:
            c.validity = np.packbits(rng.random(n) < 0.8, bitorder="little")
    return cols


@pytest.mark.parametrize("algo", ["hash", "sort"])
@pytest.mark.parametrize("jt,ojt", [("inner", orc.INNER), ("left", orc.LEFT), ("right", orc.RIGHT), ("outer", orc.FULL_OUTER)])
@pytest.mark.parametrize("keys", [[0], [0, 2], [3], [3, 0]])
def test_join_tables_all_types(api, ctx, algo, jt, ojt, keys):
    rng = np.random.default_rng(len(keys) * 5 + ojt)
    l, r = _rand_table(rng, 2000, 300), _rand_table(rng, 1500, 300)
    out = from_orc(api, ctx, l).join(from_orc(api, ctx, r), jt, algo, keys, keys, "l_", "r_")
    exp = orc.join(l, r, keys, keys, ojt, orc.HASH, "l_", "r_")
    assert_same_rows(to_orc(out), exp)
    assert out.column_names == [c.name for c in exp]


@pytest.mark.parametrize("jt,ojt", [("inner", orc.INNER), ("outer", orc.FULL_OUTER)])
@pytest.mark.parametrize("algo", ["hash", "sort"])
def test_join_drops_source_nulls_like_the_reference(api, ctx, jt, ojt, algo):
    """copy_array_by_indices (util/copy_arrray.cpp:39-47) appends a null only for index -1: a null of the INPUT comes out
    as a valid row holding its value slot. The oracle's take() restates that and is pinned to the compiled reference on
    null inputs (tests/test_oracle_vs_reference.py); the key column carries null bits too (A6: hashing and matching read
    the value slot whatever the bit says)."""
    rng = np.random.default_rng(8)
    l, r = _rand_table(rng, 500, 50, nulls=True), _rand_table(rng, 400, 50, nulls=True)
    for t in (l, r):  # null bits on the key column as well
        t[0].validity = np.packbits(rng.random(t[0].length) > 0.3, bitorder="little")
    out = to_orc(from_orc(api, ctx, l).join(from_orc(api, ctx, r), jt, algo, [0], [0]))
    exp = orc.join(l, r, [0], [0], ojt, orc.HASH)
    assert_same_rows(out, exp)
    if ojt == orc.INNER:
        assert all(c.validity is None or c.null_count == 0 for c in out)


@pytest.mark.parametrize("P", [1, 2, 3, 8])
@pytest.mark.parametrize("keys", [[0], [1], [3], [0, 3]])
def test_hash_partition_tables(api, ctx, P, keys):
    rng = np.random.default_rng(P)
    t = _rand_table(rng, 5000, 1000, nulls=(P == 3))
    parts = from_orc(api, ctx, t).hash_partition(keys, P)
    pid, _ = orc.hash_partition([t[i] for i in keys], P)
    perm, offs = orc.split_indices(pid, P)
    for p in range(P):
        exp = [c.take(perm[offs[p]:offs[p + 1]]) for c in t]
        for e, c in zip(exp, t):
            if c.validity is not None:
                e.validity = np.packbits(c.valid_mask()[perm[offs[p]:offs[p + 1]]], bitorder="little")
        got = to_orc(parts[p])
        assert got[0].length == offs[p + 1] - offs[p]
        # stable: same ORDER, not just same multiset
        for g, e in zip(got, exp):
            assert g.values() == e.values()


@pytest.mark.parametrize("name,op", [("union", orc.UNION), ("subtract", orc.SUBTRACT), ("intersect", orc.INTERSECT)])
def test_set_ops_with_strings(api, ctx, name, op):
    rng = np.random.default_rng(op)
    def tab(n):
        k = rng.integers(0, 200, n).astype(np.int64)
        return [orc.col(k), orc.col([b"w%d" % (v % 13) for v in k]), orc.col((k % 5).astype(np.float64))]
    l, r = tab(3000), tab(2500)
    out = getattr(from_orc(api, ctx, l), name)(from_orc(api, ctx, r))
    assert_same_rows(to_orc(out), orc.set_op(l, r, op))


@pytest.mark.parametrize("world", [1])
def test_groupby_totals(api, ctx, world):
    t = [orc.col(np.array([0, 0, 1, 1, 2, 2, 3, 3, 4, 4], dtype=np.int64), name="col0"),
         orc.col(np.array([0, 0, 1, 1, 2, 2, 3, 3, 4, 4], dtype=np.float64), name="col1")]
    dt = from_orc(api, ctx, t)
    out = to_orc(dt.groupby(0, [1], ["sum"], distributed=True))
    assert int(out[0].data.sum()) == 10 and float(out[1].data.sum()) == 20.0 and out[1].name == "sum_col1"
    out = to_orc(dt.groupby(0, [1], ["count"], distributed=True))
    assert int(out[1].data.sum()) == 10 and out[1].name == "count_col1"


def test_groupby_table_matches_oracle(api, ctx):
    rng = np.random.default_rng(12)
    n = 50000
    t = [orc.col(rng.integers(0, 700, n).astype(np.int64), name="g"), orc.col([b"x%d" % v for v in rng.integers(0, 3, n)], name="h"),
         orc.col(rng.random(n), name="v"), orc.col(rng.integers(-100, 100, n).astype(np.int32), name="w")]
    out = from_orc(api, ctx, t).groupby([0, 1], [2, 2, 2, 2, 3, 3], ["sum", "count", "min", "max", "sum", "max"])
    exp = orc.hash_groupby(t, [0, 1], [(2, orc.SUM), (2, orc.COUNT), (2, orc.MIN), (2, orc.MAX), (3, orc.SUM), (3, orc.MAX)])
    got = to_orc(out)
    assert out.column_names == [c.name for c in exp]
    # same group order (first-seen), so columns compare positionally; double sums within 1e-12 relative
    for g, e in zip(got, exp):
        if g.name == "sum_v":
            assert np.allclose(g.data, e.data, rtol=1e-12, atol=0)
        else:
            assert g.values() == e.values(), g.name


def test_error_codes(api, ctx):
    t = from_orc(api, ctx, [orc.col(np.arange(4, dtype=np.int64)), orc.col(np.arange(4, dtype=np.float64))])
    with pytest.raises(api.CylonError) as e:
        t.join(t, "inner", "hash", [5], [0])
    assert e.value.code == 7  # IndexError
    with pytest.raises(api.CylonError) as e:
        t.join(t, "inner", "hash", [0], [1])
    assert e.value.code == 4  # Invalid: key types differ (hash_join.cpp:195-197)
    u = from_orc(api, ctx, [orc.col(np.arange(4, dtype=np.int64))])
    with pytest.raises(api.CylonError) as e:
        t.union(u)
    assert e.value.code == 4  # VerifyTableSchema


def test_empty_tables(api, ctx):
    e = [orc.col(np.zeros(0, dtype=np.int64)), orc.col(np.zeros(0, dtype=np.float64))]
    f = [orc.col(np.arange(5, dtype=np.int64)), orc.col(np.arange(5, dtype=np.float64))]
    te, tf = from_orc(api, ctx, e), from_orc(api, ctx, f)
    assert te.join(tf, "inner", "hash", [0], [0]).row_count == 0
    assert te.join(tf, "right", "sort", [0], [0]).row_count == 5
    assert tf.join(te, "left", "hash", [0], [0]).row_count == 5
    assert te.union(tf).row_count == 5 and tf.subtract(te).row_count == 5 and te.intersect(tf).row_count == 0
    assert te.unique([0]).row_count == 0
    assert te.groupby(0, [1], ["sum"]).row_count == 0
