"""Pins the CPU oracle (oracle/) to the reference: its golden files, its known-answer vectors and,
when present, its own compiled code under oracle/_ref/. No GPU needed."""
import ctypes
import json
import os
import struct

import numpy as np
import pytest

from oracle import pyoracle as orc
from tests.helpers import GOLDEN, assert_same_rows, read_csv

KNOWN = json.load(open(os.path.join(GOLDEN, "known_answers.json")))
F64 = [np.int64, np.float64]


def test_murmur3_vectors_from_reference_build():
    for v in KNOWN["murmur3_x86_32"]:
        assert orc.murmur3_x86_32(bytes.fromhex(v["hex"]), v["seed"]) == v["hash"]


def test_murmur3_survey_known_answers():
    # computed from util/murmur3.cpp
    d = lambda x: struct.pack("<d", x)
    assert orc.murmur3_x86_32(d(10.0)) == 918068555
    assert orc.murmur3_x86_32(d(11.0)) == 2776153155
    assert orc.murmur3_x86_32(d(21.0)) == 1809509047
    assert orc.murmur3_x86_32(d(0.394)) == 3545836594
    assert orc.murmur3_x86_32(d(1.0)) == 4152582287
    assert orc.murmur3_x86_32(struct.pack("<f", 1.5)) == 376679366
    assert orc.murmur3_x86_32(b"") == 0
    assert orc.murmur3_x86_32(b"a") == 1009084850
    assert orc.murmur3_x86_32(b"abc") == 3017643002
    assert orc.murmur3_x86_32(b"hello world") == 1586663183


def test_reference_murmur_library_if_built():
    path = os.path.join(os.path.dirname(GOLDEN), "..", "oracle", "_ref", "libref_murmur3.so")
    if not os.path.exists(path):
        pytest.skip("oracle/_ref not built")
    ref = ctypes.CDLL(path)
    ref.ref_murmur3_x86_32.restype = ctypes.c_uint32
    ref.ref_murmur3_x86_32.argtypes = [ctypes.c_char_p, ctypes.c_int, ctypes.c_uint32]
    rng = np.random.default_rng(7)
    for n in list(range(0, 40)) + [100, 1000]:
        b = rng.integers(0, 256, n, dtype=np.uint8).tobytes()
        assert orc.murmur3_x86_32(b, 0) == ref.ref_murmur3_x86_32(b, n, 0)


def test_row_hash_known_answers():
    # composite (int64 123456789012, int32 7) -> 344598643 ; P=8 -> 3, P=6 -> 1
    cols = [orc.col(np.array([123456789012], dtype=np.int64)), orc.col(np.array([7], dtype=np.int32))]
    assert orc.row_hash(cols)[0] == 344598643
    assert orc.hash_partition(cols, 8)[0][0] == 3
    assert orc.hash_partition(cols, 6)[0][0] == 1
    # row (int64 26, double 0.394) -> 3545837400
    cols = [orc.col(np.array([26], dtype=np.int64)), orc.col(np.array([0.394]))]
    assert orc.row_hash(cols)[0] == 3545837400
    # int64 -5 -> 4294967291 -> P=8:3, P=3:2
    c = [orc.col(np.array([-5], dtype=np.int64))]
    assert orc.row_hash(c)[0] == 4294967291
    assert orc.hash_partition(c, 8)[0][0] == 3
    assert orc.hash_partition(c, 3)[0][0] == 2
    # small signed ints are sign-extended by static_cast<uint32_t>
    assert orc.row_hash([orc.col(np.array([-1], dtype=np.int8))])[0] == 0xFFFFFFFF
    assert orc.row_hash([orc.col(np.array([255], dtype=np.uint8))])[0] == 255
    # strings
    assert list(orc.row_hash([orc.col(["", "a", "abc", "hello world"])])) == [0, 1009084850, 3017643002, 1586663183]


@pytest.mark.parametrize("world", [1, 2, 3, 4, 8])
def test_partition_identities(world):
    # cpp/test/partition_test.cpp:32-79 with CreateTable(rows): col0 int32 i, col1 double i+10
    rows = 37
    c0 = orc.col(np.arange(rows, dtype=np.int32))
    c1 = orc.col(np.arange(rows, dtype=np.float64) + 10.0)
    pid, hist = orc.hash_partition([c0], world)
    assert list(pid) == [i % world for i in range(rows)]
    assert hist.sum() == rows
    pid, hist = orc.hash_partition([c1], world)
    assert list(pid) == [orc.murmur3_x86_32(struct.pack("<d", i + 10.0)) % world for i in range(rows)]
    perm, offs = orc.split_indices(pid, world)
    assert list(np.diff(offs)) == list(hist)
    for p in range(world):
        rows_p = perm[offs[p]:offs[p + 1]]
        assert all(pid[rows_p] == p) and list(rows_p) == sorted(rows_p)  # stable


@pytest.mark.parametrize("world", [1, 2, 4])
@pytest.mark.parametrize("algo", [orc.SORT, orc.HASH])
def test_join_inner_golden(world, algo):
    # cpp/test/join_test.cpp:20-35 -- every rank compares with its own expected file
    lefts = [read_csv("csv1_%d.csv" % r, F64) for r in range(world)]
    rights = [read_csv("csv2_%d.csv" % r, F64) for r in range(world)]
    outs = orc.distributed_join(lefts, rights, [0], [0], orc.INNER, algo)
    for r in range(world):
        exp = read_csv("join_inner_%d_%d.csv" % (world, r), [np.int64, np.float64, np.int64, np.float64])
        assert_same_rows(outs[r], exp, "join_inner world %d rank %d" % (world, r))


@pytest.mark.parametrize("world", [1, 2, 4])
@pytest.mark.parametrize("algo", [orc.SORT, orc.HASH])
def test_multi_join_golden(world, algo):
    # cpp/test/join_test.cpp:37-53 -- every rank reads the SAME two files
    lefts = [read_csv("multi_join1.csv") for _ in range(world)]
    rights = [read_csv("multi_join2.csv") for _ in range(world)]
    outs = orc.distributed_join(lefts, rights, [0, 1], [0, 1], orc.INNER, algo, "l_", "r_")
    for r in range(world):
        exp = read_csv("multi_join_%d_%d.csv" % (world, r), [np.int64] * 6)
        assert_same_rows(outs[r], exp, "multi_join world %d rank %d" % (world, r))
    assert [c.name for c in outs[0]] == ["l_a", "l_b", "l_c", "r_a", "r_b", "r_d"]


@pytest.mark.parametrize("world", [1, 2, 4])
@pytest.mark.parametrize("op,name", [(orc.UNION, "union"), (orc.SUBTRACT, "subtract"), (orc.INTERSECT, "intersect")])
def test_set_op_golden(world, op, name):
    # cpp/test/set_op_test.cpp:22-79
    lefts = [read_csv("csv1_%d.csv" % r, F64) for r in range(world)]
    rights = [read_csv("csv2_%d.csv" % r, F64) for r in range(world)]
    outs = orc.distributed_set_op(lefts, rights, op)
    for r in range(world):
        exp = read_csv("%s_%d_%d.csv" % (name, world, r), F64)
        assert_same_rows(outs[r], exp, "%s world %d rank %d" % (name, world, r))


def test_unique_known_answer():
    # cpp/test/table_op_test.cpp:49-86
    t = read_csv("duplicate_data_0.csv")
    out = orc.unique(t, [0])
    assert sorted(out[3].values()) == KNOWN["unique_col3"]


def test_dist_rl_row_counts():
    # python/test/test_dist_rl.py:76-100 (double join key -> murmur3 partitioning)
    k = KNOWN["test_dist_rl"]
    tabs = [read_csv("user_usage_tm_%d.csv" % (r + 1)) for r in range(4)]
    assert [c.type for c in tabs[0]] == [orc.DOUBLE, orc.DOUBLE, orc.DOUBLE, orc.INT64]
    joined = orc.distributed_join(tabs, tabs, [0], [0], orc.INNER, orc.HASH)
    assert [j[0].length for j in joined] == k["join_rows"] and len(joined[0]) == k["join_cols"]
    assert [t[0].length for t in orc.distributed_set_op(tabs, tabs, orc.UNION)] == k["union_rows"]
    assert [t[0].length for t in orc.distributed_set_op(tabs, tabs, orc.INTERSECT)] == k["intersect_rows"]
    assert [t[0].length for t in orc.distributed_set_op(tabs, tabs, orc.SUBTRACT)] == k["subtract_rows"]


@pytest.mark.parametrize("world", [1, 2, 4])
def test_groupby_totals(world):
    # cpp/test/groupby_test.cpp:30-110: the same 10-row table on every rank
    def table():
        return [orc.col(np.array([0, 0, 1, 1, 2, 2, 3, 3, 4, 4], dtype=np.int64), name="col0"),
                orc.col(np.array([0, 0, 1, 1, 2, 2, 3, 3, 4, 4], dtype=np.float64), name="col1")]
    tabs = [table() for _ in range(world)]
    outs = orc.distributed_hash_groupby(tabs, [0], [1], [orc.SUM])
    assert sum(int(o[0].data.sum()) for o in outs) == 10
    assert sum(float(o[1].data.sum()) for o in outs) == 2 * 10.0 * world
    assert outs[0][1].name == ("sum_col1" if world == 1 else "sum_sum_col1")  # SURVEY.md G1 naming quirk
    outs = orc.distributed_hash_groupby(tabs, [0], [1], [orc.COUNT])
    assert sum(int(o[0].data.sum()) for o in outs) == 10
    assert sum(int(o[1].data.sum()) for o in outs) == 5 * 2 * world
    assert outs[0][1].name == "count_col1"
    outs = orc.distributed_hash_groupby(tabs, [0], [1, 1], [orc.MIN, orc.MAX])
    for o in outs:
        for key, mn, mx in zip(o[0].data, o[1].data, o[2].data):
            assert mn == float(key) and mx == max(float(key), np.finfo(np.float64).tiny)  # MAX init quirk (G3)


def test_max_init_quirk():
    # aggregate_kernels.hpp:477: MAX starts from numeric_limits<double>::min() == DBL_MIN
    t = [orc.col(np.array([1, 1], dtype=np.int64)), orc.col(np.array([-3.0, -2.0]))]
    out = orc.hash_groupby(t, [0], [(1, orc.MAX), (1, orc.MIN), (1, orc.SUM), (1, orc.COUNT)])
    assert out[1].data[0] == np.finfo(np.float64).tiny
    assert out[2].data[0] == -3.0 and out[3].data[0] == -5.0 and out[4].data[0] == 2


@pytest.mark.parametrize("jt", [orc.INNER, orc.LEFT, orc.RIGHT, orc.FULL_OUTER])
def test_hash_and_sort_join_agree(jt):
    # SORT == HASH as multisets (pinned for INNER by multi_join goldens; restated for the outer types)
    rng = np.random.default_rng(jt + 11)
    lk = rng.integers(0, 50, 200).astype(np.int64)
    rk = rng.integers(25, 75, 150).astype(np.int64)
    l = [orc.col(lk), orc.col(rng.random(200))]
    r = [orc.col(rk), orc.col(rng.random(150))]
    a = orc.join(l, r, [0], [0], jt, orc.HASH)
    b = orc.join(l, r, [0], [0], jt, orc.SORT)
    assert_same_rows(a, b)
    # brute force
    exp = []
    for i in range(200):
        m = [j for j in range(150) if rk[j] == lk[i]]
        exp += [(i, j) for j in m]
        if not m and jt in (orc.LEFT, orc.FULL_OUTER):
            exp.append((i, -1))
    if jt in (orc.RIGHT, orc.FULL_OUTER):
        exp += [(-1, j) for j in range(150) if rk[j] not in set(lk)]
    li, ri = orc.hash_join_indices([l[0]], [r[0]], jt)
    assert sorted(zip(li.tolist(), ri.tolist())) == sorted(exp)
