"""Pins the oracle's Sort / MapToSortPartitions / DistributedSort (oracle/pyoracle.py) against the reference's own compiled code
run on 1-4 ranks (oracle/_ref, ops 7-9 of ref_dist_op). The reference samples at random unless num_samples equals the rank's row
count (arrow_partition_kernels.cpp:375-381), so the pinned cases use the whole column as the sample and equal row counts per rank
(the reference multiplies its own num_samples by the world size). Reference tests followed: cpp/test/partition_test.cpp:81-153."""
import numpy as np
import pytest

from oracle import pyoracle as orc
from oracle import pyref

pytestmark = pytest.mark.skipif(not (pyref.available() and pyref.dist_available()), reason="oracle/_ref/libcylon_ref.so not built")

TYPES = [("int8", -100, 100), ("uint8", 0, 255), ("int16", -3000, 3000), ("uint16", 0, 60000), ("int32", -50, 1000), ("uint32", 0, 2 ** 32 - 1),
         ("int64", -2 ** 40, 2 ** 40), ("uint64", 0, 2 ** 63), ("float32", None, None), ("float64", None, None)]


def _column(rng, kind, lo, hi, n):
    if lo is None:
        return (rng.random(n) * 1000 - 300).astype(kind)
    return rng.integers(lo, hi, n, dtype=np.uint64 if kind == "uint64" else np.int64).astype(kind)


@pytest.mark.parametrize("kind,lo,hi", TYPES)
@pytest.mark.parametrize("world", [2, 3, 4])
def test_range_partition_ids_match_reference(kind, lo, hi, world):
    """get_bin_pos in each column type's own arithmetic (int8's range wraps, 32-bit types divide unsigned, floats add in float) and
    the float quantile split: target partition of every row on every rank, both directions, several bin counts."""
    rng = np.random.default_rng(world * 100 + len(kind))
    n = 200
    for num_bins in (world, 0, 7, 64):
        for ascending in (1, 0):
            tabs = [[orc.col(_column(rng, kind, lo, hi, n)), orc.col(rng.random(n))] for _ in range(world)]
            ref = pyref.dist_op(pyref.MAP_SORT_PARTITIONS, tabs, p1=[0], p2=[ascending], join_type=num_bins, algorithm=-1)
            mine, hist = orc.map_to_sort_partitions(tabs, 0, bool(ascending), n, num_bins)
            for r in range(world):
                assert np.array_equal(ref[r][0].data, mine[r]), (kind, world, num_bins, ascending, r)
                assert np.array_equal(hist[r], np.bincount(mine[r], minlength=world))


def test_range_partition_is_monotone_on_sorted_input():
    """partition_test.cpp:81-103: on a sorted column the target partitions never decrease and the histogram sums to the row count"""
    v = np.sort(np.random.default_rng(0).integers(0, 10 ** 6, 5000)).astype(np.int32)
    pids, hist = orc.map_to_sort_partitions([[orc.col(v)]], 0, True, len(v), 4)
    # world 1 -> one partition; the same column split four ways by hand:
    pos = orc.sort_bin_pos(v, v.min(), v.max(), 64)
    p = orc.sort_bin_to_partition(np.bincount(pos, minlength=66), len(v), 4)[pos]
    assert np.all(np.diff(p.astype(np.int64)) >= 0) and np.bincount(p, minlength=4).sum() == len(v)
    assert hist[0].sum() == len(v) and np.all(pids[0] == 0)


@pytest.mark.parametrize("world", [1, 2, 4])
@pytest.mark.parametrize("cols,dirs", [([0], [1]), ([0], [0]), ([1], [1]), ([0, 1], [1, 0]), ([2, 0], [0, 1])])
def test_distributed_sort_matches_reference(world, cols, dirs):
    """Per rank: the same row multiset as the reference's rank and the same values in the sort columns, position by position
    (rows that tie on every sort column may come in any order). partition_test.cpp:105-153."""
    rng = np.random.default_rng(world * 10 + len(cols) + sum(dirs))
    n = 300
    tabs = [[orc.col(rng.integers(-20, 60, n).astype(np.int32)), orc.col(rng.random(n)), orc.col(rng.integers(0, 5, n).astype(np.int64))]
            for _ in range(world)]
    ref = pyref.dist_op(pyref.DIST_SORT, tabs, p1=cols, p2=dirs, join_type=world * 4, algorithm=-1)
    mine = orc.distributed_sort(tabs, cols, [bool(d) for d in dirs], n, world * 4)
    assert sum(len(t[0].data) for t in mine) == n * world
    for r in range(world):
        assert orc.table_rows(ref[r]) == orc.table_rows(mine[r])
        for c in cols:
            assert np.array_equal(ref[r][c].data, mine[r][c].data)
    # ranks hold consecutive ranges of the first sort column
    first = [t[cols[0]].data for t in mine if len(t[cols[0]].data)]
    for a, b in zip(first, first[1:]):
        assert (a.max() <= b.min()) if dirs[0] else (a.min() >= b.max())


@pytest.mark.parametrize("cols,dirs", [([0], [1]), ([0], [0]), ([1, 0], [1, 1]), ([2, 1], [0, 0]), ([2, 0, 1], [1, 0, 1])])
def test_local_sort_matches_reference(cols, dirs):
    rng = np.random.default_rng(len(cols) * 7 + sum(dirs))
    n = 2000
    # no uint32 / uint64 column among several sort columns: the reference's multi-column comparator takes the sign of an unsigned
    # a - b (arrow_comparator.cpp:115-121), which is no ordering at all and sends std::sort out of bounds (it crashed here)
    tab = [orc.col(rng.integers(-9, 9, n).astype(np.int16)), orc.col(rng.standard_normal(n).astype(np.float32)), orc.col(rng.integers(0, 4, n).astype(np.int64))]
    ref = pyref.dist_op(pyref.LOCAL_SORT, [tab], p1=cols, p2=dirs)[0]
    mine = orc.sort_table(tab, cols, [bool(d) for d in dirs])
    assert orc.table_rows(ref) == orc.table_rows(mine)
    for c in cols:
        assert np.array_equal(ref[c].data, mine[c].data)


def test_sort_keeps_source_validity_and_sorts_by_value_slot():
    """arrow Take carries the null bits (unlike the join's gather); the sorter compares value slots whatever the validity says"""
    v = np.array([5, 1, 4, 2, 3], dtype=np.int64)
    c = orc.col(v)
    c.validity = np.packbits(np.array([1, 0, 1, 1, 0], dtype=bool), bitorder="little")
    ref = pyref.dist_op(pyref.LOCAL_SORT, [[c, orc.col(np.arange(5.0))]], p1=[0], p2=[1])[0]
    mine = orc.sort_table([c, orc.col(np.arange(5.0))], [0], [True])
    assert np.array_equal(ref[1].data, mine[1].data)
    assert ref[0].values() == mine[0].values() == [None, 2, None, 4, 5]
