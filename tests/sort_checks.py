"""Sort / MapToSortPartitions / DistributedSort checks shared by tests/test_gpu_sort.py (one rank), tests/thread_worker.py (ranks as
threads on one GPU) and tests/dist_worker.py (one process per GPU). Every result is compared with the oracle (oracle/pyoracle.py
sort_table / map_to_sort_partitions / distributed_sort, pinned against the compiled reference in
tests/test_oracle_sort_vs_reference.py): the sort columns position by position (bit patterns), the rows as a multiset."""
import numpy as np

from oracle import pyoracle as orc
from tests.helpers import assert_same_rows, from_orc, to_orc

KINDS = ["int8", "uint8", "int16", "uint16", "int32", "uint32", "int64", "uint64", "float32", "float64"]


def _bits(a):
    a = np.ascontiguousarray(a)
    return a.view("u%d" % a.dtype.itemsize)


def column(rng, kind, n, narrow=False):
    if kind.startswith("float"):
        v = (rng.standard_normal(n) * (10 if narrow else 1e6)).astype(kind)
        return np.round(v) if narrow else v
    info = np.iinfo(kind)
    if narrow:
        return rng.integers(max(info.min, -40), min(info.max, 40), n).astype(kind)
    return rng.integers(info.min, info.max, n, dtype=kind, endpoint=True)


def check_sorted_table(got, exp, sort_cols, what, rows=True):
    assert len(got) == len(exp), what
    for c in sort_cols:
        g, e = _bits(got[c].data), _bits(exp[c].data)
        assert g.shape == e.shape, "%s: column %d has %d rows, expected %d" % (what, c, len(g), len(e))
        bad = np.nonzero(g != e)[0]
        assert len(bad) == 0, "%s: sort column %d differs at %s: %s vs %s" % (what, c, bad[:4], got[c].data[bad[:4]], exp[c].data[bad[:4]])
    if rows:
        assert_same_rows(got, exp, what)


def local_sort_single(api, ctx):
    checks = 0
    rng = np.random.default_rng(11)
    # ---- single sort column of every type, both directions, wide and narrow value ranges, payload carried along
    for kind in KINDS:
        for n, narrow in ((0, False), (1, False), (5000, False), (30000, True)):
            tab = [orc.col(column(rng, kind, n, narrow), name="k"), orc.col(np.arange(n, dtype=np.float64), name="v")]
            T = from_orc(api, ctx, tab)
            for asc in (True, False):
                out = T.sort(0, asc)
                assert out.column_names == ["k", "v"]
                check_sorted_table(to_orc(out), orc.sort_table(tab, [0], [asc]), [0, 1], "sort %s n=%d asc=%s" % (kind, n, asc), rows=False)  # ties keep input order: both columns position by position
                checks += 1
    return checks


def local_sort_specials(api, ctx):
    checks = 0
    rng = np.random.default_rng(12)
    # ---- float specials: IEEE total order (-inf < ... < -0.0 < +0.0 < ... < +inf < NaN); compared as bit patterns
    for kind in ("float32", "float64"):
        v = column(rng, kind, 300, True)
        v[:8] = [0.0, -0.0, np.inf, -np.inf, np.nan, 1.5, -1.5, 0.0]
        tab = [orc.col(v)]
        for asc in (True, False):
            got = to_orc(from_orc(api, ctx, tab).sort([0], asc))
            exp = orc.sort_table(tab, [0], [asc])
            assert np.array_equal(_bits(got[0].data), _bits(exp[0].data)), "float specials %s asc=%s" % (kind, asc)
            checks += 1
    return checks


def local_sort_multi(api, ctx):
    checks = 0
    rng = np.random.default_rng(13)
    # ---- several sort columns with mixed directions and types; many ties on the leading column
    n = 12000
    tab = [orc.col(rng.integers(-6, 6, n).astype(np.int16), name="a"), orc.col(np.round(rng.standard_normal(n) * 3).astype(np.float32), name="b"),
           orc.col(rng.integers(0, 2 ** 40, n).astype(np.uint64), name="c"), orc.col([b"s%d" % (i % 97) for i in range(n)], name="s")]
    T = from_orc(api, ctx, tab)
    for cols, dirs in (([0, 1], [True, False]), ([1, 0, 2], [False, False, True]), ([2], [False]), ([0, 2], [True, True])):
        got = to_orc(T.sort(cols, dirs))
        check_sorted_table(got, orc.sort_table(tab, cols, dirs), cols, "multi-column sort %s %s" % (cols, dirs))
        checks += 1
    # ---- a string sort column is refused, not mis-sorted
    try:
        T.sort(3)
        raise AssertionError("sorting by a string column should be NotImplemented")
    except api.CylonError:
        checks += 1
    return checks


def local_sort_nulls(api, ctx):
    checks = 0
    rng = np.random.default_rng(14)
    # ---- validity rides along with its row (arrow Take), the value slot decides the position
    v = rng.integers(0, 1000, 3000).astype(np.int64)
    c0, c1 = orc.col(v, name="k"), orc.col(rng.random(3000), name="v")
    c0.validity = np.packbits(rng.random(3000) < 0.8, bitorder="little")
    c1.validity = np.packbits(rng.random(3000) < 0.5, bitorder="little")
    got = to_orc(from_orc(api, ctx, [c0, c1]).sort(0))
    exp = orc.sort_table([c0, c1], [0], [True])
    assert np.array_equal(got[0].data, exp[0].data) and got[0].values() == exp[0].values() and got[1].values() == exp[1].values(), "sort with nulls"
    checks += 1
    return checks


def local_range_partition(api, ctx):
    checks = 0
    rng = np.random.default_rng(15)
    # ---- range partitioner on one rank into 2..7 partitions: every row's target partition and the histogram
    for kind in KINDS:
        for n, samples, narrow in ((20000, 0, False), (20000, 20000, True), (3001, 500, False), (50, 0, True)):
            v = column(rng, kind, n, narrow)
            T = from_orc(api, ctx, [orc.col(v)])
            for parts, nb, asc in ((2, 0, True), (4, 4, False), (7, 100, True), (3, 5000, False)):
                pid, hist = T.map_to_sort_partitions(0, parts, asc, samples, nb)
                epid, ehist = orc.map_to_sort_partitions([[orc.col(v)]], 0, asc, samples, nb, num_partitions=parts)
                bad = np.nonzero(pid != epid[0])[0]
                assert len(bad) == 0, "range partition %s n=%d samples=%d parts=%d bins=%d asc=%s: rows %s values %s got %s expected %s" % (
                    kind, n, samples, parts, nb, asc, bad[:4], v[bad[:4]], pid[bad[:4]], epid[0][bad[:4]])
                assert np.array_equal(hist, ehist[0]), "range partition histogram %s" % kind
                checks += 1
    return checks


LOCAL_SECTIONS = [local_sort_single, local_sort_specials, local_sort_multi, local_sort_nulls, local_range_partition]


def run_local_checks(api, ctx):
    """One rank: cylon::Sort and the range partitioner against the oracle."""
    return sum(section(api, ctx) for section in LOCAL_SECTIONS)


def run_dist_checks(api, ctx, rank, world, shared):
    """Every rank: MapToSortPartitions and DistributedSort against the oracle's result for this rank."""
    checks = 0

    def tables(seed, rows_of):
        def make():
            out = []
            for r in range(world):
                rng = np.random.default_rng(seed * 100 + r)
                n = rows_of(r)
                out.append([orc.col(rng.integers(-5000, 5000, n).astype(np.int32), name="k"), orc.col(np.round(rng.standard_normal(n), 1), name="d"),
                            orc.col(rng.integers(0, 7, n).astype(np.int64), name="g"), orc.col(rng.integers(0, 2 ** 40, n).astype(np.uint64), name="u")])
            return out
        return shared.get(("sort tables", seed), make)

    cases = [(1, lambda r: 6000, 0, 0), (2, lambda r: 4000 + 900 * r, 0, 0), (3, lambda r: 3000, 3000, world), (4, lambda r: 0 if r == 0 else 2500, 200, 64)]
    for seed, rows_of, samples, nb in cases:
        tabs = tables(seed, rows_of)
        T = from_orc(api, ctx, tabs[rank])
        for col, asc in ((0, True), (1, False), (3, True)):
            pid, hist = T.map_to_sort_partitions(col, world, asc, samples, nb)
            epid, ehist = shared.get(("sort pids", seed, col, asc), lambda: orc.map_to_sort_partitions(tabs, col, asc, samples, nb))
            assert np.array_equal(pid, epid[rank]), "MapToSortPartitions case %d column %d rank %d: %d rows differ" % (seed, col, rank, int((pid != epid[rank]).sum()))
            assert np.array_equal(hist, ehist[rank])
            checks += 1
        for cols, dirs in (([0], [True]), ([1, 0], [False, True]), ([2, 3], [True, False]), ([3], [False])):
            out = T.distributed_sort(cols, dirs, nb, samples)
            exp = shared.get(("dist sort", seed, tuple(cols), tuple(dirs)), lambda: orc.distributed_sort(tabs, cols, dirs, samples, nb))[rank]
            check_sorted_table(to_orc(out), exp, cols, "DistributedSort case %d %s %s rank %d" % (seed, cols, dirs, rank))
            assert out.column_names == ["k", "d", "g", "u"]
            checks += 1
    return checks


if __name__ == "__main__":  # every local section on cuda:0, each reported on its own
    import sys
    import traceback

    from cylon_b200 import api as _api
    _ctx = _api.CylonContext(device=0)
    _failed = 0
    for _section in LOCAL_SECTIONS:
        try:
            print("%s: %d checks ok" % (_section.__name__, _section(_api, _ctx)), flush=True)
        except BaseException:
            _failed += 1
            print("%s FAILED\n%s" % (_section.__name__, traceback.format_exc()[-1500:]), flush=True)
    sys.exit(1 if _failed else 0)

