"""Pins the oracle against the reference's OWN compiled code (oracle/_ref/libcylon_ref.so, built in place from
/root/reference by oracle/build_ref.sh): partition ids and row hashes for every key type, ArrayIndexHashJoin for all
four join types, and whole-table HashJoin / SortJoin (single and multi-column keys). Skipped where the checker
library has not been built. No GPU needed."""
import numpy as np
import pytest

from oracle import pyoracle as orc
from oracle import pyref
from tests.helpers import assert_same_rows

pytestmark = pytest.mark.skipif(not pyref.available(), reason="oracle/_ref/libcylon_ref.so not built")

KINDS = ["int8", "uint8", "int16", "uint16", "int32", "uint32", "int64", "uint64", "float32", "float64", "str"]


def rand_col(rng, n, kind, dom=None):
    if kind == "str":
        return orc.col([b"k%d" % v for v in rng.integers(0, dom or 10 ** 6, n)])
    if kind.startswith("float"):
        if dom:
            return orc.col(rng.integers(0, dom, n).astype(kind) * 0.5)
        return orc.col(rng.standard_normal(n).astype(kind))
    info = np.iinfo(kind)
    if dom:
        return orc.col(rng.integers(0, min(dom, int(info.max)), n).astype(kind))
    return orc.col(rng.integers(info.min, info.max, n, dtype=kind, endpoint=True))


@pytest.mark.parametrize("kind", KINDS)
def test_row_hash_and_partition_match_reference_kernels(kind):
    rng = np.random.default_rng(len(kind))
    c = rand_col(rng, 20011, kind)
    assert np.array_equal(orc.row_hash([c]), pyref.row_hash([c]))
    for p in (1, 2, 3, 8, 12):
        a, ah = orc.hash_partition([c], p)
        b, bh = pyref.hash_partition([c], p)
        assert np.array_equal(a, b) and np.array_equal(ah, bh)


def test_multi_column_partition_matches_reference():
    rng = np.random.default_rng(5)
    cols = [rand_col(rng, 9001, k) for k in ("int64", "int32", "float64", "str", "int8")]
    assert np.array_equal(orc.row_hash(cols), pyref.row_hash(cols))
    for p in (2, 7, 8):
        a, ah = orc.hash_partition(cols, p)
        b, bh = pyref.hash_partition(cols, p)
        assert np.array_equal(a, b) and np.array_equal(ah, bh)


@pytest.mark.parametrize("jt", [orc.INNER, orc.LEFT, orc.RIGHT, orc.FULL_OUTER])
@pytest.mark.parametrize("kind", ["int64", "int32", "float64", "uint8"])
@pytest.mark.parametrize("nl,nr", [(700, 500), (500, 700), (300, 300), (0, 40), (40, 0)])
def test_array_index_hash_join_matches_reference(jt, kind, nl, nr):
    rng = np.random.default_rng(jt * 100 + nl)
    l, r = rand_col(rng, nl, kind, 150), rand_col(rng, nr, kind, 150)
    a = orc.hash_join_indices([l], [r], jt)
    b = pyref.array_index_hash_join(l, r, jt)
    assert sorted(zip(a[0].tolist(), a[1].tolist())) == sorted(zip(b[0].tolist(), b[1].tolist()))


@pytest.mark.parametrize("algo", [orc.HASH, orc.SORT])
@pytest.mark.parametrize("jt", [orc.INNER, orc.LEFT, orc.RIGHT, orc.FULL_OUTER])
@pytest.mark.parametrize("keys", [[0], [0, 2], [2], [1, 0]])
def test_join_tables_match_reference(algo, jt, keys):
    # columns: int64, double (small domain so it can be a key), int32
    rng = np.random.default_rng(algo * 31 + jt * 7 + len(keys))
    def tab(n):
        return [rand_col(rng, n, "int64", 60), rand_col(rng, n, "float64", 4), rand_col(rng, n, "int32", 5)]
    l, r = tab(400), tab(350)
    exp = pyref.join_tables(l, r, keys, keys, jt, algo)
    got = orc.join(l, r, keys, keys, jt, algo)
    assert_same_rows(got, exp, "algo %d type %d keys %s" % (algo, jt, keys))


# ---- G2/G3: the reference's aggregation kernels themselves (compute/aggregate_kernels.hpp compiled in place) ----------
_AGG_KINDS = ["int8", "int16", "int32", "int64", "uint8", "uint16", "uint32", "uint64", "float32", "float64"]


@pytest.mark.skipif(not pyref.aggregate_available(), reason="oracle/_ref/libref_aggregate.so not built")
@pytest.mark.parametrize("kind", _AGG_KINDS)
@pytest.mark.parametrize("op", [orc.SUM, orc.MIN, orc.MAX, orc.COUNT])
def test_aggregate_matches_reference_kernels(kind, op):
    """SUM in the column's own C type (integers wrap), COUNT as int64, MIN from numeric_limits<T>::max(), MAX from
    numeric_limits<T>::min() -- for float/double the smallest positive normal, so an all-negative group reports it."""
    rng = np.random.default_rng(_AGG_KINDS.index(kind) * 10 + op)
    n, groups = 5000, 37
    dt = np.dtype(kind)
    if dt.kind == "f":
        vals = ((rng.random(n) - 0.5) * 1e6).astype(dt)
    else:
        info = np.iinfo(dt)
        vals = rng.integers(info.min, info.max, n, dtype=dt, endpoint=True)  # full range: sums wrap around
    gid = rng.integers(0, groups - 2, n)  # the last two groups stay empty: initial states are reported as they are
    if dt.kind == "f":
        vals[gid == 3] = -np.abs(vals[gid == 3]) - 1  # an all-negative group (G3: MAX returns DBL_MIN / FLT_MIN)
    got = orc.aggregate(orc.col(vals), gid, groups, op)
    exp = pyref.ref_aggregate(op, vals, gid, groups)
    assert got.dtype == exp.dtype
    assert np.array_equal(got, exp)  # floating-point SUM too: same sequential order of additions as the reference
    if dt.kind == "f" and op == orc.MAX:
        assert exp[3] == np.finfo(dt).tiny


# ---- the reference's Distributed* operators themselves, ranks as threads (oracle/ref_shim: in-process communicator) -----
from tests.helpers import assert_same_rows  # noqa: E402

_has_dist = pyref.available() and hasattr(pyref.lib(), "ref_dist_op")
dist = pytest.mark.skipif(not _has_dist, reason="oracle/_ref/libcylon_ref.so was built without the table operators")


def _kv_tables(world, n, seed, dom, two_keys=False):
    out = []
    for r in range(world):
        rng = np.random.default_rng(seed * 100 + r)
        cols = [orc.col(rng.integers(-dom, dom, n + 3 * r).astype(np.int64), name="c0")]
        if two_keys:
            cols.append(orc.col(rng.integers(0, 4, n + 3 * r).astype(np.int32), name="c1"))
        cols.append(orc.col(rng.integers(0, 1000, n + 3 * r).astype(np.float64) / 8, name="c%d" % len(cols)))
        out.append(cols)
    return out


@dist
@pytest.mark.parametrize("world", [1, 2, 3, 4])
@pytest.mark.parametrize("jt", [orc.INNER, orc.LEFT, orc.RIGHT, orc.FULL_OUTER])
@pytest.mark.parametrize("algo", [orc.SORT, orc.HASH])
@pytest.mark.parametrize("two_keys", [False, True])
def test_distributed_join_matches_reference_ranks(world, jt, algo, two_keys):
    """cylon::DistributedJoin (table.cpp:482-502) run by the reference itself on `world` ranks: hash partition of both
    tables on the key columns, the Arrow all-to-all, then the local join -- per-rank results as row multisets."""
    lefts = _kv_tables(world, 60, 1, 40, two_keys)
    rights = _kv_tables(world, 50, 2, 40, two_keys)
    keys = [0, 1] if two_keys else [0]
    exp = orc.distributed_join(lefts, rights, keys, keys, jt, algo, "l_", "r_")
    got = pyref.dist_op(pyref.DIST_JOIN, lefts, rights, p1=keys, p2=keys, join_type=jt, algorithm=algo)
    for r in range(world):
        assert [c.name for c in got[r]] == [c.name for c in exp[r]]
        assert_same_rows(got[r], exp[r], "rank %d" % r)


@dist
@pytest.mark.parametrize("world", [1, 2, 4])
@pytest.mark.parametrize("op,ref_op", [(orc.UNION, pyref.DIST_UNION), (orc.SUBTRACT, pyref.DIST_SUBTRACT), (orc.INTERSECT, pyref.DIST_INTERSECT)])
def test_distributed_set_ops_match_reference_ranks(world, op, ref_op):
    """DistributedUnion/Subtract/Intersect (table.cpp:727-785): shuffle on ALL columns, then the local set operator.
    Numeric columns only -- the reference's cross-table string comparator is broken."""
    lefts = _kv_tables(world, 80, 3, 6)   # small domains: many duplicate rows inside and across tables
    rights = _kv_tables(world, 70, 3, 6)  # same seed family: overlaps with the left tables
    for t in lefts + rights:
        t[1].data[:] = np.floor(t[1].data) % 3
    exp = orc.distributed_set_op(lefts, rights, op)
    got = pyref.dist_op(ref_op, lefts, rights)
    for r in range(world):
        assert_same_rows(got[r], exp[r], "rank %d" % r)


@dist
@pytest.mark.parametrize("world", [1, 2, 4])
@pytest.mark.parametrize("cols", [[0], [0, 1]])
def test_distributed_unique_matches_reference_ranks(world, cols):
    tables = _kv_tables(world, 90, 5, 8)
    for t in tables:
        t[1].data[:] = np.floor(t[1].data) % 2
    exp = orc.distributed_unique(tables, cols)
    got = pyref.dist_op(pyref.DIST_UNIQUE, tables, p1=cols)
    for r in range(world):
        if cols == [0, 1]:
            assert_same_rows(got[r], exp[r], "rank %d" % r)
        else:  # which duplicate survives depends on arrival order after the shuffle: compare the distinct keys
            assert sorted(got[r][0].data.tolist()) == sorted(exp[r][0].data.tolist())


@dist
@pytest.mark.parametrize("world", [2, 3, 4])
def test_shuffle_matches_reference_ranks(world):
    tables = _kv_tables(world, 100, 7, 1000)
    exp = orc.shuffle(tables, [0])
    got = pyref.dist_op(pyref.DIST_SHUFFLE, tables, p1=[0])
    assert sum(t[0].length for t in got) == sum(t[0].length for t in tables)  # cpp/test/table_op_test.cpp:38-47
    for r in range(world):
        assert_same_rows(got[r], exp[r], "rank %d" % r)


@dist
@pytest.mark.parametrize("world", [1, 2, 4])
@pytest.mark.parametrize("ops", [[orc.SUM, orc.MIN, orc.MAX], [orc.SUM, orc.COUNT, orc.MIN, orc.MAX], [orc.COUNT]])
def test_distributed_groupby_matches_reference_ranks(world, ops):
    """DistributedHashGroupBy (groupby/groupby.cpp:33-83): column names (the doubled prefix after two associative
    rounds), group placement by key hash, and the aggregate values. Values are multiples of 1/8 so sums are exact."""
    tables = _kv_tables(world, 120, 9, 10)
    for t in tables:
        t[1].data[:] -= 60.0  # negative groups exist: MAX starts from DBL_MIN (G3)
    exp = orc.distributed_hash_groupby(tables, [0], [1] * len(ops), ops)
    got = pyref.dist_op(pyref.DIST_GROUPBY, tables, p1=[0], p2=[x for op in ops for x in (1, op)])
    for r in range(world):
        assert [c.name for c in got[r]] == [c.name for c in exp[r]]
        assert_same_rows(got[r], exp[r], "rank %d" % r)


# ---- null inputs: what the reference really does with them (no reference test covers this; the compiled reference does) --
def _with_nulls(tables, seed):
    for r, t in enumerate(tables):
        rng = np.random.default_rng(seed * 31 + r)
        for c in t:
            c.validity = np.packbits(rng.random(c.length) >= 0.2, bitorder="little")
    return tables


@dist
@pytest.mark.parametrize("world", [1, 2, 3])
@pytest.mark.parametrize("jt", [orc.INNER, orc.LEFT, orc.RIGHT, orc.FULL_OUTER])
@pytest.mark.parametrize("algo", [orc.SORT, orc.HASH])
def test_join_with_null_inputs_matches_reference(world, jt, algo):
    """Keys under a null bit still hash, partition and compare by their value slot (A6); the gather does not carry source
    validity (copy_arrray.cpp:24-50), only the -1 fill makes nulls; after a shuffle every input null bit is gone (B2)."""
    lefts, rights = _with_nulls(_kv_tables(world, 60, 11, 20), 1), _with_nulls(_kv_tables(world, 50, 12, 20), 2)
    exp = orc.distributed_join(lefts, rights, [0], [0], jt, algo, "l_", "r_")
    got = pyref.dist_op(pyref.DIST_JOIN, lefts, rights, p1=[0], p2=[0], join_type=jt, algorithm=algo)
    for r in range(world):
        assert_same_rows(got[r], exp[r], "rank %d" % r)


@dist
@pytest.mark.parametrize("world", [1, 2])
def test_set_ops_unique_groupby_shuffle_with_null_inputs_match_reference(world):
    lefts, rights = _with_nulls(_kv_tables(world, 80, 13, 6), 3), _with_nulls(_kv_tables(world, 70, 13, 6), 4)
    for t in lefts + rights:
        t[1].data[:] = np.floor(t[1].data) % 3
    for op, ref_op in ((orc.UNION, pyref.DIST_UNION), (orc.SUBTRACT, pyref.DIST_SUBTRACT), (orc.INTERSECT, pyref.DIST_INTERSECT)):
        exp, got = orc.distributed_set_op(lefts, rights, op), pyref.dist_op(ref_op, lefts, rights)
        for r in range(world):
            assert_same_rows(got[r], exp[r], "set op %d rank %d" % (op, r))
    exp, got = orc.distributed_unique(lefts, [0, 1]), pyref.dist_op(pyref.DIST_UNIQUE, lefts, p1=[0, 1])
    for r in range(world):
        assert_same_rows(got[r], exp[r], "unique rank %d" % r)
    ops = [orc.SUM, orc.COUNT, orc.MIN, orc.MAX]  # COUNT counts rows under a null bit too (G3)
    exp = orc.distributed_hash_groupby(lefts, [0], [1] * 4, ops)
    got = pyref.dist_op(pyref.DIST_GROUPBY, lefts, p1=[0], p2=[x for o in ops for x in (1, o)])
    for r in range(world):
        assert_same_rows(got[r], exp[r], "group-by rank %d" % r)
    if world > 1:
        exp, got = orc.shuffle(lefts, [0]), pyref.dist_op(pyref.DIST_SHUFFLE, lefts, p1=[0])
        for r in range(world):
            assert all(c.validity is None for c in got[r])  # B2: Split appends values only
            assert_same_rows(got[r], exp[r], "shuffle rank %d" % r)


@dist
@pytest.mark.parametrize("world", [1, 2, 3])
def test_string_payload_through_join_and_shuffle_matches_reference(world):
    """Strings as PAYLOAD are safe in the reference (only cross-table string comparisons are broken, E3): the string
    splitter (arrow_kernels.cpp:115-157), the all-to-all of offset + data buffers and the binary gather
    (copy_arrray.cpp:52-85), including empty strings."""
    words = ["", "a", "bb", "cylon", "hash-partition", "x" * 40]

    def tables(seed):
        out = []
        for r in range(world):
            rng = np.random.default_rng(seed * 10 + r)
            n = 70 + 5 * r
            out.append([orc.col(rng.integers(0, 25, n).astype(np.int64), name="c0"),
                        orc.col([words[i] for i in rng.integers(0, len(words), n)], name="c1"),
                        orc.col(rng.integers(0, 100, n).astype(np.float64), name="c2")])
        return out

    lefts, rights = tables(1), tables(2)
    for jt in (orc.INNER, orc.FULL_OUTER):
        exp = orc.distributed_join(lefts, rights, [0], [0], jt, orc.HASH, "l_", "r_")
        got = pyref.dist_op(pyref.DIST_JOIN, lefts, rights, p1=[0], p2=[0], join_type=jt, algorithm=orc.HASH)
        for r in range(world):
            assert_same_rows(got[r], exp[r], "join %d rank %d" % (jt, r))
    if world > 1:
        exp, got = orc.shuffle(lefts, [0]), pyref.dist_op(pyref.DIST_SHUFFLE, lefts, p1=[0])
        for r in range(world):
            assert_same_rows(got[r], exp[r], "shuffle rank %d" % r)


# ---- floating-point key identity (NaN, -0.0 / +0.0, +-inf): what the reference really does, pinned --------------------
def _special_tables(extra_key=False):
    nan2 = np.array([0x7FF8000000000001], dtype=np.uint64).view(np.float64)[0]
    pool = np.array([np.nan, nan2, -0.0, 0.0, np.inf, -np.inf, 1.0, 2.5, -3.0, 1e300], dtype=np.float64)

    def one(n, seed, base):
        rng = np.random.default_rng(seed)
        cols = [orc.col(pool[rng.integers(0, len(pool), n)], name="k")]
        if extra_key:
            cols.append(orc.col((np.arange(n) % 2).astype(np.int64), name="z"))
        cols.append(orc.col(np.arange(n, dtype=np.int64) + base, name="v"))
        return cols
    return one(400, 1, 0), one(300, 2, 10_000)


def _bit_rows(table):
    cols = []
    for c in table:
        vals = c.values()
        if c.type == orc.DOUBLE:
            vals = [None if v is None else b for v, b in zip(vals, c.data.view(np.uint64).tolist())]
        cols.append(vals)
    return sorted(zip(*cols), key=lambda r: tuple((v is None, v if v is not None else 0) for v in r))


@pytest.mark.skipif(not pyref.dist_available(), reason="table operators of the reference not built")
@pytest.mark.parametrize("jt", [orc.INNER, orc.LEFT, orc.RIGHT, orc.FULL_OUTER])
@pytest.mark.parametrize("two_keys", [False, True])
def test_float_key_identity_hash_join(jt, two_keys):
    """Single key: operator== (a NaN joins nothing) on keys whose murmur3-of-bits hashes agree (-0.0 and +0.0 never meet).
    Two keys: the three-way compare chain, under which NaNs with identical bits are equal."""
    l, r = _special_tables(two_keys)
    keys = [0, 1] if two_keys else [0]
    ref = pyref.join_tables(l, r, keys, keys, jt, orc.HASH)
    assert _bit_rows(orc.join(l, r, keys, keys, jt, orc.HASH)) == _bit_rows(ref)
    if not two_keys and jt == orc.INNER:
        assert not np.isnan(ref[0].data).any()
        assert np.array_equal(ref[0].data.view(np.uint64), ref[2].data.view(np.uint64))


@pytest.mark.skipif(not pyref.dist_available(), reason="table operators of the reference not built")
@pytest.mark.parametrize("ncols", [1, 2])
def test_float_key_identity_row_sets(ncols):
    l, r = _special_tables(ncols == 2)
    l, r = l[:ncols], r[:ncols]
    for oop, rop in ((orc.UNION, pyref.DIST_UNION), (orc.INTERSECT, pyref.DIST_INTERSECT), (orc.SUBTRACT, pyref.DIST_SUBTRACT)):
        assert _bit_rows(orc.set_op(l, r, oop)) == _bit_rows(pyref.dist_op(rop, [l], [r])[0])
    keys = list(range(ncols))
    assert _bit_rows(orc.unique(l, keys)) == _bit_rows(pyref.dist_op(pyref.DIST_UNIQUE, [l], p1=keys)[0])
    full, _ = _special_tables(False)
    ref = pyref.dist_op(pyref.DIST_GROUPBY, [full], p1=[0], p2=[1, orc.SUM, 1, orc.COUNT])[0]
    assert _bit_rows(orc.hash_groupby(full, [0], [(1, orc.SUM), (1, orc.COUNT)])) == _bit_rows(ref)


@pytest.mark.skipif(not pyref.dist_available(), reason="table operators of the reference not built")
@pytest.mark.parametrize("dt", ["float64", "float32", "int64", "int32", "uint16", "int8", "uint8"])
def test_mean_var_stddev_match_reference(dt):
    """MeanKernel / VarianceKernel (aggregate_kernels.hpp:344-405) through the reference's HashGroupBy: result types, ddof = 1,
    states in the column's own type (the small integer types wrap)."""
    rng = np.random.default_rng(len(dt))
    k = rng.integers(0, 50, 5000).astype(np.int64)
    v = rng.integers(0, 100, 5000).astype(dt)
    t = [orc.col(k, name="c0"), orc.col(v, name="c1")]
    for op in (orc.MEAN, orc.VAR, orc.STDDEV):
        a = orc.hash_groupby(t, [0], [(1, op)])
        b = pyref.dist_op(pyref.DIST_GROUPBY, [t], p1=[0], p2=[1, op])[0]
        assert [c.name for c in a] == [c.name for c in b]
        assert np.array_equal(a[0].data, b[0].data) and a[1].data.dtype == b[1].data.dtype
        assert np.allclose(a[1].data, b[1].data, rtol=1e-12, atol=0, equal_nan=True)
