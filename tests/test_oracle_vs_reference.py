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
