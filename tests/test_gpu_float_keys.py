"""Floating-point keys with NaN, -0.0 / +0.0 and +-inf: the CUDA path against the compiled reference (oracle/_ref) and
the oracle, on the fused join, the general join, multi-column joins, the set operators, Unique and GroupBy.

What the reference does (run here, see tests/test_oracle_vs_reference.py::test_float_key_identity_*): it hashes the
BITS of a double (murmur3) and its containers only compare keys that hashed alike, so -0.0 and +0.0 are two keys
everywhere on the hash path and NaNs with identical bits are one key in Union/Unique/GroupBy and multi-column joins;
the single-key hash join compares with operator== (arrow_comparator.cpp:240-243), so there a NaN joins nothing."""
import numpy as np
import pytest

from oracle import pyoracle as orc
from oracle import pyref
from tests.helpers import from_orc, to_orc

pytestmark = pytest.mark.gpu
JT = {orc.INNER: "inner", orc.LEFT: "left", orc.RIGHT: "right", orc.FULL_OUTER: "fullouter"}
NAN2 = np.array([0x7FF8000000000001], dtype=np.uint64).view(np.float64)[0]  # a NaN with another payload


@pytest.fixture(scope="module")
def api():
    from cylon_b200 import api as a
    return a


@pytest.fixture(scope="module")
def ctx(api):
    c = api.CylonContext(device=0)
    yield c
    c.finalize()


def special_keys(n, seed):
    rng = np.random.default_rng(seed)
    pool = np.array([np.nan, NAN2, -0.0, 0.0, np.inf, -np.inf, 1.0, 2.5, -3.0, 1e300], dtype=np.float64)
    return pool[rng.integers(0, len(pool), n)]


def bit_rows(table):
    """row multiset with floating-point columns compared by bit pattern (NaN-safe)"""
    cols = []
    for c in table:
        vals = c.values()
        if c.type == orc.DOUBLE:
            bits = c.data.view(np.uint64).tolist()
            vals = [None if v is None else b for v, b in zip(vals, bits)]
        elif c.type == orc.FLOAT:
            bits = c.data.view(np.uint32).tolist()
            vals = [None if v is None else b for v, b in zip(vals, bits)]
        cols.append(vals)
    return sorted(zip(*cols), key=lambda r: tuple((v is None, v if v is not None else 0) for v in r))


def tables(n_l=400, n_r=300, extra_key=False):
    def one(n, seed, base):
        cols = [orc.col(special_keys(n, seed), name="k")]
        if extra_key:
            cols.append(orc.col((np.arange(n) % 2).astype(np.int64), name="z"))
        cols.append(orc.col(np.arange(n, dtype=np.int64) + base, name="v"))
        return cols
    return one(n_l, 1, 0), one(n_r, 2, 10_000)


@pytest.mark.parametrize("fused", ["1", "0"])
@pytest.mark.parametrize("jt", [orc.INNER, orc.LEFT, orc.RIGHT, orc.FULL_OUTER])
def test_single_key_hash_join(api, ctx, fused, jt):
    l, r = tables()
    ctx.add_config("cylon.join.fused", fused)
    try:
        out = to_orc(from_orc(api, ctx, l).join(from_orc(api, ctx, r), JT[jt], "hash", [0], [0], "l_", "r_"))
    finally:
        ctx.add_config("cylon.join.fused", "1")
    exp = orc.join(l, r, [0], [0], jt, orc.HASH, "l_", "r_")
    assert bit_rows(out) == bit_rows(exp)
    if pyref.dist_available():
        ref = pyref.join_tables(l, r, [0], [0], jt, orc.HASH)
        assert bit_rows(out) == bit_rows(ref)
    if jt == orc.INNER:  # no NaN ever matched, and the two zeros never met
        assert not np.isnan(out[0].data).any()
        assert np.array_equal(out[0].data.view(np.uint64), out[2].data.view(np.uint64))


@pytest.mark.parametrize("jt", [orc.INNER, orc.FULL_OUTER])
def test_two_key_hash_join(api, ctx, jt):
    l, r = tables(extra_key=True)
    out = to_orc(from_orc(api, ctx, l).join(from_orc(api, ctx, r), JT[jt], "hash", [0, 1], [0, 1], "l_", "r_"))
    assert bit_rows(out) == bit_rows(orc.join(l, r, [0, 1], [0, 1], jt, orc.HASH, "l_", "r_"))
    if pyref.dist_available():
        assert bit_rows(out) == bit_rows(pyref.join_tables(l, r, [0, 1], [0, 1], jt, orc.HASH))


@pytest.mark.parametrize("ncols", [1, 2])
def test_set_operators_and_unique(api, ctx, ncols):
    l, r = tables(extra_key=(ncols == 2))
    l, r = l[:ncols], r[:ncols]
    L, R = from_orc(api, ctx, l), from_orc(api, ctx, r)
    ops = (("union", orc.UNION, pyref.DIST_UNION), ("intersect", orc.INTERSECT, pyref.DIST_INTERSECT), ("subtract", orc.SUBTRACT, pyref.DIST_SUBTRACT))
    for name, oop, rop in ops:
        out = to_orc(getattr(L, name)(R))
        assert bit_rows(out) == bit_rows(orc.set_op(l, r, oop)), name
        if pyref.dist_available():
            assert bit_rows(out) == bit_rows(pyref.dist_op(rop, [l], [r])[0]), name
    keys = list(range(ncols))
    out = to_orc(L.unique(keys))
    assert bit_rows(out) == bit_rows(orc.unique(l, keys))
    if pyref.dist_available():
        assert bit_rows(out) == bit_rows(pyref.dist_op(pyref.DIST_UNIQUE, [l], p1=keys)[0])


def test_groupby_on_a_double_key(api, ctx):
    l, _ = tables()
    out = to_orc(from_orc(api, ctx, l).groupby(0, [1, 1], ["sum", "count"]))
    exp = orc.hash_groupby(l, [0], [(1, orc.SUM), (1, orc.COUNT)])
    assert bit_rows(out) == bit_rows(exp)
    if pyref.dist_available():
        ref = pyref.dist_op(pyref.DIST_GROUPBY, [l], p1=[0], p2=[1, orc.SUM, 1, orc.COUNT])[0]
        assert bit_rows(out) == bit_rows(ref)
