"""The fused group-by (cyl_groupby_fused: radix partition + one CTA per bin with its group table in shared memory)
through the C-ABI against the oracle's HashGroupBy (hash_groupby.cpp:92-201 restated) and, at table level, against the
general path and the compiled reference: keys exact, COUNT / MIN / MAX exact (MAX from numeric_limits<T>::min()),
integer SUM wrapping, double SUM within the 1e-12 relative tolerance of BASELINE.json."""
import numpy as np
import pytest

from oracle import pyoracle as orc
from oracle import pyref
from tests.helpers import assert_same_rows, from_orc, to_orc

pytestmark = pytest.mark.gpu
OPS = [orc.SUM, orc.COUNT, orc.MIN, orc.MAX]


@pytest.fixture(scope="module")
def cy():
    from cylon_b200 import capi
    capi.init(0)
    return capi


def _check(cy, key, vals, aggs, capacity=None):
    """aggs: list of (value index, op)"""
    dk = cy.DCol.from_numpy(key)
    dv = [cy.DCol.from_numpy(v) for v in vals]
    res = cy.groupby_fused(dk, dv, [a[0] for a in aggs], [a[1] for a in aggs], capacity)
    assert res is not None
    gk, gaggs, groups, attempts = res
    table = [orc.col(key)] + [orc.col(v) for v in vals]
    exp = orc.hash_groupby(table, [0], [(1 + a[0], a[1]) for a in aggs])
    assert groups == exp[0].length
    og, oe = np.argsort(gk, kind="stable"), np.argsort(exp[0].data, kind="stable")
    assert np.array_equal(gk[og].view(np.uint64), exp[0].data[oe].view(np.uint64))
    for (vi, op), got, e in zip(aggs, gaggs, exp[1:]):
        a, b = got[og], e.data[oe]
        if op == orc.SUM and a.dtype.kind == "f":
            assert np.allclose(a, b, rtol=1e-12, atol=0)
        else:
            assert np.array_equal(a, b), (op, a[:5], b[:5])
    return attempts


@pytest.mark.parametrize("n,dom", [(1, 1), (1000, 37), (100_000, 5000), (300_007, 300_007), (2_500_000, 20_000), (2_500_000, 3_000_000)])
def test_int64_key_double_value_all_ops(cy, n, dom):
    rng = np.random.default_rng(n + dom)
    key = rng.integers(-dom // 2, dom - dom // 2, n).astype(np.int64) * 7919
    val = rng.random(n) - 0.25
    _check(cy, key, [val], [(0, op) for op in OPS])


def test_max_starts_from_dbl_min_and_min_from_dbl_max(cy):
    rng = np.random.default_rng(3)
    key = rng.integers(0, 50, 5000).astype(np.int64)
    val = -rng.random(5000) - 1.0  # every group all negative: MAX must come out as DBL_MIN (aggregate_kernels.hpp:477)
    dk, dv = cy.DCol.from_numpy(key), [cy.DCol.from_numpy(val)]
    gk, gaggs, groups, _ = cy.groupby_fused(dk, dv, [0, 0], [orc.MAX, orc.MIN])
    assert groups == 50 and np.all(gaggs[0] == np.finfo(np.float64).tiny)
    _check(cy, key, [val], [(0, orc.MAX), (0, orc.MIN)])


@pytest.mark.parametrize("kdt,vdt", [(np.uint64, np.int64), (np.float64, np.uint64), (np.int64, np.int64)])
def test_key_and_value_types_integer_sum_wraps(cy, kdt, vdt):
    rng = np.random.default_rng(9)
    n = 40_000
    key = rng.integers(0, 300, n).astype(kdt)
    if kdt == np.float64:
        key = key * 0.5 - 30.0
        key[::97] = -0.0
        key[1::97] = 0.0
        key[2::97] = np.nan  # NaNs with identical bits are one group, the two zeros are two (bit-pattern identity)
    info = np.iinfo(vdt)
    val = rng.integers(info.min // 2, info.max // 2, n, dtype=vdt) * vdt(3)  # sums overflow and wrap
    if kdt == np.float64:
        dk, dv = cy.DCol.from_numpy(key), [cy.DCol.from_numpy(val)]
        gk, gaggs, groups, _ = cy.groupby_fused(dk, dv, [0, 0, 0], [orc.SUM, orc.MIN, orc.COUNT])
        assert groups == len(np.unique(key.view(np.uint64)))
        order = np.argsort(gk.view(np.uint64), kind="stable")
        uk, inv = np.unique(key.view(np.uint64), return_inverse=True)
        with np.errstate(over="ignore"):
            sums = np.zeros(len(uk), dtype=vdt)
            np.add.at(sums, inv, val)
        assert np.array_equal(gk.view(np.uint64)[order], uk)
        assert np.array_equal(gaggs[0][order], sums)
        assert np.array_equal(gaggs[2][order], np.bincount(inv))
    else:
        _check(cy, key, [val], [(0, orc.SUM), (0, orc.MIN), (0, orc.MAX), (0, orc.COUNT)])


def test_all_ones_key_and_two_value_columns(cy):
    rng = np.random.default_rng(4)
    n = 30_000
    key = rng.integers(-3, 40, n).astype(np.int64)  # -1 = the table's empty marker as a real key
    v0, v1 = rng.random(n), rng.integers(-1000, 1000, n).astype(np.int64)
    _check(cy, key, [v0, v1], [(0, orc.SUM), (1, orc.SUM), (1, orc.MAX), (0, orc.COUNT)])


def test_hot_key_and_exact_capacity_retry(cy):
    rng = np.random.default_rng(5)
    n = 600_000
    key = rng.integers(0, 200_000, n).astype(np.int64)
    key[: n // 2] = 42  # one group holds half of the rows: its bin is long but has few groups
    val = rng.random(n)
    assert _check(cy, key, [val], [(0, op) for op in OPS]) == 1
    assert _check(cy, key, [val], [(0, orc.COUNT)], capacity=100) == 2  # too small: exact count reported, second call fits


def test_unsupported_shapes_fall_through(cy):
    k32 = cy.DCol.from_numpy(np.arange(8, dtype=np.int32))
    k64 = cy.DCol.from_numpy(np.arange(8, dtype=np.int64))
    v32 = cy.DCol.from_numpy(np.arange(8, dtype=np.float32))
    v64 = cy.DCol.from_numpy(np.arange(8, dtype=np.float64))
    assert cy.groupby_fused(k32, [v64], [0], [orc.SUM]) is None
    assert cy.groupby_fused(k64, [v32], [0], [orc.SUM]) is None
    assert cy.groupby_fused(k64, [v64], [0], [orc.SUM]) is not None


@pytest.mark.parametrize("ops", [["sum", "min", "max"], ["sum", "count", "min", "max"], ["count"]])
def test_table_groupby_fused_equals_general_and_reference(ops):
    from cylon_b200 import api
    ctx = api.CylonContext(device=0)
    rng = np.random.default_rng(len(ops))
    n = 250_000
    t = [orc.col(rng.integers(-4000, 4000, n).astype(np.int64), name="c0"), orc.col(np.floor(rng.random(n) * 1000) / 8 - 60.0, name="c1")]
    T = from_orc(api, ctx, t)
    fused = T.groupby(0, [1] * len(ops), ops)
    ctx.add_config("cylon.groupby.fused", "0")
    general = T.groupby(0, [1] * len(ops), ops)
    ctx.add_config("cylon.groupby.fused", "1")
    assert fused.column_names == general.column_names
    assert_same_rows(to_orc(fused), to_orc(general))  # values are multiples of 1/8: sums exact whatever the order
    if pyref.dist_available():
        code = {"sum": orc.SUM, "min": orc.MIN, "max": orc.MAX, "count": orc.COUNT}
        exp = pyref.dist_op(pyref.DIST_GROUPBY, [t], p1=[0], p2=[x for o in ops for x in (1, code[o])])[0]
        assert fused.column_names == [c.name for c in exp]
        assert_same_rows(to_orc(fused), exp)
    ctx.finalize()
