"""MEAN / VAR / STDDEV group-by aggregates (compute/aggregate_kernels.hpp:344-405) on the GPU against the oracle and the
compiled reference: MEAN keeps the column's type (integer division for integers), VAR / STDDEV are doubles with ddof = 1,
sums and sums of squares accumulate in the column's own type (so small integer types wrap exactly as the reference's)."""
import numpy as np
import pytest

from oracle import pyoracle as orc
from oracle import pyref
from tests.helpers import from_orc, to_orc

pytestmark = pytest.mark.gpu
OPS = [("mean", orc.MEAN), ("var", orc.VAR), ("std", orc.STDDEV)]


@pytest.fixture(scope="module")
def api():
    from cylon_b200 import api as a
    return a


@pytest.fixture(scope="module")
def ctx(api):
    c = api.CylonContext(device=0)
    yield c
    c.finalize()


def _compare(got, exp):
    assert [c.name for c in got] == [c.name for c in exp]
    og, oe = np.argsort(got[0].data, kind="stable"), np.argsort(exp[0].data, kind="stable")
    assert np.array_equal(got[0].data[og], exp[0].data[oe])
    for g, e in zip(got[1:], exp[1:]):
        assert g.data.dtype == e.data.dtype, (g.name, g.data.dtype, e.data.dtype)
        a, b = g.data[og], e.data[oe]
        if a.dtype.kind == "f":
            assert np.allclose(a, b, rtol=1e-9, atol=1e-12, equal_nan=True), g.name  # float sums in another order (float32 states included)
        else:
            assert np.array_equal(a, b), g.name


@pytest.mark.parametrize("dt", [np.float64, np.float32, np.int64, np.int32, np.uint16, np.int8])
def test_moments_match_oracle_and_reference(api, ctx, dt):
    rng = np.random.default_rng(np.dtype(dt).itemsize)
    n = 20_000
    k = rng.integers(0, 300, n).astype(np.int64)
    k[:5] = 10_000 + np.arange(5)  # single-row groups: VAR = STDDEV = 0
    v = rng.integers(0, 12, n).astype(dt) if np.dtype(dt).kind != "f" else (rng.integers(0, 64, n).astype(dt) * dt(0.25))
    t = [orc.col(k, name="c0"), orc.col(v, name="c1")]
    out = to_orc(from_orc(api, ctx, t).groupby(0, [1, 1, 1, 1], ["mean", "var", "std", "sum"]))
    exp = orc.hash_groupby(t, [0], [(1, orc.MEAN), (1, orc.VAR), (1, orc.STDDEV), (1, orc.SUM)])
    _compare(out, exp)
    if pyref.dist_available():
        ref = pyref.dist_op(pyref.DIST_GROUPBY, [t], p1=[0], p2=[1, orc.MEAN, 1, orc.VAR, 1, orc.STDDEV, 1, orc.SUM])[0]
        _compare(out, ref)


def test_nunique_and_quantile_are_refused(api, ctx):
    from cylon_b200.capi import CylonError
    t = from_orc(api, ctx, [orc.col(np.arange(10, dtype=np.int64), name="k"), orc.col(np.arange(10, dtype=np.float64), name="v")])
    for op in (6, 7):
        with pytest.raises(CylonError) as e:
            t.groupby(0, [1], [op])
        assert e.value.code == 10  # Code::NotImplemented
    with pytest.raises(CylonError) as e:
        t.groupby(0, [1], [9])
    assert e.value.code == 4  # not an AggregationOpId at all: Code::Invalid
