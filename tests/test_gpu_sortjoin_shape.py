"""BASELINE config 3's shape -- LEFT / RIGHT / FULL_OUTER SortJoin on an (int64, int32) key whose left side is
Zipf(1.1)-skewed (workloads.sortjoin_tables, the generator bench.py --config sortjoin uses) -- against the compiled
reference's SortJoin (join/sort_join.cpp:396-556) and the oracle; and SURVEY.md D6: a SORT join never touches the
caller's columns (the reference's in-place variant sorts the caller's key buffer)."""
import numpy as np
import pytest

import workloads
from oracle import pyoracle as orc
from oracle import pyref
from tests.helpers import assert_same_rows, from_orc, to_orc

pytestmark = pytest.mark.gpu
JT = {orc.LEFT: "left", orc.RIGHT: "right", orc.FULL_OUTER: "fullouter", orc.INNER: "inner"}


@pytest.fixture(scope="module")
def api():
    from cylon_b200 import api as a
    return a


@pytest.fixture(scope="module")
def ctx(api):
    c = api.CylonContext(device=0)
    yield c
    c.finalize()


def _tables(rows, drop_right=0):
    l, r = workloads.sortjoin_tables(rows, 1, 0)
    if drop_right:  # leave some left rows without a partner and some right rows unmatched
        keep = np.arange(rows) % drop_right != 1
        r = [c[keep] for c in r]
    names = ["k0", "k1", "v"]
    return [orc.col(c, name=n) for c, n in zip(l, names)], [orc.col(c, name=n) for c, n in zip(r, names)]


@pytest.mark.parametrize("jt", [orc.LEFT, orc.RIGHT, orc.FULL_OUTER])
@pytest.mark.parametrize("algo", ["sort", "hash"])
def test_zipf_two_column_key_rows(api, ctx, jt, algo):
    l, r = _tables(40_000, drop_right=7)
    assert (l[0].data == 0).mean() > 0.08  # the hot key is there
    out = from_orc(api, ctx, l).join(from_orc(api, ctx, r), JT[jt], algo, [0, 1], [0, 1], "l_", "r_")
    exp = orc.join(l, r, [0, 1], [0, 1], jt, orc.SORT, "l_", "r_")
    assert_same_rows(to_orc(out), exp, "zipf %s %s" % (JT[jt], algo))
    if pyref.dist_available():
        ref = pyref.join_tables(l, r, [0, 1], [0, 1], jt, orc.SORT)
        assert out.row_count == ref[0].length and out.checksum() == orc.table_checksum(ref)


@pytest.mark.parametrize("jt", [orc.LEFT, orc.FULL_OUTER])
def test_zipf_two_column_key_checksum_1m(api, ctx, jt):
    l, r = _tables(1_000_000, drop_right=5)
    out = from_orc(api, ctx, l).join(from_orc(api, ctx, r), JT[jt], "sort", [0, 1], [0, 1], "l_", "r_")
    li, ri = orc.hash_join_indices([l[0], l[1]], [r[0], r[1]], jt)  # same pairs as the sort join, and linear time on the hot key
    exp = [c.take(li) for c in l] + [c.take(ri) for c in r]
    assert out.row_count == len(li)
    assert out.checksum() == orc.table_checksum(exp)


@pytest.mark.parametrize("keys", [[0], [0, 1]])
@pytest.mark.parametrize("jt", [orc.INNER, orc.FULL_OUTER])
def test_sort_join_leaves_inputs_untouched(api, ctx, keys, jt):
    rng = np.random.default_rng(11)
    l = [orc.col(rng.integers(0, 500, 5000).astype(np.int64), name="k0"), orc.col(rng.integers(0, 3, 5000).astype(np.int32), name="k1"),
         orc.col(rng.random(5000), name="v")]
    r = [orc.col(rng.integers(0, 500, 4000).astype(np.int64), name="k0"), orc.col(rng.integers(0, 3, 4000).astype(np.int32), name="k1"),
         orc.col(rng.random(4000), name="v")]
    L, R = from_orc(api, ctx, l), from_orc(api, ctx, r)
    out = L.join(R, JT[jt], "sort", keys, keys)
    assert out.row_count > 0
    for t, src in ((L, l), (R, r)):
        for got, c in zip(t.to_host(), src):  # same values in the same ORDER, not just the same multiset
            assert np.array_equal(got["data"], c.data), c.name
