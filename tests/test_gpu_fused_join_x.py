"""The extended fused join (cyl_join_fused, join_fused8x.cu): LEFT / RIGHT / FULL_OUTER / INNER, one- and two-column keys,
4- and 8-byte columns, against the oracle's rows and the compiled reference; hot keys that overfill a bin (build chunks
with outer semantics), empty sides, NaN and signed-zero keys, and the table-level routing of HASH and SORT."""
import numpy as np
import pytest

from oracle import pyoracle as orc
from oracle import pyref
from tests.helpers import assert_same_rows, from_orc, to_orc

pytestmark = pytest.mark.gpu
JT = {orc.INNER: "inner", orc.LEFT: "left", orc.RIGHT: "right", orc.FULL_OUTER: "fullouter"}
ALL = [orc.INNER, orc.LEFT, orc.RIGHT, orc.FULL_OUTER]


@pytest.fixture(scope="module")
def api():
    from cylon_b200 import api as a
    return a


@pytest.fixture(scope="module")
def ctx(api):
    c = api.CylonContext(device=0)
    yield c
    c.finalize()


def _join(api, ctx, l, r, jt, algo, keys):
    return from_orc(api, ctx, l).join(from_orc(api, ctx, r), JT[jt], algo, keys, keys, "l_", "r_")


def _table(rng, n, dom, two_keys, i32_payload=False):
    cols = [orc.col(rng.integers(0, dom, n).astype(np.int64), name="k0")]
    if two_keys:
        cols.append(orc.col(rng.integers(0, 3, n).astype(np.int32), name="k1"))
    cols.append(orc.col(rng.random(n), name="v"))
    if i32_payload:
        cols.append(orc.col(rng.integers(-9, 9, n).astype(np.int32), name="w"))
    return cols


@pytest.mark.parametrize("jt", ALL)
@pytest.mark.parametrize("two_keys", [False, True])
@pytest.mark.parametrize("algo", ["hash", "sort"])
@pytest.mark.parametrize("nl,nr,dom", [(3000, 2500, 1500), (2500, 3000, 1500), (50_000, 70_000, 40_000)])
def test_rows_match_oracle(api, ctx, jt, two_keys, algo, nl, nr, dom):
    rng = np.random.default_rng(nl + 7 * jt + two_keys)
    l, r = _table(rng, nl, dom, two_keys, i32_payload=not two_keys), _table(rng, nr, dom, two_keys)
    keys = [0, 1] if two_keys else [0]
    out = _join(api, ctx, l, r, jt, algo, keys)
    exp = orc.join(l, r, keys, keys, jt, orc.HASH, "l_", "r_")
    assert out.column_names == [c.name for c in exp]
    assert_same_rows(to_orc(out), exp)


@pytest.mark.parametrize("jt", ALL)
@pytest.mark.parametrize("nl,nr", [(0, 40), (40, 0), (0, 0), (1, 1)])
def test_empty_and_tiny_sides(api, ctx, jt, nl, nr):
    rng = np.random.default_rng(5)
    l, r = _table(rng, nl, 3, True), _table(rng, nr, 3, True)
    out = _join(api, ctx, l, r, jt, "hash", [0, 1])
    assert_same_rows(to_orc(out), orc.join(l, r, [0, 1], [0, 1], jt, orc.HASH, "l_", "r_"))


@pytest.mark.parametrize("jt", ALL)
@pytest.mark.parametrize("hot_side", ["left", "right", "both"])
def test_hot_key_overfills_a_bin(api, ctx, jt, hot_side):
    """30k rows of one key on the build and / or probe side: the bin is joined in several build chunks, unmatched probe
    rows are settled after the last chunk, unmatched build rows per chunk."""
    rng = np.random.default_rng(11)
    nl, nr = 60_000, 50_000
    l, r = _table(rng, nl, 200_000, True), _table(rng, nr, 200_000, True)
    if hot_side in ("left", "both"):
        l[0].data[:30_000] = 77
        l[1].data[:30_000] = 1
    if hot_side in ("right", "both"):
        r[0].data[:30_000 if hot_side == "right" else 40] = 77
        r[1].data[:30_000 if hot_side == "right" else 40] = 1
    out = _join(api, ctx, l, r, jt, "sort", [0, 1])
    li, ri = orc.hash_join_indices([l[0], l[1]], [r[0], r[1]], jt)
    exp = [c.take(li) for c in l] + [c.take(ri) for c in r]
    assert out.row_count == len(li)
    assert out.checksum() == orc.table_checksum(exp)


@pytest.mark.parametrize("jt", ALL)
def test_large_two_pass_partition_vs_reference(api, ctx, jt):
    """2.3M x 2.1M rows (the staged two-pass partition), (int64, int32) key: row count + checksum against the compiled
    reference's SortJoin when it is built, else the oracle."""
    rng = np.random.default_rng(21)
    l, r = _table(rng, 2_300_000, 3_000_000, True), _table(rng, 2_100_000, 3_000_000, True)
    out = _join(api, ctx, l, r, jt, "sort", [0, 1])
    if pyref.dist_available():
        exp = pyref.join_tables(l, r, [0, 1], [0, 1], jt, orc.SORT)
    else:
        exp = orc.join(l, r, [0, 1], [0, 1], jt, orc.SORT)
    assert out.row_count == exp[0].length
    assert out.checksum() == orc.table_checksum(exp)


@pytest.mark.parametrize("kdt", [np.float64, np.float32, np.int32, np.uint64])
@pytest.mark.parametrize("jt", [orc.LEFT, orc.FULL_OUTER])
def test_key_types_nan_and_signed_zero(api, ctx, kdt, jt):
    rng = np.random.default_rng(3)

    def tab(n):
        k = rng.integers(0, 60, n).astype(kdt)
        if np.dtype(kdt).kind == "f":
            k = k * kdt(0.5)
            k[::13] = np.nan
            k[1::13] = -0.0
            k[2::13] = 0.0
        return [orc.col(k, name="k"), orc.col(rng.integers(0, 1000, n).astype(np.int64), name="v")]

    l, r = tab(900), tab(700)
    out = to_orc(_join(api, ctx, l, r, jt, "hash", [0]))
    exp = orc.join(l, r, [0], [0], jt, orc.HASH, "l_", "r_")

    def bits(t):  # NaN-safe canonical rows
        cols = []
        for c in t:
            vals = c.values()
            if c.data.dtype.kind == "f":
                b = c.data.view(np.uint64 if c.data.dtype.itemsize == 8 else np.uint32).tolist()
                vals = [None if v is None else x for v, x in zip(vals, b)]
            cols.append(vals)
        return sorted(zip(*cols), key=lambda row: tuple((v is None, v if v is not None else 0) for v in row))
    assert bits(out) == bits(exp)


def test_fused_and_general_paths_agree(api, ctx):
    rng = np.random.default_rng(8)
    l, r = _table(rng, 30_000, 9000, True), _table(rng, 20_000, 9000, True)
    for jt in ALL:
        fused = _join(api, ctx, l, r, jt, "sort", [0, 1])
        ctx.add_config("cylon.join.fused", "0")
        try:
            general = _join(api, ctx, l, r, jt, "sort", [0, 1])
        finally:
            ctx.add_config("cylon.join.fused", "1")
        assert fused.row_count == general.row_count and fused.checksum() == general.checksum()
