"""The radix-partitioned shared-memory join (cyl_hash_join_fused_inner) against the oracle: bit-exact row multisets."""
import numpy as np
import pytest

from oracle import pyoracle as orc
from tests.helpers import assert_same_rows, from_orc, synth_kv, to_orc

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def cy():
    from cylon_b200 import capi
    capi.init(0)
    return capi


def _to_dev(cy, cols):
    return [cy.DCol.from_numpy(c.data, type_id=c.type) for c in cols]


def _out_table(cy, cols):
    out = []
    for c in cols:
        data, _o, _v = c.to_host()
        out.append(orc.HCol(c.type, data))
    return out


def _check(cy, l, r, lkey, rkey, capacity=None, oracle_algo=orc.HASH):
    cols, m, attempts = cy.hash_join_fused(_to_dev(cy, l), lkey, _to_dev(cy, r), rkey, capacity)
    exp = orc.join(l, r, [lkey], [rkey], orc.INNER, oracle_algo)
    assert m == exp[0].length
    assert_same_rows(_out_table(cy, cols), exp)
    return attempts


@pytest.mark.parametrize("nl,nr,dom", [(1, 1, 1), (1000, 700, 300), (700, 1000, 300), (50000, 40000, 60000), (300007, 250003, 200000)])
def test_fused_join_int64_double(cy, nl, nr, dom):
    rng = np.random.default_rng(nl + nr)
    l = [orc.col(rng.integers(0, dom, nl).astype(np.int64)), orc.col(rng.random(nl))]
    r = [orc.col(rng.integers(0, dom, nr).astype(np.int64)), orc.col(rng.random(nr))]
    _check(cy, l, r, 0, 0)


@pytest.mark.parametrize("key_kind", ["int8", "uint16", "int32", "uint32", "int64", "uint64", "float32", "float64"])
def test_fused_join_key_types_and_mixed_widths(cy, key_kind):
    rng = np.random.default_rng(len(key_kind))
    def tab(n):
        if key_kind.startswith("float"):
            k = rng.integers(-40, 40, n).astype(key_kind) * 0.5
        else:
            info = np.iinfo(key_kind)
            k = rng.integers(max(info.min, -50), min(info.max, 50), n).astype(key_kind)
        return [orc.col(rng.integers(0, 100, n).astype(np.int16)), orc.col(k), orc.col(rng.random(n).astype(np.float32)),
                orc.col(rng.integers(0, 2 ** 40, n).astype(np.int64)), orc.col(rng.integers(0, 200, n).astype(np.uint8))]
    _check(cy, tab(3000), tab(2000), 1, 1)


def test_fused_join_negative_zero_and_extremes(cy):
    l = [orc.col(np.array([0.0, -0.0, 1.5, np.finfo(np.float64).max, -1.5]))]
    r = [orc.col(np.array([-0.0, 0.0, -1.5, np.finfo(np.float64).max]))]
    # -0.0 vs 0.0: the reference's HASH join hashes the bit pattern and only compares keys whose hashes agree, so the two
    # zeros never meet (run against the compiled reference: tests/test_oracle_vs_reference.py::test_float_key_identity_*).
    # Same here: a floating-point key is identified by its bits.
    _check(cy, l, r, 0, 0, oracle_algo=orc.HASH)
    i64 = np.iinfo(np.int64)
    l = [orc.col(np.array([-1, i64.min, i64.max, 0, -1], dtype=np.int64))]
    r = [orc.col(np.array([i64.max, -1, 7, i64.min], dtype=np.int64))]
    _check(cy, l, r, 0, 0)


def test_fused_join_heavy_skew_uses_build_chunks(cy):
    # one key holds 30k build rows (> shared-memory capacity of a bin) and 40 probe rows: 1.2M result rows
    rng = np.random.default_rng(3)
    lk = np.concatenate([np.full(30000, 7, np.int64), rng.integers(100, 5000, 20000).astype(np.int64)])
    rk = np.concatenate([np.full(40, 7, np.int64), rng.integers(100, 5000, 60000).astype(np.int64)])
    l = [orc.col(lk), orc.col(rng.random(len(lk)))]
    r = [orc.col(rk), orc.col(rng.random(len(rk)))]
    attempts = _check(cy, l, r, 0, 0)
    assert attempts == 2  # the optimistic capacity was too small; the second call used the exact count


def test_fused_join_capacity_retry_reports_exact_count(cy):
    rng = np.random.default_rng(5)
    l = [orc.col(rng.integers(0, 50, 2000).astype(np.int64))]
    r = [orc.col(rng.integers(0, 50, 2000).astype(np.int64))]
    assert _check(cy, l, r, 0, 0, capacity=10) == 2


def test_fused_supported_predicate(cy):
    import ctypes as C
    a = [cy.DCol.from_numpy(np.arange(4, dtype=np.int64)), cy.DCol.from_numpy(np.arange(4, dtype=np.float64))]
    s = [cy.DCol.from_numpy(np.arange(4, dtype=np.int64)), cy.DCol.from_numpy(["a", "b", "c", "d"])]
    f = cy.lib().cyl_hash_join_fused_supported
    assert f(cy.c_columns(a), 2, 0, cy.c_columns(a), 2, 0, orc.INNER) == 1
    assert f(cy.c_columns(a), 2, 0, cy.c_columns(a), 2, 0, orc.LEFT) == 0      # outer joins: cyl_join_fused (join_fused8x.cu)
    assert f(cy.c_columns(a), 2, 0, cy.c_columns(s), 2, 0, orc.INNER) == 0     # strings take the general path
    assert f(cy.c_columns(a), 2, 0, cy.c_columns(a), 2, 1, orc.INNER) == 0     # key types differ


def test_table_join_takes_fused_path_and_matches_general_path():
    from cylon_b200 import api
    ctx = api.CylonContext(device=0)
    n = 400_000
    lk, lv = synth_kv(n, n, 11)
    rk, rv = synth_kv(n, n, 12)
    l = [orc.col(lk, name="k"), orc.col(lv, name="v")]
    r = [orc.col(rk, name="k"), orc.col(rv, name="v")]
    L, R = from_orc(api, ctx, l), from_orc(api, ctx, r)
    fused = L.join(R, "inner", "hash", [0], [0], "l_", "r_")
    ctx.add_config("cylon.join.fused", "0")  # the general path: LSD radix sort + merge, index vectors, gather
    general = L.join(R, "inner", "sort", [0], [0], "l_", "r_")
    ctx.add_config("cylon.join.fused", "1")
    assert fused.row_count == general.row_count
    assert fused.checksum() == general.checksum()
    m_ref, cs_ref = orc.join_i64_f64(lk, lv, rk, rv, orc.INNER, orc.HASH)
    assert fused.row_count == m_ref and fused.checksum() == cs_ref
    assert fused.column_names == ["l_k", "l_v", "r_k", "r_v"]
    ctx.finalize()


@pytest.mark.parametrize("skew", [False, True])
def test_fused_join_large(skew):
    """7M x 7M rows (32768 bins, shared-memory histogram + global cursors; with `skew` one bin is far above the
    shared-memory capacity and is probed in build chunks): result checked against the oracle by row count +
    order-independent checksum, and against the GPU sort join (independent code path)."""
    from cylon_b200 import api
    ctx = api.CylonContext(device=0)
    n = 7_000_000
    lk, lv = synth_kv(n, n, 21)
    rk, rv = synth_kv(n, n, 22)
    if skew:
        lk[: n // 3] = 12345  # one hot key on the build side: a coarse bin far above the average, and build chunks in the probe
        rk[:50] = 12345
    l = [orc.col(lk, name="k"), orc.col(lv, name="v")]
    r = [orc.col(rk, name="k"), orc.col(rv, name="v")]
    L, R = from_orc(api, ctx, l), from_orc(api, ctx, r)
    fused = L.join(R, "inner", "hash", [0], [0])
    m_ref, cs_ref = orc.join_i64_f64(lk, lv, rk, rv, orc.INNER, orc.HASH)
    assert fused.row_count == m_ref
    assert fused.checksum() == cs_ref
    ctx.add_config("cylon.join.fused", "0")
    general = L.join(R, "inner", "sort", [0], [0])
    ctx.add_config("cylon.join.fused", "1")
    assert general.row_count == m_ref and general.checksum() == cs_ref
    ctx.finalize()


@pytest.mark.parametrize("wl,wr,fkey", [(1, 1, False), (2, 2, True), (3, 2, False), (4, 4, False), (2, 3, True)])
def test_fused_join_two_pass_partition_widths(cy, wl, wr, fkey):
    """Tables above 2^21 rows take the two-pass staged radix partition (coarse digit, then fine digit); every record
    width 1-4 and a double key, ragged sizes, against the oracle's rows."""
    rng = np.random.default_rng(wl * 10 + wr)
    nl, nr = 2_300_017, 2_100_003
    dom = 3_000_000

    def table(n, w):
        k = rng.integers(0, dom, n)
        key = orc.col(k.astype(np.float64) - 0.5 * (k % 3 == 0)) if fkey else orc.col(k.astype(np.int64) - dom // 2)
        return [key] + [orc.col(rng.random(n)) if i % 2 == 0 else orc.col(rng.integers(-2 ** 60, 2 ** 60, n).astype(np.int64)) for i in range(w - 1)]

    l, r = table(nl, wl), table(nr, wr)
    # key last on the left for wide tables: exercises the key-position select inside the record
    if wl > 1:
        l = l[1:] + l[:1]
    _check(cy, l, r, wl - 1 if wl > 1 else 0, 0)


@pytest.mark.parametrize("bits", [9, 13, 16, 17, 18])
def test_fused_join_forced_bin_counts(cy, bits):
    """The many-bin code paths with a table that fits a test: 2^9 (5+4-bit digits) .. 2^18 bins (9+9-bit digits, i.e.
    512-way passes), 2^16 (histogram counted in two half-range launches), 2^17+ (histogram by global reductions)."""
    rng = np.random.default_rng(bits)
    nl, nr, dom = 2_200_013, 2_150_001, 2_500_000
    l = [orc.col(rng.integers(0, dom, nl).astype(np.int64)), orc.col(rng.random(nl))]
    r = [orc.col(rng.integers(0, dom, nr).astype(np.int64)), orc.col(rng.random(nr))]
    cy.check(cy.lib().cyl_hash_join_fused_set_bins_log2(bits))
    try:
        _check(cy, l, r, 0, 0)
    finally:
        cy.check(cy.lib().cyl_hash_join_fused_set_bins_log2(0))
    assert cy.lib().cyl_hash_join_fused_set_bins_log2(19) != 0


@pytest.mark.parametrize("nl,nr,dom,second_stream", [(0, 5, 3, False), (1000, 700, 300, False), (300007, 250003, 200000, True), (4000, 4000, 7, True)])
def test_fused_join_two_step_matches_one_call(cy, nl, nr, dom, second_stream):
    """cyl_hash_join_fused_partition + _probe (what DistributedJoin overlaps with the second shuffle): the same row
    multiset as the oracle, also with the right table partitioned on another stream and with a capacity retry (dom=7)."""
    import torch
    rng = np.random.default_rng(nl * 3 + nr)
    l = [orc.col(rng.integers(0, dom, nl).astype(np.int64)), orc.col(rng.random(nl)), orc.col(rng.integers(0, 2 ** 50, nl).astype(np.uint64))]
    r = [orc.col(rng.random(nr)), orc.col(rng.integers(0, dom, nr).astype(np.int64))]
    other = torch.cuda.Stream() if second_stream else None
    cols, m, attempts = cy.hash_join_fused_two_step(_to_dev(cy, l), 0, _to_dev(cy, r), 1, right_stream=other.cuda_stream if other else None)
    exp = orc.join(l, r, [0], [1], orc.INNER, orc.HASH)
    assert m == exp[0].length
    assert_same_rows(_out_table(cy, cols), exp)
    if dom == 7:
        assert attempts == 2

