"""Parity of the CUDA kernels (called through the C-ABI, ctypes) against the CPU oracle. Bit-exact:
everything here is integer / byte / index work."""
import numpy as np
import pytest

from oracle import pyoracle as orc

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def cy():
    from cylon_b200 import capi
    capi.init(0)
    return capi


def rand_strings(rng, n, lo=0, hi=24, vocab=None):
    if vocab is not None:
        return [vocab[i] for i in rng.integers(0, len(vocab), n)]
    lens = rng.integers(lo, hi + 1, n)
    return [bytes(rng.integers(97, 123, l, dtype=np.uint8)) for l in lens]


def make_cols(rng, n, kinds):
    """-> (oracle cols, host value lists)"""
    ocols = []
    for k in kinds:
        if k == "str":
            ocols.append(orc.col(rand_strings(rng, n)))
        elif k in ("float32", "float64"):
            ocols.append(orc.col(rng.standard_normal(n).astype(k)))
        else:
            info = np.iinfo(k)
            ocols.append(orc.col(rng.integers(info.min, info.max, n, dtype=k, endpoint=True)))
    return ocols


def to_dev(cy, ocols):
    out = []
    for c in ocols:
        if c.type in (orc.STRING, orc.BINARY):
            out.append(cy.DCol.from_numpy(c.values(), type_id=c.type))
        else:
            out.append(cy.DCol.from_numpy(c.data, type_id=c.type))
    return out


ALL_KINDS = ["int8", "uint8", "int16", "uint16", "int32", "uint32", "int64", "uint64", "float32", "float64", "str"]


@pytest.mark.parametrize("n", [0, 1, 31, 1000, 70001])
def test_row_hash_all_types(cy, n):
    rng = np.random.default_rng(n + 1)
    for kind in ALL_KINDS:
        oc = make_cols(rng, n, [kind])
        if n == 0:
            continue
        assert np.array_equal(cy.row_hash(to_dev(cy, oc)), orc.row_hash(oc)), kind
    oc = make_cols(rng, n, ["int64", "int32", "float64", "str", "int8"])
    if n:
        assert np.array_equal(cy.row_hash(to_dev(cy, oc)), orc.row_hash(oc))


def test_row_hash_known_answers(cy):
    # SURVEY.md 8(c) vectors, on the device
    c = [cy.DCol.from_numpy(np.array([123456789012], dtype=np.int64)), cy.DCol.from_numpy(np.array([7], dtype=np.int32))]
    assert cy.row_hash(c)[0] == 344598643
    c = [cy.DCol.from_numpy(np.array([26], dtype=np.int64)), cy.DCol.from_numpy(np.array([0.394]))]
    assert cy.row_hash(c)[0] == 3545837400
    assert list(cy.row_hash([cy.DCol.from_numpy(["", "a", "abc", "hello world"])])) == [0, 1009084850, 3017643002, 1586663183]
    assert cy.row_hash([cy.DCol.from_numpy(np.array([-1], dtype=np.int8))])[0] == 0xFFFFFFFF


@pytest.mark.parametrize("P", [1, 2, 3, 4, 7, 8, 64, 1000])
@pytest.mark.parametrize("kinds", [["int64"], ["float64"], ["str"], ["int64", "int32"], ["int32", "float64", "str"]])
def test_hash_partition(cy, P, kinds):
    rng = np.random.default_rng(P * 13 + len(kinds))
    n = 50021
    oc = make_cols(rng, n, kinds)
    pid, hist = cy.hash_partition(to_dev(cy, oc), P)
    epid, ehist = orc.hash_partition(oc, P)
    assert np.array_equal(pid, epid)
    assert np.array_equal(hist, ehist)


def test_hash_partition_identities(cy):
    # cpp/test/partition_test.cpp:32-79 on the device
    import struct
    rows = 1000
    c0 = cy.DCol.from_numpy(np.arange(rows, dtype=np.int32))
    c1 = cy.DCol.from_numpy(np.arange(rows, dtype=np.float64) + 10.0)
    for w in (1, 2, 3, 4, 8):
        pid, hist = cy.hash_partition([c0], w)
        assert list(pid) == [i % w for i in range(rows)] and hist.sum() == rows
        pid, hist = cy.hash_partition([c1], w)
        assert list(pid) == [orc.murmur3_x86_32(struct.pack("<d", i + 10.0)) % w for i in range(rows)]


@pytest.mark.parametrize("n", [0, 1, 5000, 300007])
@pytest.mark.parametrize("P", [1, 2, 8, 5, 256, 1024])
def test_scatter_map_is_stable_split(cy, n, P):
    rng = np.random.default_rng(n + P)
    pid = rng.integers(0, P, n).astype(np.uint32)
    if n > 100:
        pid[: n // 3] = 0  # skew
    dest, offs = cy.scatter_map(cy.DeviceBuffer.from_numpy(pid), n, P)
    perm, eoffs = orc.split_indices(pid, P)
    assert np.array_equal(offs.to_numpy(np.int64, P + 1), eoffs)
    d = dest.to_numpy(np.int64, n)
    exp_dest = np.empty(n, dtype=np.int64)
    exp_dest[perm] = np.arange(n)
    assert np.array_equal(d, exp_dest)


def test_scatter_columns(cy):
    import ctypes as C
    rng = np.random.default_rng(5)
    n, P = 120001, 8
    oc = make_cols(rng, n, ["int64", "float64", "int32", "int8", "str"])
    dc = to_dev(cy, oc)
    pid_buf, dest, offs = cy.partition_map([dc[0]], P)
    epid, _ = orc.hash_partition([oc[0]], P)
    assert np.array_equal(pid_buf.to_numpy(np.uint32, n), epid)
    perm, eoffs = orc.split_indices(epid, P)
    assert np.array_equal(offs.to_numpy(np.int64, P + 1), eoffs)
    for o, d in zip(oc[:4], dc[:4]):
        out = cy.DeviceBuffer(max(1, n * o.byte_width))
        cy.check(cy.lib().cyl_scatter_fixed(C.c_void_p(d.data.ptr), o.byte_width, C.c_void_p(dest.ptr), C.c_int64(n),
                                            C.c_void_p(out.ptr), None))
        assert np.array_equal(out.to_numpy(o.data.dtype, n), o.data[perm])
    # variable width
    s = dc[4]
    doffs = cy.DeviceBuffer(4 * (n + 1))
    ddata = cy.DeviceBuffer(max(1, s.data.nbytes))
    cy.check(cy.lib().cyl_scatter_varwidth(C.c_void_p(s.offsets.ptr), C.c_void_p(s.data.ptr), C.c_void_p(dest.ptr),
                                           C.c_int64(n), C.c_void_p(doffs.ptr), C.c_void_p(ddata.ptr), None))
    exp = oc[4].take(perm)
    assert np.array_equal(doffs.to_numpy(np.int32, n + 1), exp.offsets)
    assert np.array_equal(ddata.to_numpy(np.uint8, len(exp.data)), exp.data)
    # validity
    valid = rng.random(n) < 0.7
    vsrc = cy.DeviceBuffer.from_numpy(np.concatenate([np.packbits(valid, bitorder="little"), np.zeros(8, np.uint8)]))
    vdst = cy.DeviceBuffer(4 * ((n + 31) // 32) + 4)
    cy.check(cy.lib().cyl_scatter_validity(C.c_void_p(vsrc.ptr), C.c_void_p(dest.ptr), C.c_int64(n), C.c_void_p(vdst.ptr), None))
    got = np.unpackbits(vdst.to_numpy(np.uint8, (n + 7) // 8), bitorder="little")[:n].astype(bool)
    assert np.array_equal(got, valid[perm])


@pytest.mark.parametrize("P", [2, 8, 3])
def test_partition_scatter_fixed_fused(cy, P):
    import ctypes as C
    rng = np.random.default_rng(17 + P)
    n = 200003
    oc = make_cols(rng, n, ["int64", "float64", "int32"])
    dc = to_dev(cy, oc)
    outs = [cy.DeviceBuffer(max(1, n * o.byte_width)) for o in oc]
    offs = cy.DeviceBuffer(8 * (P + 1))
    keys = (C.c_int32 * 2)(0, 2)
    dst = (C.c_void_p * 3)(*[o.ptr for o in outs])
    cy.check(cy.lib().cyl_partition_scatter_fixed(cy.c_columns(dc), 3, keys, 2, C.c_int64(n), C.c_uint32(P), dst,
                                                  C.c_void_p(offs.ptr), None))
    epid, _ = orc.hash_partition([oc[0], oc[2]], P)
    perm, eoffs = orc.split_indices(epid, P)
    assert np.array_equal(offs.to_numpy(np.int64, P + 1), eoffs)
    for o, b in zip(oc, outs):
        assert np.array_equal(b.to_numpy(o.data.dtype, n), o.data[perm])


@pytest.mark.parametrize("n", [0, 1, 4095, 4096, 4097, 1000003])
def test_exclusive_scan(cy, n):
    rng = np.random.default_rng(n)
    a = rng.integers(0, 1000, n).astype(np.int64)
    exp = np.zeros(n + 1, dtype=np.int64)
    np.cumsum(a, out=exp[1:])
    assert np.array_equal(cy.exclusive_scan(a), exp)


def _pairs(li, ri):
    return sorted(zip(li.tolist(), ri.tolist()))


KEYSETS = {
    "int64": lambda rng, n, dom: [orc.col(rng.integers(-dom, dom, n).astype(np.int64))],
    "int32": lambda rng, n, dom: [orc.col(rng.integers(-dom, dom, n).astype(np.int32))],
    "uint8": lambda rng, n, dom: [orc.col(rng.integers(0, min(dom, 255), n).astype(np.uint8))],
    "float64": lambda rng, n, dom: [orc.col(rng.integers(1, dom + 1, n).astype(np.float64) * 0.25)],
    "str": lambda rng, n, dom: [orc.col([b"k%d" % v for v in rng.integers(0, dom, n)])],
    "int64_int32": lambda rng, n, dom: [orc.col(rng.integers(0, dom, n).astype(np.int64)),
                                        orc.col(rng.integers(0, 3, n).astype(np.int32))],
    "str_f64": lambda rng, n, dom: [orc.col([b"key-%d" % v for v in rng.integers(0, dom, n)]),
                                    orc.col(rng.integers(0, 2, n).astype(np.float64))],
}


@pytest.mark.parametrize("algo", [orc.HASH, orc.SORT])
@pytest.mark.parametrize("jt", [orc.INNER, orc.LEFT, orc.RIGHT, orc.FULL_OUTER])
@pytest.mark.parametrize("keyset", sorted(KEYSETS))
def test_join_indices_match_oracle(cy, algo, jt, keyset):
    import zlib
    rng = np.random.default_rng(zlib.crc32(repr((algo, jt, keyset)).encode()))
    nl, nr, dom = 3001, 2503, 700
    lk, rk = KEYSETS[keyset](rng, nl, dom), KEYSETS[keyset](rng, nr, dom)
    li, ri = cy.join_indices(to_dev(cy, lk), to_dev(cy, rk), jt, algo)
    eli, eri = (orc.hash_join_indices if algo == orc.HASH else orc.sort_join_indices)(lk, rk, jt)
    assert len(li) == len(eli)
    assert _pairs(li, ri) == _pairs(eli, eri)


@pytest.mark.parametrize("algo", [orc.HASH, orc.SORT])
@pytest.mark.parametrize("jt", [orc.INNER, orc.LEFT, orc.RIGHT, orc.FULL_OUTER])
@pytest.mark.parametrize("nl,nr", [(0, 0), (0, 17), (17, 0), (1, 1), (5000, 3)])
def test_join_edge_sizes(cy, algo, jt, nl, nr):
    rng = np.random.default_rng(nl * 31 + nr)
    lk = [orc.col(rng.integers(0, 4, nl).astype(np.int64))]
    rk = [orc.col(rng.integers(0, 4, nr).astype(np.int64))]
    li, ri = cy.join_indices(to_dev(cy, lk), to_dev(cy, rk), jt, algo)
    eli, eri = orc.hash_join_indices(lk, rk, jt)
    assert _pairs(li, ri) == _pairs(eli, eri)


@pytest.mark.parametrize("algo", [orc.HASH, orc.SORT])
def test_join_heavy_duplicates_and_special_keys(cy, algo):
    # one hot key on both sides (cross product), plus -1 / INT64_MIN / INT64_MAX keys (the all-ones tag has its own slot)
    rng = np.random.default_rng(3)
    special = np.array([-1, np.iinfo(np.int64).min, np.iinfo(np.int64).max, 0], dtype=np.int64)
    lk = np.concatenate([np.full(300, 42, np.int64), special, rng.integers(0, 50, 500).astype(np.int64), special])
    rk = np.concatenate([np.full(200, 42, np.int64), special, rng.integers(25, 75, 400).astype(np.int64)])
    for jt in (orc.INNER, orc.FULL_OUTER):
        li, ri = cy.join_indices(to_dev(cy, [orc.col(lk)]), to_dev(cy, [orc.col(rk)]), jt, algo)
        eli, eri = orc.hash_join_indices([orc.col(lk)], [orc.col(rk)], jt)
        assert _pairs(li, ri) == _pairs(eli, eri)


def test_join_type_mismatch_is_invalid(cy):
    l = [cy.DCol.from_numpy(np.arange(4, dtype=np.int64))]
    r = [cy.DCol.from_numpy(np.arange(4, dtype=np.int32))]
    with pytest.raises(cy.CylonError) as e:
        cy.join_indices(l, r, orc.INNER, orc.HASH)
    assert e.value.code == 4  # Code::Invalid, join/hash_join.cpp:195-197


@pytest.mark.parametrize("kind", ["int64", "float64", "int32", "int16", "uint8", "str"])
def test_gather_matches_copy_array_by_indices(cy, kind):
    rng = np.random.default_rng(11)
    n, m = 5003, 9001
    oc = make_cols(rng, n, [kind])[0]
    idx = rng.integers(-1, n, m).astype(np.int64)
    idx[idx < 0] = -1
    got = cy.gather(to_dev(cy, [oc])[0], cy.DeviceBuffer.from_numpy(idx), m)
    data, offs, valid = got.to_host()
    exp = oc.take(idx)
    assert np.array_equal(valid, idx >= 0)
    if kind == "str":
        assert np.array_equal(offs, exp.offsets) and np.array_equal(data, exp.data)
    else:
        assert np.array_equal(data, exp.data)


def test_table_checksum_matches_oracle(cy):
    rng = np.random.default_rng(23)
    n = 40001
    oc = make_cols(rng, n, ["int64", "float64", "int32", "str", "int8"])
    assert cy.table_checksum(to_dev(cy, oc)) == orc.table_checksum(oc)
    # with nulls
    valid = rng.random(n) < 0.5
    oc[1].validity = np.packbits(valid, bitorder="little")
    dc = to_dev(cy, oc)
    dc[1] = cy.DCol.from_numpy(oc[1].data, validity=valid)
    assert cy.table_checksum(dc) == orc.table_checksum(oc)


def test_large_join_checksum_and_properties(cy):
    """1M x 1M inner hash join (BASELINE config 1 shape): pair multiset checked through a checksum of
    checksums against the oracle, and row count against the independent sort join."""
    from tests.helpers import synth_kv
    n = 1_000_000
    lk, lv = synth_kv(n, 990_000, seed=1)
    rk, rv = synth_kv(n, 990_000, seed=2)
    dl, dr = [cy.DCol.from_numpy(lk)], [cy.DCol.from_numpy(rk)]
    li, ri, m = cy.join_indices_device(dl, dr, orc.INNER, orc.HASH)
    out = [cy.gather(cy.DCol.from_numpy(lk), li, m), cy.gather(cy.DCol.from_numpy(lv), li, m),
           cy.gather(cy.DCol.from_numpy(rk), ri, m), cy.gather(cy.DCol.from_numpy(rv), ri, m)]
    for c in out:
        c.validity = None  # inner join: all valid
    m_ref, cs_ref = orc.join_i64_f64(lk, lv, rk, rv, orc.INNER, orc.HASH)
    assert m == m_ref
    assert cy.table_checksum(out) == cs_ref
    _, _, m2 = cy.join_indices_device(dl, dr, orc.INNER, orc.SORT)
    assert m2 == m_ref
