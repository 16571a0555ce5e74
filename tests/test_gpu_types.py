"""Type coverage of the reference's factories on the GPU: temporal types (hashed like integers,
arrow_partition_kernels.cpp:323-327) and FIXED_SIZE_BINARY (murmur3 over byte_width bytes, :183-241; fixed-binary splitter
arrow_kernels.cpp:74-113; copy_arrray.cpp:87-110), through hashing, partitioning, split, joins, set operations, unique
and group-by keys. Bit-exact against the oracle."""
import numpy as np
import pytest

from oracle import pyoracle as orc

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def cy():
    from cylon_b200 import capi
    capi.init(0)
    return capi


def fsb_col(rng, n, width, dom):
    vocab = rng.integers(0, 256, (dom, width), dtype=np.uint8)
    data = vocab[rng.integers(0, dom, n)].reshape(-1)
    return orc.HCol(orc.FIXED_SIZE_BINARY, data, byte_width=width)


def dev(cy, c):
    if c.type == orc.FIXED_SIZE_BINARY:
        return cy.DCol(c.type, c.length, cy.DeviceBuffer.from_numpy(c.data), byte_width=c.byte_width)
    if c.type in (orc.STRING, orc.BINARY):
        return cy.DCol.from_numpy(c.values(), type_id=c.type)
    return cy.DCol.from_numpy(c.data, type_id=c.type)


@pytest.mark.parametrize("type_id,dtype", [(orc.DATE32, np.int32), (orc.DATE64, np.int64), (orc.TIMESTAMP, np.int64),
                                           (orc.TIME32, np.int32), (orc.TIME64, np.int64)])
def test_temporal_types_hash_like_integers(cy, type_id, dtype):
    rng = np.random.default_rng(type_id)
    c = orc.HCol(type_id, rng.integers(-10 ** 9, 10 ** 9, 5003).astype(dtype))
    assert np.array_equal(cy.row_hash([dev(cy, c)]), orc.row_hash([c]))
    pid, hist = cy.hash_partition([dev(cy, c)], 6)
    epid, ehist = orc.hash_partition([c], 6)
    assert np.array_equal(pid, epid) and np.array_equal(hist, ehist)
    # and equal to the same values typed as plain integers
    plain = orc.col(c.data)
    assert np.array_equal(orc.row_hash([plain]), orc.row_hash([c]))


@pytest.mark.parametrize("width", [1, 5, 8, 12, 16, 33])
def test_fixed_size_binary_hash_partition_split(cy, width):
    rng = np.random.default_rng(width)
    n = 20011
    c = fsb_col(rng, n, width, 500)
    assert np.array_equal(cy.row_hash([dev(cy, c)]), orc.row_hash([c]))
    for P in (2, 7, 8):
        pid, hist = cy.hash_partition([dev(cy, c)], P)
        epid, ehist = orc.hash_partition([c], P)
        assert np.array_equal(pid, epid) and np.array_equal(hist, ehist)
    # stable split of the column itself (generic-width scatter)
    import ctypes as C
    d = dev(cy, c)
    pid_buf, dest, offs = cy.partition_map([d], 4)
    out = cy.DeviceBuffer(n * width)
    cy.check(cy.lib().cyl_scatter_fixed(C.c_void_p(d.data.ptr), C.c_int32(width), C.c_void_p(dest.ptr), C.c_int64(n), C.c_void_p(out.ptr), None))
    perm, _ = orc.split_indices(orc.hash_partition([c], 4)[0], 4)
    assert np.array_equal(out.to_numpy(np.uint8, n * width).reshape(n, width), c.data.reshape(n, width)[perm])


@pytest.mark.parametrize("width", [4, 12])
@pytest.mark.parametrize("algo", [orc.HASH, orc.SORT])
@pytest.mark.parametrize("jt", [orc.INNER, orc.FULL_OUTER])
def test_fixed_size_binary_join_keys(cy, width, algo, jt):
    rng = np.random.default_rng(width * 10 + algo)
    vocab_rng = np.random.default_rng(1)
    vocab = vocab_rng.integers(0, 256, (300, width), dtype=np.uint8)
    def keycol(n):
        return orc.HCol(orc.FIXED_SIZE_BINARY, vocab[rng.integers(0, 300, n)].reshape(-1), byte_width=width)
    l, r = keycol(2000), keycol(1500)
    li, ri = cy.join_indices([dev(cy, l)], [dev(cy, r)], jt, algo)
    eli, eri = orc.hash_join_indices([l], [r], jt)
    assert sorted(zip(li.tolist(), ri.tolist())) == sorted(zip(eli.tolist(), eri.tolist()))
    # gather of a fixed-size binary column (copy_arrray.cpp:87-110)
    idx = np.where(li >= 0, li, -1).astype(np.int64)
    got = cy.gather(dev(cy, l), cy.DeviceBuffer.from_numpy(idx), len(idx))
    data, _o, valid = got.to_host()
    exp = l.take(idx)
    assert np.array_equal(valid, idx >= 0) and np.array_equal(data, exp.data)


def test_fixed_size_binary_setops_unique_groupby(cy):
    rng = np.random.default_rng(9)
    vocab = rng.integers(0, 256, (200, 10), dtype=np.uint8)
    def tab(n):
        k = rng.integers(0, 200, n)
        return [orc.HCol(orc.FIXED_SIZE_BINARY, vocab[k].reshape(-1), byte_width=10), orc.col((k % 3).astype(np.int32))]
    l, r = tab(3000), tab(2500)
    for op in (orc.UNION, orc.SUBTRACT, orc.INTERSECT):
        ml, mr = cy.rowset_mask([dev(cy, c) for c in l], [dev(cy, c) for c in r], op)
        eml, emr = orc.setop_mask(l, r, op)
        assert np.array_equal(ml, eml)
        if op == orc.UNION:
            assert np.array_equal(mr, emr)
    m, _ = cy.rowset_mask([dev(cy, l[0])], None, orc.UNIQUE)
    assert np.array_equal(m, orc.unique_mask([l[0]]))
    gid, first, g = cy.groupby_hash([dev(cy, l[0])])
    egid, efirst, eg = orc.groupby_ids([l[0]])
    assert g == eg and np.array_equal(gid, egid) and np.array_equal(first, efirst)


def test_unsupported_types_are_refused(cy):
    # BOOL is bit-packed in Arrow and has no splitter in the reference either (arrow_kernels.cpp:172-195); HALF_FLOAT
    # has no hash kernel (arrow_partition_kernels.cpp:307-330): both fail loudly with ExecutionError
    import ctypes as C
    for t in (1, 10):
        c = cy.DCol(t, 8, cy.DeviceBuffer(64), byte_width=2)
        out = cy.DeviceBuffer(64)
        rc = cy.lib().cyl_row_hash(cy.c_columns([c]), 1, C.c_int64(8), C.c_void_p(out.ptr), None)
        assert rc == 42
