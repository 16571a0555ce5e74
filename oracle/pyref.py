"""ctypes binding of oracle/_ref/libcylon_ref.so: the reference's OWN code (hash partition kernels, hash join, sort
join, build_final_table) compiled in place from /root/reference against pyarrow's Arrow C++. TEST INFRASTRUCTURE
ONLY -- it pins the oracle (tests/test_oracle_vs_reference.py) and serves as the "reference" CPU baseline of bench.py.
"""
import ctypes as C
import os
import time

import numpy as np

from . import pyoracle as orc

_PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref", "libcylon_ref.so")
_lib = None


def available():
    """True when the compiled reference exists AND loads here (it links pyarrow's libarrow by rpath; on a machine where
    that path differs the tests that need it skip instead of erroring)."""
    if not os.path.exists(_PATH):
        return False
    try:
        lib()
        return True
    except OSError:
        return False


def lib():
    global _lib
    if _lib is None:
        _lib = C.CDLL(_PATH)
        _lib.ref_table_rows.restype = C.c_int64
        _lib.ref_table_col_bytes.restype = C.c_int64
    return _lib


class RefColIn(C.Structure):
    _fields_ = [("type", C.c_int32), ("width", C.c_int32), ("n", C.c_int64), ("offsets", C.c_void_p), ("data", C.c_void_p),
                ("validity", C.c_void_p)]


def _args(col):
    data = col.data if col.data.size else np.zeros(1, dtype=col.data.dtype)
    offs = col.offsets.ctypes.data if col.offsets is not None else None
    return (C.c_int32(col.type), C.c_int32(col.byte_width), C.c_int64(col.length), C.c_void_p(offs), C.c_void_p(data.ctypes.data)), data


def row_hash(cols):
    h = np.zeros(max(1, cols[0].length), dtype=np.uint32)
    for c in cols:
        a, _k = _args(c)
        rc = lib().ref_update_hash(*a, h.ctypes.data_as(C.c_void_p))
        assert rc == 0, rc
    return h[: cols[0].length]


def hash_partition(cols, num_partitions):
    n = cols[0].length
    pid = np.zeros(max(1, n), dtype=np.uint32)
    hist = np.zeros(num_partitions, dtype=np.uint32)
    for c in cols[:-1]:
        a, _k = _args(c)
        assert lib().ref_update_hash(*a, pid.ctypes.data_as(C.c_void_p)) == 0
    a, _k = _args(cols[-1])
    assert lib().ref_partition(*a, C.c_uint32(num_partitions), pid.ctypes.data_as(C.c_void_p), hist.ctypes.data_as(C.c_void_p)) == 0
    return pid[:n], hist.astype(np.uint64)


def array_index_hash_join(lkey, rkey, join_type):
    la, _k1 = _args(lkey)
    ra, _k2 = _args(rkey)
    lp, rp, m = C.c_void_p(), C.c_void_p(), C.c_int64()
    rc = lib().ref_array_index_hash_join(la[0], la[1], la[2], la[3], la[4], ra[2], ra[3], ra[4], C.c_int32(join_type), C.byref(lp),
                                         C.byref(rp), C.byref(m))
    assert rc == 0, rc
    n = m.value
    li = np.ctypeslib.as_array(C.cast(lp, C.POINTER(C.c_int64)), shape=(max(1, n),))[:n].copy()
    ri = np.ctypeslib.as_array(C.cast(rp, C.POINTER(C.c_int64)), shape=(max(1, n),))[:n].copy()
    lib().ref_free(lp)
    lib().ref_free(rp)
    return li, ri


def _table_in(cols):
    arr = (RefColIn * len(cols))()
    keep = []
    for i, c in enumerate(cols):
        data = c.data if c.data.size else np.zeros(1, dtype=c.data.dtype)
        keep.append(data)
        arr[i].type, arr[i].width, arr[i].n = c.type, c.byte_width, c.length
        arr[i].offsets = c.offsets.ctypes.data if c.offsets is not None else None
        arr[i].data = data.ctypes.data
        arr[i].validity = None
        if c.validity is not None:
            v = np.ascontiguousarray(c.validity, dtype=np.uint8)
            keep.append(v)
            arr[i].validity = v.ctypes.data
    return arr, keep


def join_tables(left, right, lkeys, rkeys, join_type, algorithm):
    """cylon::join::JoinTables of the reference -> oracle-style table (list of HCol)."""
    la, _k1 = _table_in(left)
    ra, _k2 = _table_in(right)
    lk = (C.c_int32 * len(lkeys))(*lkeys)
    rk = (C.c_int32 * len(rkeys))(*rkeys)
    h = C.c_void_p()
    rc = lib().ref_join_tables(la, len(left), ra, len(right), lk, rk, len(lkeys), C.c_int32(join_type), C.c_int32(algorithm), C.byref(h))
    assert rc == 0, rc
    n = lib().ref_table_rows(h)
    out = []
    for c, src in enumerate(list(left) + list(right)):
        nbytes = lib().ref_table_col_bytes(h, c)
        valid = np.ones(max(1, n), dtype=np.uint8)
        if src.type in (orc.STRING, orc.BINARY):
            data = np.zeros(max(1, nbytes), dtype=np.uint8)
            offs = np.zeros(n + 1, dtype=np.int32)
            lib().ref_table_copy_col(h, c, C.c_void_p(data.ctypes.data), C.c_void_p(offs.ctypes.data), C.c_void_p(valid.ctypes.data))
            col = orc.HCol(src.type, data[:nbytes], offs)
        else:
            data = np.zeros(max(1, n), dtype=src.data.dtype)
            lib().ref_table_copy_col(h, c, C.c_void_p(data.ctypes.data), None, C.c_void_p(valid.ctypes.data))
            col = orc.HCol(src.type, data[:n])
        v = valid[:n].astype(bool)
        if not v.all():
            col.validity = np.packbits(v, bitorder="little")
        out.append(col)
    lib().ref_table_free(h)
    return out


def _table_out(h):
    """ref_table handle -> oracle-style table (list of HCol with the reference's column names); frees the handle."""
    L = lib()
    L.ref_table_col_name.restype = C.c_char_p
    n, ncols = L.ref_table_rows(h), L.ref_table_cols(h)
    out = []
    for c in range(ncols):
        t = L.ref_table_col_type(h, c)
        name = L.ref_table_col_name(h, c).decode()
        nbytes = L.ref_table_col_bytes(h, c)
        valid = np.ones(max(1, n), dtype=np.uint8)
        if t in (orc.STRING, orc.BINARY):
            data = np.zeros(max(1, nbytes), dtype=np.uint8)
            offs = np.zeros(n + 1, dtype=np.int32)
            L.ref_table_copy_col(h, c, C.c_void_p(data.ctypes.data), C.c_void_p(offs.ctypes.data), C.c_void_p(valid.ctypes.data))
            col = orc.HCol(t, data[:nbytes], offs, name=name)
        else:
            data = np.zeros(max(1, n), dtype=orc._TYPE2NP[t])
            L.ref_table_copy_col(h, c, C.c_void_p(data.ctypes.data), None, C.c_void_p(valid.ctypes.data))
            col = orc.HCol(t, data[:n], name=name)
        v = valid[:n].astype(bool)
        if not v.all():
            col.validity = np.packbits(v, bitorder="little")
        out.append(col)
    L.ref_table_free(h)
    return out


DIST_JOIN, DIST_UNION, DIST_SUBTRACT, DIST_INTERSECT, DIST_UNIQUE, DIST_SHUFFLE, DIST_GROUPBY, DIST_SORT, MAP_SORT_PARTITIONS, LOCAL_SORT = range(10)


def dist_op(op, lefts, rights=None, p1=(), p2=(), join_type=0, algorithm=0, rows_only=False):
    """One Distributed* operator of the reference on len(lefts) ranks (threads inside this process, an in-memory
    Communicator/Channel in place of MPI). lefts / rights: one oracle-style table per rank. -> one table per rank."""
    world = len(lefts)
    nlc = len(lefts[0])
    flat_l = [c for t in lefts for c in t]
    la, _k1 = _table_in(flat_l)
    nrc, ra, _k2 = 0, None, None
    if rights is not None:
        nrc = len(rights[0])
        ra, _k2 = _table_in([c for t in rights for c in t])
    a1 = (C.c_int32 * max(1, len(p1)))(*p1)
    a2 = (C.c_int32 * max(1, len(p2)))(*p2)
    outs = (C.c_void_p * world)()
    t0 = time.perf_counter()
    rc = lib().ref_dist_op(C.c_int32(op), C.c_int32(world), la, C.c_int32(nlc), ra, C.c_int32(nrc), a1, C.c_int32(len(p1)), a2,
                           C.c_int32(len(p2)), C.c_int32(join_type), C.c_int32(algorithm), outs)
    seconds = time.perf_counter() - t0
    if rc != 0:
        raise RuntimeError("reference operator failed with cylon::Code %d" % rc)
    if rows_only:  # benchmark use: (rows per rank, seconds inside the reference call), results not copied out
        rows = []
        for r in range(world):
            h = C.c_void_p(outs[r])
            rows.append(lib().ref_table_rows(h))
            lib().ref_table_free(h)
        return rows, seconds
    return [_table_out(C.c_void_p(outs[r])) for r in range(world)]


def dist_available():
    return available() and hasattr(lib(), "ref_dist_op")


# ---- the reference's aggregation kernels (compute/aggregate_kernels.hpp, compiled by oracle/build_ref.sh) ------------
_AGG_PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref", "libref_aggregate.so")
_agg_lib = None
_AGG_DTYPES = {"int8": 0, "int16": 1, "int32": 2, "int64": 3, "uint8": 4, "uint16": 5, "uint32": 6, "uint64": 7, "float32": 8, "float64": 9}


def aggregate_available():
    if not os.path.exists(_AGG_PATH):
        return False
    try:
        C.CDLL(_AGG_PATH)
        return True
    except OSError:
        return False


def ref_aggregate(op, values, gid, num_groups):
    """SumKernel / MinKernel / MaxKernel / CountKernel of the reference driven as groupby/hash_groupby.cpp:142-201
    drives them. op = orc.SUM/MIN/MAX/COUNT (same values as the reference's AggregationOpId)."""
    global _agg_lib
    if _agg_lib is None:
        _agg_lib = C.CDLL(_AGG_PATH)
    values = np.ascontiguousarray(values)
    gid = np.ascontiguousarray(gid, dtype=np.int64)
    out = np.zeros(max(1, num_groups), dtype=np.int64 if op == orc.COUNT else values.dtype)
    v = values if values.size else np.zeros(1, dtype=values.dtype)
    g = gid if gid.size else np.zeros(1, dtype=np.int64)
    rc = _agg_lib.ref_aggregate(int(op), _AGG_DTYPES[values.dtype.name], v.ctypes.data_as(C.c_void_p), g.ctypes.data_as(C.c_void_p),
                                C.c_int64(len(values)), C.c_int64(num_groups), out.ctypes.data_as(C.c_void_p))
    if rc != 0:
        raise RuntimeError("ref_aggregate failed: %d" % rc)
    return out[:num_groups]
