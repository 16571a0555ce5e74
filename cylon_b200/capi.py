"""ctypes binding of libcylon_b200.so (the C-ABI declared in include/cylon_b200.h).

This is plumbing for tests, bench.py and Python callers: numpy arrays in, device buffers and kernel calls
through the C-ABI, numpy arrays out. There is no CPU fallback: if the CUDA library is missing or a call
fails, a CylonError is raised.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "lib", "libcylon_b200.so")

# arrow::Type ids (include/cylon_b200_types.h)
UINT8, INT8, UINT16, INT16, UINT32, INT32, UINT64, INT64 = 2, 3, 4, 5, 6, 7, 8, 9
FLOAT, DOUBLE, STRING, BINARY, FIXED_SIZE_BINARY = 11, 12, 13, 14, 15
DATE32, DATE64, TIMESTAMP, TIME32, TIME64 = 16, 17, 18, 19, 20
INNER, LEFT, RIGHT, FULL_OUTER = 0, 1, 2, 3
SORT, HASH = 0, 1
SUM, MIN, MAX, COUNT = 0, 1, 2, 3
UNION, SUBTRACT, INTERSECT, UNIQUE = 0, 1, 2, 3

NP2TYPE = {np.dtype(k): v for k, v in {
    "uint8": UINT8, "int8": INT8, "uint16": UINT16, "int16": INT16, "uint32": UINT32, "int32": INT32,
    "uint64": UINT64, "int64": INT64, "float32": FLOAT, "float64": DOUBLE}.items()}
TYPE2NP = {v: k for k, v in NP2TYPE.items()}
TYPE2NP.update({DATE32: np.dtype("int32"), DATE64: np.dtype("int64"), TIMESTAMP: np.dtype("int64"),
                TIME32: np.dtype("int32"), TIME64: np.dtype("int64")})


class CylonError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("cylon::Code %d: %s" % (code, msg))
        self.code = code


class CylColumn(C.Structure):
    _fields_ = [("type", C.c_int32), ("byte_width", C.c_int32), ("length", C.c_int64),
                ("null_count", C.c_int64), ("validity", C.c_void_p), ("offsets", C.c_void_p),
                ("data", C.c_void_p)]


_lib = None


def lib():
    """Loads the CUDA library; raises if it has not been built (no fallback path exists)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise CylonError(5, "CUDA library %s is missing: run `python -c 'import __graft_entry__ as g; g.build()'`"
                             % LIB_PATH)
        _lib = C.CDLL(LIB_PATH)
        _lib.cyl_last_error.restype = C.c_char_p
    return _lib


def check(rc):
    if rc != 0:
        raise CylonError(rc, lib().cyl_last_error().decode(errors="replace"))


def init(device=0):
    check(lib().cyl_init(C.c_int32(device)))


def device_count():
    n = C.c_int32()
    rc = lib().cyl_device_count(C.byref(n))
    return n.value if rc == 0 else 0


def sync(stream=None):
    check(lib().cyl_stream_sync(C.c_void_p(stream)))


class DeviceBuffer:
    """A cudaMalloc'ed buffer owned by Python."""

    def __init__(self, nbytes):
        self.nbytes = int(nbytes)
        p = C.c_void_p()
        check(lib().cyl_malloc(C.byref(p), C.c_int64(self.nbytes)))
        self.ptr = p.value

    @classmethod
    def from_numpy(cls, arr, stream=None):
        arr = np.ascontiguousarray(arr)
        buf = cls(arr.nbytes)
        if arr.nbytes:
            check(lib().cyl_memcpy_h2d(C.c_void_p(buf.ptr), C.c_void_p(arr.ctypes.data), C.c_int64(arr.nbytes),
                                       C.c_void_p(stream)))
            sync(stream)
        return buf

    def to_numpy(self, dtype, count=None, stream=None):
        dtype = np.dtype(dtype)
        if count is None:
            count = self.nbytes // dtype.itemsize
        out = np.empty(count, dtype=dtype)
        if out.nbytes:
            check(lib().cyl_memcpy_d2h(C.c_void_p(out.ctypes.data), C.c_void_p(self.ptr), C.c_int64(out.nbytes),
                                       C.c_void_p(stream)))
            sync(stream)
        return out

    def free(self):
        if self.ptr:
            lib().cyl_free(C.c_void_p(self.ptr))
            self.ptr = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


class DCol:
    """A device-resident column (Arrow layout: validity bitmap, int32 offsets, data)."""

    def __init__(self, type_id, length, data, offsets=None, validity=None, byte_width=None, null_count=0, name=""):
        self.type = int(type_id)
        self.length = int(length)
        self.data, self.offsets, self.validity = data, offsets, validity
        if self.type in (STRING, BINARY):
            self.byte_width = 0
        elif byte_width is not None:
            self.byte_width = int(byte_width)
        else:
            self.byte_width = TYPE2NP[self.type].itemsize
        self.null_count = null_count
        self.name = name

    @classmethod
    def from_numpy(cls, values, type_id=None, validity=None, name=""):
        """values: numeric numpy array, or a list of str/bytes (-> STRING)."""
        if isinstance(values, np.ndarray) and values.dtype.kind in "iuf":
            t = type_id if type_id is not None else NP2TYPE[values.dtype]
            v = None
            if validity is not None:
                v = DeviceBuffer.from_numpy(_pad4(np.packbits(np.asarray(validity, dtype=bool), bitorder="little")))
            return cls(t, len(values), DeviceBuffer.from_numpy(values), None, v, name=name)
        vals = [x.encode() if isinstance(x, str) else bytes(x) for x in values]
        offsets = np.zeros(len(vals) + 1, dtype=np.int32)
        if vals:
            offsets[1:] = np.cumsum([len(x) for x in vals])
        data = np.frombuffer(b"".join(vals), dtype=np.uint8)
        v = None
        if validity is not None:
            v = DeviceBuffer.from_numpy(_pad4(np.packbits(np.asarray(validity, dtype=bool), bitorder="little")))
        return cls(type_id if type_id is not None else STRING, len(vals), DeviceBuffer.from_numpy(data),
                   DeviceBuffer.from_numpy(offsets), v, name=name)

    def c_struct(self):
        s = CylColumn()
        s.type, s.byte_width, s.length, s.null_count = self.type, self.byte_width, self.length, self.null_count
        s.validity = self.validity.ptr if self.validity is not None else None
        s.offsets = self.offsets.ptr if self.offsets is not None else None
        s.data = self.data.ptr if self.data is not None else None
        return s

    def to_host(self):
        """-> (data ndarray, offsets ndarray | None, valid-mask ndarray | None)"""
        offs = None
        if self.type in (STRING, BINARY):
            offs = self.offsets.to_numpy(np.int32, self.length + 1)
            data = self.data.to_numpy(np.uint8, int(offs[-1]) if self.length else 0)
        elif self.type == FIXED_SIZE_BINARY:
            data = self.data.to_numpy(np.uint8, self.length * self.byte_width)
        else:
            data = self.data.to_numpy(TYPE2NP[self.type], self.length)
        valid = None
        if self.validity is not None:
            bits = self.validity.to_numpy(np.uint8, (self.length + 7) // 8)
            valid = np.unpackbits(bits, bitorder="little")[: self.length].astype(bool)
        return data, offs, valid


def _pad4(a):
    pad = (-len(a)) % 4
    return np.concatenate([a, np.zeros(pad + 4, dtype=np.uint8)])


def c_columns(cols):
    arr = (CylColumn * max(1, len(cols)))()
    for i, c in enumerate(cols):
        arr[i] = c.c_struct()
    return arr


def _p(buf):
    return C.c_void_p(buf.ptr if buf is not None else None)


# ------------------------------------------------------------------------------------------------------
# thin call wrappers (each returns numpy / DeviceBuffer objects); `stream` is a raw cudaStream_t or None
# ------------------------------------------------------------------------------------------------------
def row_hash(cols, stream=None):
    n = cols[0].length
    out = DeviceBuffer(4 * max(1, n))
    check(lib().cyl_row_hash(c_columns(cols), len(cols), C.c_int64(n), _p(out), C.c_void_p(stream)))
    return out.to_numpy(np.uint32, n)


def hash_partition(cols, num_partitions, stream=None, want_pid=True):
    n = cols[0].length
    pid = DeviceBuffer(4 * max(1, n)) if want_pid else None
    hist = DeviceBuffer(8 * num_partitions)
    check(lib().cyl_hash_partition(c_columns(cols), len(cols), C.c_int64(n), C.c_uint32(num_partitions), _p(pid),
                                   _p(hist), C.c_void_p(stream)))
    return (pid.to_numpy(np.uint32, n) if want_pid else None), hist.to_numpy(np.uint64, num_partitions)


def scatter_map(pid_buf, n, num_partitions, stream=None):
    dest = DeviceBuffer(8 * max(1, n))
    offs = DeviceBuffer(8 * (num_partitions + 1))
    check(lib().cyl_scatter_map(_p(pid_buf), C.c_int64(n), C.c_uint32(num_partitions), _p(dest), _p(offs),
                                C.c_void_p(stream)))
    return dest, offs


def partition_map(cols, num_partitions, stream=None):
    n = cols[0].length
    dest = DeviceBuffer(8 * max(1, n))
    pid = DeviceBuffer(4 * max(1, n))
    offs = DeviceBuffer(8 * (num_partitions + 1))
    check(lib().cyl_partition_map(c_columns(cols), len(cols), C.c_int64(n), C.c_uint32(num_partitions), _p(pid),
                                  _p(dest), _p(offs), C.c_void_p(stream)))
    return pid, dest, offs


def exclusive_scan(arr, stream=None):
    arr = np.ascontiguousarray(arr, dtype=np.int64)
    src = DeviceBuffer.from_numpy(arr)
    out = DeviceBuffer(8 * (len(arr) + 1))
    check(lib().cyl_exclusive_scan_i64(_p(src), C.c_int64(len(arr)), _p(out), C.c_void_p(stream)))
    return out.to_numpy(np.int64, len(arr) + 1)


def join_indices(lkeys, rkeys, join_type, algorithm=HASH, stream=None):
    """count pass + materialize pass through the C-ABI; returns (l_idx, r_idx) numpy int64 arrays."""
    li, ri, m = join_indices_device(lkeys, rkeys, join_type, algorithm, stream)
    return li.to_numpy(np.int64, m), ri.to_numpy(np.int64, m)


def join_indices_device(lkeys, rkeys, join_type, algorithm=HASH, stream=None):
    nl, nr = lkeys[0].length, rkeys[0].length
    plan = C.c_void_p()
    m = C.c_int64()
    fn = lib().cyl_hash_join_count if algorithm == HASH else lib().cyl_sort_join_count
    check(fn(c_columns(lkeys), c_columns(rkeys), len(lkeys), C.c_int64(nl), C.c_int64(nr), C.c_int32(join_type),
             C.byref(plan), C.byref(m), C.c_void_p(stream)))
    li = DeviceBuffer(8 * max(1, m.value))
    ri = DeviceBuffer(8 * max(1, m.value))
    check(lib().cyl_join_materialize(plan, _p(li), _p(ri), C.c_void_p(stream)))
    return li, ri, m.value


def gather(col, idx_buf, m, stream=None):
    """copy_array_by_indices on the device -> DCol with a validity bitmap."""
    validity = DeviceBuffer(4 * ((m + 31) // 32) + 4)
    if col.type in (STRING, BINARY):
        offs = DeviceBuffer(4 * (m + 1))
        total = C.c_int64()
        check(lib().cyl_gather_varwidth_offsets(_p(col.offsets), _p(idx_buf), C.c_int64(m), _p(offs), C.byref(total),
                                                C.c_void_p(stream)))
        data = DeviceBuffer(max(1, total.value))
        check(lib().cyl_gather_varwidth_data(_p(col.offsets), _p(col.data), _p(idx_buf), C.c_int64(m), _p(offs),
                                             _p(data), C.c_void_p(stream)))
        check(lib().cyl_index_validity(_p(idx_buf), C.c_int64(m), _p(validity), C.c_void_p(stream)))
        return DCol(col.type, m, data, offs, validity, name=col.name)
    data = DeviceBuffer(max(1, m * col.byte_width))
    check(lib().cyl_gather_fixed(_p(col.data), C.c_int32(col.byte_width), _p(idx_buf), C.c_int64(m), _p(data),
                                 _p(validity), C.c_void_p(stream)))
    return DCol(col.type, m, data, None, validity, byte_width=col.byte_width, name=col.name)


def table_checksum(cols, stream=None):
    cs = C.c_uint64()
    check(lib().cyl_table_checksum(c_columns(cols), len(cols), C.c_int64(cols[0].length if cols else 0), C.byref(cs),
                                   C.c_void_p(stream)))
    return cs.value


def groupby_hash_device(keys, stream=None):
    n = keys[0].length
    gid = DeviceBuffer(8 * max(1, n))
    first = DeviceBuffer(max(1, n))
    g = C.c_int64()
    check(lib().cyl_groupby_hash(c_columns(keys), len(keys), C.c_int64(n), _p(gid), _p(first), C.byref(g),
                                 C.c_void_p(stream)))
    return gid, first, g.value


def groupby_hash(keys, stream=None):
    gid, first, g = groupby_hash_device(keys, stream)
    n = keys[0].length
    return gid.to_numpy(np.int64, n), first.to_numpy(np.uint8, n), g


def aggregate(col, gid_buf, num_groups, op, stream=None):
    dt = np.dtype("int64") if op == COUNT else TYPE2NP[col.type]
    out = DeviceBuffer(max(1, num_groups) * dt.itemsize)
    cs = col.c_struct()
    check(lib().cyl_aggregate(C.byref(cs), _p(gid_buf), C.c_int64(num_groups), C.c_int32(op), _p(out), C.c_void_p(stream)))
    return out.to_numpy(dt, num_groups)


def groupby_fused(key, values, agg_value, agg_op, capacity=None, stream=None):
    """cyl_groupby_fused through the C-ABI -> (key ndarray, [aggregate ndarrays], groups, attempts); None if unsupported."""
    av = (C.c_int32 * len(agg_value))(*agg_value)
    ao = (C.c_int32 * len(agg_op))(*agg_op)
    kc = key.c_struct()
    if not lib().cyl_groupby_fused_supported(C.byref(kc), c_columns(values), len(values), av, ao, len(agg_op)):
        return None
    n = key.length
    if capacity is None:
        capacity = max(1, min(n, n // 4 + (1 << 20)))
    attempts = 0
    while True:
        attempts += 1
        out_key = DeviceBuffer(8 * max(1, capacity))
        outs = [DeviceBuffer(8 * max(1, capacity)) for _ in agg_op]
        ptrs = (C.c_void_p * len(outs))(*[o.ptr for o in outs])
        g = C.c_int64()
        check(lib().cyl_groupby_fused(C.byref(kc), c_columns(values), len(values), av, ao, len(agg_op), C.c_int64(capacity), _p(out_key), ptrs,
                                      C.byref(g), C.c_void_p(stream)))
        if g.value <= capacity:
            aggs = []
            for a, o in enumerate(outs):
                dt = np.dtype("int64") if agg_op[a] == COUNT else TYPE2NP[values[agg_value[a]].type]
                aggs.append(o.to_numpy(dt, g.value))
            return out_key.to_numpy(TYPE2NP[key.type], g.value), aggs, g.value, attempts
        capacity = g.value


def rowset_mask(lcols, rcols, op, stream=None):
    nl = lcols[0].length
    nr = rcols[0].length if rcols else 0
    ml, mr = DeviceBuffer(max(1, nl)), DeviceBuffer(max(1, nr))
    check(lib().cyl_rowset_mask(c_columns(lcols), c_columns(rcols if rcols else lcols), len(lcols), C.c_int64(nl),
                                C.c_int64(nr), C.c_int32(op), _p(ml), _p(mr), C.c_void_p(stream)))
    return ml.to_numpy(np.uint8, nl), mr.to_numpy(np.uint8, nr)


def mask_to_indices(mask, stream=None):
    mask = np.ascontiguousarray(mask, dtype=np.uint8)
    n = len(mask)
    mb = DeviceBuffer.from_numpy(mask)
    idx = DeviceBuffer(8 * max(1, n))
    cnt = C.c_int64()
    check(lib().cyl_mask_to_indices(_p(mb), C.c_int64(n), _p(idx), C.byref(cnt), C.c_void_p(stream)))
    return idx.to_numpy(np.int64, cnt.value)


def pinned_array(shape, dtype):
    """numpy array over cudaHostAlloc'ed (page-locked) memory; freed when the array is garbage collected."""
    dtype = np.dtype(dtype)
    n = int(np.prod(shape)) if not isinstance(shape, int) else int(shape)
    nbytes = max(16, n * dtype.itemsize)
    p = C.c_void_p()
    check(lib().cyl_host_alloc_pinned(C.byref(p), C.c_int64(nbytes)))
    buf = (C.c_uint8 * nbytes).from_address(p.value)
    arr = np.frombuffer(buf, dtype=dtype, count=n)
    _PINNED[id(buf)] = (buf, p.value)
    import weakref
    weakref.finalize(arr, _free_pinned, id(buf))
    return arr


_PINNED = {}


def _free_pinned(key):
    ent = _PINNED.pop(key, None)
    if ent is not None:
        try:
            lib().cyl_host_free_pinned(C.c_void_p(ent[1]))
        except Exception:
            pass


def launch_count():
    lib().cyl_launch_count.restype = C.c_longlong
    return int(lib().cyl_launch_count())


def profile_enable(on=True):
    check(lib().cyl_profile_enable(C.c_int32(1 if on else 0)))


def profile_report():
    """-> {kernel: (launches, total_ms)}; synchronises the device and clears the records."""
    buf = C.create_string_buffer(1 << 16)
    check(lib().cyl_profile_report(buf, C.c_int64(len(buf))))
    out = {}
    for line in buf.value.decode().splitlines():
        name, n, ms = line.rsplit(",", 2)
        out[name] = (int(n), float(ms))
    return out


def hash_join_fused(lcols, lkey, rcols, rkey, capacity=None, stream=None):
    """cyl_hash_join_fused_inner through the C-ABI -> (list of DCol output columns, rows, attempts)."""
    nl, nr = lcols[0].length, rcols[0].length
    if capacity is None:
        capacity = max(nl, nr) + max(nl, nr) // 8 + 1024
    attempts = 0
    while True:
        attempts += 1
        outs = [DeviceBuffer(max(1, capacity * c.byte_width)) for c in list(lcols) + list(rcols)]
        ptrs = (C.c_void_p * len(outs))(*[o.ptr for o in outs])
        m = C.c_int64()
        check(lib().cyl_hash_join_fused_inner(c_columns(lcols), len(lcols), C.c_int32(lkey), c_columns(rcols), len(rcols), C.c_int32(rkey),
                                              C.c_int64(nl), C.c_int64(nr), C.c_int64(capacity), ptrs, C.byref(m), C.c_void_p(stream)))
        if m.value <= capacity:
            cols = [DCol(c.type, m.value, o, byte_width=c.byte_width, name=c.name) for c, o in zip(list(lcols) + list(rcols), outs)]
            return cols, m.value, attempts
        capacity = m.value


def hash_join_fused_two_step(lcols, lkey, rcols, rkey, capacity=None, stream=None, right_stream=None):
    """cyl_hash_join_fused_partition (each table, possibly on its own stream) + cyl_hash_join_fused_probe
    -> (list of DCol output columns, rows, attempts). The caller orders the two streams."""
    L = lib()
    L.cyl_hash_join_fused_side_free.restype = None
    nl, nr = lcols[0].length, rcols[0].length
    if capacity is None:
        capacity = max(nl, nr) + max(nl, nr) // 8 + 1024
    ls, rs = C.c_void_p(), C.c_void_p()
    try:
        check(L.cyl_hash_join_fused_partition(c_columns(lcols), len(lcols), C.c_int32(lkey), C.c_int64(nl), None, C.c_int64(nl), C.byref(ls),
                                              C.c_void_p(stream)))
        check(L.cyl_hash_join_fused_partition(c_columns(rcols), len(rcols), C.c_int32(rkey), C.c_int64(nr), ls, C.c_int64(0), C.byref(rs),
                                              C.c_void_p(right_stream if right_stream is not None else stream)))
        if right_stream is not None:
            sync(right_stream)
        attempts = 0
        while True:
            attempts += 1
            outs = [DeviceBuffer(max(1, capacity * 8)) for _ in list(lcols) + list(rcols)]
            ptrs = (C.c_void_p * len(outs))(*[o.ptr for o in outs])
            m = C.c_int64()
            check(L.cyl_hash_join_fused_probe(ls, rs, C.c_int64(capacity), ptrs, C.byref(m), C.c_void_p(stream)))
            if m.value <= capacity:
                cols = [DCol(c.type, m.value, o, byte_width=8, name=c.name) for c, o in zip(list(lcols) + list(rcols), outs)]
                return cols, m.value, attempts
            capacity = m.value
    finally:
        sync(stream)
        L.cyl_hash_join_fused_side_free(ls)
        L.cyl_hash_join_fused_side_free(rs)

