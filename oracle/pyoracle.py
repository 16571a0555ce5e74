"""ctypes binding + numpy glue for the CPU oracle (oracle/oracle.c). TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import this.
The product package (cylon_b200/) never does.

Tables here are plain lists of HCol (host columns backed by numpy). The distributed_* helpers
simulate W ranks of the reference in one process: partition (MapToHashPartitions) -> Split ->
route partition p to rank p (table.cpp:88-96) -> concatenate -> local operator, exactly the
sequence of table.cpp:133-179 / :482-502 / :727-767 / :984-999 and groupby/groupby.cpp:33-83.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "_build", "liboracle.so")

# arrow::Type ids (include/cylon_b200_types.h)
UINT8, INT8, UINT16, INT16, UINT32, INT32, UINT64, INT64 = 2, 3, 4, 5, 6, 7, 8, 9
FLOAT, DOUBLE, STRING, BINARY, FIXED_SIZE_BINARY = 11, 12, 13, 14, 15
DATE32, DATE64, TIMESTAMP, TIME32, TIME64 = 16, 17, 18, 19, 20
INNER, LEFT, RIGHT, FULL_OUTER = 0, 1, 2, 3
SORT, HASH = 0, 1
SUM, MIN, MAX, COUNT = 0, 1, 2, 3
MEAN, VAR, STDDEV = 4, 5, 8  # AggregationOpId (aggregate_kernels.hpp:40-50); NUNIQUE = 6 and QUANTILE = 7 are not restated
UNION, SUBTRACT, INTERSECT, UNIQUE = 0, 1, 2, 3

_NP2TYPE = {np.dtype(k): v for k, v in {
    "uint8": UINT8, "int8": INT8, "uint16": UINT16, "int16": INT16, "uint32": UINT32, "int32": INT32,
    "uint64": UINT64, "int64": INT64, "float32": FLOAT, "float64": DOUBLE}.items()}
_TYPE2NP = {v: k for k, v in _NP2TYPE.items()}
_TYPE2NP.update({DATE32: np.dtype("int32"), DATE64: np.dtype("int64"), TIMESTAMP: np.dtype("int64"),
                 TIME32: np.dtype("int32"), TIME64: np.dtype("int64")})
AGG_PREFIX = {SUM: "sum_", MIN: "min_", MAX: "max_", COUNT: "count_", MEAN: "mean_", VAR: "var_", STDDEV: "std_"}  # aggregate_kernels.hpp:186-260


class CylColumn(C.Structure):
    _fields_ = [("type", C.c_int32), ("byte_width", C.c_int32), ("length", C.c_int64),
                ("null_count", C.c_int64), ("validity", C.c_void_p), ("offsets", C.c_void_p),
                ("data", C.c_void_p)]


def build(force=False):
    """Compile oracle.c (gcc). Building the checker is not using it."""
    if force or not os.path.exists(_LIB_PATH) or \
            os.path.getmtime(_LIB_PATH) < os.path.getmtime(os.path.join(_HERE, "oracle.c")):
        subprocess.check_call(["make", "-C", _HERE, "-s"])
    return _LIB_PATH


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = C.CDLL(_LIB_PATH)
        _lib.orc_murmur3_x86_32.restype = C.c_uint32
        _lib.orc_murmur3_x86_32.argtypes = [C.c_void_p, C.c_int, C.c_uint32]
    return _lib


class HCol:
    """One host column: numpy data (+ int32 offsets for STRING/BINARY, + validity bitmap)."""

    def __init__(self, type_id, data, offsets=None, validity=None, length=None, byte_width=None, name=""):
        self.type = int(type_id)
        self.data = np.ascontiguousarray(data)
        self.offsets = None if offsets is None else np.ascontiguousarray(offsets, dtype=np.int32)
        self.validity = None if validity is None else np.ascontiguousarray(validity, dtype=np.uint8)
        if self.type in (STRING, BINARY):
            self.length = len(self.offsets) - 1
            self.byte_width = 0
        elif self.type == FIXED_SIZE_BINARY:
            self.byte_width = int(byte_width)
            self.length = self.data.size // self.byte_width if length is None else length
        else:
            self.length = len(self.data) if length is None else length
            self.byte_width = self.data.dtype.itemsize
        self.name = name

    @property
    def null_count(self):
        if self.validity is None:
            return 0
        bits = np.unpackbits(self.validity, bitorder="little")[: self.length]
        return int(self.length - bits.sum())

    def valid_mask(self):
        if self.validity is None:
            return np.ones(self.length, dtype=bool)
        return np.unpackbits(self.validity, bitorder="little")[: self.length].astype(bool)

    def values(self):
        """python list of values (None for nulls); bytes for binary types."""
        vm = self.valid_mask()
        if self.type in (STRING, BINARY):
            raw = self.data.tobytes()
            o = self.offsets
            vals = [raw[o[i]:o[i + 1]] for i in range(self.length)]
        elif self.type == FIXED_SIZE_BINARY:
            raw = self.data.tobytes()
            w = self.byte_width
            vals = [raw[i * w:(i + 1) * w] for i in range(self.length)]
        else:
            vals = self.data[: self.length].tolist()
        if self.validity is not None:
            vals = [v if ok else None for v, ok in zip(vals, vm)]
        return vals

    def take(self, idx):
        """copy_array_by_indices (util/copy_arrray.cpp:24-110): -1 -> null, source validity not propagated."""
        idx = np.asarray(idx, dtype=np.int64)
        neg = idx < 0
        safe = np.where(neg, 0, idx)
        validity = None
        if neg.any():
            validity = np.packbits(~neg, bitorder="little")
        if self.type in (STRING, BINARY):
            o = self.offsets.astype(np.int64)
            lens = np.where(neg, 0, o[safe + 1] - o[safe]) if len(idx) else np.zeros(0, np.int64)
            new_off = np.zeros(len(idx) + 1, dtype=np.int64)
            np.cumsum(lens, out=new_off[1:])
            out = np.zeros(int(new_off[-1]), dtype=np.uint8)
            src = self.data.view(np.uint8)
            if len(idx) and new_off[-1] > 0:
                starts = np.repeat(o[safe] - new_off[:-1], lens)
                out[:] = src[np.arange(new_off[-1]) + starts]
            return HCol(self.type, out, new_off.astype(np.int32), validity, name=self.name)
        if self.type == FIXED_SIZE_BINARY:
            w = self.byte_width
            src = self.data.view(np.uint8).reshape(-1, w)
            out = src[safe].copy() if len(src) else np.zeros((len(idx), w), np.uint8)
            out[neg] = 0
            return HCol(self.type, out.reshape(-1), None, validity, byte_width=w, name=self.name)
        if self.length == 0:
            out = np.zeros(len(idx), dtype=self.data.dtype)
        else:
            out = self.data[safe].copy()
            out[neg] = 0
        return HCol(self.type, out, None, validity, name=self.name)

    def filter(self, mask):
        """arrow::compute::Filter with a boolean mask (keeps source validity)."""
        idx = np.nonzero(np.asarray(mask).astype(bool))[0]
        out = self.take(idx)
        if self.validity is not None:
            out.validity = np.packbits(self.valid_mask()[idx], bitorder="little")
        return out


def col(values, type_id=None, name=""):
    """Build an HCol from a numpy array or from a list of str/bytes (-> STRING/BINARY)."""
    if isinstance(values, HCol):
        return values
    if isinstance(values, np.ndarray) and values.dtype.kind in "iuf":
        return HCol(type_id if type_id is not None else _NP2TYPE[values.dtype], values, name=name)
    vals = [v.encode() if isinstance(v, str) else bytes(v) for v in values]
    offsets = np.zeros(len(vals) + 1, dtype=np.int32)
    if vals:
        offsets[1:] = np.cumsum([len(v) for v in vals])
    data = np.frombuffer(b"".join(vals), dtype=np.uint8).copy()
    return HCol(type_id if type_id is not None else STRING, data, offsets, name=name)


def concat_cols(cols):
    first = cols[0]
    any_null = any(c.validity is not None for c in cols)
    validity = None
    if any_null:
        validity = np.packbits(np.concatenate([c.valid_mask() for c in cols]), bitorder="little")
    if first.type in (STRING, BINARY):
        datas = [c.data.view(np.uint8) for c in cols]
        offs = [np.zeros(1, np.int64)]
        base = 0
        for c in cols:
            offs.append(c.offsets[1:].astype(np.int64) - int(c.offsets[0]) + base)
            base += int(c.offsets[-1]) - int(c.offsets[0])
        return HCol(first.type, np.concatenate(datas) if datas else np.zeros(0, np.uint8),
                    np.concatenate(offs).astype(np.int32), validity, name=first.name)
    if first.type == FIXED_SIZE_BINARY:
        return HCol(first.type, np.concatenate([c.data for c in cols]), None, validity,
                    byte_width=first.byte_width, name=first.name)
    return HCol(first.type, np.concatenate([c.data[:c.length] for c in cols]), None, validity, name=first.name)


def _to_c(cols):
    arr = (CylColumn * max(1, len(cols)))()
    keep = []
    for i, c in enumerate(cols):
        arr[i].type = c.type
        arr[i].byte_width = c.byte_width
        arr[i].length = c.length
        arr[i].null_count = c.null_count
        arr[i].validity = c.validity.ctypes.data if c.validity is not None else None
        arr[i].offsets = c.offsets.ctypes.data if c.offsets is not None else None
        d = c.data if c.data.size else np.zeros(1, dtype=c.data.dtype)
        keep.append(d)
        arr[i].data = d.ctypes.data
    return arr, keep


def _check(rc, what):
    if rc != 0:
        raise RuntimeError("oracle %s failed with cylon::Code %d" % (what, rc))


def _nrows(cols):
    return cols[0].length if cols else 0


def murmur3_x86_32(b, seed=0):
    buf = (C.c_uint8 * max(1, len(b))).from_buffer_copy(bytes(b) or b"\0")
    return int(lib().orc_murmur3_x86_32(buf, len(b), seed))


def row_hash(cols):
    n = _nrows(cols)
    h = np.zeros(max(1, n), dtype=np.uint32)
    arr, _k = _to_c(cols)
    _check(lib().orc_row_hash(arr, len(cols), C.c_int64(n), h.ctypes.data_as(C.c_void_p)), "row_hash")
    return h[:n]


def hash_partition(cols, num_partitions):
    n = _nrows(cols)
    pid = np.zeros(max(1, n), dtype=np.uint32)
    hist = np.zeros(num_partitions, dtype=np.uint64)
    arr, _k = _to_c(cols)
    _check(lib().orc_hash_partition(arr, len(cols), C.c_int64(n), C.c_uint32(num_partitions),
                                    pid.ctypes.data_as(C.c_void_p), hist.ctypes.data_as(C.c_void_p)),
           "hash_partition")
    return pid[:n], hist


def split_indices(pid, num_partitions):
    n = len(pid)
    pid = np.ascontiguousarray(pid, dtype=np.uint32)
    perm = np.zeros(max(1, n), dtype=np.int64)
    offs = np.zeros(num_partitions + 1, dtype=np.int64)
    _check(lib().orc_split_indices(pid.ctypes.data_as(C.c_void_p), C.c_int64(n), C.c_uint32(num_partitions),
                                   perm.ctypes.data_as(C.c_void_p), offs.ctypes.data_as(C.c_void_p)),
           "split_indices")
    return perm[:n], offs


def _join_idx(fn, lkeys, rkeys, join_type):
    nl, nr = _nrows(lkeys), _nrows(rkeys)
    la, _k1 = _to_c(lkeys)
    ra, _k2 = _to_c(rkeys)
    lp, rp, m = C.c_void_p(), C.c_void_p(), C.c_int64()
    _check(fn(la, ra, len(lkeys), C.c_int64(nl), C.c_int64(nr), int(join_type), C.byref(lp), C.byref(rp),
              C.byref(m)), "join")
    n = m.value
    li = np.ctypeslib.as_array(C.cast(lp, C.POINTER(C.c_int64)), shape=(max(1, n),))[:n].copy()
    ri = np.ctypeslib.as_array(C.cast(rp, C.POINTER(C.c_int64)), shape=(max(1, n),))[:n].copy()
    lib().orc_free(lp)
    lib().orc_free(rp)
    return li, ri


def hash_join_indices(lkeys, rkeys, join_type):
    return _join_idx(lib().orc_hash_join, lkeys, rkeys, join_type)


def sort_join_indices(lkeys, rkeys, join_type):
    return _join_idx(lib().orc_sort_join, lkeys, rkeys, join_type)


def groupby_ids(keys):
    n = _nrows(keys)
    gid = np.zeros(max(1, n), dtype=np.int64)
    first = np.zeros(max(1, n), dtype=np.uint8)
    g = C.c_int64()
    arr, _k = _to_c(keys)
    _check(lib().orc_groupby_ids(arr, len(keys), C.c_int64(n), gid.ctypes.data_as(C.c_void_p),
                                 first.ctypes.data_as(C.c_void_p), C.byref(g)), "groupby_ids")
    return gid[:n], first[:n], g.value


def aggregate(column, gid, num_groups, op):
    dt = np.dtype("int64") if op == COUNT else (np.dtype("float64") if op in (VAR, STDDEV) else _TYPE2NP[column.type])
    out = np.zeros(max(1, num_groups), dtype=dt)
    arr, _k = _to_c([column])
    gid = np.ascontiguousarray(gid, dtype=np.int64)
    g = gid if len(gid) else np.zeros(1, np.int64)
    _check(lib().orc_aggregate(arr, g.ctypes.data_as(C.c_void_p), C.c_int64(num_groups), int(op),
                               out.ctypes.data_as(C.c_void_p)), "aggregate")
    return out[:num_groups]


def setop_mask(lcols, rcols, op):
    nl, nr = _nrows(lcols), _nrows(rcols)
    ml = np.zeros(max(1, nl), dtype=np.uint8)
    mr = np.zeros(max(1, nr), dtype=np.uint8)
    la, _k1 = _to_c(lcols)
    ra, _k2 = _to_c(rcols)
    _check(lib().orc_setop_mask(la, ra, len(lcols), C.c_int64(nl), C.c_int64(nr), int(op),
                                ml.ctypes.data_as(C.c_void_p), mr.ctypes.data_as(C.c_void_p)), "setop")
    return ml[:nl], mr[:nr]


def unique_mask(cols):
    n = _nrows(cols)
    m = np.zeros(max(1, n), dtype=np.uint8)
    arr, _k = _to_c(cols)
    _check(lib().orc_unique_mask(arr, len(cols), C.c_int64(n), m.ctypes.data_as(C.c_void_p)), "unique")
    return m[:n]


def table_checksum(cols):
    arr, _k = _to_c(cols)
    cs = C.c_uint64()
    _check(lib().orc_table_checksum(arr, len(cols), C.c_int64(_nrows(cols)), C.byref(cs)), "checksum")
    return cs.value


def join_i64_f64(lk, lv, rk, rv, join_type=INNER, algorithm=HASH):
    """Timed CPU baseline entry: returns (rows_out, checksum)."""
    m, cs = C.c_int64(), C.c_uint64()
    _check(lib().orc_join_i64_f64(lk.ctypes.data_as(C.c_void_p), lv.ctypes.data_as(C.c_void_p), C.c_int64(len(lk)),
                                  rk.ctypes.data_as(C.c_void_p), rv.ctypes.data_as(C.c_void_p), C.c_int64(len(rk)),
                                  int(join_type), int(algorithm), C.byref(m), C.byref(cs)), "join_i64_f64")
    return m.value, cs.value


# ---------------------------------------------------------------------------------------------
# table-level operators (a table = list of HCol)
# ---------------------------------------------------------------------------------------------
def table_rows(table):
    """Canonical form for multiset comparison: sorted list of row tuples (None = null)."""
    cols = [c.values() for c in table]
    rows = list(zip(*cols)) if cols else []
    return sorted(rows, key=lambda r: tuple((v is None, v if v is not None else 0) for v in r))


def take_table(table, idx):
    return [c.take(idx) for c in table]


def filter_table(table, mask):
    return [c.filter(mask) for c in table]


def concat_tables(tables):
    return [concat_cols([t[c] for t in tables]) for c in range(len(tables[0]))]


def join(left, right, lkeys, rkeys, join_type=INNER, algorithm=SORT, lprefix="", rprefix=""):
    """cylon::Join (table.cpp:428-475 -> join/join.cpp:80-100 -> build_final_table join_utils.cpp:126-181)."""
    lk = [left[i] for i in lkeys]
    rk = [right[i] for i in rkeys]
    fn = hash_join_indices if algorithm == HASH else sort_join_indices
    li, ri = fn(lk, rk, join_type)
    out = take_table(left, li) + take_table(right, ri)
    for i, (c, src) in enumerate(zip(out, list(left) + list(right))):
        c.name = (lprefix if i < len(left) else rprefix) + src.name  # "suffixes" are prefixes: join_utils.cpp:137-143
    return out


def split(table, pid, num_partitions):
    perm, offs = split_indices(pid, num_partitions)
    parts = []
    for p in range(num_partitions):
        idx = perm[offs[p]:offs[p + 1]]
        part = take_table(table, idx)
        for c in part:
            c.validity = None  # B2: Split drops validity (UnsafeAppend(Value(i)))
        parts.append(part)
    return parts


def shuffle(tables_per_rank, key_cols):
    """shuffle_table_by_hashing + all_to_all_arrow_tables (table.cpp:67-157) for every rank at once."""
    world = len(tables_per_rank)
    parts = []
    for t in tables_per_rank:
        pid, _hist = hash_partition([t[i] for i in key_cols], world)
        parts.append(split(t, pid, world))
    return [concat_tables([parts[src][dst] for src in range(world)]) for dst in range(world)]


def distributed_join(lefts, rights, lkeys, rkeys, join_type=INNER, algorithm=SORT, lprefix="", rprefix=""):
    if len(lefts) == 1:  # table.cpp:487
        return [join(lefts[0], rights[0], lkeys, rkeys, join_type, algorithm, lprefix, rprefix)]
    ls, rs = shuffle(lefts, lkeys), shuffle(rights, rkeys)
    return [join(l, r, lkeys, rkeys, join_type, algorithm, lprefix, rprefix) for l, r in zip(ls, rs)]


def set_op(left, right, op):
    ml, mr = setop_mask(left, right, op)
    if op == UNION:
        return concat_tables([filter_table(left, ml), filter_table(right, mr)])
    return filter_table(left, ml)


def distributed_set_op(lefts, rights, op):
    if len(lefts) < 2:  # table.cpp:741
        return [set_op(lefts[0], rights[0], op)]
    allc = list(range(len(lefts[0])))
    ls, rs = shuffle(lefts, allc), shuffle(rights, allc)
    return [set_op(l, r, op) for l, r in zip(ls, rs)]


def unique(table, cols):
    return filter_table(table, unique_mask([table[i] for i in cols]))


def distributed_unique(tables, cols):
    if len(tables) == 1:  # table.cpp:987
        return [unique(tables[0], cols)]
    return [unique(t, cols) for t in shuffle(tables, cols)]


def hash_groupby(table, idx_cols, aggs):
    """HashGroupBy (groupby/hash_groupby.cpp:250-320). aggs = [(col, op), ...]."""
    gid, first, g = groupby_ids([table[i] for i in idx_cols])
    out = [table[i].filter(first) for i in idx_cols]
    for ci, op in aggs:
        vals = aggregate(table[ci], gid, g, op)
        c = col(vals)
        if op not in (COUNT, VAR, STDDEV):
            c.type = table[ci].type
        c.name = AGG_PREFIX[op] + table[ci].name
        out.append(c)
    return out


def distributed_hash_groupby(tables, idx_cols, agg_cols, agg_ops):
    """DistributedHashGroupBy (groupby/groupby.cpp:33-83), including the double aggregation (and the
    doubled name prefix) when every op is associative."""
    world = len(tables)
    proj = [[t[i] for i in list(idx_cols) + list(agg_cols)] for t in tables]
    k = len(idx_cols)
    idx_after = list(range(k))
    aggs = [(k + i, op) for i, op in enumerate(agg_ops)]
    associative = all(op in (SUM, MIN, MAX) for op in agg_ops)
    if world == 1 or associative:
        local = [hash_groupby(t, idx_after, aggs) for t in proj]
    else:
        local = proj
    if world > 1:
        shuffled = shuffle(local, idx_after)
        return [hash_groupby(t, idx_after, aggs) for t in shuffled]
    return local


# ---- Sort / DistributedSort (table.cpp:291-382) and its range partitioning (arrow_partition_kernels.cpp:334-455) -------------
def sort_sample_positions(n, num_samples):
    """Rows the range partitioner looks at. The reference draws a random sample (util::SampleArray) unless num_samples equals the
    row count, in which case it uses the whole column (arrow_partition_kernels.cpp:375-381); the product takes the rows
    floor(i * n / num_samples) so that results repeat, and the whole column when num_samples >= n."""
    if num_samples >= n:
        return np.arange(n, dtype=np.int64)
    return (np.arange(num_samples, dtype=np.int64) * n) // max(1, num_samples)


def sort_bin_pos(vals, mn, mx, num_bins):
    """get_bin_pos (arrow_partition_kernels.cpp:436-440) with the promotions C++ applies for each column type:
    (val >= min) + ((val >= min) & (val < max)) * ((val - min) * num_bins / range) + (val >= max) * num_bins, range = (T)(max - min).
    Where the reference is undefined (range == 0, NaN) the quotient counts as 0 and NaN goes to the last bin; the result is clamped
    to the num_bins + 2 entries of bin_to_partition."""
    vals = np.asarray(vals)
    dt = vals.dtype
    nb = int(num_bins)
    ge_min, ge_max = vals >= mn, vals >= mx
    mask = ge_min & (vals < mx)
    with np.errstate(all="ignore"):
        if dt.kind == "f":
            f = dt.type
            rng = f(f(mx) - f(mn))
            q = (vals - f(mn)) * f(nb) / rng
            q = np.where(mask, q, f(0))
            total = (ge_min.astype(dt) + q) + (ge_max.astype(np.uint32) * np.uint32(nb)).astype(dt)
            pos = np.where(np.isnan(vals), nb + 1, np.trunc(total)).astype(np.int64)
        elif dt.itemsize <= 4:
            # int / unsigned int arithmetic: the product with the uint32 num_bins and the division are unsigned 32-bit
            wide = np.int64 if dt.kind == "i" else np.uint64
            rng_t = np.array(int(mx) - int(mn)).astype(dt)          # narrowed to T like `range = max - min`
            b = int(np.array(rng_t).astype(np.int32 if dt.kind == "i" else np.uint32).astype(np.uint32)) if dt.itemsize < 4 or dt.kind == "i" else int(rng_t)
            a = ((vals.astype(wide) - wide(mn)).astype(np.int64) & 0xFFFFFFFF).astype(np.uint64)
            a = (a * np.uint64(nb)) & np.uint64(0xFFFFFFFF)
            q = a // np.uint64(b) if b else np.zeros_like(a)
            q = np.where(mask, q, 0).astype(np.uint64)
            pos = ((ge_min.astype(np.uint64) + q + ge_max.astype(np.uint64) * np.uint64(nb)) & np.uint64(0xFFFFFFFF)).astype(np.int64)
        elif dt.kind == "i":
            rng = np.int64(np.array(int(mx) - int(mn)).astype(np.int64)) if abs(int(mx) - int(mn)) < 2 ** 63 else np.array(int(mx) - int(mn) - 2 ** 64).astype(np.int64)
            a = (vals.view(np.uint64) - np.uint64(np.int64(mn).view(np.uint64))) * np.uint64(nb)
            a = a.view(np.int64)
            q = np.zeros_like(a) if int(rng) == 0 else (np.abs(a) // abs(int(rng))) * np.sign(a) * (1 if int(rng) > 0 else -1)  # C division truncates
            q = np.where(mask, q, 0)
            pos = ge_min.astype(np.int64) + q + ge_max.astype(np.int64) * nb
        else:
            rng = np.uint64((int(mx) - int(mn)) % 2 ** 64)
            a = (vals - np.uint64(mn)) * np.uint64(nb)
            q = a // rng if int(rng) else np.zeros_like(a)
            q = np.where(mask, q, 0)
            pos = (ge_min.astype(np.uint64) + q + ge_max.astype(np.uint64) * np.uint64(nb)).astype(np.int64)
    return np.clip(pos, 0, nb + 1)


def sort_bin_to_partition(global_counts, total_samples, num_partitions):
    """Quantile split of the sample histogram, in float like the reference (float_t; arrow_partition_kernels.cpp:412-427)."""
    quantile = np.float32(1.0 / num_partitions)
    prefix, target, curr = np.float32(0), quantile, 0
    out = []
    for c in global_counts:
        out.append(curr)
        freq = np.float32(np.float32(int(c)) / np.float32(int(total_samples))) if total_samples else np.float32(0)
        prefix = np.float32(prefix + freq)
        if prefix > target:
            curr += 1 if curr < num_partitions - 1 else 0
            target = np.float32(target + quantile)
    return np.asarray(out, dtype=np.uint32)


def map_to_sort_partitions(tables, col, ascending=True, num_samples=0, num_bins=0, num_partitions=None):
    """MapToSortPartitions (partition/partition.cpp:168-198) + RangePartitionKernel on every rank at once -> (pid per rank, hist per rank).
    num_samples == 0: 1 % of the rank's rows; num_bins == 0: 16 per partition. total_samples is the number of rows actually sampled
    over all ranks (the reference multiplies ITS OWN num_samples by the world size, identical when the ranks hold equal row counts)."""
    world = num_partitions if num_partitions else len(tables)  # partitions; the ranks are len(tables)
    nb = num_bins if num_bins else world * 16
    samples = []
    for t in tables:
        v = t[col].data
        n = len(v)
        k = num_samples if num_samples else min(int(n * 0.01), n)
        samples.append(v[sort_sample_positions(n, k)])
    nonempty = [s for s in samples if len(s)]
    total = sum(len(s) for s in samples)
    if not nonempty:  # nothing sampled: every row falls below "min" -> bin 0
        pids = [np.full(len(t[col].data), 0 if ascending else world - 1, dtype=np.uint32) for t in tables]
        return pids, [np.bincount(p, minlength=world).astype(np.uint32) for p in pids]
    mn = min(s.min() for s in nonempty)
    mx = max(s.max() for s in nonempty)
    counts = np.zeros(nb + 2, dtype=np.uint64)
    for s in nonempty:
        counts += np.bincount(sort_bin_pos(s, mn, mx, nb), minlength=nb + 2).astype(np.uint64)
    b2p = sort_bin_to_partition(counts, total, world)
    pids = []
    for t in tables:
        p = b2p[sort_bin_pos(t[col].data, mn, mx, nb)]
        pids.append((p if ascending else (world - 1 - p)).astype(np.uint32))
    return pids, [np.bincount(p, minlength=world).astype(np.uint32) for p in pids]


def _sort_key(c, ascending):
    """order-preserving unsigned key; validity is ignored (the reference's sorters compare value slots)"""
    v = c.data
    if v.dtype.kind == "f":
        u = v.view(np.uint32 if v.dtype.itemsize == 4 else np.uint64)
        top = u.dtype.type(1) << u.dtype.type(u.dtype.itemsize * 8 - 1)
        k = np.where(u & top, ~u, u | top)
    elif v.dtype.kind == "i":
        u = v.view("u%d" % v.dtype.itemsize)
        k = u ^ (u.dtype.type(1) << u.dtype.type(u.dtype.itemsize * 8 - 1))
    else:
        k = v
    return k if ascending else ~k


def sort_table(table, sort_cols, directions):
    """cylon::Sort (table.cpp:291-328): argsort by the sort columns (first column most significant), then Take -- which keeps the
    source validity. Ties are in input order here (the reference's std::sort leaves them unspecified)."""
    n = _nrows(table)
    idx = np.arange(n, dtype=np.int64)
    for c, asc in reversed(list(zip(sort_cols, directions))):
        k = _sort_key(table[c], asc)
        idx = idx[np.argsort(k[idx], kind="stable")]
    out = take_table(table, idx)
    for o, c in zip(out, table):
        if c.validity is not None:
            o.validity = np.packbits(c.valid_mask()[idx], bitorder="little")
    return out


def distributed_sort(tables, sort_cols, directions, num_samples=0, num_bins=0):
    """cylon::DistributedSort (table.cpp:338-382): range-partition on the FIRST sort column, Split (drops validity, B2), all-to-all,
    local sort by all sort columns."""
    world = len(tables)
    if world == 1:
        return [sort_table(tables[0], sort_cols, directions)]
    pids, _ = map_to_sort_partitions(tables, sort_cols[0], directions[0], num_samples, num_bins)
    parts = [split(t, p, world) for t, p in zip(tables, pids)]
    return [sort_table(concat_tables([parts[src][dst] for src in range(world)]), sort_cols, directions) for dst in range(world)]
