"""Python mirror of the Cylon table API over the C facade (include/cylon_b200_api.h).

Same operator names and argument meaning as pycylon's Table (python/pycylon/data/table.pyx:314-380,538,642,714,757):
join / distributed_join, union / subtract / intersect and their distributed_ forms, unique, shuffle, groupby.
Tables live in HBM; every call goes through the C++ host layer in libcylon_b200.so -- there is no Python or CPU
compute path here, and a missing library raises.
"""
import ctypes as C

import numpy as np

from . import capi
from .capi import (BINARY, COUNT, DOUBLE, FIXED_SIZE_BINARY, FLOAT, FULL_OUTER, HASH, INNER, INTERSECT, LEFT, MAX, MIN, RIGHT,  # noqa: F401
                   SORT, STRING, SUBTRACT, SUM, UNION, CylColumn, CylonError, check, lib)

_JOIN_TYPES = {"inner": INNER, "left": LEFT, "right": RIGHT, "outer": FULL_OUTER, "fullouter": FULL_OUTER, "full_outer": FULL_OUTER}
_ALGOS = {"sort": SORT, "hash": HASH}
_AGGS = {"sum": SUM, "min": MIN, "max": MAX, "count": COUNT, "mean": 4, "var": 5, "std": 8, "stddev": 8}  # AggregationOpId values


def nccl_unique_id():
    buf = C.create_string_buffer(128)
    check(lib().cylon_nccl_unique_id(buf))
    return buf.raw


class CylonContext:
    """CylonContext(config=None, distributed=False) of pycylon; distributed contexts take the NCCL bootstrap."""

    def __init__(self, device=0, distributed=False, rank=0, world_size=1, nccl_id=None, thread_group=None):
        self._h = C.c_void_p()
        self._stream = None
        if thread_group is not None:  # ranks as threads of this process (include/cylon_b200_api.h: cylon_context_init_threads)
            check(lib().cylon_context_init_threads(str(thread_group).encode(), C.c_int32(rank), C.c_int32(world_size), C.c_int32(device),
                                                   C.byref(self._h)))
        elif distributed:
            if nccl_id is None or len(nccl_id) != 128:
                raise ValueError("distributed context needs the 128-byte NCCL unique id created on rank 0")
            check(lib().cylon_context_init_distributed(C.c_int32(rank), C.c_int32(world_size), C.c_int32(device),
                                                       C.c_char_p(nccl_id), C.byref(self._h)))
        else:
            check(lib().cylon_context_init(C.c_int32(device), C.byref(self._h)))

    def set_stream(self, cuda_stream):
        """All operators of tables created on this context run (and synchronise) on `cuda_stream`."""
        check(lib().cylon_context_set_stream(self._h, C.c_void_p(cuda_stream)))
        self._stream = cuda_stream

    def add_config(self, key, value):
        """CylonContext::AddConfig; see include/cylon_b200_api.h for the keys this library reads."""
        check(lib().cylon_context_add_config(self._h, str(key).encode(), str(value).encode()))

    def get_rank(self):
        return lib().cylon_context_rank(self._h)

    def get_world_size(self):
        return lib().cylon_context_world_size(self._h)

    def barrier(self):
        check(lib().cylon_context_barrier(self._h))

    def finalize(self):
        if self._h:
            lib().cylon_context_finalize(self._h)
            self._h = None


def _host_column(values, type_id=None, validity=None):
    """-> (CylColumn with host pointers, keepalive list)"""
    keep = []
    s = CylColumn()
    if isinstance(values, np.ndarray) and values.dtype.kind in "iuf":
        a = np.ascontiguousarray(values)
        keep.append(a)
        s.type = type_id if type_id is not None else capi.NP2TYPE[a.dtype]
        s.byte_width = a.dtype.itemsize
        s.length = len(a)
        s.data = a.ctypes.data if a.size else None
        if a.size == 0:
            z = np.zeros(1, dtype=a.dtype)
            keep.append(z)
            s.data = z.ctypes.data
    elif isinstance(values, tuple):  # (offsets int32, bytes uint8) -- Arrow string layout
        offs = np.ascontiguousarray(values[0], dtype=np.int32)
        data = np.ascontiguousarray(values[1], dtype=np.uint8)
        if data.size == 0:
            data = np.zeros(1, dtype=np.uint8)
        keep += [offs, data]
        s.type = type_id if type_id is not None else STRING
        s.byte_width = 0
        s.length = len(offs) - 1
        s.offsets = offs.ctypes.data
        s.data = data.ctypes.data
    else:
        vals = [x.encode() if isinstance(x, str) else bytes(x) for x in values]
        offs = np.zeros(len(vals) + 1, dtype=np.int32)
        if vals:
            offs[1:] = np.cumsum([len(x) for x in vals])
        data = np.frombuffer(b"".join(vals) or b"\0", dtype=np.uint8).copy()
        keep += [offs, data]
        s.type = type_id if type_id is not None else STRING
        s.byte_width = 0
        s.length = len(vals)
        s.offsets = offs.ctypes.data
        s.data = data.ctypes.data
    if validity is not None:
        bits = np.packbits(np.asarray(validity, dtype=bool), bitorder="little")
        bits = np.concatenate([bits, np.zeros(8, dtype=np.uint8)])
        keep.append(bits)
        s.validity = bits.ctypes.data
        s.null_count = int(len(validity) - np.count_nonzero(validity))
    return s, keep


class Table:
    def __init__(self, handle, ctx):
        self._h = handle
        self.ctx = ctx

    # ------------------------------------------------------------------ construction / inspection
    @classmethod
    def from_numpy(cls, ctx, columns, names=None, types=None, validity=None):
        """columns: list of numeric ndarrays / lists of str|bytes / (offsets, bytes) tuples. Copies host -> HBM."""
        n = len(columns)
        arr = (CylColumn * max(1, n))()
        keep = []
        for i, col in enumerate(columns):
            s, k = _host_column(col, None if types is None else types[i], None if validity is None else validity[i])
            arr[i] = s
            keep.append(k)
        cnames = None
        if names is not None:
            cnames = (C.c_char_p * n)(*[str(x).encode() for x in names])
        h = C.c_void_p()
        check(lib().cylon_table_from_host(ctx._h, arr, cnames, C.c_int32(n), C.byref(h)))
        capi.sync(ctx._stream)  # the host arrays may be reused as soon as this returns
        return cls(h, ctx)

    @classmethod
    def from_arrow(cls, ctx, arrow_table):
        """Table::FromArrowTable (table.cpp:202-211) for a pyarrow.Table in host memory: every column's Arrow buffers
        (validity, offsets, data) go to HBM as they are -- numeric, date/time/timestamp, string and binary columns;
        chunked columns are combined first. Anything else raises CylonError(Code::NotImplemented)."""
        import pyarrow as pa

        t = arrow_table.combine_chunks()
        columns, types, validity, names = [], [], [], []
        for name, chunked in zip(t.column_names, t.columns):
            arr = chunked.chunk(0) if chunked.num_chunks else pa.array([], type=chunked.type)
            tid, n, off = arr.type.id, len(arr), arr.offset
            bufs = arr.buffers()
            if tid in capi.TYPE2NP:
                dt = capi.TYPE2NP[tid]
                data = np.frombuffer(bufs[1], dtype=dt)[off:off + n] if n else np.zeros(0, dtype=dt)
                columns.append(data)
            elif tid in (STRING, BINARY):
                offs = np.frombuffer(bufs[1], dtype=np.int32)[off:off + n + 1] if n else np.zeros(1, dtype=np.int32)
                data = np.frombuffer(bufs[2], dtype=np.uint8) if bufs[2] is not None and bufs[2].size else np.zeros(0, dtype=np.uint8)
                columns.append((offs, data))
            else:
                raise capi.CylonError(10, "from_arrow: column %r has unsupported Arrow type %s" % (name, arr.type))
            v = None
            if arr.null_count:
                v = np.unpackbits(np.frombuffer(bufs[0], dtype=np.uint8), bitorder="little")[off:off + n].astype(bool)
            types.append(tid)
            validity.append(v)
            names.append(name)
        return cls.from_numpy(ctx, columns, names=names, types=types, validity=validity)

    def to_arrow(self):
        """Table::ToArrowTable (table.cpp:477-480): a pyarrow.Table in host memory with the same column names, types
        (by arrow::Type id) and nulls."""
        import pyarrow as pa

        by_id = {capi.UINT8: pa.uint8(), capi.INT8: pa.int8(), capi.UINT16: pa.uint16(), capi.INT16: pa.int16(), capi.UINT32: pa.uint32(),
                 capi.INT32: pa.int32(), capi.UINT64: pa.uint64(), capi.INT64: pa.int64(), FLOAT: pa.float32(), DOUBLE: pa.float64(),
                 STRING: pa.string(), BINARY: pa.binary(),
                 capi.DATE32: pa.date32(), capi.DATE64: pa.date64(), capi.TIMESTAMP: pa.timestamp("us"), capi.TIME32: pa.time32("ms"),
                 capi.TIME64: pa.time64("us")}
        arrays, names = [], []
        for c in self.to_host():
            if c["type"] not in by_id:
                raise capi.CylonError(10, "to_arrow: column %r has a type without an Arrow mapping here (%d)" % (c["name"], c["type"]))
            n = len(c["offsets"]) - 1 if c["offsets"] is not None else len(c["data"])
            vbuf, nulls = None, 0
            if c["valid"] is not None:
                nulls = int(n - np.count_nonzero(c["valid"]))
                vbuf = pa.py_buffer(np.packbits(c["valid"], bitorder="little").tobytes())
            if c["type"] in (STRING, BINARY):
                bufs = [vbuf, pa.py_buffer(c["offsets"].tobytes()), pa.py_buffer(c["data"].tobytes())]
            else:
                bufs = [vbuf, pa.py_buffer(np.ascontiguousarray(c["data"]).tobytes())]
            arrays.append(pa.Array.from_buffers(by_id[c["type"]], n, bufs, null_count=nulls))
            names.append(c["name"])
        return pa.Table.from_arrays(arrays, names=names)

    @classmethod
    def from_device(cls, ctx, dcols, names=None):
        """Zero-copy view over capi.DCol device columns (they must outlive the table)."""
        n = len(dcols)
        arr = capi.c_columns(dcols)
        cnames = None
        if names is not None:
            cnames = (C.c_char_p * n)(*[str(x).encode() for x in names])
        h = C.c_void_p()
        check(lib().cylon_table_from_device(ctx._h, arr, cnames, C.c_int32(n), C.byref(h)))
        t = cls(h, ctx)
        t._keep = dcols
        return t

    @property
    def row_count(self):
        lib().cylon_table_rows.restype = C.c_int64
        return lib().cylon_table_rows(self._h)

    @property
    def column_count(self):
        return lib().cylon_table_columns(self._h)

    @property
    def column_names(self):
        lib().cylon_table_column_name.restype = C.c_char_p
        return [lib().cylon_table_column_name(self._h, C.c_int32(i)).decode() for i in range(self.column_count)]

    def column_desc(self, i):
        d = CylColumn()
        nbytes = C.c_int64()
        check(lib().cylon_table_column(self._h, C.c_int32(i), C.byref(d), C.byref(nbytes)))
        return d, nbytes.value

    def to_host(self):
        """-> list of dicts {type, byte_width, data, offsets, valid, name} (numpy; device -> host copies)."""
        out = []
        names = self.column_names
        for i in range(self.column_count):
            d, nbytes = self.column_desc(i)
            n = d.length
            offs = None
            if d.type in (STRING, BINARY):
                data = np.zeros(max(1, nbytes), dtype=np.uint8)
                offs = np.zeros(n + 1, dtype=np.int32)
            elif d.type == FIXED_SIZE_BINARY:
                data = np.zeros(max(1, nbytes), dtype=np.uint8)
            else:
                data = np.zeros(max(1, n), dtype=capi.TYPE2NP[d.type])
            vbits = np.zeros((n + 7) // 8 + 8, dtype=np.uint8) if d.validity else None
            check(lib().cylon_table_column_to_host(self._h, C.c_int32(i), C.c_void_p(data.ctypes.data),
                                                   C.c_void_p(offs.ctypes.data) if offs is not None else None,
                                                   C.c_void_p(vbits.ctypes.data) if vbits is not None else None))
            valid = None
            if vbits is not None:
                valid = np.unpackbits(vbits, bitorder="little")[:n].astype(bool)
            if d.type in (STRING, BINARY, FIXED_SIZE_BINARY):
                data = data[:nbytes]
            else:
                data = data[:n]
            out.append({"type": d.type, "byte_width": d.byte_width, "data": data, "offsets": offs, "valid": valid, "name": names[i]})
        return out

    def checksum(self):
        cs = C.c_uint64()
        check(lib().cylon_table_checksum(self._h, C.byref(cs)))
        return cs.value

    def retain_memory(self, retain):
        check(lib().cylon_table_retain_memory(self._h, C.c_int32(1 if retain else 0)))

    def free(self):
        if self._h:
            lib().cylon_table_free(self._h)
            self._h = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass

    # ------------------------------------------------------------------ operators
    def _join(self, table, join_type, algorithm, left_on, right_on, left_prefix, right_prefix, distributed):
        jt = _JOIN_TYPES[join_type] if isinstance(join_type, str) else int(join_type)
        al = _ALGOS[algorithm] if isinstance(algorithm, str) else int(algorithm)
        lo = (C.c_int32 * len(left_on))(*left_on)
        ro = (C.c_int32 * len(right_on))(*right_on)
        h = C.c_void_p()
        check(lib().cylon_join(self._h, table._h, C.c_int32(jt), C.c_int32(al), lo, ro, C.c_int32(len(left_on)),
                               left_prefix.encode(), right_prefix.encode(), C.c_int32(1 if distributed else 0), C.byref(h)))
        return Table(h, self.ctx)

    def join(self, table, join_type="inner", algorithm="sort", left_on=(0,), right_on=(0,), left_prefix="", right_prefix=""):
        return self._join(table, join_type, algorithm, list(left_on), list(right_on), left_prefix, right_prefix, False)

    def distributed_join(self, table, join_type="inner", algorithm="sort", left_on=(0,), right_on=(0,), left_prefix="", right_prefix=""):
        return self._join(table, join_type, algorithm, list(left_on), list(right_on), left_prefix, right_prefix, True)

    def _setop(self, table, op, distributed):
        h = C.c_void_p()
        check(lib().cylon_set_operation(self._h, table._h, C.c_int32(op), C.c_int32(1 if distributed else 0), C.byref(h)))
        return Table(h, self.ctx)

    def union(self, table):
        return self._setop(table, UNION, False)

    def subtract(self, table):
        return self._setop(table, SUBTRACT, False)

    def intersect(self, table):
        return self._setop(table, INTERSECT, False)

    def distributed_union(self, table):
        return self._setop(table, UNION, True)

    def distributed_subtract(self, table):
        return self._setop(table, SUBTRACT, True)

    def distributed_intersect(self, table):
        return self._setop(table, INTERSECT, True)

    def unique(self, columns, distributed=False):
        cols = (C.c_int32 * len(columns))(*columns)
        h = C.c_void_p()
        check(lib().cylon_unique(self._h, cols, C.c_int32(len(columns)), C.c_int32(1 if distributed else 0), C.byref(h)))
        return Table(h, self.ctx)

    def distributed_unique(self, columns):
        return self.unique(columns, True)

    def sort(self, columns, ascending=True, distributed=False, num_bins=0, num_samples=0):
        """cylon::Sort / DistributedSort (table.hpp:353-404): columns[0] most significant; ascending: bool or one per column"""
        columns = [columns] if isinstance(columns, int) else list(columns)
        dirs = [ascending] * len(columns) if isinstance(ascending, bool) else list(ascending)
        cols = (C.c_int32 * len(columns))(*columns)
        asc = (C.c_int32 * len(columns))(*[1 if d else 0 for d in dirs])
        h = C.c_void_p()
        check(lib().cylon_sort(self._h, cols, asc, C.c_int32(len(columns)), C.c_int32(1 if distributed else 0), C.c_uint32(num_bins),
                               C.c_uint64(num_samples), C.byref(h)))
        return Table(h, self.ctx)

    def distributed_sort(self, columns, ascending=True, num_bins=0, num_samples=0):
        return self.sort(columns, ascending, True, num_bins, num_samples)

    def map_to_sort_partitions(self, column, num_partitions, ascending=True, num_samples=0, num_bins=0):
        """cylon::MapToSortPartitions -> (target partition per row, rows per partition) as numpy arrays; collective"""
        pid = np.zeros(self.row_count, dtype=np.uint32)
        hist = np.zeros(num_partitions, dtype=np.uint32)
        check(lib().cylon_map_to_sort_partitions(self._h, C.c_int32(column), C.c_uint32(num_partitions), C.c_int32(1 if ascending else 0),
                                                 C.c_uint64(num_samples), C.c_uint32(num_bins), pid.ctypes.data_as(C.c_void_p),
                                                 hist.ctypes.data_as(C.c_void_p)))
        return pid, hist

    def shuffle(self, hash_columns):
        cols = (C.c_int32 * len(hash_columns))(*hash_columns)
        h = C.c_void_p()
        check(lib().cylon_shuffle(self._h, cols, C.c_int32(len(hash_columns)), C.byref(h)))
        return Table(h, self.ctx)

    def hash_partition(self, hash_columns, num_partitions):
        cols = (C.c_int32 * len(hash_columns))(*hash_columns)
        outs = (C.c_void_p * num_partitions)()
        check(lib().cylon_hash_partition(self._h, cols, C.c_int32(len(hash_columns)), C.c_int32(num_partitions), outs))
        return [Table(C.c_void_p(outs[p]), self.ctx) for p in range(num_partitions)]

    def groupby(self, index, agg_cols, agg_ops, distributed=False):
        """DistributedHashGroupBy(table, index_cols, aggregate_cols, aggregate_ops) when distributed, else HashGroupBy."""
        index = [index] if isinstance(index, int) else list(index)
        ops = [_AGGS[o] if isinstance(o, str) else int(o) for o in agg_ops]
        ic = (C.c_int32 * len(index))(*index)
        ac = (C.c_int32 * len(agg_cols))(*agg_cols)
        oc = (C.c_int32 * len(ops))(*ops)
        h = C.c_void_p()
        check(lib().cylon_groupby(self._h, ic, C.c_int32(len(index)), ac, oc, C.c_int32(len(ops)), C.c_int32(1 if distributed else 0),
                                  C.byref(h)))
        return Table(h, self.ctx)
