"""Host-side half of Table.from_arrow: which buffers, types, names and null masks reach the C facade
for chunked, sliced and null-carrying pyarrow tables. No GPU: the upload itself is replaced by a recorder."""
import numpy as np
import pyarrow as pa
import pytest

from cylon_b200 import api, capi


class _Recorder(api.Table):
    seen = {}

    @classmethod
    def from_numpy(cls, ctx, columns, names=None, types=None, validity=None):
        cls.seen = dict(columns=columns, names=names, types=types, validity=validity)
        return None


def test_sliced_table_with_nulls_and_strings():
    t = pa.table({"k": pa.array([1, 2, None, 4], type=pa.int64()), "s": pa.array(["a", None, "ccc", ""]),
                  "d": pa.array([0.5, 1.5, 2.5, 3.5])}).slice(1, 3)
    _Recorder.from_arrow(None, t)
    s = _Recorder.seen
    assert s["names"] == ["k", "s", "d"] and s["types"] == [capi.INT64, capi.STRING, capi.DOUBLE]
    assert s["columns"][0].tolist()[0] == 2 and s["columns"][0].tolist()[2] == 4
    assert s["validity"][0].tolist() == [True, False, True]
    offs, data = s["columns"][1]
    assert offs.tolist() == [1, 1, 4, 4] and bytes(data) == b"accc"  # offsets keep the slice's base: FromHost rebases
    assert s["validity"][1].tolist() == [False, True, True]
    assert s["columns"][2].tolist() == [1.5, 2.5, 3.5] and s["validity"][2] is None


def test_chunked_columns_are_combined_and_temporal_types_keep_their_ids():
    a = pa.chunked_array([pa.array([1, 2], type=pa.int32()), pa.array([3], type=pa.int32())])
    ts = pa.chunked_array([pa.array([10, 20, 30], type=pa.timestamp("us"))])
    _Recorder.from_arrow(None, pa.table({"a": a, "ts": ts}))
    s = _Recorder.seen
    assert s["types"] == [capi.INT32, capi.TIMESTAMP]
    assert s["columns"][0].tolist() == [1, 2, 3] and s["columns"][1].dtype == np.int64


def test_empty_table_and_unsupported_type():
    _Recorder.from_arrow(None, pa.table({"k": pa.array([], type=pa.int64()), "s": pa.array([], type=pa.string())}))
    s = _Recorder.seen
    assert len(s["columns"][0]) == 0 and s["columns"][1][0].tolist() == [0]
    with pytest.raises(capi.CylonError):
        _Recorder.from_arrow(None, pa.table({"l": pa.array([[1, 2], [3]])}))
