"""Shared test helpers: CSV fixtures -> oracle tables, seeded synthetic tables, multiset compare."""
import csv
import os

import numpy as np

from oracle import pyoracle as orc

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
REF_DATA = os.path.join(GOLDEN, "ref_data")


def read_csv(name, types=None):
    """Arrow-CSV-like inference: a column is int64 if every field parses as int, else double, else string.
    `types` forces numpy dtypes per column (the reference tests do that for set ops:
    cpp/test/test_utils.hpp:71-76)."""
    with open(os.path.join(REF_DATA, name), newline="") as f:
        rows = list(csv.reader(f))
    header, body = rows[0], [r for r in rows[1:] if r]
    cols = []
    for c, name_c in enumerate(header):
        raw = [r[c] for r in body]
        if types is not None:
            arr = np.array([types[c](v) for v in raw], dtype=types[c])
        else:
            try:
                arr = np.array([int(v) for v in raw], dtype=np.int64)
            except ValueError:
                try:
                    arr = np.array([float(v) for v in raw], dtype=np.float64)
                except ValueError:
                    arr = raw
        if isinstance(arr, np.ndarray) and arr.size == 0:
            arr = np.zeros(0, dtype=np.int64 if types is None else types[c])
        cols.append(orc.col(arr, name=name_c))
    return cols


def splitmix64(x):
    x = (x + np.uint64(0x9E3779B97F4A7C15)) & np.uint64(0xFFFFFFFFFFFFFFFF)
    z = x
    z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
    z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
    return z ^ (z >> np.uint64(31))


def synth_kv(nrows, key_range, seed, rank=0):
    """(int64 key uniform in [0, key_range), double payload in [DBL_MIN, 1)) -- SURVEY.md 8(d)."""
    with np.errstate(over="ignore"):
        i = np.arange(nrows, dtype=np.uint64)
        # one independent stream per (seed, rank, column)
        kbase = splitmix64(np.array([seed * 4096 + rank * 2 + 0], dtype=np.uint64))[0]
        vbase = splitmix64(np.array([seed * 4096 + rank * 2 + 1], dtype=np.uint64))[0]
        k = splitmix64(i + kbase)
        v = splitmix64(i + vbase)
    keys = (k % np.uint64(key_range)).astype(np.int64)
    vals = (v >> np.uint64(11)).astype(np.float64) * (1.0 / 9007199254740992.0)
    vals = np.maximum(vals, np.finfo(np.float64).tiny)
    return keys, vals


def assert_same_rows(got, expected, what=""):
    g, e = orc.table_rows(got), orc.table_rows(expected)
    assert len(g) == len(e), "%s: %d rows, expected %d" % (what, len(g), len(e))
    for a, b in zip(g, e):
        assert a == b, "%s: row %r != expected %r" % (what, a, b)


def to_orc(table):
    """cylon_b200.api.Table -> oracle table (list of HCol), for multiset comparison."""
    cols = []
    for c in table.to_host():
        validity = None
        if c["valid"] is not None:
            validity = np.packbits(c["valid"], bitorder="little")
        if c["type"] in (orc.STRING, orc.BINARY):
            cols.append(orc.HCol(c["type"], c["data"], c["offsets"], validity, name=c["name"]))
        elif c["type"] == orc.FIXED_SIZE_BINARY:
            cols.append(orc.HCol(c["type"], c["data"], None, validity, byte_width=c["byte_width"], name=c["name"]))
        else:
            cols.append(orc.HCol(c["type"], c["data"], None, validity, name=c["name"]))
    return cols


def from_orc(api, ctx, ocols):
    """oracle table -> device table through the public API (host buffers in, H2D inside)."""
    columns, types, validity, names = [], [], [], []
    any_valid = False
    for c in ocols:
        if c.type in (orc.STRING, orc.BINARY):
            columns.append((c.offsets, c.data))
        else:
            columns.append(c.data)
        types.append(c.type)
        names.append(c.name)
        if c.validity is not None:
            validity.append(c.valid_mask())
            any_valid = True
        else:
            validity.append(None)
    return api.Table.from_numpy(ctx, columns, names=names, types=types, validity=validity if any_valid else None)
