"""cylon_b200/csrc/sort_bins.cuh compiled for the host (g++): the order-preserving sort keys and the range partitioner's bin
expression -- the arithmetic the CUDA kernels of sort.cu inline -- against the oracle (which tests/test_oracle_sort_vs_reference.py
pins against the compiled reference). No GPU needed."""
import ctypes as C
import os
import shutil
import subprocess

import numpy as np
import pytest

from oracle import pyoracle as orc

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
TYPE_ID = {"uint8": 2, "int8": 3, "uint16": 4, "int16": 5, "uint32": 6, "int32": 7, "uint64": 8, "int64": 9, "float32": 11, "float64": 12}


@pytest.fixture(scope="module")
def host():
    if shutil.which("g++") is None:
        pytest.skip("needs g++")
    out = os.path.join(ROOT, "build", "tests")
    os.makedirs(out, exist_ok=True)
    so = os.path.join(out, "libsort_bins_host.so")
    subprocess.run(["g++", "-std=c++17", "-O2", "-shared", "-fPIC", "-o", so, os.path.join(ROOT, "tests", "cpp", "sort_bins_host.cpp")], check=True)
    lib = C.CDLL(so)
    lib.host_order_key.restype = C.c_uint64
    lib.host_order_key_inverse.restype = C.c_uint64
    return lib


def _key(lib, kind, v):
    a = np.array([v], dtype=kind)
    bits = int(a.view("u%d" % a.dtype.itemsize)[0])
    return lib.host_order_key(TYPE_ID[kind], a.dtype.itemsize, C.c_uint64(bits))


def _values(rng, kind, n):
    if kind.startswith("float"):
        return (rng.standard_normal(n) * 1000).astype(kind)
    info = np.iinfo(kind)
    wide = rng.integers(info.min, info.max, n, dtype=kind, endpoint=True)
    narrow = rng.integers(max(info.min, -100), min(info.max, 100), n).astype(kind)
    return np.where(rng.random(n) < 0.5, wide, narrow).astype(kind)


@pytest.mark.parametrize("kind", sorted(TYPE_ID))
def test_order_keys_sort_like_values_and_invert(host, kind):
    rng = np.random.default_rng(len(kind))
    v = _values(rng, kind, 4000)
    if kind.startswith("float"):
        v[:4] = [0.0, -0.0, np.inf, -np.inf]
    keys = np.array([_key(host, kind, x) for x in v], dtype=np.uint64)
    order = np.argsort(keys, kind="stable")
    assert np.all(np.diff(v[order].astype(np.float64 if kind.startswith("float") else object)) >= 0)
    w = v.dtype.itemsize
    assert np.all(keys < (1 << (8 * w)) if w < 8 else True)
    bits = v.view("u%d" % w).astype(np.uint64)
    back = np.array([host.host_order_key_inverse(TYPE_ID[kind], w, C.c_uint64(int(k))) for k in keys], dtype=np.uint64)
    assert np.array_equal(back, bits)


@pytest.mark.parametrize("kind", sorted(TYPE_ID))
def test_bin_positions_match_oracle(host, kind):
    """Full-range and narrow-range columns, several bin counts, min/max taken from a sample (so values fall outside [min, max] too)"""
    rng = np.random.default_rng(7 + len(kind))
    for trial in range(12):
        v = _values(rng, kind, 3000)
        if trial % 3 == 0 and not kind.startswith("float"):
            v = (v % 50).astype(kind)  # small range, many ties
        sample = v[: 1 + trial * 100]
        mn, mx = sample.min(), sample.max()
        for nb in (1, 2, 7, 32, 128, 5000):
            out = np.zeros(len(v), dtype=np.uint32)
            rc = host.host_bin_pos(TYPE_ID[kind], v.ctypes.data_as(C.c_void_p), C.c_int64(len(v)), C.c_uint64(_key(host, kind, mn)),
                                   C.c_uint64(_key(host, kind, mx)), C.c_uint32(nb), out.ctypes.data_as(C.c_void_p))
            assert rc == 0
            exp = orc.sort_bin_pos(v, mn, mx, nb)
            bad = np.nonzero(out.astype(np.int64) != exp)[0]
            assert len(bad) == 0, (kind, trial, nb, v[bad[:3]], out[bad[:3]], exp[bad[:3]], mn, mx)


def test_quantile_split_matches_oracle():
    """cyl_range_quantile_split (host-only entry point of the library) against the oracle's float restatement of
    build_bin_to_partition, on random, skewed, empty and single-bin histograms"""
    so = os.path.join(ROOT, "cylon_b200", "lib", "libcylon_b200.so")
    if not os.path.exists(so):
        pytest.skip("library not built")
    lib = C.CDLL(so)
    rng = np.random.default_rng(3)
    for trial in range(300):
        nb = int(rng.choice([1, 2, 7, 16, 64, 128, 1000]))
        parts = int(rng.choice([1, 2, 3, 4, 8, 13, 64]))
        kind = trial % 4
        if kind == 0:
            counts = rng.integers(0, 1000, nb + 2)
        elif kind == 1:
            counts = (rng.pareto(1.2, nb + 2) * 50).astype(np.int64)
        elif kind == 2:
            counts = np.zeros(nb + 2, dtype=np.int64)
            counts[rng.integers(0, nb + 2)] = 12345
        else:
            counts = np.zeros(nb + 2, dtype=np.int64)
        counts = counts.astype(np.uint64)
        total = int(counts.sum()) if trial % 5 else int(counts.sum()) + 17  # the reference's total need not equal the sum
        out = np.zeros(nb + 2, dtype=np.uint32)
        rc = lib.cyl_range_quantile_split(counts.ctypes.data_as(C.c_void_p), C.c_uint32(nb), C.c_uint64(total), C.c_uint32(parts),
                                          out.ctypes.data_as(C.c_void_p))
        assert rc == 0
        assert np.array_equal(out, orc.sort_bin_to_partition(counts, total, parts)), (nb, parts, kind)
