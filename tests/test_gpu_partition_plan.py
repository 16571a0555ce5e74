"""cyl_partition_plan_create / _scatter (the two-step fused partition + scatter behind the peer-memory shuffle) on ONE
GPU: every "peer" destination is a separate local buffer, so the exact kernels that store into other GPUs' columns over
NVLink are checked here against the oracle's stable split (bit-exact, including row order inside a partition)."""
import ctypes as C

import numpy as np
import pytest

from oracle import pyoracle as orc

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def cy():
    from cylon_b200 import capi
    capi.init(0)
    return capi


def _run(cy, ocols, key_idx, P, row_base_shift=0):
    n = ocols[0].length
    dcols = [cy.DCol.from_numpy(c.data, type_id=c.type) for c in ocols]
    plan = C.c_void_p()
    rows = (C.c_int64 * P)()
    keys = (C.c_int32 * len(key_idx))(*key_idx)
    cy.check(cy.lib().cyl_partition_plan_create(cy.c_columns(dcols), len(dcols), keys, len(key_idx), C.c_int64(n), C.c_uint32(P),
                                                C.byref(plan), rows, None))
    rows = list(rows)
    # destination p: its own buffers, rows placed after `row_base_shift + p` foreign rows (as if lower ranks wrote first)
    bufs, ptrs, bases = [], [], []
    for p in range(P):
        base = row_base_shift + p
        bases.append(base)
        for c in ocols:
            b = cy.DeviceBuffer(max(1, (rows[p] + base + 1) * c.byte_width))
            cy.check(cy.lib().cyl_memset(C.c_void_p(b.ptr), 0, C.c_int64(b.nbytes), None))
            bufs.append(b)
            ptrs.append(b.ptr)
    dst = (C.c_void_p * len(ptrs))(*ptrs)
    cy.check(cy.lib().cyl_partition_plan_scatter(plan, dst, (C.c_int64 * P)(*bases), None))
    cy.sync()
    cy.lib().cyl_partition_plan_free(plan)
    out = []
    for p in range(P):
        part = []
        for ci, c in enumerate(ocols):
            arr = bufs[p * len(ocols) + ci].to_numpy(c.data.dtype, rows[p] + bases[p] + 1)
            assert not arr[:bases[p]].any() and not arr[bases[p] + rows[p]:].any()  # nothing outside the assigned range
            part.append(arr[bases[p]:bases[p] + rows[p]])
        out.append(part)
    return rows, out


@pytest.mark.parametrize("P", [1, 2, 3, 4, 8, 32, 33, 100])
@pytest.mark.parametrize("n", [0, 1, 700, 200003])
def test_plan_scatter_matches_stable_split(cy, P, n):
    rng = np.random.default_rng(P * 7 + n)
    ocols = [orc.col(rng.integers(-10 ** 12, 10 ** 12, n).astype(np.int64)), orc.col(rng.random(n))]
    rows, parts = _run(cy, ocols, [0], P, row_base_shift=5)
    pid, hist = orc.hash_partition([ocols[0]], P)
    perm, offs = orc.split_indices(pid, P)
    assert rows == [int(h) for h in hist]
    for p in range(P):
        idx = perm[offs[p]:offs[p + 1]]
        for got, c in zip(parts[p], ocols):
            assert np.array_equal(got, c.data[idx])


@pytest.mark.parametrize("P", [2, 8])
def test_plan_scatter_mixed_widths_and_wide_rows(cy, P):
    rng = np.random.default_rng(P)
    n = 50021
    narrow = [orc.col(rng.integers(0, 1000, n).astype(np.int32)), orc.col(rng.integers(0, 100, n).astype(np.int8)),
              orc.col(rng.random(n).astype(np.float32)), orc.col(rng.integers(0, 9, n).astype(np.int16))]
    wide = narrow + [orc.col(rng.integers(0, 2 ** 50, n).astype(np.int64)), orc.col(rng.random(n)), orc.col(rng.random(n))]  # > 24 B/row: direct path
    for ocols, keys in ((narrow, [0, 3]), (wide, [4])):
        rows, parts = _run(cy, ocols, keys, P)
        pid, hist = orc.hash_partition([ocols[k] for k in keys], P)
        perm, offs = orc.split_indices(pid, P)
        for p in range(P):
            idx = perm[offs[p]:offs[p + 1]]
            for got, c in zip(parts[p], ocols):
                assert np.array_equal(got, c.data[idx])


@pytest.mark.parametrize("P", [2, 3, 5, 8, 16])
@pytest.mark.parametrize("layout", ["k", "kv", "vk", "kvv", "vkv", "vvk", "dk_v", "two_keys"])
def test_plan_scatter_all_8_byte_variants(cy, P, layout):
    """Every instantiation of the all-8-byte staged kernel (1-3 columns x key position, int64 'modulo' key, double
    'murmur' key, two-column key = generic row hash), on row bases of
    both parities and ragged tails."""
    rng = np.random.default_rng(P * 31 + len(layout))
    n = 100_003
    ik = lambda: orc.col(rng.integers(-2 ** 40, 2 ** 40, n).astype(np.int64))  # noqa: E731
    dv = lambda: orc.col(rng.random(n))  # noqa: E731
    if layout == "dk_v":
        ocols, keys = [orc.col(rng.integers(0, 5000, n).astype(np.float64)), ik()], [0]
    elif layout == "two_keys":
        ocols, keys = [ik(), ik(), dv()], [0, 1]
    else:
        ocols = [ik() if ch == "k" else dv() for ch in layout]
        keys = [layout.index("k")]
    for shift in (0, 1):
        rows, parts = _run(cy, ocols, keys, P, row_base_shift=shift)
        pid, hist = orc.hash_partition([ocols[k] for k in keys], P)
        perm, offs = orc.split_indices(pid, P)
        assert rows == [int(h) for h in hist]
        for p in range(P):
            idx = perm[offs[p]:offs[p + 1]]
            for got, c in zip(parts[p], ocols):
                assert np.array_equal(got, c.data[idx])
