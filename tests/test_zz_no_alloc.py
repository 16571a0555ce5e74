"""The allocation-free scatter map (cyl_scatter_map_static) against the pooled one (cyl_scatter_map).

Some users of this library cannot let it allocate: an engine that owns the whole GPU through its own arena, or a
service that bans allocator calls on its request path. For them the library carries one entry point whose scratch is a
fixed array inside the library image. This file checks, through the C-ABI, that it returns exactly what
cyl_scatter_map returns (integer indices: bit-exact, the tolerance of every index test here), on two calls in a row
that share that array -- the second smaller than the first, so anything the first call left behind would show -- and
that an input beyond the array's capacity is refused instead of overrunning it. It is named test_zz_* so that pytest
collects it last."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def cy():
    from cylon_b200 import capi
    capi.init(0)
    return capi


def _pooled(cy, pid, n, P):
    dest, offs = cy.scatter_map(cy.DeviceBuffer.from_numpy(pid), n, P)
    return dest.to_numpy(np.int64, n), offs.to_numpy(np.int64, P + 1)


def _static(cy, pid, n, P):
    pid_buf = cy.DeviceBuffer.from_numpy(pid)
    dest = cy.DeviceBuffer(8 * max(1, n))
    offs = cy.DeviceBuffer(8 * (P + 1))
    cy.check(cy.lib().cyl_scatter_map_static(C.c_void_p(pid_buf.ptr), C.c_int64(n), C.c_uint32(P), C.c_void_p(dest.ptr),
                                             C.c_void_p(offs.ptr), None))
    return dest.to_numpy(np.int64, n), offs.to_numpy(np.int64, P + 1)


def test_static_scatter_map_equals_pooled_on_two_calls_in_a_row(cy):
    rng = np.random.default_rng(2024)
    # call 1: large and skewed; call 2: smaller, other partition count -> reuses (and must not trust) the same scratch
    for n, P in [(3_000_017, 8), (250_003, 5)]:
        pid = rng.integers(0, P, n).astype(np.uint32)
        pid[: n // 4] = P - 1
        d_ref, o_ref = _pooled(cy, pid, n, P)
        d_got, o_got = _static(cy, pid, n, P)
        assert np.array_equal(o_got, o_ref)
        assert np.array_equal(d_got, d_ref)
        # and it is the stable split: a permutation, rows of one partition keep their input order
        assert np.array_equal(np.sort(d_got), np.arange(n))
        for p in range(P):
            rows = np.nonzero(pid == p)[0]
            assert np.array_equal(d_got[rows], o_got[p] + np.arange(len(rows)))


def test_static_scatter_map_many_partitions_and_empty(cy):
    rng = np.random.default_rng(7)
    n, P = 100_001, 300
    pid = rng.integers(0, P, n).astype(np.uint32)
    d_ref, o_ref = _pooled(cy, pid, n, P)
    d_got, o_got = _static(cy, pid, n, P)
    assert np.array_equal(d_got, d_ref) and np.array_equal(o_got, o_ref)
    d0, o0 = _static(cy, np.zeros(0, dtype=np.uint32), 0, 4)
    assert len(d0) == 0 and not o0.any()


def test_static_scatter_map_refuses_what_does_not_fit(cy):
    n, P = 40_000_000, 8  # 39063 tiles x 8 partitions > 262144 counters
    pid_buf = cy.DeviceBuffer(4 * n)
    dest = cy.DeviceBuffer(8 * n)
    offs = cy.DeviceBuffer(8 * (P + 1))
    rc = cy.lib().cyl_scatter_map_static(C.c_void_p(pid_buf.ptr), C.c_int64(n), C.c_uint32(P), C.c_void_p(dest.ptr),
                                         C.c_void_p(offs.ptr), None)
    assert rc == 6  # cylon::Code::CapacityError
    msg = cy.lib().cyl_last_error().decode()
    assert "static scratch" in msg and "cyl_scatter_map" in msg
