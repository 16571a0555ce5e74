#!/usr/bin/env python
"""One-GPU timing of cyl_partition_plan_create + cyl_partition_plan_scatter with every destination LOCAL: the floor of
the peer shuffle's kernels without NVLink in the way (compare with partition_scatter_peer in the multi-GPU bench)."""
import ctypes as C
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from cylon_b200 import capi as cy  # noqa: E402


def main():
    cy.init(0)
    n = int(os.environ.get("ROWS", 100_000_000))
    rng = np.random.default_rng(1)
    k = rng.integers(0, n, n, dtype=np.int64)
    v = rng.random(n)
    dcols = [cy.DCol.from_numpy(k), cy.DCol.from_numpy(v)]
    out = []
    for P in (2, 4, 8):
        keys = (C.c_int32 * 1)(0)
        rows = (C.c_int64 * P)()
        bufs = None
        cy.lib().cyl_profile_enable(1)
        for it in range(6):
            plan = C.c_void_p()
            cy.check(cy.lib().cyl_partition_plan_create(cy.c_columns(dcols), 2, keys, 1, C.c_int64(n), C.c_uint32(P), C.byref(plan), rows, None))
            if bufs is None:
                bufs = [cy.DeviceBuffer(max(16, (rows[p] + 1) * 8)) for p in range(P) for _ in range(2)]
            dst = (C.c_void_p * (2 * P))(*[b.ptr for b in bufs])
            if it == 1:
                cy.sync()
                cy.profile_report()
            cy.check(cy.lib().cyl_partition_plan_scatter(plan, dst, (C.c_int64 * P)(*([0] * P)), None))
            cy.sync()
            cy.lib().cyl_partition_plan_free(plan)
        rep = cy.profile_report()
        out.append({"P": P, "rows": n, "kernels_ms": {k_: round(ms / max(1, c_), 4) for k_, (c_, ms) in rep.items()}})
        del bufs
    print(json.dumps({"bulk": os.environ.get("CYLON_B200_PEER_BULK", "1"), "results": out}))


if __name__ == "__main__":
    main()
