"""BASELINE.json config 2 at its full single-GPU size (100M x 100M rows, int64 key + double payload): the oracle
cannot finish that in seconds, so parity is checked through size-independent properties -- the fused shared-memory
join and the independent radix-sort/merge join must agree on the row count and on the order-independent checksum of
the whole result table, every output row must carry equal keys, and the count must match the closed form computed
from the key histograms with numpy (sum over keys of count_left * count_right). The oracle's C join is run at the full
size too and its row count and checksum must equal the GPU's."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def test_join_100m_fused_equals_sort_merge_and_closed_form():
    import bench
    from cylon_b200 import api
    ctx = api.CylonContext(device=0)
    n = 100_000_000
    lk, lv = bench.synth_table(n, n, 1, 0)
    rk, rv = bench.synth_table(n, n, 2, 0)
    expected = int(np.dot(np.bincount(lk, minlength=n).astype(np.int64), np.bincount(rk, minlength=n).astype(np.int64)))
    L = api.Table.from_numpy(ctx, [lk, lv], names=["k", "v"])
    R = api.Table.from_numpy(ctx, [rk, rv], names=["k", "v"])
    fused = L.distributed_join(R, "inner", "hash", [0], [0], "l_", "r_")
    assert fused.row_count == expected
    cs = fused.checksum()
    cols = fused.to_host()
    assert np.array_equal(cols[0]["data"], cols[2]["data"])          # l_k == r_k on every output row
    assert fused.column_names == ["l_k", "l_v", "r_k", "r_v"]
    del cols
    fused.free()
    ctx.add_config("cylon.join.fused", "0")  # the independent path: LSD radix sort + merge, index vectors, gather
    merged = L.join(R, "inner", "sort", [0], [0], "l_", "r_")
    ctx.add_config("cylon.join.fused", "1")
    assert merged.row_count == expected and merged.checksum() == cs
    merged.free()
    # ... and the oracle itself at the bench's full shape (oracle/oracle.c orc_join_i64_f64: hash join + materialise +
    # checksum in C, about a minute on one core): row count and order-independent checksum of the whole result table
    from oracle import pyoracle as orc
    m_ref, cs_ref = orc.join_i64_f64(lk, lv, rk, rv, orc.INNER, orc.HASH)
    assert m_ref == expected and cs_ref == cs
    # joining is symmetric: swapping the sides swaps the column groups, nothing else
    swapped = R.join(L, "inner", "hash", [0], [0], "r_", "l_")
    assert swapped.row_count == expected
    ctx.finalize()
