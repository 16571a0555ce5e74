"""The CUDA path against the REFERENCE ITSELF (not the restatement): every operator of the hot path on one GPU compared
with the reference's own code compiled into oracle/_ref/libcylon_ref.so (prebuilt; /root/reference is not read at run
time). World > 1 is emulated on one GPU the way tests/test_gpu_tables.py does (GPU hash partition per simulated rank,
partitions routed on the host, local operator per rank) and compared with the reference's Distributed* operator run on
that many ranks (threads) -- per-rank row multisets. Real multi-GPU transport: tests/test_gpu_distributed.py."""
import numpy as np
import pytest

from oracle import pyoracle as orc
from oracle import pyref
from tests.helpers import assert_same_rows, from_orc, to_orc

pytestmark = [pytest.mark.gpu, pytest.mark.skipif(not pyref.dist_available(), reason="oracle/_ref/libcylon_ref.so (table operators) not built")]

JT = {orc.INNER: "inner", orc.LEFT: "left", orc.RIGHT: "right", orc.FULL_OUTER: "fullouter"}


@pytest.fixture(scope="module")
def api():
    from cylon_b200 import api as a
    return a


@pytest.fixture(scope="module")
def ctx(api):
    c = api.CylonContext(device=0)
    yield c
    c.finalize()


def _tables(world, n, seed, dom, two_keys=False):
    out = []
    for r in range(world):
        rng = np.random.default_rng(seed * 1000 + r)
        m = n + 17 * r
        cols = [orc.col(rng.integers(-dom, dom, m).astype(np.int64), name="c0")]
        if two_keys:
            cols.append(orc.col(rng.integers(0, 4, m).astype(np.int32), name="c1"))
        cols.append(orc.col(rng.integers(0, 1000, m).astype(np.float64) / 8, name="c%d" % len(cols)))
        out.append(cols)
    return out


def _gpu_shuffle(api, ctx, tables, key_cols):
    world = len(tables)
    if world == 1:
        return tables
    parts = [from_orc(api, ctx, t).hash_partition(key_cols, world) for t in tables]
    return [orc.concat_tables([to_orc(parts[src][dst]) for src in range(world)]) for dst in range(world)]


@pytest.mark.parametrize("world", [1, 2, 4])
@pytest.mark.parametrize("jt", [orc.INNER, orc.LEFT, orc.RIGHT, orc.FULL_OUTER])
@pytest.mark.parametrize("algo", ["sort", "hash"])
@pytest.mark.parametrize("two_keys", [False, True])
def test_join_vs_reference(api, ctx, world, jt, algo, two_keys):
    lefts, rights = _tables(world, 3000, 1, 1500, two_keys), _tables(world, 2500, 2, 1500, two_keys)
    keys = [0, 1] if two_keys else [0]
    exp = pyref.dist_op(pyref.DIST_JOIN, lefts, rights, p1=keys, p2=keys, join_type=jt, algorithm=orc.SORT if algo == "sort" else orc.HASH)
    ls, rs = _gpu_shuffle(api, ctx, lefts, keys), _gpu_shuffle(api, ctx, rights, keys)
    for r in range(world):
        out = from_orc(api, ctx, ls[r]).join(from_orc(api, ctx, rs[r]), JT[jt], algo, keys, keys, "l_", "r_")
        assert out.column_names == [c.name for c in exp[r]]
        assert_same_rows(to_orc(out), exp[r], "rank %d" % r)


@pytest.mark.parametrize("nl,nr,dom", [(200_000, 150_000, 120_000), (2_300_000, 2_200_000, 3_000_000)])
def test_fused_inner_hash_join_vs_reference(api, ctx, nl, nr, dom):
    """The benchmark's path (fused radix join; the larger case takes the two-pass staged partition) against the
    reference's HashJoin on the same (int64, double) tables."""
    rng = np.random.default_rng(nl)
    l = [orc.col(rng.integers(0, dom, nl).astype(np.int64), name="c0"), orc.col(rng.random(nl), name="c1")]
    r = [orc.col(rng.integers(0, dom, nr).astype(np.int64), name="c0"), orc.col(rng.random(nr), name="c1")]
    exp = pyref.dist_op(pyref.DIST_JOIN, [l], [r], p1=[0], p2=[0], join_type=orc.INNER, algorithm=orc.HASH)[0]
    out = from_orc(api, ctx, l).join(from_orc(api, ctx, r), "inner", "hash", [0], [0], "l_", "r_")
    assert out.row_count == exp[0].length
    assert out.checksum() == orc.table_checksum(exp)


@pytest.mark.parametrize("world", [1, 2, 4])
@pytest.mark.parametrize("name,ref_op", [("union", pyref.DIST_UNION), ("subtract", pyref.DIST_SUBTRACT), ("intersect", pyref.DIST_INTERSECT)])
def test_set_ops_vs_reference(api, ctx, world, name, ref_op):
    lefts, rights = _tables(world, 4000, 3, 30), _tables(world, 3500, 3, 30)
    for t in lefts + rights:
        t[1].data[:] = np.floor(t[1].data) % 5
    exp = pyref.dist_op(ref_op, lefts, rights)
    ls, rs = _gpu_shuffle(api, ctx, lefts, [0, 1]), _gpu_shuffle(api, ctx, rights, [0, 1])
    for r in range(world):
        out = getattr(from_orc(api, ctx, ls[r]), name)(from_orc(api, ctx, rs[r]))
        assert_same_rows(to_orc(out), exp[r], "rank %d" % r)


@pytest.mark.parametrize("world", [1, 2, 4])
def test_unique_vs_reference(api, ctx, world):
    tables = _tables(world, 5000, 5, 40)
    for t in tables:
        t[1].data[:] = np.floor(t[1].data) % 3
    exp = pyref.dist_op(pyref.DIST_UNIQUE, tables, p1=[0, 1])
    ts = _gpu_shuffle(api, ctx, tables, [0, 1])
    for r in range(world):
        assert_same_rows(to_orc(from_orc(api, ctx, ts[r]).unique([0, 1])), exp[r], "rank %d" % r)


@pytest.mark.parametrize("world", [2, 4])
def test_hash_partition_vs_reference_shuffle(api, ctx, world):
    tables = _tables(world, 6000, 7, 100000)
    exp = pyref.dist_op(pyref.DIST_SHUFFLE, tables, p1=[0])
    got = _gpu_shuffle(api, ctx, tables, [0])
    for r in range(world):
        assert_same_rows(got[r], exp[r], "rank %d" % r)


@pytest.mark.parametrize("ops", [["sum", "min", "max"], ["sum", "count", "min", "max"]])
def test_groupby_vs_reference(api, ctx, ops):
    """HashGroupBy on one GPU against the reference's (world 1): names, first-seen group order, aggregate values
    (multiples of 1/8: sums exact; negative groups: MAX starts from DBL_MIN)."""
    t = _tables(1, 20000, 9, 300)[0]
    t[1].data[:] -= 60.0
    code = {"sum": orc.SUM, "min": orc.MIN, "max": orc.MAX, "count": orc.COUNT}
    exp = pyref.dist_op(pyref.DIST_GROUPBY, [t], p1=[0], p2=[x for o in ops for x in (1, code[o])])[0]
    out = from_orc(api, ctx, t).groupby(0, [1] * len(ops), ops)
    assert out.column_names == [c.name for c in exp]
    assert_same_rows(to_orc(out), exp)
