"""GPU parity for group-by, aggregation, set operations and unique (C-ABI vs CPU oracle)."""
import numpy as np
import pytest

from oracle import pyoracle as orc
from tests.test_gpu_kernels import KEYSETS, make_cols, to_dev

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def cy():
    from cylon_b200 import capi
    capi.init(0)
    return capi


@pytest.mark.parametrize("keyset", sorted(KEYSETS))
@pytest.mark.parametrize("n,dom", [(1, 5), (5000, 100), (200003, 50000), (100000, 10 ** 9)])
def test_groupby_ids_first_seen_order(cy, keyset, n, dom):
    rng = np.random.default_rng(n + dom % 97)
    keys = KEYSETS[keyset](rng, n, dom)
    gid, first, g = cy.groupby_hash(to_dev(cy, keys))
    egid, efirst, eg = orc.groupby_ids(keys)
    assert g == eg
    assert np.array_equal(first, efirst)
    assert np.array_equal(gid, egid)  # same numbering: groups in first-seen order (hash_groupby.cpp:113-121)


def test_groupby_empty(cy):
    k = [cy.DCol.from_numpy(np.zeros(0, dtype=np.int64))]
    gid, first, g = cy.groupby_hash(k)
    assert g == 0 and len(gid) == 0


INT_KINDS = ["int8", "uint8", "int16", "uint16", "int32", "uint32", "int64", "uint64"]


@pytest.mark.parametrize("kind", INT_KINDS + ["float32", "float64"])
@pytest.mark.parametrize("op", [orc.SUM, orc.MIN, orc.MAX, orc.COUNT])
def test_aggregate_matches_reference_kernels(cy, kind, op):
    rng = np.random.default_rng(len(kind) * 7 + op)
    n, groups = 60001, 997
    keys = [orc.col(rng.integers(0, groups, n).astype(np.int64))]
    vals = make_cols(rng, n, [kind])[0]
    if kind.startswith("float"):
        vals = orc.col((rng.random(n) * 2 - 0.5).astype(kind))
    gidb, _first, g = cy.groupby_hash_device(to_dev(cy, keys))
    got = cy.aggregate(to_dev(cy, [vals])[0], gidb, g, op)
    egid, _ef, eg = orc.groupby_ids(keys)
    exp = orc.aggregate(vals, egid, eg, op)
    assert g == eg
    if kind.startswith("float") and op == orc.SUM:
        # atomicAdd order != row order: BASELINE.json tolerance 1e-12 (relative) for double; float32 sums are
        # compared at float32 resolution
        rtol = 1e-12 if kind == "float64" else 2e-5
        scale = np.maximum(np.abs(exp.astype(np.float64)), 1.0)
        assert np.all(np.abs(got.astype(np.float64) - exp.astype(np.float64)) <= rtol * scale * (64 if kind == "float64" else 1))
    else:
        assert np.array_equal(got, exp)


def test_max_init_quirk_on_device(cy):
    # aggregate_kernels.hpp:477: MAX starts from numeric_limits<double>::min() == DBL_MIN
    keys = [cy.DCol.from_numpy(np.array([1, 1, 2], dtype=np.int64))]
    vals = cy.DCol.from_numpy(np.array([-3.0, -2.0, 5.0]))
    gidb, _f, g = cy.groupby_hash_device(keys)
    assert list(cy.aggregate(vals, gidb, g, orc.MAX)) == [np.finfo(np.float64).tiny, 5.0]
    assert list(cy.aggregate(vals, gidb, g, orc.MIN)) == [-3.0, 5.0]
    assert list(cy.aggregate(vals, gidb, g, orc.SUM)) == [-5.0, 5.0]
    assert list(cy.aggregate(vals, gidb, g, orc.COUNT)) == [2, 1]


def _table(rng, n, dom, with_str):
    k = rng.integers(0, dom, n).astype(np.int64)
    cols = [orc.col(k), orc.col((k % 7).astype(np.float64) * 0.5)]
    if with_str:
        cols.append(orc.col([b"w%d" % (v % 13) for v in k]))
    return cols


@pytest.mark.parametrize("op", [orc.UNION, orc.SUBTRACT, orc.INTERSECT])
@pytest.mark.parametrize("with_str", [False, True])
@pytest.mark.parametrize("nl,nr,dom", [(0, 0, 5), (0, 50, 5), (50, 0, 5), (4000, 3000, 500), (150001, 120001, 10 ** 6)])
def test_setop_masks(cy, op, with_str, nl, nr, dom):
    rng = np.random.default_rng(nl + 3 * nr + op)
    l, r = _table(rng, nl, dom, with_str), _table(rng, nr, dom, with_str)
    ml, mr = cy.rowset_mask(to_dev(cy, l), to_dev(cy, r), op)
    eml, emr = orc.setop_mask(l, r, op)
    assert np.array_equal(ml, eml)
    if op == orc.UNION:
        assert np.array_equal(mr, emr)


@pytest.mark.parametrize("keyset", sorted(KEYSETS))
def test_unique_mask(cy, keyset):
    rng = np.random.default_rng(99)
    keys = KEYSETS[keyset](rng, 70001, 3000)
    m, _ = cy.rowset_mask(to_dev(cy, keys), None, orc.UNIQUE)
    assert np.array_equal(m, orc.unique_mask(keys))


def test_single_column_setops_exact_mode(cy):
    rng = np.random.default_rng(4)
    special = np.array([-1, 0, np.iinfo(np.int64).min], dtype=np.int64)
    l = [orc.col(np.concatenate([rng.integers(0, 100, 1000).astype(np.int64), special]))]
    r = [orc.col(np.concatenate([special[:1], rng.integers(50, 150, 800).astype(np.int64)]))]
    for op in (orc.UNION, orc.SUBTRACT, orc.INTERSECT):
        ml, mr = cy.rowset_mask(to_dev(cy, l), to_dev(cy, r), op)
        eml, emr = orc.setop_mask(l, r, op)
        assert np.array_equal(ml, eml)
        if op == orc.UNION:
            assert np.array_equal(mr, emr)


def test_schema_mismatch_is_invalid(cy):
    l = [cy.DCol.from_numpy(np.arange(4, dtype=np.int64))]
    r = [cy.DCol.from_numpy(np.arange(4, dtype=np.float64))]
    with pytest.raises(cy.CylonError) as e:
        cy.rowset_mask(l, r, orc.UNION)
    assert e.value.code == 4


@pytest.mark.parametrize("n", [0, 1, 33, 100001])
def test_mask_to_indices(cy, n):
    rng = np.random.default_rng(n)
    mask = (rng.random(n) < 0.3).astype(np.uint8)
    assert np.array_equal(cy.mask_to_indices(mask), np.nonzero(mask)[0])
