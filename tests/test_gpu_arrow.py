"""Table.from_arrow / to_arrow on the GPU (Arrow host tables in, Arrow host tables out)."""
import numpy as np
import pyarrow as pa
import pytest

from oracle import pyoracle as orc
from tests.helpers import assert_same_rows, to_orc

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def api():
    from cylon_b200 import api as a
    return a


@pytest.fixture(scope="module")
def ctx(api):
    c = api.CylonContext(device=0)
    yield c
    c.finalize()


def test_arrow_round_trip_keeps_values_nulls_names_and_types(api, ctx):
    rng = np.random.default_rng(0)
    n = 5003
    t = pa.table({
        "k": pa.array(rng.integers(-10 ** 12, 10 ** 12, n), type=pa.int64(), mask=rng.random(n) < 0.1),
        "f": pa.array(rng.random(n).astype(np.float32)),
        "s": pa.array([None if x % 7 == 0 else "w%d" % (x % 50) for x in range(n)]),
        "u": pa.array(rng.integers(0, 200, n).astype(np.uint8)),
        "ts": pa.array(rng.integers(0, 10 ** 15, n), type=pa.timestamp("us")),
    })
    for tab in (t, t.slice(17, 4000)):
        back = api.Table.from_arrow(ctx, tab).to_arrow()
        assert back.schema.names == tab.schema.names
        assert [f.type for f in back.schema] == [f.type for f in tab.schema]
        assert back.equals(tab.combine_chunks())


def test_join_of_arrow_tables_matches_oracle(api, ctx):
    rng = np.random.default_rng(1)
    nl, nr = 30000, 25000
    lk, rk = rng.integers(0, 20000, nl), rng.integers(0, 20000, nr)
    lv, rv = rng.random(nl), rng.random(nr)
    L = api.Table.from_arrow(ctx, pa.table({"k": lk, "v": lv}))
    R = api.Table.from_arrow(ctx, pa.table({"k": rk, "v": rv}))
    out = L.join(R, "inner", "hash", [0], [0], "l_", "r_")
    exp = orc.join([orc.col(lk, name="k"), orc.col(lv, name="v")], [orc.col(rk, name="k"), orc.col(rv, name="v")], [0], [0],
                   orc.INNER, orc.HASH, "l_", "r_")
    assert_same_rows(to_orc(out), exp)
    at = out.to_arrow()
    assert at.schema.names == ["l_k", "l_v", "r_k", "r_v"] and at.num_rows == exp[0].length
    assert np.array_equal(at.column("l_k").to_numpy(), at.column("r_k").to_numpy())
