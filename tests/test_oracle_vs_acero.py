"""CPU: independent cross-check of the oracle with pyarrow's Acero engine and exactly-rounded sums (math.fsum)
on the synthetic TPC-H-shaped workload."""
import math

import numpy as np
import pyarrow as pa
import pyarrow.compute as pc
import pytest

from util import assert_rows_match

Q1_F = ["sum_qty", "sum_base_price", "sum_disc_price", "sum_charge", "avg_qty", "avg_price", "avg_disc"]


@pytest.mark.parametrize("n,parts", [(1, 1), (1000, 1), (200_000, 4)])
def test_q1_vs_acero(O, n, parts):
    from ballista_b200 import tpch
    cols = O.gen_lineitem(0.1, 64, n)
    got = O.q1(cols, n_partitions=parts)
    t = pa.Table.from_batches([O.to_arrow(cols, tpch.Q1_COLUMNS)])
    t = t.filter(pc.less_equal(t["l_shipdate"], pa.scalar(O.DATE_1998_09_02, pa.int32()).cast(pa.date32())))
    one = pa.scalar(1.0)
    disc_price = pc.multiply(t["l_extendedprice"], pc.subtract(one, t["l_discount"]))
    t = t.append_column("disc_price", disc_price).append_column("charge", pc.multiply(disc_price, pc.add(one, t["l_tax"])))
    r = t.group_by(["l_returnflag", "l_linestatus"]).aggregate(
        [("l_quantity", "sum"), ("l_extendedprice", "sum"), ("disc_price", "sum"), ("charge", "sum"),
         ("l_quantity", "mean"), ("l_extendedprice", "mean"), ("l_discount", "mean"), ("l_quantity", "count")]).to_pylist()
    want = [{"l_returnflag": x["l_returnflag"], "l_linestatus": x["l_linestatus"], "sum_qty": x["l_quantity_sum"],
             "sum_base_price": x["l_extendedprice_sum"], "sum_disc_price": x["disc_price_sum"],
             "sum_charge": x["charge_sum"], "avg_qty": x["l_quantity_mean"], "avg_price": x["l_extendedprice_mean"],
             "avg_disc": x["l_discount_mean"], "count_order": x["l_quantity_count"]} for x in r]
    assert_rows_match(got, want, ["l_returnflag", "l_linestatus"], exact=["count_order"], approx=Q1_F)


def test_q6_vs_fsum(O):
    cols = O.gen_lineitem(0.2, 0, 500_000)
    rev, passing = O.q6(cols, n_partitions=3)
    lo, hi = 0.06 - 0.01, 0.06 + 0.01
    m = ((cols["l_shipdate"] >= O.DATE_1994_01_01) & (cols["l_shipdate"] < O.DATE_1995_01_01)
         & (cols["l_discount"] >= lo) & (cols["l_discount"] <= hi) & (cols["l_quantity"] < 24.0))
    assert passing == int(m.sum()) > 0
    exact = math.fsum((cols["l_extendedprice"][m] * cols["l_discount"][m]).tolist())
    assert math.isclose(rev, exact, rel_tol=1e-12)
    assert not np.any(m & (cols["l_discount"] == 0.07))  # f64-folded upper bound excludes 0.07


def test_inner_join_vs_acero(O):
    sf = 0.01
    o = O.gen_orders(sf, 0, O.tpch_num_rows("orders", sf))
    li = O.gen_lineitem(sf, 0, O.tpch_num_rows("lineitem", sf))
    keep = o["o_orderdate"] < O.DATE_1995_03_15
    okeys = o["o_orderkey"][keep]
    lidx, ridx = O.hash_join([okeys], [li["l_orderkey"]])
    got = sorted(zip(okeys[lidx].tolist(), ridx.tolist()))
    left = pa.table({"k": okeys})
    right = pa.table({"k": li["l_orderkey"], "row": np.arange(len(li["l_orderkey"]))})
    j = left.join(right, keys="k", join_type="inner")
    want = sorted(zip(j["k"].to_pylist(), j["row"].to_pylist()))
    assert got == want and len(got) > 0


def test_generator_shape(O):
    """4 lines per order on average, the four Q1 groups, value ranges."""
    sf = 0.02
    n = O.tpch_num_rows("lineitem", sf)
    assert n == 4 * O.tpch_num_rows("orders", sf)
    c = O.gen_lineitem(sf, 0, n)
    groups = set(zip(c["l_returnflag"].to_list(), c["l_linestatus"].to_list()))
    assert groups == {("A", "F"), ("N", "F"), ("N", "O"), ("R", "F")}
    assert c["l_quantity"].min() == 1 and c["l_quantity"].max() == 50
    assert set(np.round(c["l_discount"] * 100).astype(int)) == set(range(11))
    assert np.all(np.diff(c["l_orderkey"]) >= 0)
    # any row range is reproducible from the row index alone
    sub = O.gen_lineitem(sf, 1000, 77)
    assert np.array_equal(sub["l_extendedprice"], c["l_extendedprice"][1000:1077])


@pytest.mark.parametrize("how", ["Inner", "Left", "Right"])
def test_two_key_outer_joins_with_nulls_vs_acero(O, how):
    """The oracle's outer-join restatement over TWO key columns with NULLs in either (the checker of the GPU's
    Int32 multi-key join) against pyarrow's Acero join on the same data."""
    rng = np.random.default_rng(41)
    nl, nr = 1500, 4001
    la, lb = rng.integers(0, 200, nl).astype(np.int32), rng.integers(0, 5, nl).astype(np.int32)
    ra, rb = rng.integers(50, 250, nr).astype(np.int32), rng.integers(0, 5, nr).astype(np.int32)
    lma, lmb = rng.random(nl) < 0.05, rng.random(nl) < 0.05
    rma, rmb = rng.random(nr) < 0.05, rng.random(nr) < 0.05
    li, ri = O.hash_join_outer([la, lb], [ra, rb], how, ~(lma | lmb), ~(rma | rmb))
    left = pa.table({"a": pa.array(la, mask=lma), "b": pa.array(lb, mask=lmb), "lrow": np.arange(nl, dtype=np.int64)})
    right = pa.table({"a2": pa.array(ra, mask=rma), "b2": pa.array(rb, mask=rmb), "rrow": np.arange(nr, dtype=np.int64)})
    jt = {"Inner": "inner", "Left": "left outer", "Right": "right outer"}[how]
    ac = left.join(right, keys=["a", "b"], right_keys=["a2", "b2"], join_type=jt, coalesce_keys=False)
    want = sorted((-1 if l is None else l, -1 if r is None else r)
                  for l, r in zip(ac["lrow"].to_pylist(), ac["rrow"].to_pylist()))
    got = sorted(zip(li.tolist(), ri.tolist()))
    assert got == want and len(got) > 500
