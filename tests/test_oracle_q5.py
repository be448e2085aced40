"""CPU: the oracle's Q5 (hash joins in the reference's left-deep order) against a join-free numpy evaluation of the same
SQL (direct indexing by the dense keys of the synthetic tables, exactly rounded sums)."""
import math

import numpy as np


def test_oracle_q5_against_direct_evaluation(O):
    sf = 0.2
    got = O.q5(sf, threads=2, range_orders=1 << 15)
    no = O.tpch_num_rows("orders", sf)
    o, l = O.gen_orders(sf, 0, no), O.gen_lineitem(sf, 0, 4 * no)
    c = O.gen_customer(sf, 0, O.tpch_num_rows("customer", sf))
    s = O.gen_supplier(sf, 0, O.tpch_num_rows("supplier", sf))
    n = O.gen_nation()
    om = (o["o_orderdate"] >= O.DATE_1994_01_01) & (o["o_orderdate"] < O.DATE_1995_01_01)
    cn = c["c_nationkey"][o["o_custkey"] - 1]            # c_custkey = row + 1
    idx = np.searchsorted(o["o_orderkey"], l["l_orderkey"])  # order row of every lineitem row
    sn = s["s_nationkey"][l["l_suppkey"] - 1]            # s_suppkey = row + 1
    m = om[idx] & (sn == cn[idx]) & (n["n_regionkey"][sn] == 2)  # region 2 = ASIA
    rev = l["l_extendedprice"] * (1.0 - l["l_discount"])
    names = n["n_name"].to_list()
    want = {names[k]: math.fsum(rev[m & (sn == k)].tolist()) for k in range(25) if (m & (sn == k)).any()}
    assert set(got) == set(want) and len(want) == 5
    for k, v in want.items():
        assert abs(got[k] - v) <= 1e-9 * abs(v)
    # one range or many: the same groups, sums within rounding
    one = O.q5(sf, threads=1, range_orders=1 << 30)
    assert set(one) == set(got) and all(abs(one[k] - got[k]) <= 1e-12 * abs(got[k]) for k in got)
