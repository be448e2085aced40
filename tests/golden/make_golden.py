#!/usr/bin/env python
"""Builds tests/golden/ref_testdata.json from the reference's own table fixtures
(/root/reference/rust/scheduler/testdata/*: the 10-row dbgen-format .tbl files its planner tests read,
rust/scheduler/src/test_utils.rs:31-116) plus known answers for the hot-path plans over them.

The reference asserts no query outputs on these files (rust/scheduler/src/planner.rs:428-458 checks plan shape
only) and cannot be built here (no cargo; DataFusion un-vendored), so the expected values are derived
independently of both the oracle and the CUDA code: exact decimal arithmetic (fractions.Fraction) on the
fixture's text, rounded once to float64 at the end.  A Float64 plan result must agree to ~1e-12 relative
(the plans add <= 20 terms); integer / count results exactly.

Run in the build container (needs /root/reference):  python tests/golden/make_golden.py
"""
import datetime
import json
import os
from fractions import Fraction

REF = "/root/reference/rust/scheduler/testdata"
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "ref_testdata.json")
EPOCH = datetime.date(1970, 1, 1)


def days(s):
    return (datetime.date.fromisoformat(s) - EPOCH).days


def read_tbl(path):
    with open(path) as f:
        return [line.rstrip("\n").split("|") for line in f if line.strip()]


def main():
    # ---- inputs (column order: rust/benchmarks/tpch/src/main.rs:303-343) ----
    parts = [read_tbl(f"{REF}/lineitem/partition{p}.tbl") for p in (0, 1)]
    lineitem = {"partitions": []}
    for rows in parts:
        lineitem["partitions"].append({
            "l_orderkey": [int(r[0]) for r in rows], "l_partkey": [int(r[1]) for r in rows],
            "l_suppkey": [int(r[2]) for r in rows], "l_linenumber": [int(r[3]) for r in rows],
            "l_quantity": [r[4] for r in rows], "l_extendedprice": [r[5] for r in rows],
            "l_discount": [r[6] for r in rows], "l_tax": [r[7] for r in rows],
            "l_returnflag": [r[8] for r in rows], "l_linestatus": [r[9] for r in rows],
            "l_shipdate": [days(r[10]) for r in rows], "l_commitdate": [days(r[11]) for r in rows],
            "l_receiptdate": [days(r[12]) for r in rows],
        })
    orows = read_tbl(f"{REF}/orders/orders.tbl")
    orders = {"o_orderkey": [int(r[0]) for r in orows], "o_custkey": [int(r[1]) for r in orows],
              "o_orderdate": [days(r[4]) for r in orows], "o_shippriority": [int(r[7]) for r in orows]}
    crows = read_tbl(f"{REF}/customer/customer.tbl")
    customer = {"c_custkey": [int(r[0]) for r in crows], "c_nationkey": [int(r[3]) for r in crows],
                "c_mktsegment": [r[6] for r in crows]}

    allrows = [r for rows in parts for r in rows]
    F = Fraction

    # ---- Q1 (queries/q1.sql): l_shipdate <= 1998-09-02, group by (returnflag, linestatus) ----
    groups = {}
    for r in allrows:
        if days(r[10]) > days("1998-09-02"):
            continue
        g = groups.setdefault((r[8], r[9]), {"n": 0, "qty": F(0), "price": F(0), "disc_price": F(0), "charge": F(0), "disc": F(0)})
        q, p, d, t = F(r[4]), F(r[5]), F(r[6]), F(r[7])
        g["n"] += 1
        g["qty"] += q
        g["price"] += p
        g["disc_price"] += p * (1 - d)
        g["charge"] += p * (1 - d) * (1 + t)
        g["disc"] += d
    q1 = []
    for (rf, ls), g in sorted(groups.items()):
        q1.append({"l_returnflag": rf, "l_linestatus": ls, "sum_qty": float(g["qty"]), "sum_base_price": float(g["price"]),
                   "sum_disc_price": float(g["disc_price"]), "sum_charge": float(g["charge"]),
                   "avg_qty": float(g["qty"] / g["n"]), "avg_price": float(g["price"] / g["n"]),
                   "avg_disc": float(g["disc"] / g["n"]), "count_order": g["n"]})

    # ---- Q6-shaped (queries/q6.sql) with the date range widened to [1994-01-01, 1997-01-01) and quantity < 50 so
    #      that fixture rows pass; the two l_discount == 0.07 rows inside the range must NOT pass, because the
    #      reference folds `0.06 + 0.01` in f64 to 0.06999999999999999 ----
    lo, hi = 0.06 - 0.01, 0.06 + 0.01  # IEEE f64 folding, as the reference evaluates literals at run time
    q6_sum, q6_n = F(0), 0
    for r in allrows:
        if days("1994-01-01") <= days(r[10]) < days("1997-01-01") and lo <= float(r[6]) <= hi and float(r[4]) < 50:
            q6_sum += F(r[5]) * F(r[6])
            q6_n += 1

    # ---- joins: orders x lineitem on orderkey (inner), customer x orders on custkey ----
    okeys = {int(r[0]) for r in orows}
    j_ol = sorted((int(r[0]), int(r[3])) for r in allrows if int(r[0]) in okeys)  # (orderkey, linenumber)
    ckeys = {int(r[0]) for r in crows}
    j_co = sorted((int(r[1]), int(r[0])) for r in orows if int(r[1]) in ckeys)  # (custkey, orderkey)

    # ---- Q3-shaped (queries/q3.sql) without the segment/date filters that would empty the 10-row fixtures:
    #      orders x lineitem, group by (l_orderkey, o_orderdate, o_shippriority), sum(price*(1-disc)) ----
    odate = {int(r[0]): (days(r[4]), int(r[7])) for r in orows}
    q3 = {}
    for r in allrows:
        k = int(r[0])
        if k in odate:
            key = (k, odate[k][0], odate[k][1])
            q3[key] = q3.get(key, F(0)) + F(r[5]) * (1 - F(r[6]))
    q3_rows = [{"l_orderkey": k[0], "o_orderdate": k[1], "o_shippriority": k[2], "revenue": float(v)}
               for k, v in sorted(q3.items())]

    out = {"source": "rust/scheduler/testdata/{lineitem/partition0.tbl,lineitem/partition1.tbl,orders/orders.tbl,customer/customer.tbl}",
           "lineitem": lineitem, "orders": orders, "customer": customer,
           "expected": {"q1": q1, "q6_wide": {"revenue": float(q6_sum), "rows_passing": q6_n, "date_lo": days("1994-01-01"),
                                              "date_hi": days("1997-01-01"), "quantity_lt": 50},
                        "join_orders_lineitem": j_ol, "join_customer_orders": j_co, "q3_nofilter": q3_rows}}
    with open(OUT, "w") as f:
        json.dump(out, f, indent=1)
    print("wrote", OUT, "q1 groups:", [(r["l_returnflag"], r["l_linestatus"], r["count_order"], r["sum_qty"], r["sum_base_price"]) for r in q1])
    print("q6:", float(q6_sum), q6_n, "joins:", len(j_ol), len(j_co), "q3 groups:", len(q3_rows))


if __name__ == "__main__":
    main()
