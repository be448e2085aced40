"""GPU: the CUDA operators, called through the C ABI with host Arrow batches, against the golden vectors derived
from the reference's own table fixtures (tests/golden/ref_testdata.json)."""
import math

import numpy as np
import pyarrow as pa
import pytest

import golden_util as G
from util import assert_rows_match

pytestmark = pytest.mark.gpu
Q1_F = ["sum_qty", "sum_base_price", "sum_disc_price", "sum_charge", "avg_qty", "avg_price", "avg_disc"]


@pytest.fixture(scope="module")
def gold():
    return G.load()


def lineitem_batches(gold, O, names):
    return [[O.to_arrow(G.lineitem_columns(p, O.Utf8), names)] for p in gold["lineitem"]["partitions"]]


@pytest.mark.parametrize("fused", [True, False])
def test_q1_fixture(ctx, O, gold, fused, monkeypatch):
    """Two scan partitions (partition0.tbl, partition1.tbl) -> Partial per partition -> Merge -> Final."""
    if not fused:
        monkeypatch.setenv("B2_DISABLE_FUSION", "1")
    from ballista_b200 import MemoryExec, tpch
    scan = tpch.scan(ctx, lineitem_batches(gold, O, tpch.Q1_COLUMNS))
    assert scan.output_partitioning() == ("UnknownPartitioning", 2)
    partial = tpch.q1_partial(scan)
    assert partial.schema().names == ["l_returnflag", "l_linestatus", "sum_qty[sum]", "sum_base_price[sum]",
                                      "sum_disc_price[sum]", "sum_charge[sum]", "avg_qty[count]", "avg_qty[sum]",
                                      "avg_price[count]", "avg_price[sum]", "avg_disc[count]", "avg_disc[sum]",
                                      "count_order[count]"]  # SURVEY A.2 partial-state schema
    states = [s for p in range(2) for s in partial.collect_device(p)]
    final = tpch.q1_final(MemoryExec(ctx, partial.schema(), [states]), scan.schema())
    got = final.collect().to_pylist()
    assert_rows_match(got, gold["expected"]["q1"], ["l_returnflag", "l_linestatus"], exact=["count_order"], approx=Q1_F)
    t = final.schema()
    assert str(t.field("count_order").type) == "uint64" and str(t.field("avg_disc").type) == "double"


def test_q6_fixture_excludes_discount_007(ctx, O, gold):
    from ballista_b200 import (CoalesceBatchesExec, FilterExec, HashAggregateExec, MemoryExec, col, count, date, lit,
                               tpch)
    e = gold["expected"]["q6_wide"]
    pred = ((col("l_shipdate") >= date("1994-01-01")) & (col("l_shipdate") < date("1997-01-01"))
            & col("l_discount").between(lit(0.06) - lit(0.01), lit(0.06) + lit(0.01)) & (col("l_quantity") < lit(50)))
    scan = tpch.scan(ctx, lineitem_batches(gold, O, tpch.Q6_COLUMNS))
    partial = HashAggregateExec("Partial", [], tpch.q6_aggregates() + [count().alias("n")],
                                CoalesceBatchesExec(FilterExec(pred, scan)))
    states = [s for p in range(2) for s in partial.collect_device(p)]
    final = HashAggregateExec("Final", [], tpch.q6_aggregates() + [count().alias("n")],
                              tpch.MergeExec(MemoryExec(ctx, partial.schema(), [states])), input_schema=scan.schema())
    got = final.collect().to_pylist()
    assert got[0]["n"] == e["rows_passing"] == 4
    assert math.isclose(got[0]["revenue"], e["revenue"], rel_tol=1e-9)


def test_join_and_q3_shaped_fixture(ctx, O, gold):
    from ballista_b200 import CoalesceBatchesExec, HashAggregateExec, HashJoinExec, MemoryExec, tpch
    li = G.concat_lineitem(gold["lineitem"]["partitions"], O.Utf8)
    orders = pa.RecordBatch.from_arrays(
        [pa.array(np.array(gold["orders"][k], dtype=np.int32)) if k != "o_orderdate"
         else pa.array(np.array(gold["orders"][k], dtype=np.int32)).cast(pa.date32()) for k in tpch.Q3_ORDERS_COLUMNS],
        names=tpch.Q3_ORDERS_COLUMNS)
    lrb = O.to_arrow(li, ["l_orderkey", "l_linenumber", "l_extendedprice", "l_discount"])
    o_scan, l_scan = tpch.scan(ctx, [[orders]]), tpch.scan(ctx, [[lrb]])
    join = CoalesceBatchesExec(HashJoinExec(o_scan, l_scan, [("o_orderkey", "l_orderkey")]))
    # output schema: all left fields, then right fields (the right key has a different name, so it is kept)
    assert join.schema().names == tpch.Q3_ORDERS_COLUMNS + ["l_orderkey", "l_linenumber", "l_extendedprice", "l_discount"]
    out = join.collect()
    got = sorted(zip(out["o_orderkey"].to_pylist(), out["l_linenumber"].to_pylist()))
    assert got == [tuple(x) for x in gold["expected"]["join_orders_lineitem"]]
    assert out["o_orderkey"].to_pylist() == out["l_orderkey"].to_pylist()
    # Q3-shaped aggregate on top of the join (unfused path: input is a join)
    partial = HashAggregateExec("Partial", tpch.q3_groups(), tpch.q3_aggregates(), join)
    states = partial.collect_device(0)
    final = tpch.q3_final(MemoryExec(ctx, partial.schema(), [states]), join.schema())
    rows = final.collect().to_pylist()
    for r in rows:
        r["o_orderdate"] = (r["o_orderdate"] - __import__("datetime").date(1970, 1, 1)).days
    assert_rows_match(rows, gold["expected"]["q3_nofilter"], ["l_orderkey", "o_orderdate", "o_shippriority"],
                      approx=["revenue"])


def test_empty_join_fixture(ctx, gold):
    from ballista_b200 import HashJoinExec, tpch
    cust = pa.RecordBatch.from_arrays([pa.array(np.array(gold["customer"]["c_custkey"], dtype=np.int32)),
                                       pa.array(gold["customer"]["c_mktsegment"], pa.utf8())],
                                      names=["c_custkey", "c_mktsegment"])
    orders = pa.RecordBatch.from_arrays([pa.array(np.array(gold["orders"][k], dtype=np.int32))
                                         for k in ("o_orderkey", "o_custkey")], names=["o_orderkey", "o_custkey"])
    j = HashJoinExec(tpch.scan(ctx, [[cust]]), tpch.scan(ctx, [[orders]]), [("c_custkey", "o_custkey")])
    out = j.collect()
    assert out.num_rows == 0 and out.schema.names == ["c_custkey", "c_mktsegment", "o_orderkey", "o_custkey"]


def test_q3_partial_state_schema_and_two_partition_final(ctx, O, gold):
    """The state schema the Partial stage of Q3 ships to the Final stage (group columns, then one `name[sum]` state per
    SUM: SURVEY A.2), and a Final over the states of TWO Partial tasks (the fixture's two lineitem partitions joined
    against the same orders) against the exact-arithmetic answer."""
    import datetime
    from ballista_b200 import CoalesceBatchesExec, HashAggregateExec, HashJoinExec, MemoryExec, tpch
    orders = pa.RecordBatch.from_arrays(
        [pa.array(np.array(gold["orders"][k], dtype=np.int32)) if k != "o_orderdate"
         else pa.array(np.array(gold["orders"][k], dtype=np.int32)).cast(pa.date32()) for k in tpch.Q3_ORDERS_COLUMNS],
        names=tpch.Q3_ORDERS_COLUMNS)
    lparts = [[O.to_arrow(G.lineitem_columns(p, O.Utf8), ["l_orderkey", "l_linenumber", "l_extendedprice", "l_discount"])]
              for p in gold["lineitem"]["partitions"]]
    join = CoalesceBatchesExec(HashJoinExec(tpch.scan(ctx, [[orders]]), tpch.scan(ctx, lparts),
                                            [("o_orderkey", "l_orderkey")]))
    assert join.output_partitioning()[1] == 2  # the probe side's partitions
    partial = HashAggregateExec("Partial", tpch.q3_groups(), tpch.q3_aggregates(), join)
    ps = partial.schema()
    assert ps.names == ["l_orderkey", "o_orderdate", "o_shippriority", "revenue[sum]"]
    assert [str(ps.field(n).type) for n in ps.names] == ["int32", "date32[day]", "int32", "double"]
    states = [s for p in range(2) for s in partial.collect_device(p)]
    assert len(states) == 2
    final = tpch.q3_final(MemoryExec(ctx, ps, [states]), join.schema())
    rows = final.collect().to_pylist()
    for r in rows:
        r["o_orderdate"] = (r["o_orderdate"] - datetime.date(1970, 1, 1)).days
    assert_rows_match(rows, gold["expected"]["q3_nofilter"], ["l_orderkey", "o_orderdate", "o_shippriority"],
                      approx=["revenue"])


@pytest.mark.parametrize("n_out", [2, 3])
def test_repartition_fixture_keeps_arrival_order(ctx, O, gold, n_out):
    """RepartitionExec(Hash(l_orderkey), n) over the fixture's two lineitem partitions: every input row comes out once;
    an order's lines share one output partition; inside an output partition the rows appear in arrival order (input
    partition 0 before 1, file order within each) - what the reference's per-input-batch `take` + channel send
    produces for a single consumer draining the inputs in order."""
    from ballista_b200 import RepartitionExec, col, tpch
    names = ["l_orderkey", "l_linenumber", "l_extendedprice"]
    parts = [[O.to_arrow(G.lineitem_columns(p, O.Utf8), names)] for p in gold["lineitem"]["partitions"]]
    rep = RepartitionExec(tpch.scan(ctx, parts), "Hash", n_out, [col("l_orderkey")])
    arrival = [(k, ln) for p in gold["lineitem"]["partitions"] for k, ln in zip(p["l_orderkey"], p["l_linenumber"])]
    seen, owner = [], {}
    for d in range(n_out):
        t = rep.collect(d)
        rows = list(zip(t["l_orderkey"].to_pylist(), t["l_linenumber"].to_pylist()))
        for k, _ in rows:
            assert owner.setdefault(k, d) == d
        assert rows == [r for r in arrival if owner.get(r[0]) == d]  # a subsequence of the arrival order
        seen += rows
    assert sorted(seen) == sorted(arrival) and len(seen) == 20
