import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import ballista_b200 as b2
from ballista_b200 import tpch, col, lit, date, FilterExec, CoalesceBatchesExec, HashJoinExec, HashAggregateExec
sf = float(sys.argv[1]) if len(sys.argv) > 1 else 0.02
ctx = b2.Context(0)
def stage(name, plan):
    try:
        t = plan.collect(); ctx.synchronize(); print("ok", name, t.num_rows, flush=True)
    except Exception as e:
        print("FAIL", name, e, flush=True); sys.exit(1)
cust = lambda: tpch.scan(ctx, [[ctx.tpch("customer", sf, columns=tpch.Q3_CUSTOMER_COLUMNS)]])
orders = lambda: tpch.scan(ctx, [[ctx.tpch("orders", sf, columns=tpch.Q3_ORDERS_COLUMNS)]])
line = lambda: tpch.scan(ctx, [[ctx.tpch("lineitem", sf, columns=tpch.Q3_LINEITEM_COLUMNS)]])
stage("cust scan", cust()); stage("orders scan", orders()); stage("line scan", line())
c = lambda: CoalesceBatchesExec(FilterExec(col("c_mktsegment") == lit("BUILDING"), cust()))
o = lambda: CoalesceBatchesExec(FilterExec(col("o_orderdate") < date("1995-03-15"), orders()))
l = lambda: CoalesceBatchesExec(FilterExec(col("l_shipdate") > date("1995-03-15"), line()))
stage("cust filter", c()); stage("orders filter", o()); stage("line filter", l())
j1 = lambda: CoalesceBatchesExec(HashJoinExec(c(), o(), [("c_custkey", "o_custkey")]))
stage("j1", j1())
j2 = lambda: CoalesceBatchesExec(HashJoinExec(j1(), l(), [("o_orderkey", "l_orderkey")]))
stage("j2", j2())
p = HashAggregateExec("Partial", tpch.q3_groups(), tpch.q3_aggregates(), j2())
stage("partial", p); print(p.last_path())
