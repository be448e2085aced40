"""A few steps of Q1 / Q6 / Q3 for an ncu launch list: python profiles/tools/steps_for_ncu.py <sf> <steps>"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import ballista_b200 as b2
from ballista_b200 import tpch, MemoryExec
ctx = b2.Context(0)
sf = float(sys.argv[1]) if len(sys.argv) > 1 else 10.0
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
which = sys.argv[3] if len(sys.argv) > 3 else "q1,q6,q3"
if "q1" in which:
    li = ctx.tpch("lineitem", sf, columns=tpch.Q1_COLUMNS)
    scan = tpch.scan(ctx, [[li]])
    q = tpch.TwoStage(ctx, tpch.q1_partial(scan), lambda s: tpch.q1_final(s, scan.schema()))
    for _ in range(steps): q.run()
    del li, scan, q
if "q6" in which:
    li = ctx.tpch("lineitem", sf, columns=tpch.Q6_COLUMNS)
    scan = tpch.scan(ctx, [[li]])
    q = tpch.TwoStage(ctx, tpch.q6_partial(scan), lambda s: tpch.q6_final(s, scan.schema()))
    for _ in range(steps): q.run()
    del li, scan, q
if "q3" in which:
    cu = ctx.tpch("customer", sf, columns=tpch.Q3_CUSTOMER_COLUMNS)
    od = ctx.tpch("orders", sf, columns=tpch.Q3_ORDERS_COLUMNS)
    li = ctx.tpch("lineitem", sf, columns=tpch.Q3_LINEITEM_COLUMNS)
    for _ in range(steps):
        c, o, l = (tpch.scan(ctx, [[x]]) for x in (cu, od, li))
        partial = tpch.q3_partial(c, o, l)
        states = partial.collect_device(0)
        final = tpch.q3_final(MemoryExec(ctx, partial.schema(), [states]), partial.children()[0].schema())
        n = sum(b.num_rows for b in final.collect_device(0))
    print("q3 groups", n)
ctx.synchronize()
