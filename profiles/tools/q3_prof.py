import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import ballista_b200 as b2
from ballista_b200 import tpch, MemoryExec
ctx = b2.Context(0)
sf = float(sys.argv[1]) if len(sys.argv) > 1 else 10
c = ctx.tpch("customer", sf, columns=tpch.Q3_CUSTOMER_COLUMNS)
o = ctx.tpch("orders", sf, columns=tpch.Q3_ORDERS_COLUMNS)
l = ctx.tpch("lineitem", sf, columns=tpch.Q3_LINEITEM_COLUMNS)
ctx.synchronize()
T = {}
def tick(name, t0):
    ctx.synchronize(); T[name] = T.get(name, 0) + time.perf_counter() - t0
N = 20
for it in range(N + 3):
    if it == 3: T.clear()
    t = time.perf_counter()
    cust, orders, line = (tpch.scan(ctx, [[x]]) for x in (c, o, l))
    partial = tpch.q3_partial(cust, orders, line); tick("build partial plan", t)
    t = time.perf_counter(); states = partial.collect_device(0); tick("partial execute", t)
    t = time.perf_counter(); final = tpch.q3_final(MemoryExec(ctx, partial.schema(), [states]), partial.children()[0].schema()); tick("build final plan", t)
    t = time.perf_counter(); out = final.collect_device(0); tick("final execute", t)
    t = time.perf_counter(); del cust, orders, line, partial, states, final, out; tick("teardown", t)
for k, v in T.items(): print(f"{k:24s} {v / N * 1e3:8.3f} ms")
if os.environ.get("B2_TRACE"): print(ctx.trace_report() if hasattr(ctx, "trace_report") else "")
