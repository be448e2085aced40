import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import ballista_b200 as b2
from ballista_b200 import tpch, MemoryExec
ctx = b2.Context(0)
sf = float(sys.argv[1]) if len(sys.argv) > 1 else 10
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 5
c = ctx.tpch("customer", sf, columns=tpch.Q3_CUSTOMER_COLUMNS)
o = ctx.tpch("orders", sf, columns=tpch.Q3_ORDERS_COLUMNS)
l = ctx.tpch("lineitem", sf, columns=tpch.Q3_LINEITEM_COLUMNS)
ctx.synchronize()
def step():
    cust, orders, line = (tpch.scan(ctx, [[x]]) for x in (c, o, l))
    partial = tpch.q3_partial(cust, orders, line)
    states = partial.collect_device(0)
    final = tpch.q3_final(MemoryExec(ctx, partial.schema(), [states]), partial.children()[0].schema())
    return final.collect_device(0)
for _ in range(2): out = step()
ctx.set_profiling(2); ctx.synchronize(); l0 = ctx.kernel_launches(); t = time.perf_counter()
for _ in range(steps): out = step()
ctx.synchronize(); dt = (time.perf_counter() - t) / steps
rows = c.num_rows + o.num_rows + l.num_rows
scan_bytes = c.num_bytes + o.num_bytes + l.num_bytes
print(f"Q3 SF{sf:g}: {dt*1e3:.2f} ms/step, {rows/dt/1e9:.3f} G input rows/s, scan-only {scan_bytes/dt/1e9:.1f} GB/s, "
      f"groups {sum(b.num_rows for b in out)}, launches/step {(ctx.kernel_launches()-l0)/steps}")
del out, c, o, l
import gc; gc.collect()
ctx.close()
