"""Where a Q6 / Q1 step spends its time: Python plan construction vs the two executes, wall clock per phase."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import ballista_b200 as b2
from ballista_b200 import tpch, MemoryExec

ctx = b2.Context(0)
sf = float(sys.argv[1]) if len(sys.argv) > 1 else 10.0
q = sys.argv[2] if len(sys.argv) > 2 else "q6"
cols = tpch.Q6_COLUMNS if q == "q6" else tpch.Q1_COLUMNS
li = ctx.tpch("lineitem", sf, columns=cols)
mk_partial = tpch.q6_partial if q == "q6" else tpch.q1_partial
mk_final = tpch.q6_final if q == "q6" else tpch.q1_final
ctx.synchronize()
T = {}
def tick(name, t0):
    T[name] = T.get(name, 0.0) + time.perf_counter() - t0
N = 200
for it in range(N + 20):
    if it == 20:
        T.clear()
        ctx.synchronize()
        t_all = time.perf_counter()
    t0 = time.perf_counter(); scan = tpch.scan(ctx, [[li]]); part = mk_partial(scan); tick("build partial plan", t0)
    t0 = time.perf_counter(); states = part.collect_device(0); tick("execute partial (to device batch)", t0)
    t0 = time.perf_counter(); sch = part.schema(); tick("partial.schema()", t0)
    t0 = time.perf_counter(); fin = mk_final(MemoryExec(ctx, sch, [states]), scan.schema()); tick("build final plan", t0)
    t0 = time.perf_counter(); tab = fin.collect(); tick("execute final + D2H + to Table", t0)
    t0 = time.perf_counter(); del scan, part, states, fin, tab; tick("destroy", t0)
ctx.synchronize()
total = time.perf_counter() - t_all
print(f"{q} SF{sf:g}: {total / N * 1e3:.3f} ms/step wall")
for k, v in T.items():
    print(f"  {k:40s} {v / N * 1e6:8.1f} us")
