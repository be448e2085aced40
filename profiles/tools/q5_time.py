"""Q5 on one GPU, timed: python profiles/tools/q5_time.py <sf> <steps>"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import ballista_b200 as b2
from ballista_b200 import tpch
ctx = b2.Context(0)
sf = float(sys.argv[1]) if len(sys.argv) > 1 else 10
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 10
batches = [ctx.tpch(t, sf, columns=c) for t, c in tpch.Q5_TABLES]
def step():
    scans = [tpch.scan(ctx, [[b]]) for b in batches]
    partial = tpch.q5_partial(*scans)
    q = tpch.TwoStage(ctx, partial, lambda s: tpch.q5_final(s, partial.children()[0].schema()))
    return q.run()
for _ in range(3): out = step()
ctx.synchronize(); l0 = ctx.kernel_launches(); t = time.perf_counter()
for _ in range(steps): out = step()
ctx.synchronize(); dt = (time.perf_counter() - t) / steps
print(f"Q5 SF{sf:g}: {dt*1e3:.2f} ms/step, launches/step {(ctx.kernel_launches()-l0)/steps}, "
      f"first row {out.to_pylist()[0]}")
