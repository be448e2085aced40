import sys, os, time, cProfile, pstats
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import ballista_b200 as b2
from ballista_b200 import tpch, MemoryExec
ctx = b2.Context(0)
li = ctx.tpch("lineitem", 0.01, columns=tpch.Q1_COLUMNS); ctx.synchronize()
def build():
    scan = tpch.scan(ctx, [[li]]); partial = tpch.q1_partial(scan)
    final = tpch.q1_final(MemoryExec(ctx, partial.schema(), [[]]), scan.schema())
    return partial, final
for _ in range(50): build()
t = time.perf_counter()
for _ in range(500): build()
print("plan build (partial+final) ms:", (time.perf_counter() - t) / 500 * 1e3)
pr = cProfile.Profile(); pr.enable()
for _ in range(500): build()
pr.disable()
pstats.Stats(pr).sort_stats("tottime").print_stats(18)
