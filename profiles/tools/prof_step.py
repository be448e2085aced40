import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import ballista_b200 as b2
from ballista_b200 import tpch, MemoryExec
ctx = b2.Context(0)
sf = float(sys.argv[1]) if len(sys.argv) > 1 else 10
li = ctx.tpch("lineitem", sf, columns=tpch.Q1_COLUMNS); ctx.synchronize()
T = {}
def tick(name, t0):
    ctx.synchronize(); T[name] = T.get(name, 0) + time.perf_counter() - t0
N = 20
for it in range(N + 3):
    if it == 3: T.clear(); ctx.set_profiling(2)
    t = time.perf_counter(); scan = tpch.scan(ctx, [[li]]); partial = tpch.q1_partial(scan); tick("build partial plan", t)
    t = time.perf_counter(); states = partial.collect_device(0); tick("partial execute", t)
    t = time.perf_counter(); final = tpch.q1_final(MemoryExec(ctx, partial.schema(), [states]), scan.schema()); tick("build final plan", t)
    t = time.perf_counter(); tab = final.collect(); tick("final execute+D2H", t)
    t = time.perf_counter(); del scan, partial, states, final; tick("teardown", t)
for k, v in T.items(): print(f"{k:24s} {v / N * 1e3:8.3f} ms")
print("launches/step", ctx.kernel_launches() / (N + 3))

del li
import gc; gc.collect()
ctx.close()
