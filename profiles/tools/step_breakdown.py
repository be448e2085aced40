"""Where a cached-plan Q6 / Q1 step spends its host time (wall clock per phase; the GPU runs asynchronously)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import ballista_b200 as b2
from ballista_b200 import tpch, MemoryExec

ctx = b2.Context(0)
sf = float(sys.argv[1]) if len(sys.argv) > 1 else 10.0
q = sys.argv[2] if len(sys.argv) > 2 else "q6"
cols = tpch.Q6_COLUMNS if q == "q6" else tpch.Q1_COLUMNS
li = ctx.tpch("lineitem", sf, columns=cols)
mk_partial = tpch.q6_partial if q == "q6" else tpch.q1_partial
mk_final = tpch.q6_final if q == "q6" else tpch.q1_final
scan = tpch.scan(ctx, [[li]])
part = mk_partial(scan)
leaf = MemoryExec(ctx, part.schema(), [[]])
fin = mk_final(leaf, scan.schema())
ctx.synchronize()
T = {}
def tick(name, t0):
    T[name] = T.get(name, 0.0) + time.perf_counter() - t0
N = 300
for it in range(N + 20):
    if it == 20:
        T.clear()
        ctx.synchronize()
        t_all = time.perf_counter()
    t0 = time.perf_counter(); states = part.collect_device(0); tick("partial: execute + next_device (async)", t0)
    t0 = time.perf_counter(); leaf.clear(); [leaf.add(0, s) for s in states]; tick("re-feed final leaf", t0)
    t0 = time.perf_counter(); st = fin.execute(0); tick("final: execute()", t0)
    t0 = time.perf_counter(); it_ = iter(st); rb = next(it_); tick("final: stream_next (kernels + D2H + sync) + import", t0)
    t0 = time.perf_counter(); rest = list(it_); import pyarrow as pa; tab = pa.Table.from_batches([rb], schema=fin.schema()); tick("eos + Table", t0)
    t0 = time.perf_counter(); del states, st, it_, rb, rest, tab; tick("destroy", t0)
ctx.synchronize()
total = time.perf_counter() - t_all
print(f"{q} SF{sf:g}: {total / N * 1e3:.3f} ms/step wall")
for k, v in T.items():
    print(f"  {k:56s} {v / N * 1e6:8.1f} us")
