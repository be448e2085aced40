import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import ballista_b200 as b2
from ballista_b200 import tpch, MemoryExec
ctx = b2.Context(0)
sf = float(sys.argv[1]); steps = int(sys.argv[2])
batches = [ctx.tpch(t, sf, columns=c) for t, c in tpch.Q5_TABLES]
for _ in range(steps):
    scans = [tpch.scan(ctx, [[b]]) for b in batches]
    partial = tpch.q5_partial(*scans)
    q = tpch.TwoStage(ctx, partial, lambda s: tpch.q5_final(s, partial.children()[0].schema()))
    print(q.run().to_pylist()[0])
ctx.synchronize()
