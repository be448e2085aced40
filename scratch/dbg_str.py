import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import pyarrow as pa, numpy as np
import ballista_b200 as b2
from ballista_b200 import tpch, MemoryExec, HashAggregateExec, col, sum as sum_
ctx = b2.Context(0)
b = pa.RecordBatch.from_arrays([pa.array(["INDONESIA", "ABCDEFGHIJKLMN", "INDONESIA", "XY"]), pa.array([1.0, 2.0, 3.0, 4.0])], names=["s", "v"])
for env in ({}, {"B2_DISABLE_SPECULATION": "1"}):
    os.environ.update(env)
    p = HashAggregateExec("Partial", [col("s")], [sum_(col("v")).alias("t")], tpch.scan(ctx, [[b]]))
    st = p.collect_device(0)
    print(env, "partial:", st[0].to_arrow().to_pylist(), p.last_path())
    f = HashAggregateExec("Final", [col("s")], [sum_(col("v")).alias("t")], MemoryExec(ctx, p.schema(), [st]), input_schema=b.schema)
    print("final:", f.collect().to_pylist(), f.last_path())
