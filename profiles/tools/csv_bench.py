"""Throughput of the GPU .tbl decode (host text in, device columns out) beside pyarrow's multi-threaded CSV reader."""
import io, sys, time, os
sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", "..")); sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", "..", "tests"))
import pyarrow as pa, pyarrow.csv
from test_gpu_csv import tbl_text, LINEITEM, arrow_read
import ballista_b200 as b2
chunk = tbl_text(20_000, seed=3)
reps = int(sys.argv[1]) if len(sys.argv) > 1 else 60
text = chunk * reps
names = ["l_quantity", "l_extendedprice", "l_discount", "l_tax", "l_returnflag", "l_linestatus", "l_shipdate"]
ctx = b2.Context(0)
for _ in range(2):
    b = ctx.read_csv(text, LINEITEM, names); n = b.num_rows; del b
ts = []
for _ in range(5):
    t0 = time.perf_counter(); b = ctx.read_csv(text, LINEITEM, names); ctx.synchronize() if hasattr(ctx, "synchronize") else None
    ts.append(time.perf_counter() - t0); del b
tg = min(ts)
t0 = time.perf_counter(); t = arrow_read(text, LINEITEM, names); ta = time.perf_counter() - t0
print(f"text {len(text)/1e6:.0f} MB rows {n}: gpu decode {tg*1e3:.1f} ms = {len(text)/tg/1e9:.2f} GB/s text, {n/tg/1e6:.0f} Mrows/s; "
      f"pyarrow ({pa.cpu_count()} threads) {ta*1e3:.0f} ms = {len(text)/ta/1e9:.2f} GB/s")
import gc; del t; gc.collect(); ctx.close()
