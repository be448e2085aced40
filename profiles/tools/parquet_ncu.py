"""One decode of a pyarrow-default (snappy + dictionary) lineitem file, for an ncu launch list."""
import io, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import pyarrow as pa, pyarrow.parquet as pq
import ballista_b200 as b2
sf = float(sys.argv[1]) if len(sys.argv) > 1 else 1.0
cols = sys.argv[2].split(",") if len(sys.argv) > 2 else None
ctx = b2.Context(0)
t = pa.Table.from_batches([ctx.tpch("lineitem", sf).to_arrow()])
bio = io.BytesIO(); pq.write_table(t, bio)
f = b2.ParquetFile(bio.getvalue(), pinned_by=ctx)
d = f.read(ctx, cols)
ctx.synchronize()
print(d.num_rows, d.num_bytes)
