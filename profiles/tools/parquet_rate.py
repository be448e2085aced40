"""Parquet decode rate on the GPU against pyarrow's reader on the host: lineitem SF1 written by pyarrow."""
import io, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import pyarrow as pa, pyarrow.parquet as pq
import ballista_b200 as b2

sf = float(sys.argv[1]) if len(sys.argv) > 1 else 1.0
ctx = b2.Context(0)
t = pa.Table.from_batches([ctx.tpch("lineitem", sf).to_arrow()])
print("rows", t.num_rows, "arrow bytes", t.nbytes, flush=True)
for name, kw in (("snappy+dict (pyarrow default)", {}), ("uncompressed+dict", dict(compression="none")),
                 ("uncompressed plain", dict(compression="none", use_dictionary=False)),
                 ("snappy plain", dict(compression="snappy", use_dictionary=False))):
    bio = io.BytesIO()
    t0 = time.time(); pq.write_table(t, bio, **kw); tw = time.time() - t0
    data = bio.getvalue()
    f = b2.ParquetFile(data, pinned_by=ctx if os.environ.get("PINNED", "1") == "1" else None)
    for proj, label in ((None, f"all {t.num_columns} columns"), ([t.schema.names.index(c) for c in b2.tpch.Q1_COLUMNS], "Q1's 7 columns")):
        best = 1e9
        for it in range(4):
            ctx.synchronize(); t0 = time.time(); d = f.read(ctx, proj); ctx.synchronize(); best = min(best, time.time() - t0)
        out_bytes = d.num_bytes
        del d
        cpu = {}
        for threads in (False, True):
            b = 1e9
            for it in range(2):
                t0 = time.time()
                r = pq.read_table(io.BytesIO(data), columns=None if proj is None else [t.schema.names[i] for i in proj], use_threads=threads)
                b = min(b, time.time() - t0)
            cpu[threads] = b
        print(f"{name:32s} {label:16s} file {len(data)/1e6:8.1f} MB  gpu {best*1e3:8.2f} ms = {out_bytes/best/1e9:6.2f} GB/s decoded"
              f"   pyarrow 1 thread {cpu[False]*1e3:8.1f} ms, all threads {cpu[True]*1e3:8.1f} ms  (write {tw:.1f} s)", flush=True)
    print(f.describe().split("row_group 0")[1].split("row_group 1")[0][:1400])
