"""GPU: SortExec / GlobalLimitExec / LocalLimitExec (the operators after the Final aggregate: q1.sql ORDER BY flags,
q3.sql ORDER BY revenue DESC, o_orderdate LIMIT 10) against numpy / Python sorts; bit-exact (a sort only moves rows)."""
import numpy as np
import pyarrow as pa
import pytest

pytestmark = pytest.mark.gpu


def test_sort_multi_key_types_and_stability(ctx):
    from ballista_b200 import SortExec, col, tpch
    rng = np.random.default_rng(11)
    n = 70_001  # several tiles, ragged tail
    f = rng.integers(-5, 6, n).astype(np.float64) / 4.0  # many ties, negative values, -0.0 / 0.0
    f[::97] = -0.0
    i = rng.integers(-2**31, 2**31 - 1, n).astype(np.int32)
    s = np.array(["", "A", "AB", "B", "BUILDING", "AUTOMOBILE", "Z" * 16], dtype=object)[rng.integers(0, 7, n)]
    rb = pa.RecordBatch.from_arrays([pa.array(f), pa.array(i), pa.array(list(s)), pa.array(np.arange(n, dtype=np.int64))],
                                    names=["f", "i", "s", "row"])
    parts = [[rb.slice(0, 30_000), rb.slice(30_000, 1)], ]
    parts[0].append(rb.slice(30_001))
    plan = SortExec([(col("s"), True), (col("f"), False), (col("i"), True)], tpch.scan(ctx, parts))
    got = plan.collect()
    rows = list(zip(s, f, i, range(n)))
    # python sort is stable; -0.0 == 0.0 compare equal in python but differ in IEEE total order: normalise like the kernel
    def fkey(x):
        b = np.float64(x).view(np.uint64)
        return int(~b & 0xFFFFFFFFFFFFFFFF) if b >> 63 else int(b | (1 << 63))
    want = sorted(rows, key=lambda r: (r[0].encode(), -fkey(r[1]), int(r[2])))
    assert got["row"].to_pylist() == [r[3] for r in want]
    assert got["s"].to_pylist() == [r[0] for r in want]
    assert plan.output_partitioning()[1] == 1


def test_sort_rejects_multiple_partitions_and_long_strings(ctx):
    from ballista_b200 import BallistaError, SortExec, col, tpch
    rb = pa.RecordBatch.from_arrays([pa.array(["x" * 17, "y"])], names=["s"])
    with pytest.raises(BallistaError):
        SortExec([col("s")], tpch.scan(ctx, [[rb]])).collect()
    ok = pa.RecordBatch.from_arrays([pa.array(["b", "a"])], names=["s"])
    with pytest.raises(BallistaError):
        SortExec([col("s")], tpch.scan(ctx, [[ok], [ok]])).collect()


def test_q1_order_by_and_q3_top10(ctx, O):
    import datetime
    from ballista_b200 import MemoryExec, tpch
    sf, n = 0.05, 200_000
    li = ctx.tpch("lineitem", sf, 0, n, columns=tpch.Q1_COLUMNS)
    scan = tpch.scan(ctx, [[li]])
    partial = tpch.q1_partial(scan)
    states = partial.collect_device(0)
    final = tpch.q1_final(MemoryExec(ctx, partial.schema(), [states]), scan.schema())
    got = tpch.q1_sorted(final).collect().to_pylist()
    keys = [(r["l_returnflag"], r["l_linestatus"]) for r in got]
    assert keys == sorted(keys) and len(keys) == 4

    sf3 = 0.1
    cust = tpch.scan(ctx, [[ctx.tpch("customer", sf3, columns=tpch.Q3_CUSTOMER_COLUMNS)]])
    orders = tpch.scan(ctx, [[ctx.tpch("orders", sf3, columns=tpch.Q3_ORDERS_COLUMNS)]])
    line = tpch.scan(ctx, [[ctx.tpch("lineitem", sf3, columns=tpch.Q3_LINEITEM_COLUMNS)]])
    partial = tpch.q3_partial(cust, orders, line)
    states = partial.collect_device(0)
    def final():
        return tpch.q3_final(MemoryExec(ctx, partial.schema(), [states]), partial.children()[0].schema())
    everything = final().collect().to_pylist()
    top = tpch.q3_top(final(), 10).collect().to_pylist()
    want = sorted(everything, key=lambda r: (-r["revenue"], r["o_orderdate"]))[:10]
    assert [(r["l_orderkey"], r["revenue"]) for r in top] == [(r["l_orderkey"], r["revenue"]) for r in want]
    assert len(top) == 10


def test_limits(ctx):
    from ballista_b200 import GlobalLimitExec, LocalLimitExec, tpch
    bs = [ctx.tpch("orders", 0.01, i * 100, i * 100 + 100) for i in range(5)]
    g = GlobalLimitExec(tpch.scan(ctx, [bs]), 250).collect()
    assert g.num_rows == 250
    assert g["o_orderkey"].to_pylist() == tpch.scan(ctx, [bs]).collect()["o_orderkey"].to_pylist()[:250]
    l = LocalLimitExec(tpch.scan(ctx, [bs[:2], bs[2:]]), 150)
    assert l.collect(0).num_rows == 150 and l.collect(1).num_rows == 150
    assert GlobalLimitExec(tpch.scan(ctx, [bs]), 0).collect().num_rows == 0
