"""Parity of the fused and unfused scan->filter->aggregate paths (TPC-H Q1, Q6) against the oracle."""
import os

import numpy as np
import pytest

from util import assert_rows_match

pytestmark = pytest.mark.gpu

Q1_F = ["sum_qty", "sum_base_price", "sum_disc_price", "sum_charge", "avg_qty", "avg_price", "avg_disc"]


def run_q1(ctx, batches, merge_parts=1):
    from ballista_b200 import tpch
    scan = tpch.scan(ctx, batches)
    partial = tpch.q1_partial(scan)
    states = [s for p in range(len(batches)) for s in partial.collect_device(p)]
    from ballista_b200 import MemoryExec
    final = tpch.q1_final(MemoryExec(ctx, partial.schema(), [states]), scan.schema())
    return final.collect().to_pylist(), partial


@pytest.mark.parametrize("n", [1, 5, 1023, 1 << 16, 300_000])
def test_generator_matches_host(ctx, O, n):
    sf = 0.1
    dev = ctx.tpch("lineitem", sf, 7 * 32, 7 * 32 + n).to_arrow()
    host = O.gen_lineitem(sf, 7 * 32, n)
    for name in ["l_orderkey", "l_suppkey", "l_quantity", "l_extendedprice", "l_discount", "l_tax", "l_shipdate",
                 "l_commitdate", "l_receiptdate"]:
        got = dev.column(name).to_numpy(zero_copy_only=False)
        if "date" in name:
            got = dev.column(name).cast("int32").to_numpy()
        assert np.array_equal(got.view(np.uint8), np.ascontiguousarray(host[name]).view(np.uint8)), name
    assert dev.column("l_returnflag").to_pylist() == host["l_returnflag"].to_list()
    assert dev.column("l_linestatus").to_pylist() == host["l_linestatus"].to_list()


@pytest.mark.parametrize("tma", [True, False])
@pytest.mark.parametrize("n,parts", [(1, 1), (3, 1), (959, 1), (960, 1), (961, 1), (4097, 1), (1 << 16, 1), (300_000, 3)])
def test_q1_fused_matches_oracle(ctx, O, n, parts, tma, monkeypatch):
    """tma=True: the TMA-fed bandwidth kernel (agg_tma.cuh); False: the register-load kernel it falls back to."""
    if not tma:
        monkeypatch.setenv("B2_DISABLE_TMA", "1")
    sf = 0.05
    bounds = [n * p // parts for p in range(parts + 1)]
    batches = [[ctx.tpch("lineitem", sf, bounds[p], bounds[p + 1])] for p in range(parts)]
    got, partial = run_q1(ctx, batches)
    assert partial.last_path().startswith("fused:"), partial.last_path()
    assert ("[tma" in partial.last_path()) == tma, partial.last_path()
    want = O.q1(O.gen_lineitem(sf, 0, n), n_partitions=parts)
    assert_rows_match(got, want, ["l_returnflag", "l_linestatus"], exact=["count_order"], approx=Q1_F)


def test_q1_unfused_matches_oracle(ctx, O, monkeypatch):
    monkeypatch.setenv("B2_DISABLE_FUSION", "1")
    sf, n = 0.05, 100_000
    got, partial = run_q1(ctx, [[ctx.tpch("lineitem", sf, 0, n)]])
    assert partial.last_path().startswith("unfused"), partial.last_path()
    want = O.q1(O.gen_lineitem(sf, 0, n))
    assert_rows_match(got, want, ["l_returnflag", "l_linestatus"], exact=["count_order"], approx=Q1_F)


def test_q1_host_batches_through_arrow(ctx, O):
    """Host RecordBatches (32,768-row batches like the reference's CsvExec) pushed through the C ABI."""
    sf, n = 0.05, 150_000
    cols = O.gen_lineitem(sf, 0, n)
    from ballista_b200 import tpch
    rb = O.to_arrow(cols, tpch.Q1_COLUMNS)
    batches = [[rb.slice(o, 32768) for o in range(0, n, 32768)]]
    got, partial = run_q1(ctx, batches)
    want = O.q1(cols)
    assert_rows_match(got, want, ["l_returnflag", "l_linestatus"], exact=["count_order"], approx=Q1_F)


@pytest.mark.parametrize("fused", [True, False])
def test_q6_matches_oracle(ctx, O, fused, monkeypatch):
    if not fused:
        monkeypatch.setenv("B2_DISABLE_FUSION", "1")
    from ballista_b200 import MemoryExec, tpch
    sf, n = 0.1, 600_000
    li = ctx.tpch("lineitem", sf, 0, n, columns=tpch.Q6_COLUMNS)
    scan = tpch.scan(ctx, [[li]])
    partial = tpch.q6_partial(scan)
    states = partial.collect_device(0)
    assert partial.last_path().startswith("fused:" if fused else "unfused"), partial.last_path()
    got = tpch.q6_final(MemoryExec(ctx, partial.schema(), [states]), scan.schema()).collect().to_pylist()
    rev, passing = O.q6(O.gen_lineitem(sf, 0, n))
    assert passing > 0
    assert abs(got[0]["revenue"] - rev) <= 1e-9 * abs(rev)
    # the f64-folded BETWEEN bounds exclude l_discount == 0.07: count through a COUNT aggregate
    from ballista_b200 import CoalesceBatchesExec, FilterExec, HashAggregateExec, count
    cnt = HashAggregateExec("Partial", [], [count().alias("c")],
                            CoalesceBatchesExec(FilterExec(tpch.q6_predicate(), tpch.scan(ctx, [[li]]))))
    assert cnt.collect().to_pylist()[0]["c[count]"] == passing  # bit-exact


@pytest.mark.parametrize("mapping,expect", [
    ({"R": "Ret", "A": "Acc", "N": "N"}, "[tma,narrow]/small"),           # <= 3 characters: narrow 8-byte keys
    ({"R": "Return", "A": "Accept", "N": "N"}, "/small+widekeys"),           # overflow -> redone with 15-byte keys
])
def test_q1_long_flag_strings(ctx, O, mapping, expect):
    """Utf8 group keys of several characters: the narrow-key kernel handles <= 3, longer ones fall back to wide keys."""
    import pyarrow as pa
    from ballista_b200 import tpch
    sf, n = 0.05, 70_000
    cols = O.gen_lineitem(sf, 0, n)
    want = O.q1(cols)
    rb = O.to_arrow(cols, tpch.Q1_COLUMNS)
    rf = pa.array([mapping[v] for v in rb.column("l_returnflag").to_pylist()], pa.utf8())
    rb = rb.set_column(rb.schema.get_field_index("l_returnflag"), "l_returnflag", rf)
    got, partial = run_q1(ctx, [[rb]])
    assert partial.last_path().endswith(expect), partial.last_path()
    for r in got:
        assert r["l_returnflag"] in mapping.values()
        r["l_returnflag"] = r["l_returnflag"][0]
    assert_rows_match(got, want, ["l_returnflag", "l_linestatus"], exact=["count_order"], approx=Q1_F)


def test_q1_key_strings_beyond_packed_budget_are_rejected(ctx, O):
    """Two Utf8 keys share 15 packed bytes (6 characters each): longer values are NotImplemented, loudly."""
    import pyarrow as pa
    from ballista_b200 import NotImplementedOnGpu, tpch
    cols = O.gen_lineitem(0.05, 0, 5000)
    rb = O.to_arrow(cols, tpch.Q1_COLUMNS)
    rf = pa.array(["Returned!" for _ in range(rb.num_rows)], pa.utf8())
    rb = rb.set_column(rb.schema.get_field_index("l_returnflag"), "l_returnflag", rf)
    with pytest.raises(NotImplementedOnGpu):
        run_q1(ctx, [[rb]])


@pytest.mark.parametrize("n", [3000, 200_000])
def test_fused_high_cardinality_falls_back_to_global_table(ctx, O, n):
    """group by l_orderkey: the per-CTA directory overflows (> 128 groups), the launch aborts and the batch is redone
    on the HBM hash table (agg_global_kernel) — regression for the directory's out-of-ids sentinel."""
    from ballista_b200 import HashAggregateExec, MemoryExec, col, sum as sum_, tpch
    sf = 0.05
    li = ctx.tpch("lineitem", sf, 0, n, columns=["l_orderkey", "l_quantity"])
    partial = HashAggregateExec("Partial", [col("l_orderkey")], [sum_(col("l_quantity")).alias("q")], tpch.scan(ctx, [[li]]))
    got = partial.collect().to_pylist()
    assert partial.last_path().startswith("fused:") and partial.last_path().endswith("/global"), partial.last_path()
    host = O.gen_lineitem(sf, 0, n)
    rep, sums, counts = O.hash_aggregate([host["l_orderkey"]], [host["l_quantity"]])
    want = [{"l_orderkey": int(host["l_orderkey"][r]), "q[sum]": float(sums[g, 0])} for g, r in enumerate(rep)]
    assert_rows_match(got, want, ["l_orderkey"], approx=["q[sum]"])


def test_high_cardinality_over_growing_batches_regrows_the_group_table(ctx, O):
    """group by l_orderkey over batches of 4 k, 40 k and 260 k rows in ONE partition: the HBM group table is re-sized
    between batches (table_merge_kernel moves the groups claimed so far, listing them behind one reservation per CTA
    round) and every group must come out once with the sum over all batches."""
    from ballista_b200 import HashAggregateExec, col, count, sum as sum_, tpch
    sf, cuts = 0.1, [0, 4_000, 44_000, 304_000]
    bs = [ctx.tpch("lineitem", sf, a, b, columns=["l_orderkey", "l_quantity"]) for a, b in zip(cuts, cuts[1:])]
    partial = HashAggregateExec("Partial", [col("l_orderkey")],
                                [sum_(col("l_quantity")).alias("q"), count(col("l_quantity")).alias("c")],
                                tpch.scan(ctx, [bs]))
    got = partial.collect().to_pylist()
    host = O.gen_lineitem(sf, 0, cuts[-1])
    rep, sums, counts = O.hash_aggregate([host["l_orderkey"]], [host["l_quantity"]])
    want = [{"l_orderkey": int(host["l_orderkey"][r]), "q[sum]": float(sums[g, 0]), "c[count]": int(counts[g, 0])}
            for g, r in enumerate(rep)]
    assert len(got) == len(want) > 70_000
    assert_rows_match(got, want, ["l_orderkey"], exact=["c[count]"], approx=["q[sum]"])
