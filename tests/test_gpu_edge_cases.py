"""GPU: edge cases of the fused join / repartition / sort paths (empty and ragged inputs, filters that reject everything,
extreme key values, duplicate build keys under a fused filter, fused vs unfused agreement)."""
import os

import numpy as np
import pyarrow as pa
import pytest

pytestmark = pytest.mark.gpu


def rb(**cols):
    return pa.RecordBatch.from_arrays([pa.array(v) for v in cols.values()], names=list(cols))


def join_rows(ctx, left, right, on, lpred=None, rpred=None):
    from ballista_b200 import CoalesceBatchesExec, FilterExec, HashJoinExec, tpch
    l = tpch.scan(ctx, left)
    r = tpch.scan(ctx, right)
    if lpred is not None:
        l = CoalesceBatchesExec(FilterExec(lpred, l))
    if rpred is not None:
        r = CoalesceBatchesExec(FilterExec(rpred, r))
    j = HashJoinExec(l, r, on)
    out = []
    for p in range(j.output_partitioning()[1]):
        t = j.collect(p)
        out += list(zip(*[t[c].to_pylist() for c in t.schema.names]))
    return sorted(out), j


def test_join_extreme_keys_empty_sides_and_ragged_batches(ctx):
    from ballista_b200 import col, lit
    i32 = np.int32
    lk = np.array([-1, -2**31, 2**31 - 1, 0, 7, 7, -1], dtype=i32)  # -1 = 0xFFFFFFFF (the table's EMPTY key bits), dups
    rk = np.array([7, -1, 5, -2**31, 2**31 - 1, 0, 0, 9], dtype=i32)
    left = rb(k=lk, lv=np.arange(len(lk), dtype=np.int64))
    right = rb(k2=rk, rv=np.arange(len(rk), dtype=np.int64))
    want = sorted((int(a), i, int(b), j) for i, a in enumerate(lk) for j, b in enumerate(rk) if a == b)
    got, j = join_rows(ctx, [[left]], [[right.slice(0, 3), right.slice(3, 0)], [right.slice(3)]], [("k", "k2")])
    assert got == want and "k32" in j.last_path()
    # empty build side, empty probe side
    got, _ = join_rows(ctx, [[left.slice(0, 0)]], [[right]], [("k", "k2")])
    assert got == []
    got, _ = join_rows(ctx, [[left]], [[right.slice(0, 0)]], [("k", "k2")])
    assert got == []
    # fused filters: reject everything on either side, then a selective one with duplicate build keys
    got, _ = join_rows(ctx, [[left]], [[right]], [("k", "k2")], lpred=col("lv") < lit(0))
    assert got == []
    got, _ = join_rows(ctx, [[left]], [[right]], [("k", "k2")], rpred=col("rv") > lit(100))
    assert got == []
    got, j = join_rows(ctx, [[left]], [[right]], [("k", "k2")], lpred=col("lv") >= lit(4), rpred=col("rv") <= lit(1))
    want = sorted((int(a), i, int(b), jx) for i, a in enumerate(lk) for jx, b in enumerate(rk) if a == b and i >= 4 and jx <= 1)
    assert got == want and len(got) == 3 and "fused" in j.last_path()


def test_join_fused_equals_unfused_on_random_data(ctx):
    from ballista_b200 import col, lit
    rng = np.random.default_rng(3)
    n_l, n_r = 50_000, 300_001
    left = rb(k=rng.integers(0, 40_000, n_l).astype(np.int32), a=rng.random(n_l))
    right = rb(k2=rng.integers(0, 60_000, n_r).astype(np.int32), b=rng.integers(0, 100, n_r).astype(np.int32),
               s=pa.array([["x", "yy", "BUILDING"][i % 3] for i in range(n_r)]))
    parts = [[right.slice(0, 100_000), right.slice(100_000, 7)], [right.slice(100_007)]]
    args = dict(lpred=col("a") < lit(0.5), rpred=(col("b") >= lit(10)) & (col("b") < lit(60)))
    fused, j = join_rows(ctx, [[left]], parts, [("k", "k2")], **args)
    assert "fused" in j.last_path()
    os.environ["B2_DISABLE_FUSION"] = "1"
    try:
        plain, j2 = join_rows(ctx, [[left]], parts, [("k", "k2")], **args)
    finally:
        del os.environ["B2_DISABLE_FUSION"]
    assert "fused" not in j2.last_path()
    assert fused == plain and len(fused) > 10_000
    # a string-equality predicate is evaluated as a byte mask and still fused
    m1, j3 = join_rows(ctx, [[left]], parts, [("k", "k2")], rpred=col("s") == lit("BUILDING"))
    lk, la = left["k"].to_numpy(), left["a"].to_numpy()
    rkk = right["k2"].to_numpy()
    from collections import defaultdict
    idx = defaultdict(list)
    for i, k in enumerate(lk):
        idx[int(k)].append(i)
    want = sum(len(idx.get(int(k), ())) for i, k in enumerate(rkk) if i % 3 == 2)
    assert len(m1) == want


def test_repartition_fused_filter_empty_and_all_rejected(ctx):
    from ballista_b200 import CoalesceBatchesExec, FilterExec, RepartitionExec, col, lit, tpch
    n = 100_003
    rng = np.random.default_rng(9)
    b = rb(k=rng.integers(-1000, 1000, n).astype(np.int32), v=rng.random(n), s=pa.array([str(i % 11) for i in range(n)]))
    def run(pred, batches):
        src = tpch.scan(ctx, batches)
        if pred is not None:
            src = CoalesceBatchesExec(FilterExec(pred, src))
        rep = RepartitionExec(src, "Hash", 5, [col("k")])
        rows, owner = [], {}
        for p in range(5):
            t = rep.collect(p)
            for r in zip(t["k"].to_pylist(), t["v"].to_pylist(), t["s"].to_pylist()):
                assert owner.setdefault(r[0], p) == p
                rows.append(r)
        return sorted(rows), rep
    allrows = sorted(zip(b["k"].to_pylist(), b["v"].to_pylist(), b["s"].to_pylist()))
    got, rep = run(None, [[b.slice(0, 40_000)], [b.slice(40_000, 0), b.slice(40_000)]])
    assert got == allrows
    got, rep = run(col("v") < lit(0.25), [[b]])
    assert got == [r for r in allrows if r[1] < 0.25] and "fused-filter" in rep.last_path()
    got, _ = run(col("v") < lit(-1.0), [[b]])
    assert got == []
    got, _ = run(None, [[b.slice(0, 0)]])
    assert got == []


def test_sort_empty_single_row_and_presorted(ctx):
    from ballista_b200 import SortExec, col, tpch
    b = rb(x=np.array([3, 1, 2], dtype=np.int32), y=np.array([0.5, -0.5, 0.0]))
    assert SortExec([col("x")], tpch.scan(ctx, [[b.slice(0, 0)]])).collect().num_rows == 0
    assert SortExec([col("x")], tpch.scan(ctx, [[b.slice(1, 1)]])).collect()["x"].to_pylist() == [1]
    assert SortExec([(col("y"), False)], tpch.scan(ctx, [[b]])).collect()["x"].to_pylist() == [3, 2, 1]
    big = rb(x=np.arange(300_000, dtype=np.int32))
    assert SortExec([(col("x"), False)], tpch.scan(ctx, [[big]])).collect()["x"].to_pylist() == list(range(299_999, -1, -1))


@pytest.mark.parametrize("dups", [False, True])
def test_join_long_clustered_probe_side_against_numpy(ctx, dups):
    """9 M probe rows clustered on the key in runs of 1..7 (the lineitem shape): more warp tiles than the single-CTA
    scan takes (chunked scan), runs that inherit their head's lookup, Bloom false positives that the table rejects, and
    with `dups` a build side whose keys repeat (chain walk, count + write passes).  Checked against a numpy join:
    same pairs, probe order kept."""
    from ballista_b200 import HashJoinExec, tpch
    rng = np.random.default_rng(11 + dups)
    n_keys = 2_300_000
    runs = rng.integers(1, 8, n_keys)
    pk = np.repeat(np.arange(n_keys, dtype=np.int32) * 3 + 1, runs)[:9_000_000]
    bk = rng.choice(np.arange(n_keys, dtype=np.int32) * 3 + 1, 40_000, replace=False)
    if dups:
        bk = np.concatenate([bk, bk[:5000], bk[:50]])
    bk = np.concatenate([bk, np.array([2, 5, 8], dtype=np.int32)])  # keys the probe side never has
    left = rb(k=bk, lrow=np.arange(len(bk), dtype=np.int64))
    right = rb(k2=pk, rrow=np.arange(len(pk), dtype=np.int64))
    j = HashJoinExec(tpch.scan(ctx, [[left]]), tpch.scan(ctx, [[right]]), [("k", "k2")])
    t = j.collect(0)
    lrow, rrow = t["lrow"].to_numpy(), t["rrow"].to_numpy()
    # numpy statement of the same join: for every probe row in order, every build row with that key
    order = np.argsort(bk, kind="stable")
    sk = bk[order]
    lo, hi = np.searchsorted(sk, pk, "left"), np.searchsorted(sk, pk, "right")
    cnt = hi - lo
    assert len(rrow) == int(cnt.sum())
    want_r = np.repeat(np.arange(len(pk), dtype=np.int64), cnt)
    assert np.array_equal(rrow, want_r)  # probe order, every row as often as its key occurs on the build side
    assert np.array_equal(bk[lrow], pk[rrow])
    # each (probe row, build row) pair exactly once
    starts = np.repeat(lo, cnt)
    within = np.arange(len(want_r)) - np.repeat(np.cumsum(cnt) - cnt, cnt)
    want_l = np.sort((order[starts + within]).reshape(-1))
    assert np.array_equal(np.sort(lrow), want_l)
    pair = rrow * (len(bk) + 1) + lrow
    assert len(np.unique(pair)) == len(pair)


def test_join_two_column_and_string_keys_general_path(ctx):
    """The packed-key join (any key shape that is not one Int32 column): Q5's two-column key
    (l_suppkey = s_suppkey AND c_nationkey = s_nationkey, q5.sql:11-12), an Int64 key and a Utf8 key, with duplicate
    build keys; pairs and probe order against a dictionary join."""
    from ballista_b200 import HashJoinExec, tpch
    rng = np.random.default_rng(23)
    nb, npr = 3_000, 40_000
    b_supp = rng.integers(0, 500, nb).astype(np.int32)
    b_nat = rng.integers(0, 25, nb).astype(np.int32)
    p_supp = rng.integers(0, 600, npr).astype(np.int32)
    p_nat = rng.integers(0, 25, npr).astype(np.int32)

    def check(left, right, on, lkey, rkey, expect="key128"):
        j = HashJoinExec(tpch.scan(ctx, [[left]]), tpch.scan(ctx, [[right.slice(0, 17_001), right.slice(17_001)]]), on)
        assert expect in (j.collect_device(0) and j.last_path()), j.last_path()
        t = j.collect(0)
        build = {}
        for i, k in enumerate(lkey):
            build.setdefault(k, []).append(i)
        want = [(i, r) for r, k in enumerate(rkey) for i in build.get(k, [])]
        got = list(zip(t["lrow"].to_pylist(), t["rrow"].to_pylist()))
        assert [r for _, r in got] == [r for _, r in want]          # probe order
        assert sorted(got) == sorted(want)                            # every pair exactly once

    left = rb(s_suppkey=b_supp, s_nationkey=b_nat, lrow=np.arange(nb, dtype=np.int64))
    right = rb(l_suppkey=p_supp, c_nationkey=p_nat, rrow=np.arange(npr, dtype=np.int64))
    # all-Int32 key columns: 32-bit table on the first pair, the others verified on the candidate pairs ...
    check(left, right, [("s_suppkey", "l_suppkey"), ("s_nationkey", "c_nationkey")],
          list(zip(b_supp.tolist(), b_nat.tolist())), list(zip(p_supp.tolist(), p_nat.tolist())), expect="+verify")
    # ... also with the low-cardinality column first (long chains on the table's key)
    check(left, right, [("s_nationkey", "c_nationkey"), ("s_suppkey", "l_suppkey")],
          list(zip(b_nat.tolist(), b_supp.tolist())), list(zip(p_nat.tolist(), p_supp.tolist())), expect="+verify")
    # ... and the same join through the packed-key table
    import os
    os.environ["B2_DISABLE_K32_VERIFY"] = "1"
    try:
        check(left, right, [("s_suppkey", "l_suppkey"), ("s_nationkey", "c_nationkey")],
              list(zip(b_supp.tolist(), b_nat.tolist())), list(zip(p_supp.tolist(), p_nat.tolist())))
    finally:
        del os.environ["B2_DISABLE_K32_VERIFY"]
    # Int64 key with values beyond 32 bits
    b64 = (b_supp.astype(np.int64) << 33) + b_nat
    p64 = (p_supp.astype(np.int64) << 33) + p_nat
    check(rb(k=b64, lrow=np.arange(nb, dtype=np.int64)), rb(k2=p64, rrow=np.arange(npr, dtype=np.int64)), [("k", "k2")],
          b64.tolist(), p64.tolist())
    # Utf8 key
    bs = [f"n{a}-{b}" for a, b in zip(b_supp.tolist(), b_nat.tolist())]
    ps = [f"n{a}-{b}" for a, b in zip(p_supp.tolist(), p_nat.tolist())]
    check(rb(k=bs, lrow=np.arange(nb, dtype=np.int64)), rb(k2=ps, rrow=np.arange(npr, dtype=np.int64)), [("k", "k2")], bs, ps)


def test_final_over_one_unique_partial_output_is_the_identity(ctx):
    """A Final aggregate whose whole input is one Partial output (every key once) returns the states unchanged; two
    Partial outputs with overlapping keys, or an AVG, take the merging path.  Same rows either way."""
    from ballista_b200 import HashAggregateExec, MemoryExec, MergeExec, avg, col, sum as sum_, tpch
    rng = np.random.default_rng(3)
    n = 50_000
    b = rb(k=rng.integers(0, 20_000, n).astype(np.int32), d=rng.integers(0, 3, n).astype(np.int32),
           v=rng.random(n))
    groups = [col("k"), col("d")]

    def run(aggs, parts):
        states, schema = [], None
        for p in parts:
            partial = HashAggregateExec("Partial", groups, aggs, tpch.scan(ctx, [[p]]))
            states += partial.collect_device(0)
            schema = partial.schema()
        final = HashAggregateExec("Final", groups, aggs, MergeExec(MemoryExec(ctx, schema, [states])),
                                  input_schema=b.schema)
        t = final.collect()
        rows = sorted(zip(*[t[c].to_pylist() for c in t.schema.names]))
        return rows, final.last_path()

    want = {}
    for k, d, v in zip(b["k"].to_pylist(), b["d"].to_pylist(), b["v"].to_pylist()):
        want.setdefault((k, d), []).append(v)
    one, path1 = run([sum_(col("v")).alias("s")], [b])
    assert path1.startswith("identity"), path1
    two, path2 = run([sum_(col("v")).alias("s")], [b.slice(0, n // 2), b.slice(n // 2)])
    assert not path2.startswith("identity"), path2
    withavg, path3 = run([sum_(col("v")).alias("s"), avg(col("v")).alias("a")], [b])
    assert not path3.startswith("identity"), path3
    for rows in (one, two):
        assert [r[:2] for r in rows] == sorted(want)
        for r in rows:
            assert abs(r[2] - sum(want[r[:2]])) <= 1e-9 * abs(sum(want[r[:2]]))
    assert [r[:3] for r in withavg] == [tuple(r) for r in one] or all(
        abs(x[2] - y[2]) <= 1e-12 * abs(y[2]) for x, y in zip(withavg, one))
    for r in withavg:
        vs = want[r[:2]]
        assert abs(r[3] - sum(vs) / len(vs)) <= 1e-9 * abs(sum(vs) / len(vs))
