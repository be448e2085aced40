"""GPU: NULL semantics, Left / Right outer joins (from_proto.rs:268-272), and the wider expression
vocabulary (CASE, LIKE, Utf8 ordering, IS [NOT] NULL, scalar functions, integer SUM; ballista.proto:14-45,80-114)
through the C ABI, against the oracle's restatements and pyarrow (an independent engine)."""
import math

import numpy as np
import pyarrow as pa
import pyarrow.compute as pc
import pytest

pytestmark = pytest.mark.gpu


def rb(**cols):
    return pa.RecordBatch.from_arrays([v if isinstance(v, pa.Array) else pa.array(v) for v in cols.values()],
                                      names=list(cols))


def scan(ctx, parts):
    from ballista_b200 import tpch
    return tpch.scan(ctx, parts)


# ---------------------------------------------------------------------------------------------------------
# the silent-wrong-value hole of round 1: ungrouped aggregate over no rows = (SUM NULL, COUNT 0), not 0.0
# ---------------------------------------------------------------------------------------------------------
def test_ungrouped_aggregate_over_rejected_rows_is_null(ctx, O):
    from ballista_b200 import (CoalesceBatchesExec, FilterExec, HashAggregateExec, MemoryExec, avg, col, count, lit, sum,
                               tpch)
    n = 50_000
    li = ctx.tpch("lineitem", 0.1, 0, n, columns=tpch.Q6_COLUMNS)
    aggs = [sum(col("l_extendedprice") * col("l_discount")).alias("revenue"), count().alias("n"),
            avg(col("l_quantity")).alias("aq")]
    for fused in (True, False):
        src = scan(ctx, [[li]])
        # quantity is in [1, 50]: nothing passes
        filt = CoalesceBatchesExec(FilterExec(col("l_quantity") < lit(0.5), src))
        part = HashAggregateExec("Partial", [], aggs, filt)
        if not fused:
            import os
            os.environ["B2_DISABLE_FUSION"] = "1"
        try:
            states = part.collect_device(0)
        finally:
            if not fused:
                del os.environ["B2_DISABLE_FUSION"]
        st = states[0].to_arrow().to_pylist()
        assert st == [{"revenue[sum]": None, "n[count]": 0, "aq[count]": 0, "aq[sum]": None}], st
        fin = HashAggregateExec("Final", [], aggs, MemoryExec(ctx, part.schema(), [states]), input_schema=src.schema())
        assert fin.collect().to_pylist() == [{"revenue": None, "n": 0, "aq": None}]
    # the oracle agrees: Q6 over rows none of which pass is NULL
    cols = O.gen_lineitem(0.1, 0, 1000)
    assert O.q6(cols, quantity_lt=0.5) == (None, 0)


def test_final_merges_an_empty_partition_with_a_non_empty_one(ctx, O):
    from ballista_b200 import MemoryExec, tpch
    n = 40_000
    li = ctx.tpch("lineitem", 0.1, 0, n, columns=tpch.Q6_COLUMNS)
    empty = ctx.tpch("lineitem", 0.1, 0, 0, columns=tpch.Q6_COLUMNS)
    rev, passing = O.q6(O.gen_lineitem(0.1, 0, n))
    assert passing > 0
    for order in ([[empty], [li]], [[li], [empty]]):
        src = scan(ctx, order)
        part = tpch.q6_partial(src)
        states = [s for p in range(2) for s in part.collect_device(p)]
        assert sorted(s.to_arrow().to_pylist()[0]["revenue[sum]"] is None for s in states) == [False, True]
        fin = tpch.q6_final(MemoryExec(ctx, part.schema(), [states]), src.schema())
        got = fin.collect().to_pylist()
        assert len(got) == 1 and abs(got[0]["revenue"] - rev) <= 1e-9 * abs(rev)
    # both partitions empty: NULL
    src = scan(ctx, [[empty], [empty]])
    part = tpch.q6_partial(src)
    states = [s for p in range(2) for s in part.collect_device(p)]
    fin = tpch.q6_final(MemoryExec(ctx, part.schema(), [states]), src.schema())
    assert fin.collect().to_pylist() == [{"revenue": None}]


# ---------------------------------------------------------------------------------------------------------
# validity bitmaps across the C Data Interface; null predicate = false; IS [NOT] NULL
# ---------------------------------------------------------------------------------------------------------
def nullable_batch(n=10_007, seed=5):
    rng = np.random.default_rng(seed)
    k = rng.integers(0, 7, n).astype(np.int32)
    x = rng.random(n) * 100
    y = rng.integers(-50, 50, n).astype(np.int64)
    s = np.array(["ab", "abc", "b", "", "zz%", "a_c"])[rng.integers(0, 6, n)]
    xm, ym, sm, km = (rng.random(n) < p for p in (0.2, 0.3, 0.1, 0.05))
    return rb(k=pa.array(k, mask=km), x=pa.array(x, mask=xm), y=pa.array(y, mask=ym), s=pa.array(s, mask=sm),
              i=pa.array(np.arange(n, dtype=np.int32)))


def test_validity_round_trip_and_sliced_import(ctx):
    b = nullable_batch()
    assert ctx.upload(b).to_arrow().equals(b)
    sl = b.slice(13, 4001)  # bitmap offset not a multiple of 8
    assert ctx.upload(sl).to_arrow().equals(sl)
    assert ctx.upload(b).to_arrow().column("x").null_count == b.column("x").null_count


def test_null_predicate_is_false_and_is_null_filters(ctx):
    from ballista_b200 import FilterExec, col, lit
    b = nullable_batch()
    t = pa.Table.from_batches([b])
    cases = [
        (col("x") < lit(50.0), pc.less(t["x"], 50.0)),
        ((col("x") < lit(50.0)) & (col("y") > lit(0)), pc.and_(pc.less(t["x"], 50.0), pc.greater(t["y"], 0))),
        # arrow 4.0's `or` is not Kleene: NULL OR true = NULL -> the row is dropped
        ((col("x") < lit(50.0)) | (col("y") > lit(0)), pc.or_(pc.less(t["x"], 50.0), pc.greater(t["y"], 0))),
        (col("x").is_null(), pc.is_null(t["x"])),
        (col("s").is_not_null() & col("y").is_null(), pc.and_(pc.is_valid(t["s"]), pc.is_null(t["y"]))),
        (~(col("y") >= lit(10)), pc.invert(pc.greater_equal(t["y"], 10))),
    ]
    for pred, want_mask in cases:
        got = FilterExec(pred, scan(ctx, [[b.slice(0, 5000)], [b.slice(5000)]])).collect()
        want = t.filter(want_mask)  # pyarrow drops NULL mask entries too
        assert got.equals(want), pred.display()


# ---------------------------------------------------------------------------------------------------------
# SUM / AVG / COUNT skip NULLs (A.2); SUM of nothing is NULL; integer SUM -> Int64
# ---------------------------------------------------------------------------------------------------------
def test_grouped_aggregates_skip_nulls(ctx, O):
    from ballista_b200 import HashAggregateExec, MemoryExec, avg, col, count, sum
    b = nullable_batch(30_011, seed=11)
    # group keys must not be NULL (refused loudly): group by the dense row id modulo 9, computed on the host
    g = pa.array((np.arange(b.num_rows) % 9).astype(np.int32))
    b = b.set_column(0, "k", g)
    # group 8: every x is NULL -> SUM(x) NULL, AVG(x) NULL, COUNT(x) 0
    xs = b.column("x").to_pylist()
    xs = [None if i % 9 == 8 else v for i, v in enumerate(xs)]
    b = b.set_column(1, "x", pa.array(xs, type=pa.float64()))
    aggs = [sum(col("x")).alias("sx"), avg(col("x")).alias("ax"), count(col("x")).alias("cx"), count().alias("n"),
            sum(col("y")).alias("sy"), count(col("s")).alias("cs")]
    parts = [[b.slice(0, 9000)], [b.slice(9000, 4), b.slice(9004)]]
    src = scan(ctx, parts)
    partial = HashAggregateExec("Partial", [col("k")], aggs, src)
    states = [s for p in range(2) for s in partial.collect_device(p)]
    final = HashAggregateExec("Final", [col("k")], aggs, MemoryExec(ctx, partial.schema(), [states]),
                              input_schema=src.schema())
    got = {r["k"]: r for r in final.collect().to_pylist()}
    assert final.schema().field("sy").type == pa.int64()
    ks, ys, ss = b.column("k").to_pylist(), b.column("y").to_pylist(), b.column("s").to_pylist()
    want_x = O.aggregate_skip_nulls(ks, [0.0 if v is None else v for v in xs], [v is not None for v in xs])
    want_y = O.aggregate_skip_nulls(ks, [0 if v is None else v for v in ys], [v is not None for v in ys])
    assert sorted(got) == sorted(want_x)
    for k, (rows, cnt, tot) in want_x.items():
        r = got[k]
        assert r["n"] == rows and r["cx"] == cnt
        assert r["cs"] == builtin_sum(1 for kk, v in zip(ks, ss) if kk == k and v is not None)
        if cnt == 0:
            assert r["sx"] is None and r["ax"] is None
        else:
            assert math.isclose(r["sx"], tot, rel_tol=1e-9) and math.isclose(r["ax"], tot / cnt, rel_tol=1e-9)
        assert r["sy"] == want_y[k][2]  # integer sums are exact
    # independent engine
    t = pa.Table.from_batches([b]).group_by("k").aggregate([("x", "sum"), ("x", "count"), ("y", "sum")])
    for r in t.to_pylist():
        assert got[r["k"]]["cx"] == r["x_count"] and got[r["k"]]["sy"] == r["y_sum"]
        assert (got[r["k"]]["sx"] is None) == (r["x_sum"] is None)


def builtin_sum(it):
    import builtins
    return builtins.sum(it)


def test_null_group_keys_are_refused(ctx):
    from ballista_b200 import HashAggregateExec, NotImplementedOnGpu, col, sum
    b = nullable_batch(1000)
    with pytest.raises(NotImplementedOnGpu):
        HashAggregateExec("Partial", [col("k")], [sum(col("x")).alias("s")], scan(ctx, [[b]])).collect_device(0)


# ---------------------------------------------------------------------------------------------------------
# joins: null keys never match (A.3); Left / Right joins emit null-extended rows
# ---------------------------------------------------------------------------------------------------------
def join_table(ctx, left, right, on, how, lparts=None, rparts=None):
    from ballista_b200 import HashJoinExec
    j = HashJoinExec(scan(ctx, lparts or [[left]]), scan(ctx, rparts or [[right]]), on, how)
    return j.collect(), j


def canon_rows(t):
    key = lambda r: tuple((v is None, v) for v in r)
    return sorted((tuple(r[c] for c in t.schema.names) for r in t.to_pylist()), key=key)


@pytest.mark.parametrize("how", ["Inner", "Left", "Right"])
@pytest.mark.parametrize("keytype", ["int32", "int64", "utf8"])
def test_outer_joins_and_null_keys_match_oracle_and_acero(ctx, O, how, keytype):
    rng = np.random.default_rng(21)
    nl, nr = 3000, 7001
    lk = rng.integers(0, 2500, nl)
    rk = rng.integers(1500, 4000, nr)
    conv = {"int32": lambda a: a.astype(np.int32), "int64": lambda a: a.astype(np.int64) * 3_000_000_000,
            "utf8": lambda a: np.array([f"k{v:05d}" for v in a])}[keytype]
    lm, rm = rng.random(nl) < 0.05, rng.random(nr) < 0.05
    left = rb(k=pa.array(conv(lk), mask=lm), lv=pa.array(rng.random(nl), mask=rng.random(nl) < 0.1),
              ls=pa.array([f"L{i % 13}" for i in range(nl)]))
    right = rb(k2=pa.array(conv(rk), mask=rm), rv=np.arange(nr, dtype=np.int64))
    got, j = join_table(ctx, left, right, [("k", "k2")], how, rparts=[[right.slice(0, 3000)], [right.slice(3000)]]
                        if how != "Left" else None)
    assert ("k32" in j.last_path()) == (keytype == "int32")
    # oracle restatement: pairs + null-extended rows
    okl = [O.Utf8.from_list(list(conv(lk)))] if keytype == "utf8" else [conv(lk)]
    okr = [O.Utf8.from_list(list(conv(rk)))] if keytype == "utf8" else [conv(rk)]
    li, ri = O.hash_join_outer(okl, okr, how, ~lm, ~rm)
    lrows, rrows = left.to_pylist(), right.to_pylist()
    nul_l, nul_r = {c: None for c in left.schema.names}, {c: None for c in right.schema.names}
    want = []
    for a, b in zip(li, ri):
        row = dict(lrows[a] if a >= 0 else nul_l)
        row.update(rrows[b] if b >= 0 else nul_r)
        want.append(tuple(row[c] for c in got.schema.names))
    key = lambda r: tuple((v is None, v) for v in r)
    assert canon_rows(got) == sorted(want, key=key)
    # independent engine (pyarrow Acero); it coalesces nothing: keep both key columns
    jt = {"Inner": "inner", "Left": "left outer", "Right": "right outer"}[how]
    ac = pa.Table.from_batches([left]).join(pa.Table.from_batches([right]), keys="k", right_keys="k2", join_type=jt,
                                            coalesce_keys=False)
    assert canon_rows(ac.select(got.schema.names)) == canon_rows(got)
    assert got.num_rows > 1000


@pytest.mark.parametrize("how", ["Inner", "Left", "Right"])
def test_two_int32_key_columns_with_nulls_match_oracle(ctx, O, how):
    """Two Int32 key columns (32-bit table on the first, the second verified on the candidate pairs): NULLs in either
    column never match, outer joins null-extend the rows left without a VERIFIED partner."""
    rng = np.random.default_rng(33)
    nl, nr = 2500, 6001
    la, lb = rng.integers(0, 300, nl).astype(np.int32), rng.integers(0, 6, nl).astype(np.int32)
    ra, rb_ = rng.integers(100, 400, nr).astype(np.int32), rng.integers(0, 6, nr).astype(np.int32)
    lma, lmb = rng.random(nl) < 0.04, rng.random(nl) < 0.04
    rma, rmb = rng.random(nr) < 0.04, rng.random(nr) < 0.04
    left = rb(a=pa.array(la, mask=lma), b=pa.array(lb, mask=lmb), lv=np.arange(nl, dtype=np.int64))
    right = rb(a2=pa.array(ra, mask=rma), b2=pa.array(rb_, mask=rmb), rv=np.arange(nr, dtype=np.int64))
    got, j = join_table(ctx, left, right, [("a", "a2"), ("b", "b2")], how)
    assert "k32" in j.last_path() and "+verify" in j.last_path(), j.last_path()
    li, ri = O.hash_join_outer([la, lb], [ra, rb_], how, ~(lma | lmb), ~(rma | rmb))
    lrows, rrows = left.to_pylist(), right.to_pylist()
    nul_l, nul_r = {c: None for c in left.schema.names}, {c: None for c in right.schema.names}
    want = []
    for a, b in zip(li, ri):
        row = dict(lrows[a] if a >= 0 else nul_l)
        row.update(rrows[b] if b >= 0 else nul_r)
        want.append(tuple(row[c] for c in got.schema.names))
    key = lambda r: tuple((v is None, v) for v in r)
    assert canon_rows(got) == sorted(want, key=key)
    assert got.num_rows > 1000


def test_right_join_drops_the_left_duplicate_key_column(ctx):
    left = rb(k=np.array([1, 2, 3], dtype=np.int32), a=np.array([1.0, 2.0, 3.0]))
    right = rb(k=np.array([2, 3, 4, 4], dtype=np.int32), b=np.array([20, 30, 40, 41], dtype=np.int64))
    got, _ = join_table(ctx, left, right, [("k", "k")], "Right")
    assert got.schema.names == ["a", "k", "b"]  # build_join_schema: Right keeps the right key
    assert canon_rows(got) == canon_rows(pa.table({"a": [2.0, 3.0, None, None], "k": [2, 3, 4, 4], "b": [20, 30, 40, 41]}))
    got, _ = join_table(ctx, left, right, [("k", "k")], "Left")
    assert got.schema.names == ["k", "a", "b"]
    assert canon_rows(got) == canon_rows(pa.table({"k": [1, 2, 3], "a": [1.0, 2.0, 3.0], "b": [None, 20, 30]}))


def test_left_join_feeds_a_null_skipping_aggregate(ctx):
    """customers LEFT JOIN orders, COUNT(o_orderkey) per customer: the q13 shape (rows without orders count 0)."""
    from ballista_b200 import HashAggregateExec, HashJoinExec, col, count, sum
    rng = np.random.default_rng(2)
    cust = rb(c=np.arange(500, dtype=np.int32))
    ok = rng.integers(0, 400, 3000).astype(np.int32)  # customers 400..499 have no orders
    orders = rb(oc=ok, amount=rng.random(3000))
    j = HashJoinExec(scan(ctx, [[cust]]), scan(ctx, [[orders]]), [("c", "oc")], "Left")
    agg = HashAggregateExec("Partial", [col("c")], [count(col("oc")).alias("n"), sum(col("amount")).alias("s")], j)
    got = {r["c"]: (r["n[count]"], r["s[sum]"]) for b in agg.collect_device(0) for r in b.to_arrow().to_pylist()}
    cnt = np.bincount(ok, minlength=500)
    assert len(got) == 500
    for c in range(500):
        assert got[c][0] == cnt[c]
        assert (got[c][1] is None) == (cnt[c] == 0)


# ---------------------------------------------------------------------------------------------------------
# expression vocabulary: CASE, LIKE, Utf8 ordering, scalar functions, casts, integer division
# ---------------------------------------------------------------------------------------------------------
def project(ctx, batch, exprs):
    from ballista_b200 import ProjectionExec
    return ProjectionExec(exprs, scan(ctx, [[batch]])).collect()


def test_case_expressions(ctx):
    from ballista_b200 import case, col, lit
    b = nullable_batch(5003, seed=3)
    t = pa.Table.from_batches([b])
    x, y, s = (t[c].to_pylist() for c in ("x", "y", "s"))
    got = project(ctx, b, [
        # q12's shape: CASE WHEN s = 'ab' OR s = 'abc' THEN 1 ELSE 0 END (Int64)
        case([((col("s") == lit("ab")) | (col("s") == lit("abc")), lit(1))], otherwise=lit(0)).alias("hi"),
        # no ELSE: NULL when no branch fires; a NULL condition is false
        case([(col("x") < lit(25.0), col("x") * lit(2.0)), (col("x") < lit(50.0), lit(-1.0))]).alias("c2"),
        # CASE base WHEN value ...; Int64 and Float64 branches widen to Float64
        case([(lit(0), lit(100)), (lit(1), col("x"))], otherwise=lit(0.5), base=col("y")).alias("c3"),
    ])
    for i in range(b.num_rows):
        hi = 1 if s[i] in ("ab", "abc") else 0  # NULL = 'ab' is NULL -> false -> ELSE
        assert got["hi"][i].as_py() == hi
        c2 = None if x[i] is None else (x[i] * 2.0 if x[i] < 25.0 else (-1.0 if x[i] < 50.0 else None))
        assert got["c2"][i].as_py() == c2
        c3 = 100.0 if y[i] == 0 else (x[i] if y[i] == 1 else 0.5)
        assert got["c3"][i].as_py() == c3
    assert got.schema.field("hi").type == pa.int64() and got.schema.field("c3").type == pa.float64()


def test_case_producing_strings(ctx):
    """CASE whose THEN / ELSE values are Utf8 (literals and columns): first condition that holds (a NULL condition does
    not), NULL when nothing fires and there is no ELSE, a NULL string in the chosen branch stays NULL."""
    from ballista_b200 import case, col, lit
    b = nullable_batch(4003, seed=9)
    t = pa.Table.from_batches([b])
    x, y, s = (t[c].to_pylist() for c in ("x", "y", "s"))
    got = project(ctx, b, [
        case([(col("x") < lit(25.0), lit("low")), (col("x") < lit(50.0), col("s"))], otherwise=lit("")).alias("a"),
        case([(col("y") == lit(0), lit("zero")), (col("y") == lit(1), lit("one-and-a-bit longer"))]).alias("b"),
        case([(lit(1), col("s"))], otherwise=lit("other"), base=col("y")).alias("c"),
    ])
    for i in range(b.num_rows):
        a = "low" if (x[i] is not None and x[i] < 25.0) else (s[i] if (x[i] is not None and x[i] < 50.0) else "")
        assert got["a"][i].as_py() == a, i
        bb = "zero" if y[i] == 0 else ("one-and-a-bit longer" if y[i] == 1 else None)
        assert got["b"][i].as_py() == bb, i
        assert got["c"][i].as_py() == (s[i] if y[i] == 1 else "other"), i
    assert got.schema.field("a").type == pa.utf8()
    assert project(ctx, b.slice(0, 0), [case([(col("x") < lit(1.0), lit("q"))]).alias("e")]).num_rows == 0


def test_like_utf8_ordering_and_scalar_functions(ctx):
    from ballista_b200 import FilterExec, col, lit, scalar_fn
    words = ["PROMO BRUSHED", "promo", "STANDARD POLISHED", "SMALL PROMO", "", "a%b", "häßlich", "MEDIUM", "x_y", "PRO"]
    n = 5000
    rng = np.random.default_rng(8)
    s = [words[i] for i in rng.integers(0, len(words), n)]
    f = rng.random(n) * 20 - 10
    b = rb(s=pa.array(s, mask=rng.random(n) < 0.1), f=f, i=np.arange(n, dtype=np.int32))
    t = pa.Table.from_batches([b])
    for pat in ["PROMO%", "%PROMO%", "%PROMO", "PRO", "%", "", "_ROMO%", "%_", "h__lich", "%O%O%", "a%%b", "x_y"]:
        got = FilterExec(col("s").like(pat), scan(ctx, [[b]])).collect()
        want = t.filter(pc.match_like(t["s"], pat))
        assert got.equals(want), pat
    for op, fn in (("__lt__", pc.less), ("__le__", pc.less_equal), ("__gt__", pc.greater), ("__ge__", pc.greater_equal)):
        got = FilterExec(getattr(col("s"), op)(lit("PROMO")), scan(ctx, [[b]])).collect()
        assert got.equals(t.filter(fn(t["s"], "PROMO"))), op
    fns = {"sqrt": np.sqrt, "exp": np.exp, "floor": np.floor, "ceil": np.ceil, "trunc": np.trunc, "abs": np.abs,
           "sin": np.sin, "log": np.log, "round": lambda v: np.sign(v) * np.floor(np.abs(v) + 0.5),
           "signum": lambda v: np.where(np.signbit(v), -1.0, 1.0)}
    got = project(ctx, b, [scalar_fn(k, col("f")).alias(k) for k in fns] + [scalar_fn("octetlength", col("s")).alias("len")])
    with np.errstate(invalid="ignore", divide="ignore"):
        for k, fn in fns.items():
            want = fn(f)
            g = got[k].to_numpy()
            ok = np.isclose(g, want, rtol=1e-12, atol=0, equal_nan=True)
            assert ok.all(), (k, g[~ok][:3], want[~ok][:3])
    lens = [None if v is None else len(v.encode()) for v in t["s"].to_pylist()]
    assert got["len"].to_pylist() == lens
    ascii_rows = rb(s=pa.array([w for w in words if w.isascii()]))
    up = project(ctx, ascii_rows, [scalar_fn("upper", col("s")).alias("u"), scalar_fn("lower", col("s")).alias("l")])
    assert up["u"].to_pylist() == [w.upper() for w in ascii_rows["s"].to_pylist()]
    assert up["l"].to_pylist() == [w.lower() for w in ascii_rows["s"].to_pylist()]


def test_trim_functions(ctx):
    """trim / ltrim / rtrim (ScalarFunction TRIM / LTRIM / RTRIM, ballista.proto:101-103).  The reference's DataFusion
    revision lies between Rust's str::trim* (all Unicode white space) and the Postgres forms (U+0020 only) and cannot be
    read here: spaces are trimmed - where both agree - and a row whose trimmed edge is another white-space character, or
    a non-ASCII one, is refused instead of guessed.  NULLs stay NULL."""
    from ballista_b200 import NotImplementedOnGpu, col, scalar_fn
    words = ["", " ", "  a", "a  ", "  a b  ", "x\ty", "no-space", "   ", " q ", "inner  space", " 7 ", "x" * 40 + "  ",
             "  a\u00e9b  "]
    rng = np.random.default_rng(5)
    n = 4001
    s = [words[i] for i in rng.integers(0, len(words), n)]
    b = rb(s=pa.array(s, mask=rng.random(n) < 0.1))
    got = project(ctx, b, [scalar_fn("trim", col("s")).alias("t"), scalar_fn("ltrim", col("s")).alias("l"),
                           scalar_fn("rtrim", col("s")).alias("r")])
    src = pa.Table.from_batches([b])["s"].to_pylist()
    assert got["t"].to_pylist() == [None if v is None else v.strip(" ") for v in src]
    assert got["l"].to_pylist() == [None if v is None else v.lstrip(" ") for v in src]
    assert got["r"].to_pylist() == [None if v is None else v.rstrip(" ") for v in src]
    # an edge that the two readings would treat differently: loud refusal (only on the side that is trimmed)
    for bad in [" \tx", "x\n ", " \u00a0x "]:
        with pytest.raises(NotImplementedOnGpu):
            project(ctx, rb(s=pa.array([bad, "ok"])), [scalar_fn("trim", col("s")).alias("t")])
    assert project(ctx, rb(s=pa.array(["\tx  "])), [scalar_fn("rtrim", col("s")).alias("r")])["r"].to_pylist() == ["\tx"]
    # a NULL row's bytes are not looked at
    nul = rb(s=pa.array(["\tz", " a "], mask=np.array([True, False])))
    assert project(ctx, nul, [scalar_fn("trim", col("s")).alias("t")])["t"].to_pylist() == [None, "a"]
    assert project(ctx, rb(s=pa.array([], pa.utf8())), [scalar_fn("trim", col("s")).alias("t")]).num_rows == 0


def test_concat(ctx):
    """concat(s1, s2, ...) (ScalarFunction CONCAT, ballista.proto:98) over Utf8 columns and literals.  Rows without a NULL
    argument are concatenated (every reading of the pinned DataFusion revision agrees on them); a NULL argument is
    refused, because that revision's rule (NULL result, or the argument skipped) cannot be read here."""
    from ballista_b200 import NotImplementedOnGpu, col, lit, scalar_fn
    rng = np.random.default_rng(6)
    words = np.array(["", "a", "PROMO", "h\u00e4\u00df", "x" * 33])
    n = 3001
    a, c = words[rng.integers(0, 5, n)], words[rng.integers(0, 5, n)]
    b = rb(a=pa.array(a), c=pa.array(c))
    got = project(ctx, b, [scalar_fn("concat", col("a"), lit("-"), col("c")).alias("j"),
                           scalar_fn("concat", col("a")).alias("one")])
    assert got["j"].to_pylist() == [f"{x}-{y}" for x, y in zip(a.tolist(), c.tolist())]
    assert got["one"].to_pylist() == a.tolist()
    withnull = rb(a=pa.array(a, mask=rng.random(n) < 0.01), c=pa.array(c))
    with pytest.raises(NotImplementedOnGpu, match="NULL argument"):
        project(ctx, withnull, [scalar_fn("concat", col("a"), col("c")).alias("j")])
    with pytest.raises(NotImplementedOnGpu):
        project(ctx, b, [scalar_fn("concat", lit("x"), lit("y")).alias("j")])


def test_md5(ctx):
    """md5(Utf8) -> 32 lower-case hex digits (ScalarFunction MD5, ballista.proto:108), NULL for NULL; lengths around the
    55 / 56 / 64-byte padding boundaries and multi-byte characters, against hashlib."""
    import hashlib
    from ballista_b200 import col, scalar_fn
    vals = ["", "a", "abc", "message digest", "h\u00e4\u00dflich"] + ["x" * k for k in (54, 55, 56, 57, 63, 64, 65, 119, 120, 128, 1000)]
    rng = np.random.default_rng(4)
    s = [vals[i] for i in rng.integers(0, len(vals), 3001)]
    b = rb(s=pa.array(s, mask=rng.random(len(s)) < 0.1))
    got = project(ctx, b, [scalar_fn("md5", col("s")).alias("h")])["h"].to_pylist()
    src = pa.Table.from_batches([b])["s"].to_pylist()
    assert got == [None if v is None else hashlib.md5(v.encode()).hexdigest() for v in src]
    assert hashlib.md5(b"").hexdigest() in got and got.count(None) > 100


def test_nullif(ctx):
    """NULLIF(a, b) (ScalarFunction NULLIF, ballista.proto:106; DataFusion nullif_func = eq + arrow's nullif): NULL where
    a = b holds, else a; a NULL on either side leaves a as it is.  Same-typed arguments only."""
    from ballista_b200 import NotImplementedOnGpu, col, lit, scalar_fn
    rng = np.random.default_rng(12)
    n = 3001
    a_i = rng.integers(0, 5, n).astype(np.int32)
    b_i = rng.integers(0, 5, n).astype(np.int32)
    a_f = rng.integers(0, 4, n).astype(np.float64) / 2
    a_l = rng.integers(-2, 3, n).astype(np.int64) * 3_000_000_000
    ma, mb = rng.random(n) < 0.1, rng.random(n) < 0.1
    b = rb(ai=pa.array(a_i, mask=ma), bi=pa.array(b_i, mask=mb), af=pa.array(a_f, mask=ma), al=a_l)
    got = project(ctx, b, [scalar_fn("nullif", col("ai"), col("bi")).alias("ii"),
                           scalar_fn("nullif", col("af"), lit(0.5)).alias("f"),
                           scalar_fn("nullif", col("al"), lit(3_000_000_000)).alias("l")])
    t = pa.Table.from_batches([b])
    ai, bi, af, al = (t[c].to_pylist() for c in ("ai", "bi", "af", "al"))
    assert got["ii"].to_pylist() == [None if (x is not None and y is not None and x == y) else x for x, y in zip(ai, bi)]
    assert got["f"].to_pylist() == [None if x == 0.5 else x for x in af]
    assert got["l"].to_pylist() == [None if x == 3_000_000_000 else x for x in al]
    assert got.schema.field("ii").type == pa.int32() and got.schema.field("l").type == pa.int64()
    with pytest.raises(NotImplementedOnGpu):  # Int32 against an Int64 literal: the reference would coerce, not restated
        project(ctx, b, [scalar_fn("nullif", col("ai"), lit(1)).alias("x")])


def test_integer_division_by_zero_fails_and_casts(ctx):
    from ballista_b200 import BallistaError, col, lit
    from ballista_b200 import _capi as K
    b = rb(a=np.array([10, -7, 0, 2**40], dtype=np.int64), d=np.array([3, 2, 5, -1], dtype=np.int64),
           f=np.array([1.9, -1.9, 3e10, float("nan")]))
    got = project(ctx, b, [(col("a") / col("d")).alias("q"), col("f").cast(K.INT32).alias("fi"),
                           col("f").cast(K.INT64).alias("fl"), col("a").cast(K.FLOAT64).alias("af")])
    assert got["q"].to_pylist() == [3, -3, 0, -(2**40)]  # truncating division
    assert got["fi"].to_pylist() == [1, -1, 2**31 - 1, 0]  # saturating, NaN -> 0 (Rust `as`)
    assert got["fl"].to_pylist() == [1, -1, 30000000000, 0]
    z = rb(a=np.array([1, 2], dtype=np.int64), d=np.array([1, 0], dtype=np.int64))
    with pytest.raises(BallistaError, match="Divide by zero"):
        project(ctx, z, [(col("a") / col("d")).alias("q")])
    # a NULL divisor row is not a division
    zn = rb(a=np.array([1, 2], dtype=np.int64), d=pa.array([1, 0], mask=np.array([False, True]), type=pa.int64()))
    assert project(ctx, zn, [(col("a") / col("d")).alias("q")])["q"].to_pylist() == [1, None]


# ---------------------------------------------------------------------------------------------------------
# a bare FilterExec: every column, row order, Utf8 compaction vs the oracle's restatement of A.1
# ---------------------------------------------------------------------------------------------------------
def test_bare_filter_exec_matches_oracle(ctx, O):
    from ballista_b200 import CoalesceBatchesExec, FilterExec, col, date, lit
    n = 200_003
    names = ["l_orderkey", "l_quantity", "l_extendedprice", "l_discount", "l_shipdate", "l_returnflag", "l_linestatus"]
    cols = O.gen_lineitem(0.1, 0, n)
    li = ctx.tpch("lineitem", 0.1, 0, n, columns=names)
    pred = (col("l_shipdate") > date("1995-03-15")) & (col("l_discount") >= lit(0.05)) & (col("l_returnflag") == lit("N"))
    mask = (cols["l_shipdate"] > O.DATE_1995_03_15) & (cols["l_discount"] >= 0.05) & cols["l_returnflag"].equals_literal("N")
    want = O.filter_columns([cols[c] for c in names], mask)
    # three ragged partitions, one of them empty
    rbatch = li.to_arrow()
    parts = [[rbatch.slice(0, 70_001)], [rbatch.slice(70_001, 0)], [rbatch.slice(70_001, 50_000), rbatch.slice(120_001)]]
    f = CoalesceBatchesExec(FilterExec(pred, scan(ctx, parts)))
    got = f.collect()
    assert got.num_rows == int(mask.sum()) > 10_000
    for name, w in zip(names, want):
        g = got[name].combine_chunks()
        if isinstance(w, O.Utf8):
            assert g.to_pylist() == w.to_list(), name
        elif "date" in name:
            assert np.array_equal(g.cast(pa.int32()).to_numpy(), w), name
        else:
            assert np.array_equal(g.to_numpy(), w), name  # bit-exact copies, input order


# ---------------------------------------------------------------------------------------------------------
# re-execution after re-feeding a leaf; over-long Utf8 key in a Final aggregate
# ---------------------------------------------------------------------------------------------------------
def test_join_and_repartition_recompute_after_the_leaves_are_refed(ctx):
    from ballista_b200 import HashJoinExec, MemoryExec, RepartitionExec, col
    from ballista_b200 import _capi as K
    l1 = rb(k=np.array([1, 2, 3], dtype=np.int32), a=np.array([1.0, 2.0, 3.0]))
    l2 = rb(k=np.array([3, 4], dtype=np.int32), a=np.array([30.0, 40.0]))
    r = rb(k2=np.array([1, 3, 4, 4], dtype=np.int32), b=np.array([10, 30, 40, 41], dtype=np.int64))
    left = MemoryExec(ctx, l1.schema, [[l1]])
    right = MemoryExec(ctx, r.schema, [[r]])
    j = HashJoinExec(left, RepartitionExec(right, "Hash", 2, [col("k2")]), [("k", "k2")])
    first = sorted(tuple(x.values()) for p in range(2) for x in j.collect(p).to_pylist())
    assert first == [(1, 1.0, 1, 10), (3, 3.0, 3, 30)]
    K.check(K.lib().b2_memory_exec_clear(left._h))
    left.add(0, l2)
    K.check(K.lib().b2_memory_exec_clear(right._h))
    right.add(0, r.slice(1))
    second = sorted(tuple(x.values()) for p in range(2) for x in j.collect(p).to_pylist())
    assert second == [(3, 30.0, 3, 30), (4, 40.0, 4, 40), (4, 40.0, 4, 41)]


def test_final_aggregate_rejects_overlong_utf8_key(ctx):
    from ballista_b200 import HashAggregateExec, NotImplementedOnGpu, col, sum
    states = rb(g=pa.array(["short", "a-twenty-byte-key-xx", "short"]), s=pa.array([1.0, 2.0, 3.0]))
    in_schema = pa.schema([("g", pa.utf8()), ("v", pa.float64())])
    fin = HashAggregateExec("Final", [col("g")], [sum(col("v")).alias("s")],
                            scan(ctx, [[states.rename_columns(["g", "s[sum]"])]]), input_schema=in_schema)
    with pytest.raises(NotImplementedOnGpu, match="packed-key budget"):
        fin.collect()


def test_group_keys_of_eight_to_fourteen_characters(ctx):
    """a character at byte 7 of a packed string key sits at the 64-bit word boundary of the 128-bit key (regression:
    nvcc's __int128 lost it); through the unfused Partial, the small Final and the general Final"""
    import os
    from ballista_b200 import HashAggregateExec, MemoryExec, col, count, sum
    names = ["INDONESIA", "UNITED KINGDOM", "ABCDEFGH", "ABCDEFGI", "X", "", "SAUDI ARABIA", "ABCDEFG"]
    n = 4000
    s = [names[i % len(names)] for i in range(n)]
    b = rb(s=pa.array(s), v=np.arange(n, dtype=np.float64))
    want = {}
    for k, v in zip(s, range(n)):
        c, t = want.get(k, (0, 0.0))
        want[k] = (c + 1, t + v)
    aggs = [sum(col("v")).alias("t"), count().alias("c")]
    for env in ({}, {"B2_DISABLE_SPECULATION": "1"}):
        os.environ.update(env)
        try:
            p = HashAggregateExec("Partial", [col("s")], aggs, scan(ctx, [[b.slice(0, 1500)], [b.slice(1500)]]))
            states = [x for q in range(2) for x in p.collect_device(q)]
            f = HashAggregateExec("Final", [col("s")], aggs, MemoryExec(ctx, p.schema(), [states]), input_schema=b.schema)
            got = {r["s"]: (r["c"], r["t"]) for r in f.collect().to_pylist()}
        finally:
            for k in env:
                del os.environ[k]
        assert got == want, (env, f.last_path())
