"""GPU: the fast scan path is general.  Random predicates (conjunctions / disjunctions of < <= > >= = <> BETWEEN IN over
Int32 / Date32 / Float64 / Utf8 columns), 0-2 group keys and 1-5 SUM / AVG / COUNT arguments over + - * / are compiled at
run time into the TMA-fed bandwidth kernel (agg_tma.cuh) and checked against a numpy evaluation of the same plan
(counts bit-exact, sums <= 1e-9); `last_path` must name the TMA kernel.  compile_expr in the reference accepts any
expression (from_proto.rs:348-364)."""
import datetime
import math

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

F64 = ["l_quantity", "l_extendedprice", "l_discount", "l_tax"]
COLS = ["l_suppkey", "l_linenumber"] + F64 + ["l_returnflag", "l_linestatus", "l_shipdate", "l_commitdate"]


def random_shape(rng):
    """-> (predicate Expr or None, numpy mask fn, key names, [(kind, Expr, numpy fn)])"""
    from ballista_b200 import col, lit, date
    def d(days):
        return (datetime.date(1970, 1, 1) + datetime.timedelta(days=int(days))).isoformat()

    def term():
        k = rng.integers(0, 6)
        if k == 0:
            c, v = rng.choice(F64), None
            v = {"l_quantity": 25.0, "l_extendedprice": 30000.0, "l_discount": 0.05, "l_tax": 0.04}[c]
            op = rng.choice(["<", "<=", ">", ">="])
            return {"<": col(c) < lit(v), "<=": col(c) <= lit(v), ">": col(c) > lit(v), ">=": col(c) >= lit(v)}[op], \
                lambda t, c=c, v=v, op=op: {"<": t[c] < v, "<=": t[c] <= v, ">": t[c] > v, ">=": t[c] >= v}[op]
        if k == 1:
            c, days = rng.choice(["l_shipdate", "l_commitdate"]), int(rng.integers(8500, 10300))
            op = rng.choice(["<", ">="])
            return (col(c) < date(d(days))) if op == "<" else (col(c) >= date(d(days))), \
                lambda t, c=c, days=days, op=op: (t[c] < days) if op == "<" else (t[c] >= days)
        if k == 2:
            lo, hi = 0.02, 0.07
            return col("l_discount").between(lit(lo), lit(hi)), lambda t: (t["l_discount"] >= lo) & (t["l_discount"] <= hi)
        if k == 3:
            v = str(rng.choice(["N", "R", "A", "F", "O", "zz"]))
            c = "l_returnflag" if v in "NRA" else "l_linestatus"
            neg = bool(rng.integers(0, 2))
            return (col(c) != lit(v)) if neg else (col(c) == lit(v)), \
                lambda t, c=c, v=v, neg=neg: (t[c] != v) if neg else (t[c] == v)
        if k == 4:
            vals = [int(x) for x in rng.choice(np.arange(1, 8), 3, replace=False)]
            return col("l_linenumber").in_list([lit(v).cast(2) for v in vals]), lambda t, vals=vals: np.isin(t["l_linenumber"], vals)
        v = int(rng.integers(1, 800))
        return col("l_suppkey") > lit(v), lambda t, v=v: t["l_suppkey"] > v
    nt = int(rng.integers(0, 4))
    pred, fn = None, (lambda t: np.ones(len(t["l_quantity"]), bool))
    for _ in range(nt):
        e, f = term()
        if pred is None:
            pred, fn = e, f
        elif rng.integers(0, 3) == 0:
            pred, fn = pred | e, (lambda t, a=fn, b=f: a(t) | b(t))
        else:
            pred, fn = pred & e, (lambda t, a=fn, b=f: a(t) & b(t))
    keys = [k for k in ["l_returnflag", "l_linestatus", "l_linenumber"] if rng.integers(0, 2)][:2]

    def arg():
        a, b = rng.choice(F64, 2, replace=False)
        k = rng.integers(0, 5)
        if k == 0:
            return col(a), lambda t, a=a: t[a]
        if k == 1:
            return col(a) * (lit(1) - col(b)), lambda t, a=a, b=b: t[a] * (1.0 - t[b])
        if k == 2:
            return col(a) + col(b) * lit(0.5), lambda t, a=a, b=b: t[a] + t[b] * 0.5
        if k == 3:
            return col(a) / col("l_quantity"), lambda t, a=a: t[a] / t["l_quantity"]
        return -col(a) - lit(2.5), lambda t, a=a: -t[a] - 2.5
    aggs = []
    for i in range(int(rng.integers(1, 6))):
        e, f = arg()
        aggs.append((["sum", "avg"][int(rng.integers(0, 2))], e, f))
    return pred, fn, keys, aggs


@pytest.mark.parametrize("seed", range(6))
def test_random_shapes_run_on_the_tma_kernel_and_match_numpy(ctx, seed):
    from ballista_b200 import CoalesceBatchesExec, FilterExec, HashAggregateExec, MemoryExec, avg, col, count, sum, tpch
    rng = np.random.default_rng(1000 + seed)
    n = 300_000 + int(rng.integers(0, 1000))
    li = ctx.tpch("lineitem", 0.1, 0, n, columns=COLS)
    t = li.to_arrow()
    host = {}
    for name in COLS:
        c = t[name]
        if name in ("l_returnflag", "l_linestatus"):
            host[name] = np.array(c.to_pylist())
        elif "date" in name:
            host[name] = np.frombuffer(c.buffers()[1], dtype=np.int32, count=n)
        else:
            host[name] = c.to_numpy()
    pred, mask_fn, keys, aggs = random_shape(rng)
    src = tpch.scan(ctx, [[li]])
    inp = CoalesceBatchesExec(FilterExec(pred, src)) if pred is not None else src
    specs = [(sum if kind == "sum" else avg)(e).alias(f"a{i}") for i, (kind, e, _) in enumerate(aggs)] + [count().alias("n")]
    partial = HashAggregateExec("Partial", [col(k) for k in keys], specs, inp)
    states = partial.collect_device(0)
    final = HashAggregateExec("Final", [col(k) for k in keys], specs, MemoryExec(ctx, partial.schema(), [states]),
                              input_schema=src.schema())
    got = final.collect().to_pylist()
    path = partial.last_path()
    assert path.startswith("fused:") and "[tma" in path, path  # the bandwidth kernel, not the column-at-a-time path
    m = mask_fn(host)
    vals = [f(host)[m] for _, _, f in aggs]
    kcols = [host[k][m] for k in keys]
    want = {}
    for r in range(int(m.sum())):
        k = tuple(c[r].item() if hasattr(c[r], "item") else c[r] for c in kcols)
        w = want.setdefault(k, [0, [0.0] * len(aggs)])
        w[0] += 1
        for a in range(len(aggs)):
            w[1][a] += vals[a][r]
    if not keys and not want:
        assert got == [{**{f"a{i}": None for i in range(len(aggs))}, "n": 0}]
        return
    assert len(got) == len(want), (len(got), len(want), path)
    for r in got:
        k = tuple(r[c] for c in keys)
        cnt, sums = want[k]
        assert r["n"] == cnt
        for a, (kind, _, _) in enumerate(aggs):
            w = sums[a] if kind == "sum" else sums[a] / cnt
            assert math.isclose(r[f"a{a}"], w, rel_tol=1e-9, abs_tol=1e-9 * abs(cnt)), (a, kind, r[f"a{a}"], w, path)
