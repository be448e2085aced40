"""GPU: TPC-H Q5 (rust/benchmarks/tpch/queries/q5.sql) - six tables, five left-deep inner joins, one of them on the
two-column key (l_suppkey, c_nationkey) = (s_suppkey, s_nationkey), r_name = 'ASIA', group by n_name - through the C ABI
against the oracle's restatement (itself checked against a join-free numpy evaluation in test_oracle_q5.py)."""
import os

import pytest

pytestmark = pytest.mark.gpu


def run_q5(ctx, sf, parts=1, sort=False):
    from ballista_b200 import MemoryExec, tpch
    scans = []
    for table, cols in tpch.Q5_TABLES:
        if table == "lineitem" and parts > 1:
            n = tpch.K.lib().b2_tpch_num_rows(b"lineitem", float(sf))
            bounds = [(n * p // parts) // 32 * 32 for p in range(parts)] + [n]
            scans.append(tpch.scan(ctx, [[ctx.tpch(table, sf, bounds[p], bounds[p + 1], columns=cols)] for p in range(parts)]))
        else:
            scans.append(tpch.scan(ctx, [[ctx.tpch(table, sf, columns=cols)]]))
    partial = tpch.q5_partial(*scans)
    # the probe side of the last three joins (supplier, nation, region) has one partition: so has the plan
    states = [s for p in range(partial.output_partitioning()[1]) for s in partial.collect_device(p)]
    final = tpch.q5_final(MemoryExec(ctx, partial.schema(), [states]), partial.children()[0].schema())
    plan = tpch.q5_sorted(final) if sort else final
    return plan.collect().to_pylist(), partial


@pytest.mark.parametrize("sf,parts", [(0.01, 1), (0.1, 1), (0.3, 3)])
def test_q5_matches_oracle(ctx, O, sf, parts):
    got, partial = run_q5(ctx, sf, parts)
    want = O.q5(sf, threads=4)
    assert len(want) == 5  # the five nations of ASIA
    assert sorted(r["n_name"] for r in got) == sorted(want)  # bit-exact group keys
    for r in got:
        assert abs(r["revenue"] - want[r["n_name"]]) <= 1e-9 * abs(want[r["n_name"]]), r


def test_q5_join_side_selection_and_order_by(ctx, O):
    """the joins against nation / region build on the small right side; the result equals the collect-left plan's"""
    got, partial = run_q5(ctx, 0.1, sort=True)
    revs = [r["revenue"] for r in got]
    assert revs == sorted(revs, reverse=True)  # ORDER BY revenue DESC
    os.environ["B2_DISABLE_JOIN_SWAP"] = "1"
    try:
        plain, _ = run_q5(ctx, 0.1, sort=True)
    finally:
        del os.environ["B2_DISABLE_JOIN_SWAP"]
    assert [r["n_name"] for r in plain] == [r["n_name"] for r in got]
    for a, b in zip(plain, got):
        assert abs(a["revenue"] - b["revenue"]) <= 1e-9 * abs(a["revenue"])


def test_generated_supplier_nation_region_match_the_oracle(ctx, O):
    import numpy as np
    sf = 0.5
    s = ctx.tpch("supplier", sf).to_arrow()
    w = O.gen_supplier(sf, 0, O.tpch_num_rows("supplier", sf))
    assert np.array_equal(s["s_suppkey"].to_numpy(), w["s_suppkey"]) and np.array_equal(s["s_nationkey"].to_numpy(), w["s_nationkey"])
    n, wn = ctx.tpch("nation", sf).to_arrow(), O.gen_nation()
    assert n["n_name"].to_pylist() == wn["n_name"].to_list() and n["n_regionkey"].to_pylist() == wn["n_regionkey"].tolist()
    r, wr = ctx.tpch("region", sf).to_arrow(), O.gen_region()
    assert r["r_name"].to_pylist() == wr["r_name"].to_list() and r["r_regionkey"].to_pylist() == [0, 1, 2, 3, 4]
    # every l_suppkey has a supplier
    li = ctx.tpch("lineitem", sf, 0, 100_000, columns=["l_suppkey"]).to_arrow()
    assert 1 <= min(li["l_suppkey"].to_pylist()) and max(li["l_suppkey"].to_pylist()) <= len(w["s_suppkey"])
