"""Wire compatibility: serialized PhysicalPlanNode messages (tests/golden/plan_protos.json, encoded by Google's protobuf
runtime from the reference's own ballista.proto by tests/golden/make_plan_protos.py) decoded by the hand-written reader
in csrc/proto.cpp.  The structural checks need no GPU; the execution checks compare with the oracle."""
import base64
import datetime
import json
import os

import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
with open(os.path.join(HERE, "golden", "plan_protos.json")) as f:
    PLANS = {k: v for k, v in json.load(f).items() if not k.startswith("_")}


def plan_bytes(name):
    return base64.b64decode(PLANS[name]["bytes_b64"])


@pytest.mark.parametrize("name", sorted(PLANS))
def test_describe_matches_the_encoded_tree(name):
    from ballista_b200 import describe_plan_proto
    lines, n_leaves = describe_plan_proto(plan_bytes(name))
    assert [l.split(" [")[0] for l in lines] == PLANS[name]["nodes_preorder"]
    assert n_leaves == len(PLANS[name]["leaves"])


def test_malformed_bytes_are_rejected_not_crashed():
    from ballista_b200 import BallistaError, describe_plan_proto
    good = plan_bytes("q1_partial_stage")
    for bad in (b"", good[:17], good[:-3], b"\xff" * 40, bytes([0x7A, 0x05, 1, 2, 3])):
        with pytest.raises(BallistaError):
            describe_plan_proto(bad)
    # a truncated expression payload deep in the tree
    with pytest.raises(BallistaError):
        describe_plan_proto(good[:200] + b"\x00" * 5)


def _leaf_batches(ctx, sf, spec):
    table, cols = spec.split(":")
    return ctx.tpch(table, sf, columns=cols.split(","))


@pytest.mark.gpu
def test_q1_stage_plans_from_proto_match_oracle(ctx, O):
    from ballista_b200 import MemoryExec, plan_from_proto, tpch
    sf, n = 0.05, 200_000
    li = ctx.tpch("lineitem", sf, 0, n, columns=tpch.Q1_COLUMNS)
    scan = tpch.scan(ctx, [[li]])
    partial = plan_from_proto(ctx, plan_bytes("q1_partial_stage"), [scan])
    states = partial.collect_device(0)
    assert partial.last_path().startswith("fused:")  # the decoded plan takes the same fused kernel
    final = plan_from_proto(ctx, plan_bytes("q1_final_stage"), [MemoryExec(ctx, partial.schema(), [states])])
    got = final.collect().to_pylist()
    want = O.q1(O.gen_lineitem(sf, 0, n))
    assert [(r["l_returnflag"], r["l_linestatus"]) for r in got] == [(w["l_returnflag"], w["l_linestatus"]) for w in want]
    for g, w in zip(got, want):  # sorted by the decoded SortExec
        assert g["count_order"] == w["count_order"]
        for k in ("sum_qty", "sum_base_price", "sum_disc_price", "sum_charge", "avg_qty", "avg_price", "avg_disc"):
            assert abs(g[k] - w[k]) <= 1e-9 * abs(w[k]), (k, g[k], w[k])


@pytest.mark.gpu
def test_q6_and_q3_stage_plans_from_proto(ctx, O):
    from ballista_b200 import MemoryExec, plan_from_proto, tpch
    from test_gpu_join_q3 import oracle_q3
    sf = 0.1
    # Q6
    li6 = ctx.tpch("lineitem", sf, columns=tpch.Q6_COLUMNS)
    p6 = plan_from_proto(ctx, plan_bytes("q6_partial_stage"), [tpch.scan(ctx, [[li6]])])
    s6 = p6.collect_device(0)
    got6 = tpch.q6_final(MemoryExec(ctx, p6.schema(), [s6]), tpch.scan(ctx, [[li6]]).schema()).collect().to_pylist()
    rev, _ = O.q6(O.gen_lineitem(sf, 0, li6.num_rows))
    assert abs(got6[0]["revenue"] - rev) <= 1e-9 * abs(rev)
    # Q3: three leaves in pre-order (customer, orders, lineitem), then the final stage with ORDER BY + LIMIT 10
    leaves = [tpch.scan(ctx, [[_leaf_batches(ctx, sf, spec)]]) for spec in PLANS["q3_partial_stage"]["leaves"]]
    partial = plan_from_proto(ctx, plan_bytes("q3_partial_stage"), leaves)
    states = partial.collect_device(0)
    top = plan_from_proto(ctx, plan_bytes("q3_final_stage"), [MemoryExec(ctx, partial.schema(), [states])]).collect().to_pylist()
    no, nl, nc = (O.tpch_num_rows(t, sf) for t in ("orders", "lineitem", "customer"))
    want, _, _ = oracle_q3(O, sf, no, nl, nc)
    want.sort(key=lambda r: (-r["revenue"], r["o_orderdate"]))
    assert len(top) == 10
    for g, w in zip(top, want[:10]):
        assert g["l_orderkey"] == w["l_orderkey"]
        assert (g["o_orderdate"] - datetime.date(1970, 1, 1)).days == w["o_orderdate"]
        assert abs(g["revenue"] - w["revenue"]) <= 1e-9 * abs(w["revenue"])


@pytest.mark.gpu
def test_repartition_stage_plan_and_leaf_checks(ctx):
    from ballista_b200 import BallistaError, plan_from_proto, tpch
    sf = 0.02
    spec = PLANS["lineitem_repartition_stage"]["leaves"][0]
    li = _leaf_batches(ctx, sf, spec)
    rep = plan_from_proto(ctx, plan_bytes("lineitem_repartition_stage"), [tpch.scan(ctx, [[li]])])
    assert rep.output_partitioning() == ("Hash", 4)
    total = sum(rep.collect(p).num_rows for p in range(4))
    passing = tpch.scan(ctx, [[li]]).collect()["l_shipdate"].to_pylist()
    assert total == sum(d > datetime.date(1995, 3, 15) for d in passing)
    with pytest.raises(BallistaError):  # wrong number of leaves
        plan_from_proto(ctx, plan_bytes("lineitem_repartition_stage"), [])
    with pytest.raises(BallistaError):  # a source whose columns are not the ones the plan was built on
        plan_from_proto(ctx, plan_bytes("lineitem_repartition_stage"),
                        [tpch.scan(ctx, [[ctx.tpch("orders", sf, columns=tpch.Q3_ORDERS_COLUMNS)]])])


@pytest.mark.gpu
def test_projection_stage_with_string_functions_from_proto(ctx):
    """A ProjectionExec whose expressions are ScalarFunctionNode / CaseNode messages (trim, md5, concat over lower,
    nullif, a CASE with Utf8 branches) encoded by the protobuf runtime, decoded by proto.cpp and evaluated on the GPU."""
    import hashlib
    from ballista_b200 import plan_from_proto, tpch
    spec = PLANS["customer_projection_stage"]["leaves"][0]
    cu = _leaf_batches(ctx, 0.01, spec)
    plan = plan_from_proto(ctx, plan_bytes("customer_projection_stage"), [tpch.scan(ctx, [[cu]])])
    assert plan.schema().names == ["c_custkey", "seg", "h", "tag", "nk", "cls"]
    got = plan.collect()
    src = cu.to_arrow()
    ck, nk, seg = (src.column(src.schema.get_field_index(c)).to_pylist() for c in ("c_custkey", "c_nationkey", "c_mktsegment"))
    assert got.num_rows == len(ck) == 1500
    assert got["c_custkey"].to_pylist() == ck
    assert got["seg"].to_pylist() == [s.strip(" ") for s in seg]
    assert got["h"].to_pylist() == [hashlib.md5(s.encode()).hexdigest() for s in seg]
    assert got["tag"].to_pylist() == [s + "#" + s.lower() for s in seg]
    assert got["nk"].to_pylist() == [None if n == 3 else n for n in nk]
    assert got["cls"].to_pylist() == ["b" if s == "BUILDING" else (s if n < 5 else "other") for s, n in zip(seg, nk)]
    assert any(v is None for v in got["nk"].to_pylist()) and "other" in got["cls"].to_pylist()
