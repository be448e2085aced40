"""N-rank TPC-H Q3 with the hash-repartition shuffle (torchrun): row-range partitions of customer / orders / lineitem
per rank, both join inputs exchanged on the join key over NVLink, each rank finalises its own groups.  Rank 0 gathers
the result rows and compares them with the oracle's single-node Q3 on the same synthetic tables (bit-exact keys,
<= 1e-9 relative on the Float64 sums)."""
import datetime
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, HERE)
import torch
import torch.distributed as dist

import ballista_b200 as b2
from ballista_b200 import MemoryExec, cluster, tpch

rank, local_rank, world = cluster.env_rank()
torch.cuda.set_device(local_rank)
dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
ctx = b2.Context(local_rank)
cluster.attach_communicator(ctx, dist, rank, world)


def part(table, sf, columns, align=1):
    n = b2._capi.lib().b2_tpch_num_rows(table.encode(), float(sf))
    r0, r1 = cluster.partition_rows(n, world, rank, align=align)
    return ctx.tpch(table, sf, r0, r1, columns=columns)


def run(sf, mode):
    if mode == "nccl":
        os.environ["B2_SHUFFLE"] = "nccl"
    else:
        os.environ.pop("B2_SHUFFLE", None)
    cu = part("customer", sf, tpch.Q3_CUSTOMER_COLUMNS)
    od = part("orders", sf, tpch.Q3_ORDERS_COLUMNS)
    li = part("lineitem", sf, tpch.Q3_LINEITEM_COLUMNS, align=32)
    scans = [tpch.scan(ctx, [[x]]) for x in (cu, od, li)]
    if mode == "broadcast":  # collect-left build sides all-gathered, probe sides stay put, states exchanged by group key
        partial, in_schema, p = tpch.q3_broadcast_stages(ctx, *scans, world)
        states = partial.collect_device(p)
        final = tpch.q3_final(MemoryExec(ctx, partial.schema(), [states]), in_schema)
    else:
        partial = tpch.q3_distributed_partial(*scans, world)
        states = partial.collect_device(rank)
        final = tpch.q3_final(MemoryExec(ctx, partial.schema(), [states]), partial.children()[0].schema())
    got = final.collect().to_pylist()
    for r in got:
        r["o_orderdate"] = (r["o_orderdate"] - datetime.date(1970, 1, 1)).days
    gathered = [None] * world
    dist.all_gather_object(gathered, got)
    if rank == 0:
        from oracle import oracle_py as O
        from test_gpu_join_q3 import oracle_q3
        from util import assert_rows_match
        no, nl, nc = (O.tpch_num_rows(t, sf) for t in ("orders", "lineitem", "customer"))
        want, _, n_j2 = oracle_q3(O, sf, no, nl, nc)
        rows = [r for p in gathered for r in p]
        keys = [r["l_orderkey"] for r in rows]
        assert len(keys) == len(set(keys)), "a group appears on two ranks"
        assert_rows_match(rows, want, ["l_orderkey", "o_orderdate", "o_shippriority"], approx=["revenue"])
        print(f"ok [{mode}] sf={sf} groups={len(rows)} per-rank={[len(p) for p in gathered]} join rows={n_j2}", flush=True)
    dist.barrier()


def run_one_sided(sf):
    """A join whose BUILD side arrives through an exchange while its probe side does not (each rank probes with its own
    row-range partition): not a partition-wise join - the build partitions are all-gathered (collect-left, as the
    reference's MergeExec(left)), and no match is lost."""
    import numpy as np
    from ballista_b200 import HashJoinExec, RepartitionExec, col
    od = part("orders", sf, ["o_orderkey", "o_custkey"])
    li = part("lineitem", sf, ["l_orderkey", "l_suppkey"], align=32)
    o, l = tpch.scan(ctx, [[od]]), tpch.scan(ctx, [[li]])
    j = HashJoinExec(RepartitionExec(o, "Hash", world, [col("o_orderkey")]), l, [("o_orderkey", "l_orderkey")])
    mine = j.collect(0)  # the probe side's own (single) partition
    assert "partitioned" not in j.last_path(), j.last_path()
    rows = list(zip(mine["o_orderkey"].to_pylist(), mine["o_custkey"].to_pylist(), mine["l_suppkey"].to_pylist()))
    gathered = [None] * world
    dist.all_gather_object(gathered, rows)
    if rank == 0:
        from oracle import oracle_py as O
        no, nl = O.tpch_num_rows("orders", sf), O.tpch_num_rows("lineitem", sf)
        oo, ll = O.gen_orders(sf, 0, no), O.gen_lineitem(sf, 0, nl)
        got = sorted(r for p_ in gathered for r in p_)
        lidx, ridx = O.hash_join([oo["o_orderkey"]], [ll["l_orderkey"]])  # every lineitem row finds its order
        want = sorted(zip(oo["o_orderkey"][lidx].tolist(), oo["o_custkey"][lidx].tolist(), ll["l_suppkey"][ridx].tolist()))
        assert got == want, (len(got), len(want))
        print(f"ok [one-sided exchange: build side only] sf={sf} rows={len(got)} path={j.last_path()}", flush=True)
    dist.barrier()


run(0.02, "p2p")
run_one_sided(0.02)
run(0.02, "nccl")
run(0.2, "p2p")
run(0.02, "broadcast")
run(0.2, "broadcast")
ctx.synchronize()
if rank == 0:
    print("q3_dist: all ok")
