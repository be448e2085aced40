"""N-rank correctness check of the in-box shuffle (torchrun): the multiset of full rows is preserved bit-exactly,
equal keys land on one rank, row order inside (source, destination) is kept.  Runs both exchange paths.

    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 tests/shuffle_check_worker.py
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

import ballista_b200 as b2
from ballista_b200 import RepartitionExec, cluster, col, tpch

rank, local_rank, world = cluster.env_rank()
torch.cuda.set_device(local_rank)
dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
ctx = b2.Context(local_rank)
cluster.attach_communicator(ctx, dist, rank, world)
cols = ["l_orderkey", "l_extendedprice", "l_returnflag", "l_shipdate"]


def rows_of(table):
    return list(zip(*[table[c].to_pylist() for c in cols]))


def run(sf, rows_per_rank, mode):
    if mode == "nccl":
        os.environ["B2_SHUFFLE"] = "nccl"
    else:
        os.environ.pop("B2_SHUFFLE", None)
    # uneven partitions; the last rank also gets an empty second batch
    n = rows_per_rank + 1000 * rank
    r0 = rank * (rows_per_rank + 5000)
    part = ctx.tpch("lineitem", sf, r0, r0 + n, columns=cols)
    import pyarrow as pa
    mine = rows_of(pa.Table.from_batches([part.to_arrow()]))
    rep = RepartitionExec(tpch.scan(ctx, [[part]]), "Hash", world, [col("l_orderkey")])
    got = rep.collect(rank)
    path = rep.last_path()
    got_rows = rows_of(got)
    gathered_in, gathered_out = [None] * world, [None] * world
    dist.all_gather_object(gathered_in, mine)
    dist.all_gather_object(gathered_out, got_rows)
    if rank == 0:
        all_in = sorted(r for p in gathered_in for r in p)
        all_out = sorted(r for p in gathered_out for r in p)
        assert all_in == all_out, f"{mode}: multiset of rows changed ({len(all_in)} vs {len(all_out)})"
        owner = {}
        for p, rows in enumerate(gathered_out):
            for r in rows:
                assert owner.setdefault(r[0], p) == p, f"{mode}: key {r[0]} on two ranks"
        # order: inside one destination, rows of one source keep their source order and sources arrive in rank order
        for p, rows in enumerate(gathered_out):
            want = [r for s in range(world) for r in gathered_in[s] if owner[r[0]] == p]
            assert rows == want, f"{mode}: row order inside partition {p} differs"
        print(f"ok [{mode}] path={path} sf={sf} rows_in={len(all_in)} parts={[len(p) for p in gathered_out]}", flush=True)
    if mode == "p2p":
        assert path == "nvlink-peer-pull", path
    dist.barrier()


def run_filtered(sf, rows_per_rank, cutoff, label):
    """A FilterExec directly below the repartition rides in the split kernels.  `cutoff` picks the density of the
    survivors: most rows (dense move kernel) or under one in eight (sparse move kernel); same contract either way."""
    import datetime
    import pyarrow as pa
    from ballista_b200 import CoalesceBatchesExec, FilterExec, date
    os.environ.pop("B2_SHUFFLE", None)
    n = rows_per_rank + 777 * rank
    r0 = rank * (rows_per_rank + 5000)
    part = ctx.tpch("lineitem", sf, r0, r0 + n, columns=cols)
    day = (datetime.date.fromisoformat(cutoff) - datetime.date(1970, 1, 1)).days
    mine = [r for r in rows_of(pa.Table.from_batches([part.to_arrow()]))
            if (r[3] - datetime.date(1970, 1, 1)).days > day]
    rep = RepartitionExec(CoalesceBatchesExec(FilterExec(col("l_shipdate") > date(cutoff), tpch.scan(ctx, [[part]]))),
                          "Hash", world, [col("l_orderkey")])
    got_rows = rows_of(rep.collect(rank))
    path = rep.last_path()
    gathered_in, gathered_out = [None] * world, [None] * world
    dist.all_gather_object(gathered_in, mine)
    dist.all_gather_object(gathered_out, got_rows)
    if rank == 0:
        owner = {}
        for q, rows in enumerate(gathered_out):
            for r in rows:
                assert owner.setdefault(r[0], q) == q, f"{label}: key {r[0]} on two ranks"
        assert sorted(r for q in gathered_in for r in q) == sorted(r for q in gathered_out for r in q), label
        for q, rows in enumerate(gathered_out):
            want = [r for s in range(world) for r in gathered_in[s] if owner[r[0]] == q]
            assert rows == want, f"{label}: row order inside partition {q} differs"
        kept = sum(len(q) for q in gathered_in)
        print(f"ok [filtered:{label}] path={path} kept {kept} of ~{world * n} rows", flush=True)
    assert path == "nvlink-peer-pull+fused-filter", path
    dist.barrier()


fixed_cols = ["l_orderkey", "l_extendedprice", "l_discount", "l_shipdate"]


def run_fixed(sf, rows_per_rank, mode, label, cutoff=None, expect=None):
    """Fixed-width columns only: the PUSH form (rows stored straight into the destination's final columns over NVLink,
    outputs are windows of the receive arena).  mode "pull" forces the staged form on the same input.  `cutoff` puts a
    FilterExec below the repartition (fused into the split).  Returns the path taken."""
    import datetime
    import pyarrow as pa
    from ballista_b200 import CoalesceBatchesExec, FilterExec, date
    if mode == "pull":
        os.environ["B2_SHUFFLE"] = "pull"
    else:
        os.environ.pop("B2_SHUFFLE", None)
    n = rows_per_rank + 1111 * rank
    r0 = rank * (rows_per_rank + 5000)
    part = ctx.tpch("lineitem", sf, r0, r0 + n, columns=fixed_cols)
    t = pa.Table.from_batches([part.to_arrow()])
    mine = list(zip(*[t[c].to_pylist() for c in fixed_cols]))
    src = tpch.scan(ctx, [[part]])
    if cutoff:
        day = (datetime.date.fromisoformat(cutoff) - datetime.date(1970, 1, 1)).days
        mine = [r for r in mine if (r[3] - datetime.date(1970, 1, 1)).days > day]
        src = CoalesceBatchesExec(FilterExec(col("l_shipdate") > date(cutoff), src))
    rep = RepartitionExec(src, "Hash", world, [col("l_orderkey")])
    got = rep.collect(rank)
    path = rep.last_path()
    got_rows = list(zip(*[got[c].to_pylist() for c in fixed_cols]))
    gathered_in, gathered_out = [None] * world, [None] * world
    dist.all_gather_object(gathered_in, mine)
    dist.all_gather_object(gathered_out, got_rows)
    if rank == 0:
        owner = {}
        for q, rows in enumerate(gathered_out):
            for r in rows:
                assert owner.setdefault(r[0], q) == q, f"{label}: key {r[0]} on two ranks"
        assert sorted(r for q in gathered_in for r in q) == sorted(r for q in gathered_out for r in q), label
        for q, rows in enumerate(gathered_out):
            want = [r for s in range(world) for r in gathered_in[s] if owner[r[0]] == q]
            assert rows == want, f"{label}: row order inside partition {q} differs"
        print(f"ok [fixed:{label}] path={path} rows_in={sum(len(q) for q in gathered_in)} "
              f"parts={[len(q) for q in gathered_out]}", flush=True)
    if expect:
        assert path == expect, (label, path, expect)
    os.environ.pop("B2_SHUFFLE", None)
    dist.barrier()
    return path


def check_allgather(sf, n):
    """b2_comm_allgather_batch: rank-ordered concatenation, small-slot protocol (n <= 256) and general protocol."""
    import pyarrow as pa
    part = ctx.tpch("lineitem", sf, rank * n, rank * n + n + rank, columns=cols)
    mine = rows_of(pa.Table.from_batches([part.to_arrow()]))
    got = rows_of(pa.Table.from_batches([ctx.allgather(part).to_arrow()]))
    gathered = [None] * world
    dist.all_gather_object(gathered, mine)
    want = [r for p in gathered for r in p]
    assert got == want, f"allgather({n}) differs on rank {rank}"
    if rank == 0:
        print(f"ok [allgather] rows/rank~{n} total={len(got)}", flush=True)
    dist.barrier()


sf0 = 0.01 * world  # 60 k lineitem rows per rank
for mode in ("p2p", "nccl", "p2p"):
    run(sf0, 20_000, mode)
run(0.1 * world, 150_000, "p2p")   # arena regrow
run(sf0, 3_000, "p2p")             # smaller again: arena reused
run_filtered(sf0, 20_000, "1993-01-01", "dense")    # ~85 % of the rows survive
run_filtered(sf0, 20_000, "1998-06-01", "sparse")   # ~7 %: the sparse move kernel
run_filtered(sf0, 20_000, "1999-01-01", "none")     # nothing survives
# the push form.  The receive arena's floor is shrunk so that these sizes reach its grow / too-small paths.
os.environ["B2_PUSH_ARENA_MIN_MB"] = "1"
run_fixed(sf0, 20_000, "push", "first", expect="nvlink-peer-push")
run_fixed(sf0, 20_000, "pull", "forced-pull", expect="nvlink-peer-pull")
run_fixed(sf0, 3_000, "push", "small", expect="nvlink-peer-push")
# an exchange larger than anything offered so far: this one falls back to the staged form on every rank, the arena grows,
# and the same exchange is pushed the next time
run_fixed(0.2 * world, 400_000, "push", "too-big-for-the-offer", expect="nvlink-peer-pull")
run_fixed(0.2 * world, 400_000, "push", "after-growth", expect="nvlink-peer-push")
held = []
for i in range(3):  # outputs kept alive: later exchanges land in other spans of the arena, earlier outputs stay intact
    part = ctx.tpch("lineitem", sf0, rank * 30_000, rank * 30_000 + 20_000 + i, columns=fixed_cols)
    rep = RepartitionExec(tpch.scan(ctx, [[part]]), "Hash", world, [col("l_orderkey")])
    held.append((rep, rep.collect_device(rank), part))
first = held[0][1][0].to_arrow()
run_fixed(sf0, 20_000, "push", "while-three-outputs-are-held", expect="nvlink-peer-push")
again = held[0][1][0].to_arrow()
assert first.equals(again), "an output still referenced was overwritten by a later exchange"
del held, first, again
run_fixed(sf0, 20_000, "push", "filtered-dense", cutoff="1993-01-01", expect="nvlink-peer-push+fused-filter")
run_fixed(sf0, 20_000, "push", "filtered-sparse", cutoff="1998-06-01", expect="nvlink-peer-push+fused-filter")
run_fixed(sf0, 20_000, "push", "filtered-none", cutoff="1999-01-01", expect="nvlink-peer-push+fused-filter")
run_fixed(sf0, 20_000, "pull", "filtered-sparse-pull", cutoff="1998-06-01", expect="nvlink-peer-pull+fused-filter")
os.environ.pop("B2_PUSH_ARENA_MIN_MB", None)
check_allgather(sf0, 4)
check_allgather(sf0, 256 - world)
check_allgather(sf0, 3000)
ctx.synchronize()
if rank == 0:
    print("shuffle_check: all ok")
