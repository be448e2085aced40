"""torchrun: per-stage wall time (synchronised after every stage) of the broadcast-join Q3 plan."""
import os, sys, time, gc
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch, torch.distributed as dist
import pyarrow as pa
import ballista_b200 as b2
from ballista_b200 import MemoryExec, cluster, tpch, col
from ballista_b200.plan import *
rank, local_rank, world = cluster.env_rank()
torch.cuda.set_device(local_rank)
dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
ctx = b2.Context(local_rank)
cluster.attach_communicator(ctx, dist, rank, world)
sf = float(sys.argv[1]) if len(sys.argv) > 1 else 100.0
def part(table, columns, align=1):
    n = b2._capi.lib().b2_tpch_num_rows(table.encode(), sf)
    r0, r1 = cluster.partition_rows(n, world, rank, align=align)
    return ctx.tpch(table, sf, r0, r1, columns=columns)
cu, od, li = part("customer", tpch.Q3_CUSTOMER_COLUMNS), part("orders", tpch.Q3_ORDERS_COLUMNS), part("lineitem", tpch.Q3_LINEITEM_COLUMNS, 32)
pc, po, pl = tpch._q3_predicates("BUILDING", "1995-03-15")
T = {}
def lap(name, t0):
    ctx.synchronize(); T[name] = T.get(name, 0.0) + time.perf_counter() - t0; return time.perf_counter()
def step():
    customer, orders, lineitem = (tpch.scan(ctx, [[x]]) for x in (cu, od, li))
    t = time.perf_counter()
    p1 = ProjectionExec([col("c_custkey")], CoalesceBatchesExec(FilterExec(pc, customer)))
    g = CoalesceBatchesExec(p1, (1 << 31) - 2).collect_device(0); t = lap("1 filter customer", t)
    g = [ctx.allgather(g[0])]; t = lap("1 allgather", t)
    s1 = MemoryExec(ctx, p1.schema(), [g])
    j1 = HashJoinExec(s1, CoalesceBatchesExec(FilterExec(po, orders)), [("c_custkey", "o_custkey")])
    p2 = ProjectionExec([col("o_orderkey"), col("o_orderdate"), col("o_shippriority")], CoalesceBatchesExec(j1))
    g = CoalesceBatchesExec(p2, (1 << 31) - 2).collect_device(0); t = lap("2 join orders", t)
    g = [ctx.allgather(g[0])]; t = lap("2 allgather", t)
    s2 = MemoryExec(ctx, p2.schema(), [g])
    j2 = CoalesceBatchesExec(HashJoinExec(s2, CoalesceBatchesExec(FilterExec(pl, lineitem)), [("o_orderkey", "l_orderkey")]))
    partial = HashAggregateExec("Partial", tpch.q3_groups(), tpch.q3_aggregates(), j2)
    st = partial.collect_device(0); t = lap("3 join lineitem + partial", t)
    rep = RepartitionExec(MemoryExec(ctx, partial.schema(), [st]), "Hash", world, [col("l_orderkey")])
    st = rep.collect_device(rank); t = lap("3 exchange states", t)
    final = tpch.q3_final(MemoryExec(ctx, partial.schema(), [st]), j2.schema())
    out = final.collect_device(0); t = lap("4 final", t)
    return sum(b.num_rows for b in out)
for _ in range(3): step()
T.clear(); dist.barrier()
for _ in range(10): g = step()
if rank == 0:
    for k, v in T.items(): print(f"{k:28s} {v/10*1e3:7.3f} ms", flush=True)
    print("sum", sum(T.values()) / 10 * 1e3, "groups", g, flush=True)
dist.barrier()
del cu, od, li; gc.collect(); ctx.close(); dist.destroy_process_group()
