"""torchrun: time distributed Q3 over an SF<total> database (row-range partitions per rank), with the host trace of rank 0.
usage: torchrun ... profiles/tools/q3_dist_time.py <total_sf> <steps>"""
import os, sys, time, gc
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch, torch.distributed as dist
import ballista_b200 as b2
from ballista_b200 import MemoryExec, cluster, tpch
rank, local_rank, world = cluster.env_rank()
torch.cuda.set_device(local_rank)
dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
ctx = b2.Context(local_rank)
cluster.attach_communicator(ctx, dist, rank, world)
sf = float(sys.argv[1]) if len(sys.argv) > 1 else 20.0
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 10
mode = sys.argv[3] if len(sys.argv) > 3 else "repartition"
def part(table, columns, align=1):
    n = b2._capi.lib().b2_tpch_num_rows(table.encode(), sf)
    r0, r1 = cluster.partition_rows(n, world, rank, align=align)
    return ctx.tpch(table, sf, r0, r1, columns=columns)
cu, od, li = part("customer", tpch.Q3_CUSTOMER_COLUMNS), part("orders", tpch.Q3_ORDERS_COLUMNS), part("lineitem", tpch.Q3_LINEITEM_COLUMNS, 32)
def step():
    scans = [tpch.scan(ctx, [[x]]) for x in (cu, od, li)]
    if mode == "broadcast":
        partial, in_schema, p = tpch.q3_broadcast_stages(ctx, *scans, world)
        states = partial.collect_device(p)
        final = tpch.q3_final(MemoryExec(ctx, partial.schema(), [states]), in_schema)
    else:
        partial = tpch.q3_distributed_partial(*scans, world)
        states = partial.collect_device(rank)
        final = tpch.q3_final(MemoryExec(ctx, partial.schema(), [states]), partial.children()[0].schema())
    return sum(b.num_rows for b in final.collect_device(0))
for _ in range(3): g = step()
dist.barrier(); ctx.synchronize()
if rank == 0 and os.environ.get("B2_TRACE"): ctx.set_profiling(2)
l0 = ctx.kernel_launches(); t = time.perf_counter()
for _ in range(steps): g = step()
ctx.synchronize(); dt = (time.perf_counter() - t) / steps
dist.barrier()
if rank == 0:
    print(f"Q3 distributed [{mode}] SF{sf:g} on {world} GPUs: {dt*1e3:.2f} ms/step, groups(rank0) {g}, launches/step {(ctx.kernel_launches()-l0)/steps}", flush=True)
del cu, od, li; gc.collect()
ctx.close()
dist.destroy_process_group()
