"""Hash-repartition shuffle (BASELINE config 5): RepartitionExec(Hash(l_orderkey), N) over this rank's lineitem
partition (l_orderkey, l_extendedprice, l_discount = 20 B/row), exchanged inside the box.

    python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 profiles/tools/shuffle_bench.py [sf_per_gpu] [steps]
"""
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
import torch.distributed as dist

import ballista_b200 as b2
from ballista_b200 import RepartitionExec, cluster, col, tpch

rank, local_rank, world = cluster.env_rank()
torch.cuda.set_device(local_rank)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
ctx = b2.Context(local_rank)
cluster.attach_communicator(ctx, dist, rank, world)
stream = torch.cuda.ExternalStream(ctx.cuda_stream)
sf = float(sys.argv[1]) if len(sys.argv) > 1 else 10
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 5
n_out = world if world > 1 else int(os.environ.get("N_OUT", 8))
cols = ["l_orderkey", "l_extendedprice", "l_discount"]
rows = b2._capi.lib().b2_tpch_num_rows(b"lineitem", sf)
part = ctx.tpch("lineitem", sf * world, rank * rows, (rank + 1) * rows, columns=cols)
ctx.synchronize()


def step():
    rep = RepartitionExec(tpch.scan(ctx, [[part]]), "Hash", n_out, [col("l_orderkey")])
    if world > 1:
        out = rep.collect_device(rank)
    else:
        out = [b for p in range(n_out) for b in rep.collect_device(p)]
    return sum(b.num_rows for b in out), rep.last_path()


for _ in range(2):
    n_recv, path = step()
ctx.set_profiling(2)
if world > 1:
    dist.barrier()
ctx.synchronize()
sent0 = ctx.comm_bytes_sent()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
t0 = time.perf_counter()
e0.record(stream)
for _ in range(steps):
    n_recv, path = step()
e1.record(stream)
ctx.synchronize()
torch.cuda.synchronize()
wall = (time.perf_counter() - t0) / steps * 1e3
ms = cluster.max_over_ranks(dist, torch, e0.elapsed_time(e1) / steps, world)
sent = (ctx.comm_bytes_sent() - sent0) / steps
if rank == 0:
    print(json.dumps({"workload": f"hash repartition of lineitem SF{sf:g}/GPU ({rows} rows x 20 B) -> {n_out} partitions",
                      "n_gpus": world, "path": path, "ms_per_step": ms, "host_wall_ms": wall,
                      "rows_per_s_total": rows * world / (ms / 1e3),
                      "egress_bytes_per_gpu": sent, "egress_GBps_per_gpu": sent / (ms / 1e3) / 1e9,
                      "nvlink_frac_of_900GBps": sent / (ms / 1e3) / 1e9 / 900.0,
                      "split_GBps_local": rows * 20 / (ms / 1e3) / 1e9, "rows_received_rank0": n_recv}))
del part
import gc; gc.collect()
if world > 1:
    dist.barrier()
if os.environ.get('B2_TRACE') and rank == 0:
    ctx.close()
