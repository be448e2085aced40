"""CPU, world_size 2 over gloo: the host-side multi-rank logic (partition ranges, communicator-id broadcast,
max-over-ranks timing) and the Partial -> merge -> Final decomposition checked with the oracle standing in for
the per-rank compute."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, n, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from ballista_b200 import cluster
        from oracle import oracle_py as O
        uid = cluster.broadcast_unique_id(dist, lambda: bytes([7]) * 128, rank)
        assert uid == bytes([7]) * 128
        b, e = cluster.partition_rows(n, world, rank, align=32)
        cols = O.gen_lineitem(0.05, b, e - b)
        # Partial stage on this rank: group states (sums / counts) of its partition
        m = cols["l_shipdate"] <= O.DATE_1998_09_02
        f = O.filter_columns([cols["l_returnflag"], cols["l_linestatus"], cols["l_quantity"], cols["l_extendedprice"]], m)
        rep, sums, counts = O.hash_aggregate([f[0], f[1]], [f[2], f[3]])
        rf, ls = f[0].to_list(), f[1].to_list()
        states = [(rf[r], ls[r], float(sums[g, 0]), float(sums[g, 1]), int(counts[g, 0])) for g, r in enumerate(rep)]
        gathered = [None] * world
        dist.all_gather_object(gathered, states)  # the exchange (all-gather of partial states)
        final = {}
        for part in gathered:  # Final stage: merge states in rank order
            for k1, k2, s1, s2, c in part:
                a = final.setdefault((k1, k2), [0.0, 0.0, 0])
                a[0] += s1
                a[1] += s2
                a[2] += c
        t = cluster.max_over_ranks(dist, torch, 10.0 + rank, world, device="cpu")
        if rank == 0:
            out.put((sorted((k, v) for k, v in final.items()), t, (b, e)))
    finally:
        dist.destroy_process_group()


def test_two_rank_partial_merge_final_matches_single_process(O):
    n, world = 100_000, 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, n, q)) for r in range(world)]
    for p in procs:
        p.start()
    got, tmax, bounds = q.get(timeout=240)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert tmax == 11.0
    want = O.q1(O.gen_lineitem(0.05, 0, n))
    assert len(got) == len(want)
    for (key, (s_qty, s_price, cnt)), w in zip(got, want):
        assert key == (w["l_returnflag"], w["l_linestatus"])
        assert cnt == w["count_order"]
        assert abs(s_qty - w["sum_qty"]) <= 1e-9 * abs(w["sum_qty"])
        assert abs(s_price - w["sum_base_price"]) <= 1e-9 * abs(w["sum_base_price"])


@pytest.mark.parametrize("total,world", [(0, 1), (1, 2), (600_000_000, 8), (1000, 3), (31, 4)])
def test_partition_rows_tile_the_table(total, world):
    from ballista_b200 import cluster
    ranges = [cluster.partition_rows(total, world, r, align=32) for r in range(world)]
    assert ranges[0][0] == 0 and ranges[-1][1] == total
    for (b0, e0), (b1, e1) in zip(ranges, ranges[1:]):
        assert e0 == b1 and b0 <= e0
    for b, e in ranges[:-1]:
        assert e % 32 == 0 or e == total
    with pytest.raises(ValueError):
        cluster.partition_rows(total, world, world)
