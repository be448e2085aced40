"""Host-side plumbing for one-process-per-GPU execution (torch.distributed carries the rendezvous only).

Mirrors how the reference maps work: one task per partition (rust/scheduler/src/lib.rs:381-399), a stage's
partitions are independent up to the exchange (docs/architecture.md:17-18).  Here partition p of the scan stage
lives on rank p; partial aggregate states are all-gathered over NCCL by the library (b2_comm_allgather_batch),
which replaces the MergeExec stage hop (rust/scheduler/src/planner.rs:136-148).
"""
import os


def env_rank():
    """(rank, local_rank, world_size) as torchrun exports them."""
    return (int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0)),
            int(os.environ.get("WORLD_SIZE", 1)))


def partition_rows(total_rows, world, rank, align=1):
    """Contiguous row range [begin, end) of partition `rank` of `world`; ranges tile [0, total_rows) exactly.
    `align` keeps boundaries on multiples (e.g. 32 = one lineitem row group of the synthetic generator)."""
    if not (0 <= rank < world):
        raise ValueError(f"rank {rank} outside world of {world}")
    def bound(p):
        b = total_rows * p // world
        return min(total_rows, (b // align) * align) if p < world else total_rows
    return bound(rank), bound(rank + 1)


def broadcast_unique_id(dist, make_id, rank, src=0):
    """Rank `src` creates the communicator id (b2_comm_unique_id); everyone receives the same 128 bytes."""
    box = [make_id() if rank == src else None]
    dist.broadcast_object_list(box, src=src)
    uid = box[0]
    if not isinstance(uid, (bytes, bytearray)) or len(uid) != 128:
        raise ValueError("communicator id must be 128 bytes")
    return bytes(uid)


def attach_communicator(ctx, dist, rank, world):
    """Give `ctx` its NCCL communicator (no-op for a single rank)."""
    if world > 1:
        from .plan import Context
        ctx.comm_init(rank, world, broadcast_unique_id(dist, Context.comm_unique_id, rank))
    return ctx


def max_over_ranks(dist, torch, value, world, device="cuda"):
    """Multi-GPU timings are the max over ranks of the device-side time."""
    t = torch.tensor([float(value)], dtype=torch.float64, device=device)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())
