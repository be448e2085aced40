"""Cross-box interop of stage outputs: the reference's shuffle files.

Inside one box the GPUs exchange over NVLink (csrc/shuffle.cu).  Towards CPU executors, or across boxes, a stage's
output partition is what the reference writes: an Arrow IPC *file* at `<work_dir>/<job_id>/<stage_id>/<partition>/
data.arrow` (rust/executor/src/flight_service.rs:104-111), produced by `utils::write_stream_to_disk`
(rust/core/src/utils.rs:49-84), which also returns the `PartitionStats` the scheduler records (rows, batches,
bytes).  A CPU `ShuffleReaderExec` / `FetchPartition` (flight_service.rs:193-228) reads that file back unchanged.
The IPC encoding itself is the Arrow library's (pyarrow here, arrow-rs there): host I/O, not part of the GPU path.
"""
import os
from dataclasses import dataclass

import pyarrow as pa
import pyarrow.ipc


@dataclass
class PartitionStats:
    """PartitionStats::new(num_rows, num_batches, num_bytes) (rust/core/src/serde/scheduler/mod.rs)."""
    num_rows: int
    num_batches: int
    num_bytes: int


def partition_path(work_dir, job_id, stage_id, partition):
    """The executor's layout: <work_dir>/<job_id>/<stage_id>/<partition>/data.arrow (flight_service.rs:104-111)."""
    d = os.path.join(work_dir, str(job_id), str(stage_id), str(partition))
    os.makedirs(d, exist_ok=True)
    return os.path.join(d, "data.arrow")


def write_stream_to_disk(stream, path, schema=None):
    """Drain a RecordBatchStream (anything iterable over pyarrow RecordBatches with `.schema()` / `.schema`) into an
    Arrow IPC file; returns PartitionStats.  num_bytes sums the buffer sizes of every column, like
    `array.get_array_memory_size()` in utils.rs:65-72."""
    if schema is None:
        schema = stream.schema() if callable(getattr(stream, "schema", None)) else stream.schema
    rows = batches = nbytes = 0
    try:
        sink = pa.OSFile(path, "wb")
    except OSError as e:  # BallistaError::General("Failed to create partition file at ...")
        raise OSError(f"Failed to create partition file at {path}: {e!r}") from e
    with sink, pa.ipc.new_file(sink, schema) as writer:
        for batch in stream:
            batches += 1
            rows += batch.num_rows
            nbytes += sum(col.nbytes for col in batch.columns)
            writer.write_batch(batch)
    return PartitionStats(rows, batches, nbytes)


def execute_partition_to_disk(plan, partition, work_dir, job_id, stage_id):
    """What the executor's ExecutePartition does for one partition (flight_service.rs:96-135): run it, stream the
    result to the shuffle file, report (path, PartitionStats)."""
    path = partition_path(work_dir, job_id, stage_id, partition)
    stats = write_stream_to_disk(plan.execute(partition), path, schema=plan.schema())
    return path, stats


def read_partition(path):
    """The reader side (FetchPartition / ShuffleReaderExec): all RecordBatches of one shuffle file."""
    with pa.memory_map(path, "r") as src:
        reader = pa.ipc.open_file(src)
        return [reader.get_batch(i) for i in range(reader.num_record_batches)], reader.schema
