"""Shuffle-file interop: Arrow IPC files in the executor's directory layout + PartitionStats."""
import os

import numpy as np
import pyarrow as pa
import pytest


def test_write_read_roundtrip_and_stats(tmp_path):
    from ballista_b200 import shuffle_io
    schema = pa.schema([("k", pa.int32()), ("v", pa.float64()), ("s", pa.utf8())])
    batches = [pa.RecordBatch.from_arrays([pa.array(np.arange(n, dtype=np.int32)), pa.array(np.arange(n) * 0.5),
                                           pa.array([str(i) for i in range(n)])], schema=schema) for n in (5, 0, 1000)]

    class Stream:
        def schema(self):
            return schema
        def __iter__(self):
            return iter(batches)
    path = shuffle_io.partition_path(str(tmp_path), "job7", 2, 3)
    assert path.endswith(os.path.join("job7", "2", "3", "data.arrow"))
    stats = shuffle_io.write_stream_to_disk(Stream(), path)
    assert (stats.num_rows, stats.num_batches) == (1005, 3)
    assert stats.num_bytes == sum(c.nbytes for b in batches for c in b.columns)
    got, sch = shuffle_io.read_partition(path)
    assert sch == schema and [b.num_rows for b in got] == [5, 0, 1000]
    assert pa.Table.from_batches(got).equals(pa.Table.from_batches(batches))
    with pytest.raises(OSError):
        shuffle_io.write_stream_to_disk(Stream(), os.path.join(str(tmp_path), "missing_dir", "data.arrow"))


@pytest.mark.gpu
def test_gpu_stage_output_written_as_shuffle_files(ctx, tmp_path):
    """A GPU repartition stage writes one shuffle file per output partition; a CPU reader (pyarrow) reads them back."""
    from ballista_b200 import RepartitionExec, col, shuffle_io, tpch
    li = ctx.tpch("lineitem", 0.02, columns=["l_orderkey", "l_extendedprice", "l_returnflag"])
    rep = RepartitionExec(tpch.scan(ctx, [[li]]), "Hash", 3, [col("l_orderkey")])
    total, seen = 0, {}
    for p in range(3):
        path, stats = shuffle_io.execute_partition_to_disk(rep, p, str(tmp_path), "jobA", 1)
        got, _ = shuffle_io.read_partition(path)
        assert sum(b.num_rows for b in got) == stats.num_rows
        total += stats.num_rows
        for k in set(k for b in got for k in b["l_orderkey"].to_pylist()):
            assert seen.setdefault(k, p) == p
    assert total == li.num_rows
