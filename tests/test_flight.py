"""CPU: the FetchPartition Flight endpoint serves the executor's shuffle files (flight_service.rs:193-228) and its
ticket is the reference's protobuf Action (ballista.proto:433-471), checked against the protobuf runtime."""
import os

import pyarrow as pa
import pytest


def test_ticket_matches_the_protobuf_runtime():
    """hand-encoded Action { fetch_partition: PartitionId } == what google.protobuf encodes from the same schema"""
    from ballista_b200 import flight
    try:
        from google.protobuf import descriptor_pb2, descriptor_pool, message_factory
    except Exception:
        pytest.skip("protobuf runtime not importable")
    fd = descriptor_pb2.FileDescriptorProto(name="t.proto", package="t", syntax="proto3")
    pid = fd.message_type.add(name="PartitionId")
    pid.field.add(name="job_id", number=1, type=9, label=1)
    pid.field.add(name="stage_id", number=2, type=13, label=1)
    pid.field.add(name="partition_id", number=4, type=13, label=1)
    act = fd.message_type.add(name="Action")
    act.field.add(name="fetch_partition", number=3, type=11, label=1, type_name=".t.PartitionId")
    pool = descriptor_pool.DescriptorPool()
    pool.Add(fd)
    Action = message_factory.GetMessageClass(pool.FindMessageTypeByName("t.Action"))
    for job, stage, part in [("job-7Xk", 2, 5), ("j", 0, 0), ("x" * 200, 300, 70000)]:
        m = Action()
        m.fetch_partition.job_id, m.fetch_partition.stage_id, m.fetch_partition.partition_id = job, stage, part
        assert flight.encode_fetch_partition(job, stage, part) == m.SerializeToString()
        assert flight.decode_action(m.SerializeToString()) == ("fetch_partition", job, stage, part)


def test_fetch_partition_round_trip(tmp_path):
    from ballista_b200 import flight, shuffle_io
    schema = pa.schema([("k", pa.int32()), ("s", pa.utf8()), ("v", pa.float64())])
    batches = [pa.RecordBatch.from_pydict({"k": [1, 2, 3], "s": ["a", None, "ccc"], "v": [0.5, 1.5, 2.5]}, schema=schema),
               pa.RecordBatch.from_pydict({"k": [], "s": [], "v": []}, schema=schema),
               pa.RecordBatch.from_pydict({"k": [9], "s": ["z"], "v": [-1.0]}, schema=schema)]
    path = shuffle_io.partition_path(str(tmp_path), "job1", 3, 2)
    stats = shuffle_io.write_stream_to_disk(iter(batches), path, schema=schema)
    assert (stats.num_rows, stats.num_batches) == (4, 3)
    server, port = flight.serve(str(tmp_path))
    try:
        got, got_schema = flight.fetch_partition("127.0.0.1", port, "job1", 3, 2)
        assert got_schema.equals(schema)
        assert pa.Table.from_batches(got, schema).equals(pa.Table.from_batches(batches, schema))
        with pytest.raises(pa.ArrowException, match="Failed to open partition file"):
            flight.fetch_partition("127.0.0.1", port, "job1", 3, 99)
    finally:
        server.shutdown()
