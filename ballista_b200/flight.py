"""`FetchPartition` over Arrow Flight: the cross-box half of the shuffle.

Inside one box the GPUs exchange over NVLink (csrc/shuffle.cu).  Towards CPU executors, or across boxes, a stage's output
partition is the file `shuffle_io.execute_partition_to_disk` writes, and the reference's executors fetch it with a Flight
`do_get` whose ticket is a serialized `protobuf::Action { fetch_partition: PartitionId { job_id, stage_id, partition_id } }`
(rust/core/src/client.rs:123-160; server side rust/executor/src/flight_service.rs:193-228; messages
rust/core/proto/ballista.proto:433-471).  This module serves exactly that request from the same files, so a stock
Ballista `ShuffleReaderExec` (rust/core/src/execution_plans/shuffle_reader.rs:77-99) can read what a GPU stage produced,
and offers the matching client for the opposite direction.  Host I/O only: pyarrow's Flight and IPC, no GPU work.
"""
import os
import threading

import pyarrow as pa
import pyarrow.flight as fl
import pyarrow.ipc


# ---- the two protobuf messages involved, encoded / decoded by hand (proto3 wire format) ----
def _varint(v):
    out = bytearray()
    while True:
        b = v & 0x7F
        v >>= 7
        if v:
            out.append(b | 0x80)
        else:
            out.append(b)
            return bytes(out)


def _read_varint(buf, pos):
    shift = v = 0
    while True:
        if pos >= len(buf):
            raise ValueError("truncated varint")
        b = buf[pos]
        pos += 1
        v |= (b & 0x7F) << shift
        if not b & 0x80:
            return v, pos
        shift += 7


def _fields(buf):
    """(field number, wire type, value) of one message; length-delimited values are bytes"""
    pos = 0
    while pos < len(buf):
        tag, pos = _read_varint(buf, pos)
        num, wire = tag >> 3, tag & 7
        if wire == 0:
            v, pos = _read_varint(buf, pos)
        elif wire == 2:
            n, pos = _read_varint(buf, pos)
            v = bytes(buf[pos:pos + n])
            if len(v) != n:
                raise ValueError("truncated field")
            pos += n
        elif wire == 1:
            v, pos = bytes(buf[pos:pos + 8]), pos + 8
        elif wire == 5:
            v, pos = bytes(buf[pos:pos + 4]), pos + 4
        else:
            raise ValueError(f"unsupported wire type {wire}")
        yield num, wire, v


def encode_fetch_partition(job_id, stage_id, partition_id):
    """Ticket bytes of Action::FetchPartition(PartitionId::new(job_id, stage_id, partition_id)) (client.rs:123-131):
    Action.fetch_partition = 3; PartitionId.job_id = 1, .stage_id = 2, .partition_id = 4 (ballista.proto:443,467-471)."""
    job = job_id.encode()
    pid = b"\x0a" + _varint(len(job)) + job
    if stage_id:
        pid += b"\x10" + _varint(int(stage_id))
    if partition_id:
        pid += b"\x20" + _varint(int(partition_id))
    return b"\x1a" + _varint(len(pid)) + pid


def decode_action(ticket):
    """-> ("fetch_partition", job_id, stage_id, partition_id) | ("query",) | ("execute_partition",)"""
    kind = None
    for num, wire, v in _fields(ticket):
        if num == 3 and wire == 2:
            job, stage, part = "", 0, 0
            for n2, w2, v2 in _fields(v):
                if n2 == 1 and w2 == 2:
                    job = v2.decode()
                elif n2 == 2 and w2 == 0:
                    stage = v2
                elif n2 == 4 and w2 == 0:
                    part = v2
            kind = ("fetch_partition", job, stage, part)
        elif num == 1 and wire == 2:
            kind = ("query",)
        elif num == 2 and wire == 2:
            kind = ("execute_partition",)
    if kind is None:
        raise ValueError("Action without an ActionType")
    return kind


class ExecutorFlightService(fl.FlightServerBase):
    """The executor's Flight endpoint for stage outputs: `do_get(FetchPartition)` streams
    `<work_dir>/<job_id>/<stage_id>/<partition_id>/data.arrow` (flight_service.rs:193-228).  Executing plans is the
    executor's job (it calls b2_plan_from_proto + b2_op_execute and writes the file); asking this endpoint for it is
    answered as the reference answers an unsupported action."""

    def __init__(self, work_dir, location="grpc://127.0.0.1:0", **kw):
        super().__init__(location, **kw)
        self.work_dir = work_dir

    def do_get(self, context, ticket):
        try:
            action = decode_action(ticket.ticket)
        except ValueError as e:
            raise fl.FlightServerError(f"Ballista Error: General({e!r})")
        if action[0] != "fetch_partition":
            raise fl.FlightServerError(f"Ballista Error: NotImplemented(\"{action[0]} is served by the executor proper\")")
        _, job, stage, part = action
        path = os.path.join(self.work_dir, job, str(stage), str(part), "data.arrow")
        try:
            src = pa.memory_map(path, "r")
        except OSError as e:  # BallistaError::General("Failed to open partition file at ...")
            raise fl.FlightServerError(f"Ballista Error: General(\"Failed to open partition file at {path}: {e!r}\")")
        reader = pa.ipc.open_file(src)
        schema = reader.schema

        def batches():
            for i in range(reader.num_record_batches):
                yield reader.get_batch(i)
        return fl.GeneratorStream(schema, batches())


def serve(work_dir, host="127.0.0.1", port=0):
    """Start the endpoint on a background thread; -> (server, port)."""
    server = ExecutorFlightService(work_dir, f"grpc://{host}:{port}")
    threading.Thread(target=server.serve, daemon=True).start()
    return server, server.port


def fetch_partition(host, port, job_id, stage_id, partition_id):
    """BallistaClient::fetch_partition (client.rs:123-131): -> (list of RecordBatches, schema).  Works against this
    module's endpoint and against a stock Ballista executor alike."""
    client = fl.connect(f"grpc://{host}:{port}")
    try:
        reader = client.do_get(fl.Ticket(encode_fetch_partition(job_id, stage_id, partition_id)))
        table = reader.read_all()
        return table.to_batches(), table.schema
    finally:
        client.close()
