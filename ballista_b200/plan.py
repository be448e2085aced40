"""Host-side mirror of the reference's physical-operator interface over the C ABI.

Each class corresponds to the DataFusion operator Ballista instantiates on the executor
(rust/core/src/serde/physical_plan/from_proto.rs:58-346) and exposes the `ExecutionPlan` trait as the
reference's own operators implement it (rust/core/src/execution_plans/query_stage.rs:50-84):
schema() / output_partitioning() / children() / execute(partition) -> RecordBatchStream.
All compute happens in libballista_b200.so; pyarrow is used only to hand RecordBatches across the
Arrow C Data Interface.
"""
import ctypes as C

import pyarrow as pa

from . import _capi as K
from .expr import AggExpr, Expr


def _addr(struct):
    return C.addressof(struct)


class Context:
    """One GPU (b2_ctx): stream + memory pool (+ NCCL communicator when attached)."""

    def __init__(self, device=0):
        self._h = C.c_void_p()
        K.check(K.lib().b2_ctx_create(device, C.byref(self._h)))
        self.device = device
        self.rank, self.nranks = 0, 1

    def close(self):
        if self._h:
            K.lib().b2_ctx_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def cuda_stream(self):
        return K.lib().b2_ctx_cuda_stream(self._h)

    def synchronize(self):
        K.check(K.lib().b2_ctx_synchronize(self._h))

    def kernel_launches(self):
        return K.lib().b2_ctx_kernel_launches(self._h)

    def set_profiling(self, on=True):
        K.check(K.lib().b2_ctx_set_profiling(self._h, int(on)))

    def profile(self):
        """-> (total ms, launches) of the aggregate scan kernels since the last call (CUDA events)."""
        ms, n = C.c_double(0), C.c_int64(0)
        K.check(K.lib().b2_ctx_profile(self._h, C.byref(ms), C.byref(n)))
        return ms.value, n.value

    def read_csv(self, text, schema, projection=None, delimiter="|", has_header=False):
        """Decode delimiter-separated text (bytes of whole rows, e.g. one dbgen .tbl partition file) on the GPU into a
        DeviceBatch of the projected columns: the CsvExec leaf's decode (from_proto.rs:93-110)."""
        sch = _export_schema(schema)
        proj = list(projection) if projection is not None else list(range(len(schema.names)))
        proj = [schema.names.index(p) if isinstance(p, str) else int(p) for p in proj]
        arr = (C.c_int * max(len(proj), 1))(*proj)
        h = C.c_void_p()
        data = bytes(text)
        try:
            K.check(K.lib().b2_csv_decode(self._h, C.byref(sch), arr, len(proj), delimiter.encode()[:1], int(has_header),
                                          data, len(data), C.byref(h)))
        finally:
            _release_schema(sch)
        return DeviceBatch(self, h)

    # ---- HBM-resident batches ----
    def upload(self, batch):
        """pyarrow RecordBatch (or Table with one chunk) -> DeviceBatch."""
        if isinstance(batch, pa.Table):
            batch = batch.combine_chunks().to_batches()[0] if batch.num_rows else pa.RecordBatch.from_pylist([], schema=batch.schema)
        arr, sch = K.ArrowArray(), K.ArrowSchema()
        batch._export_to_c(_addr(arr), _addr(sch))
        h = C.c_void_p()
        try:
            K.check(K.lib().b2_batch_from_arrow(self._h, C.byref(sch), C.byref(arr), C.byref(h)))
        finally:
            _release_schema(sch)
            _release_array(arr)
        return DeviceBatch(self, h)

    def tpch(self, table, sf, row_begin=0, row_end=None, columns=None):
        """Synthetic TPC-H-shaped table rows [row_begin,row_end) generated directly in HBM."""
        if row_end is None:
            row_end = K.lib().b2_tpch_num_rows(table.encode(), float(sf))
        cols = None
        if columns is not None:
            raw = [c.encode() for c in columns] + [None]
            cols = (C.c_char_p * len(raw))(*raw)
        h = C.c_void_p()
        K.check(K.lib().b2_tpch_generate(self._h, table.encode(), float(sf), row_begin, row_end, cols, C.byref(h)))
        return DeviceBatch(self, h)

    # ---- multi-GPU ----
    @staticmethod
    def comm_unique_id():
        buf = C.create_string_buffer(128)
        K.check(K.lib().b2_comm_unique_id(buf))
        return bytes(buf.raw)

    def comm_init(self, rank, nranks, unique_id):
        K.check(K.lib().b2_comm_init(self._h, rank, nranks, C.c_char_p(unique_id)))
        self.rank, self.nranks = rank, nranks

    def allgather(self, batch):
        h = C.c_void_p()
        K.check(K.lib().b2_comm_allgather_batch(self._h, batch._h, C.byref(h)))
        return DeviceBatch(self, h)

    def comm_bytes_sent(self):
        return K.lib().b2_comm_bytes_sent(self._h)


def _release_schema(s):
    if s.release:
        C.CFUNCTYPE(None, C.POINTER(K.ArrowSchema))(s.release)(C.byref(s))


def _release_array(a):
    if a.release:
        C.CFUNCTYPE(None, C.POINTER(K.ArrowArray))(a.release)(C.byref(a))


class DeviceBatch:
    """A RecordBatch resident in HBM (b2_batch)."""

    def __init__(self, ctx, handle):
        self.ctx, self._h = ctx, handle

    def __del__(self):
        try:
            if self._h:
                K.lib().b2_batch_destroy(self._h)
                self._h = None
        except Exception:
            pass

    @property
    def num_rows(self):
        return K.lib().b2_batch_num_rows(self._h)

    @property
    def num_bytes(self):
        return K.lib().b2_batch_num_bytes(self._h)

    @property
    def schema(self):
        sch = K.ArrowSchema()
        K.check(K.lib().b2_batch_schema(self._h, C.byref(sch)))
        return pa.Schema._import_from_c(_addr(sch))

    def to_arrow(self):
        arr, sch = K.ArrowArray(), K.ArrowSchema()
        K.check(K.lib().b2_batch_schema(self._h, C.byref(sch)))
        K.check(K.lib().b2_batch_to_arrow(self._h, C.byref(arr)))
        return pa.RecordBatch._import_from_c(_addr(arr), _addr(sch))


class RecordBatchStream:
    """SendableRecordBatchStream: iterate host RecordBatches, or `.device()` for HBM-resident ones."""

    def __init__(self, plan, handle):
        self.plan, self._h = plan, handle

    def __del__(self):
        try:
            if self._h:
                K.lib().b2_stream_destroy(self._h)
                self._h = None
        except Exception:
            pass

    def schema(self):
        return self.plan.schema()

    def __iter__(self):
        schema = self.plan.schema()
        lib = K.lib()
        while True:
            arr, eos = K.ArrowArray(), C.c_int(0)
            K.check(lib.b2_stream_next(self._h, C.byref(arr), C.byref(eos)))
            if eos.value:
                return
            yield pa.RecordBatch._import_from_c(_addr(arr), schema)

    def device(self):
        while True:
            h, eos = C.c_void_p(), C.c_int(0)
            K.check(K.lib().b2_stream_next_device(self._h, C.byref(h), C.byref(eos)))
            if eos.value:
                return
            yield DeviceBatch(self.plan.ctx, h)


class ExecutionPlan:
    """Arc<dyn ExecutionPlan> backed by a b2_op."""

    def __init__(self, ctx, handle, children=()):
        self.ctx, self._h, self._children = ctx, handle, list(children)

    def __del__(self):
        try:
            if self._h:
                K.lib().b2_op_destroy(self._h)
                self._h = None
        except Exception:
            pass

    def schema(self):
        # an operator's schema never changes: one trip through the C Data Interface, then the cached object
        s = getattr(self, "_schema", None)
        if s is None:
            sch = K.ArrowSchema()
            K.check(K.lib().b2_op_schema(self._h, C.byref(sch)))
            s = self._schema = pa.Schema._import_from_c(_addr(sch))
        return s

    def output_partitioning(self):
        scheme, n = C.c_int(0), C.c_int(0)
        K.check(K.lib().b2_op_output_partitioning(self._h, C.byref(scheme), C.byref(n)))
        return {K.PART_UNKNOWN: "UnknownPartitioning", K.PART_ROUND_ROBIN: "RoundRobinBatch",
                K.PART_HASH: "Hash"}[scheme.value], n.value

    def children(self):
        return list(self._children)

    def name(self):
        return K.lib().b2_op_name(self._h).decode()

    def last_path(self):
        return K.lib().b2_op_last_path(self._h).decode()

    def execute(self, partition):
        h = C.c_void_p()
        K.check(K.lib().b2_op_execute(self._h, partition, C.byref(h)))
        return RecordBatchStream(self, h)

    # conveniences (utils::collect_stream, rust/core/src/utils.rs:86-94)
    def collect(self, partition=None):
        parts = range(self.output_partitioning()[1]) if partition is None else [partition]
        batches = [b for p in parts for b in self.execute(p)]
        return pa.Table.from_batches(batches, schema=self.schema())

    def collect_device(self, partition=0):
        return list(self.execute(partition).device())

    def __repr__(self):
        return f"{self.name()}({', '.join(repr(c) for c in self._children)})"


def _names(plan_or_schema):
    s = plan_or_schema.schema() if isinstance(plan_or_schema, ExecutionPlan) else plan_or_schema
    return list(s.names)


def _export_schema(schema):
    sch = K.ArrowSchema()
    schema._export_to_c(_addr(sch))
    return sch


class MemoryExec(ExecutionPlan):
    """Leaf over in-memory partitions of host RecordBatches and/or DeviceBatches."""

    def __init__(self, ctx, schema, partitions):
        h = C.c_void_p()
        sch = _export_schema(schema)
        try:
            K.check(K.lib().b2_memory_exec_create(ctx._h, C.byref(sch), len(partitions), C.byref(h)))
        finally:
            _release_schema(sch)
        super().__init__(ctx, h)
        self._keep = []
        for p, batches in enumerate(partitions):
            for b in batches:
                self.add(p, b)

    def clear(self):
        """Drop every batch of every partition (b2_memory_exec_clear): the plans above recompute on their next execute."""
        K.check(K.lib().b2_memory_exec_clear(self._h))
        self._keep = []

    def add(self, partition, b):
        if isinstance(b, DeviceBatch):
            K.check(K.lib().b2_memory_exec_add_device_batch(self._h, partition, b._h))
            self._keep.append(b)
        else:
            arr, sch = K.ArrowArray(), K.ArrowSchema()
            b._export_to_c(_addr(arr), _addr(sch))
            _release_schema(sch)
            try:
                K.check(K.lib().b2_memory_exec_add_batch(self._h, partition, C.byref(arr)))
            finally:
                _release_array(arr)  # no-op when the library took ownership (release == NULL)


class FilterExec(ExecutionPlan):
    def __init__(self, predicate, input):
        keep, h = [], C.c_void_p()
        e = predicate.lower(_names(input), keep)
        K.check(K.lib().b2_filter_create(input.ctx._h, input._h, C.byref(e), C.byref(h)))
        super().__init__(input.ctx, h, [input])


class CoalesceBatchesExec(ExecutionPlan):
    def __init__(self, input, target_batch_size=32768):
        h = C.c_void_p()
        K.check(K.lib().b2_coalesce_batches_create(input.ctx._h, input._h, target_batch_size, C.byref(h)))
        super().__init__(input.ctx, h, [input])


class MergeExec(ExecutionPlan):
    def __init__(self, input):
        h = C.c_void_p()
        K.check(K.lib().b2_merge_create(input.ctx._h, input._h, C.byref(h)))
        super().__init__(input.ctx, h, [input])


class ProjectionExec(ExecutionPlan):
    def __init__(self, exprs, input):
        keep, h = [], C.c_void_p()
        names = _names(input)
        lowered = [(e if isinstance(e, Expr) else e[0].alias(e[1])).lower(names, keep) for e in exprs]
        arr = (K.Expr * len(lowered))(*lowered)
        K.check(K.lib().b2_projection_create(input.ctx._h, input._h, arr, len(lowered), C.byref(h)))
        super().__init__(input.ctx, h, [input])


def plan_from_proto(ctx, data, leaves):
    """Serialized `PhysicalPlanNode` (ballista.proto:294-422) -> ExecutionPlan, as the executor rebuilds a task's plan
    in from_proto.rs:58-346.  `leaves`: the operators bound, in pre-order, to the plan's scan / shuffle-reader nodes."""
    h = C.c_void_p()
    arr = (C.c_void_p * max(len(leaves), 1))(*[l._h for l in leaves])
    K.check(K.lib().b2_plan_from_proto(ctx._h, bytes(data), len(data), arr, len(leaves), C.byref(h)))
    return ExecutionPlan(ctx, h, list(leaves))


def describe_plan_proto(data):
    """(pre-order operator names, number of leaf nodes) of a serialized PhysicalPlanNode; needs no GPU."""
    buf, n = C.create_string_buffer(1 << 16), C.c_int(0)
    K.check(K.lib().b2_plan_proto_describe(bytes(data), len(data), buf, len(buf), C.byref(n)))
    return [l.strip() for l in buf.value.decode().splitlines()], n.value


class SortExec(ExecutionPlan):
    """SortExec::try_new(exprs, input); `exprs` = [expr | (expr, ascending)] (from_proto.rs:291-330)."""

    def __init__(self, exprs, input):
        keep, h = [], C.c_void_p()
        names = _names(input)
        pairs = [(e, True) if isinstance(e, Expr) else (e[0], bool(e[1])) for e in exprs]
        lowered = [e.lower(names, keep) for e, _ in pairs]
        arr = (K.Expr * len(lowered))(*lowered)
        desc = (C.c_int * len(pairs))(*[0 if asc else 1 for _, asc in pairs])
        nulls = (C.c_int * len(pairs))(*[0] * len(pairs))
        K.check(K.lib().b2_sort_create(input.ctx._h, input._h, arr, desc, nulls, len(lowered), C.byref(h)))
        super().__init__(input.ctx, h, [input])


class GlobalLimitExec(ExecutionPlan):
    def __init__(self, input, limit):
        h = C.c_void_p()
        K.check(K.lib().b2_limit_create(input.ctx._h, input._h, int(limit), 1, C.byref(h)))
        super().__init__(input.ctx, h, [input])


class LocalLimitExec(ExecutionPlan):
    def __init__(self, input, limit):
        h = C.c_void_p()
        K.check(K.lib().b2_limit_create(input.ctx._h, input._h, int(limit), 0, C.byref(h)))
        super().__init__(input.ctx, h, [input])


class HashAggregateExec(ExecutionPlan):
    """HashAggregateExec::try_new(mode, group_expr, aggr_expr, input, input_schema)."""

    def __init__(self, mode, group_exprs, aggr_exprs, input, input_schema=None):
        keep, h = [], C.c_void_p()
        mode_i = {"Partial": K.AGG_PARTIAL, "Final": K.AGG_FINAL}.get(mode, mode)
        in_names = _names(input)
        arg_schema = input_schema if input_schema is not None else input.schema()
        arg_names = list(arg_schema.names)
        groups = [g.lower(in_names, keep) for g in group_exprs]
        aggs = []
        for k, a in enumerate(aggr_exprs):
            assert isinstance(a, AggExpr)
            s = K.Agg()
            s.fn = a.fn
            s.arg = a.arg.lower(arg_names, keep, default_name=a.name or f"agg{k}")
            if a.name:
                nb = a.name.encode()
                keep.append(nb)
                s.arg.name = nb
            aggs.append(s)
        garr = (K.Expr * max(len(groups), 1))(*groups)
        aarr = (K.Agg * max(len(aggs), 1))(*aggs)
        sch = _export_schema(arg_schema)
        try:
            K.check(K.lib().b2_hash_aggregate_create(input.ctx._h, input._h, mode_i, garr, len(groups), aarr,
                                                     len(aggs), C.byref(sch), C.byref(h)))
        finally:
            _release_schema(sch)
        super().__init__(input.ctx, h, [input])
        self.mode = "Partial" if mode_i == K.AGG_PARTIAL else "Final"


class HashJoinExec(ExecutionPlan):
    """HashJoinExec::try_new(left, right, &on, &join_type); `on` = [(left_name, right_name)]."""

    def __init__(self, left, right, on, join_type="Inner"):
        h = C.c_void_p()
        jt = {"Inner": K.JOIN_INNER, "Left": K.JOIN_LEFT, "Right": K.JOIN_RIGHT}.get(join_type, join_type)
        l = (C.c_char_p * len(on))(*[a.encode() for a, _ in on])
        r = (C.c_char_p * len(on))(*[b.encode() for _, b in on])
        K.check(K.lib().b2_hash_join_create(left.ctx._h, left._h, right._h, l, r, len(on), jt, C.byref(h)))
        super().__init__(left.ctx, h, [left, right])


class RepartitionExec(ExecutionPlan):
    """RepartitionExec::try_new(input, Partitioning::Hash(exprs, n) | RoundRobinBatch(n))."""

    def __init__(self, input, partitioning, n, exprs=()):
        keep, h = [], C.c_void_p()
        scheme = {"Hash": K.PART_HASH, "RoundRobinBatch": K.PART_ROUND_ROBIN}.get(partitioning, partitioning)
        names = _names(input)
        lowered = [e.lower(names, keep) for e in exprs]
        arr = (K.Expr * max(len(lowered), 1))(*lowered)
        K.check(K.lib().b2_repartition_create(input.ctx._h, input._h, scheme, arr, len(lowered), n, C.byref(h)))
        super().__init__(input.ctx, h, [input])
