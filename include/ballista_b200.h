/*
 * ballista_b200.h — C ABI of libballista_b200: B200-native (sm_100a) replacements for the executor-side
 * columnar operators Ballista instantiates from a wire plan.
 *
 * This is the drop-in boundary a Rust `extern "C"` FFI crate would bind (INTEGRATION.md shows the shim).
 * Each entry point cites the reference interface it replaces, relative to the reference repository root.
 * The shape mirrors DataFusion's `ExecutionPlan` trait as the reference's own operators implement it
 * (rust/core/src/execution_plans/query_stage.rs:50-84, shuffle_reader.rs:51-99):
 *
 *     schema() / output_partitioning() / children() / execute(partition) -> RecordBatchStream
 *
 * Data crosses the boundary as Arrow C Data Interface structs (one struct-typed ArrowArray == one
 * RecordBatch), or stays resident in HBM as an opaque `b2_batch`.
 *
 * Rules:
 *   - every function returns 0 (B2_OK) or a negative B2_ERR_* code; the message is b2_last_error()
 *     (thread-local).  No exceptions, no abort() cross the ABI.
 *   - there is NO CPU fallback: without a CUDA device b2_ctx_create fails with B2_ERR_CUDA.
 *   - a b2_ctx owns one GPU, one stream and one stream-ordered memory pool.  A ctx and everything created
 *     from it may be driven by one thread at a time (Send, not Sync).
 *   - ownership: "moved" ArrowArray arguments are released by the library; ArrowArray / ArrowSchema
 *     outputs carry a release callback the caller must invoke; handles are destroyed by the caller with
 *     the matching *_destroy (NULL is accepted).
 */
#ifndef BALLISTA_B200_H
#define BALLISTA_B200_H

#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- Arrow C Data Interface (verbatim ABI from the Arrow specification) ---- */
#ifndef ARROW_C_DATA_INTERFACE
#define ARROW_C_DATA_INTERFACE
#define ARROW_FLAG_DICTIONARY_ORDERED 1
#define ARROW_FLAG_NULLABLE 2
#define ARROW_FLAG_MAP_KEYS_SORTED 4
struct ArrowSchema {
  const char* format;
  const char* name;
  const char* metadata;
  int64_t flags;
  int64_t n_children;
  struct ArrowSchema** children;
  struct ArrowSchema* dictionary;
  void (*release)(struct ArrowSchema*);
  void* private_data;
};
struct ArrowArray {
  int64_t length;
  int64_t null_count;
  int64_t offset;
  int64_t n_buffers;
  int64_t n_children;
  const void** buffers;
  struct ArrowArray** children;
  struct ArrowArray* dictionary;
  void (*release)(struct ArrowArray*);
  void* private_data;
};
#endif

/* ---- status codes (map onto BallistaError / DataFusionError, rust/core/src/error.rs:31-47) ---- */
enum {
  B2_OK = 0,
  B2_ERR_INVALID_ARG = -1,     /* BallistaError::General */
  B2_ERR_NOT_IMPLEMENTED = -2, /* BallistaError::NotImplemented / DataFusionError::NotImplemented */
  B2_ERR_EXECUTION = -3,       /* DataFusionError::Execution */
  B2_ERR_ARROW = -4,           /* ArrowError */
  B2_ERR_CUDA = -5,            /* BallistaError::Internal */
  B2_ERR_NCCL = -6,            /* replaces tonic/Flight transport errors (error.rs:41-43) */
  B2_ERR_OOM = -7
};

typedef struct b2_ctx b2_ctx;       /* one GPU + stream + memory pool                                     */
typedef struct b2_op b2_op;         /* a physical operator node  (Arc<dyn ExecutionPlan>)                  */
typedef struct b2_stream b2_stream; /* result of execute(partition) (SendableRecordBatchStream)            */
typedef struct b2_batch b2_batch;   /* a RecordBatch resident in HBM                                       */

const char* b2_last_error(void);
const char* b2_version(void);

/* ---- context ---- */
int b2_ctx_create(int device, b2_ctx** out);
void b2_ctx_destroy(b2_ctx* ctx);
/* the cudaStream_t all work of this ctx is enqueued on (for CUDA-event timing by the caller) */
void* b2_ctx_cuda_stream(b2_ctx* ctx);
int b2_ctx_synchronize(b2_ctx* ctx);
/* number of kernels this ctx has launched so far (bench.py's gpu_launches) */
int64_t b2_ctx_kernel_launches(b2_ctx* ctx);
/* Per-launch CUDA-event timing of the aggregate scan kernels (the dominant kernel of Q1/Q6/Q3), recorded
 * on the ctx stream.  b2_ctx_profile synchronises, returns the summed duration and launch count since the
 * last call, and resets. */
int b2_ctx_set_profiling(b2_ctx* ctx, int on);
int b2_ctx_profile(b2_ctx* ctx, double* ms_total, int64_t* launches);
/* pinned host memory for Arrow buffers that will be pushed (so H2D runs at PCIe speed) */
int b2_host_alloc(b2_ctx* ctx, size_t nbytes, void** out);
int b2_host_free(b2_ctx* ctx, void* p);

/* ---- expression IR: flat post-order array mirroring LogicalExprNode (rust/core/proto/ballista.proto:14-45)
 *      and the binary-operator table (rust/core/src/serde/logical_plan/from_proto.rs:937-957).
 *      Children of a node are the nodes immediately preceding it (stack machine order). ---- */
enum b2_expr_kind {
  B2_EXPR_COLUMN = 1,  /* column_name        : `col` = index into the input schema                        */
  B2_EXPR_LITERAL = 2, /* ScalarValue        : dtype + i64/f64/str                                        */
  B2_EXPR_BINARY = 3,  /* BinaryExprNode     : op = b2_binary_op, 2 children                              */
  B2_EXPR_NOT = 4,     /* Not                : 1 child                                                    */
  B2_EXPR_IS_NULL = 5,
  B2_EXPR_IS_NOT_NULL = 6,
  B2_EXPR_CAST = 7,     /* CastNode          : dtype = target, 1 child                                    */
  B2_EXPR_NEGATIVE = 8, /* NegativeNode      : 1 child                                                    */
  B2_EXPR_BETWEEN = 9,  /* BetweenNode       : 3 children (expr, low, high); op!=0 => NOT BETWEEN         */
  B2_EXPR_IN_LIST = 10, /* InListNode        : nchildren = 1 + list length; op!=0 => NOT IN               */
  B2_EXPR_CASE = 11,    /* CaseNode          : nchildren children = [base] (when, then)* [else];
                                               op bit 0 = base expression present, bit 1 = ELSE present        */
  B2_EXPR_SCALAR_FN = 12 /* ScalarFunctionNode: op = ScalarFunction enum value (ballista.proto:80-114),
                                               nchildren arguments                                             */
};
enum b2_binary_op {
  B2_OP_AND = 1, B2_OP_OR, B2_OP_EQ, B2_OP_NOTEQ, B2_OP_LTEQ, B2_OP_LT, B2_OP_GT, B2_OP_GTEQ,
  B2_OP_PLUS, B2_OP_MINUS, B2_OP_MULTIPLY, B2_OP_DIVIDE, B2_OP_LIKE
};
enum b2_dtype { /* the Arrow types that occur on the TPC-H path (rust/benchmarks/tpch/src/main.rs:267-360) */
  B2_BOOL = 1, B2_INT32 = 2, B2_INT64 = 3, B2_UINT64 = 4, B2_FLOAT64 = 5, B2_DATE32 = 6, B2_UTF8 = 7,
  B2_UINT8 = 8
};
typedef struct b2_expr_node {
  int32_t kind;      /* b2_expr_kind                      */
  int32_t op;        /* b2_binary_op / negation flag      */
  int32_t dtype;     /* b2_dtype for LITERAL and CAST     */
  int32_t col;       /* COLUMN: input-schema index        */
  int32_t nchildren; /* IN_LIST / CASE / SCALAR_FN; others are implied */
  int32_t reserved;
  int64_t i;         /* integer / date32 literal          */
  double f;          /* Float64 literal                   */
  const char* s;     /* Utf8 literal (NUL-terminated)     */
} b2_expr_node;
typedef struct b2_expr {
  const b2_expr_node* nodes;
  int32_t n_nodes;
  const char* name; /* output field name (group_expr_name / aggr_expr_name / expr_name) */
} b2_expr;

/* AggregateExprNode vocabulary that can cross the Ballista wire: rust/core/src/serde/physical_plan/
 * to_proto.rs:352-363 (Avg, Sum, Count only; always non-distinct, from_proto.rs:230-236). */
enum b2_agg_fn { B2_AGG_SUM = 1, B2_AGG_AVG = 2, B2_AGG_COUNT = 3 };
typedef struct b2_agg {
  int32_t fn;   /* b2_agg_fn */
  b2_expr arg;  /* argument; COUNT(*) arrives as COUNT(UInt8(1)) literal  */
} b2_agg;

enum b2_agg_mode { B2_AGG_PARTIAL = 0, B2_AGG_FINAL = 1 }; /* ballista.proto:370-373 */
enum b2_join_type { B2_JOIN_INNER = 0, B2_JOIN_LEFT = 1, B2_JOIN_RIGHT = 2 }; /* ballista.proto:266-270 */
enum b2_partitioning { /* to_proto.rs:279-295 */
  B2_PART_UNKNOWN = 0, B2_PART_ROUND_ROBIN = 1, B2_PART_HASH = 2
};

/* ---- plan construction: one function per from_proto.rs match arm on the hot path ---- */

/* Leaf holding in-memory partitions (stands where CsvExec/ParquetExec/ShuffleReaderExec feed the path,
 * from_proto.rs:93-121, shuffle_reader.rs:77-99; same role as rust/core/src/memory_stream.rs:29-93). */
int b2_memory_exec_create(b2_ctx* ctx, const struct ArrowSchema* schema, int n_partitions, b2_op** out);
/* append a host RecordBatch (moved; copied host->device when the partition is executed) */
int b2_memory_exec_add_batch(b2_op* op, int partition, struct ArrowArray* batch);
/* append an HBM-resident batch (shared: the caller keeps its handle) */
int b2_memory_exec_add_device_batch(b2_op* op, int partition, b2_batch* batch);
int b2_memory_exec_clear(b2_op* op);

/* FilterExec::try_new(predicate, input)                       from_proto.rs:81-92 */
int b2_filter_create(b2_ctx* ctx, b2_op* input, const b2_expr* predicate, b2_op** out);
/* CoalesceBatchesExec::new(input, target_batch_size)          from_proto.rs:122-128 */
int b2_coalesce_batches_create(b2_ctx* ctx, b2_op* input, int64_t target_batch_size, b2_op** out);
/* MergeExec::new(input)                                       from_proto.rs:129-132 */
int b2_merge_create(b2_ctx* ctx, b2_op* input, b2_op** out);
/* ProjectionExec::try_new(exprs, input)                       from_proto.rs:60-80 */
int b2_projection_create(b2_ctx* ctx, b2_op* input, const b2_expr* exprs, int n_exprs, b2_op** out);
/* Column decode of the CsvExec leaf (from_proto.rs:93-110: CsvReadOptions{has_header, delimiter, schema} + projection)
 * on the GPU: `text` = `len` bytes of whole rows of a delimiter-separated file (dbgen .tbl: '|', no quoting, trailing
 * delimiter allowed) in host memory; the result holds the projected columns of `file_schema` in HBM.  Int32, Int64,
 * Float64 (plain decimals of <= 15 significant digits, converted with one correctly rounded division), Date32
 * (YYYY-MM-DD) and Utf8 columns; malformed input is B2_ERR_ARROW with the line number. */
int b2_csv_decode(b2_ctx* ctx, const struct ArrowSchema* file_schema, const int* projection, int n_projection,
                  char delimiter, int has_header, const char* text, size_t len, b2_batch** out);
/* Column decode of the ParquetExec leaf (from_proto.rs:111-121: ParquetExec::try_from_files(filenames, projection, None,
 * batch_size, num_partitions); the benchmark's `--format parquet`) on the GPU.  `file` = the `len` bytes of one whole
 * Parquet file in host memory (an mmap is fine: only the footer, the page headers and the projected column chunks are
 * touched).  Flat schemas; BOOLEAN, INT32 (Int32 / Date32), INT64 (Int64 / UInt64), DOUBLE and BYTE_ARRAY (Utf8)
 * columns, REQUIRED or OPTIONAL; data pages V1 and V2; PLAIN, PLAIN_DICTIONARY / RLE_DICTIONARY (and RLE booleans);
 * UNCOMPRESSED and SNAPPY.  Anything else is B2_ERR_NOT_IMPLEMENTED naming the feature, a malformed file B2_ERR_ARROW.
 *   b2_parquet_describe: no device needed; *num_rows / *n_row_groups / *n_columns (each optional) and a text listing of
 *                        columns, row groups, codecs, encodings and page counts into `text` (optional, NUL-terminated).
 *   b2_parquet_schema:   Arrow schema of the projected columns (projection NULL = all), names and nullability as in
 *                        the file (DataFusion's ParquetExec schema).
 *   b2_parquet_decode:   row groups [row_group_begin, row_group_end) (end < 0 = to the last) of the projected columns
 *                        as ONE batch in HBM; the file bytes may be released as soon as it returns. */
int b2_parquet_describe(const uint8_t* file, size_t len, int64_t* num_rows, int* n_row_groups, int* n_columns,
                        char* text, size_t cap);
int b2_parquet_schema(const uint8_t* file, size_t len, const int* projection, int n_projection,
                      struct ArrowSchema* out);
int b2_parquet_decode(b2_ctx* ctx, const uint8_t* file, size_t len, const int* projection, int n_projection,
                      int row_group_begin, int row_group_end, b2_batch** out);
/* Wire compatibility: a serialized `PhysicalPlanNode` (rust/core/proto/ballista.proto:294-422, as the scheduler
 * stores it, rust/scheduler/src/state/mod.rs:149-162) decoded into the operator tree the way the executor does in
 * rust/core/src/serde/physical_plan/from_proto.rs:58-346.  Leaf nodes (CsvScan, ParquetScan, ShuffleReader,
 * UnresolvedShuffle, Empty) are not turned into readers: the i-th leaf in pre-order is bound to leaves[i]
 * (a MemoryExec fed by the CPU reader or by the previous stage).  The handles in `leaves` stay owned by the caller. */
int b2_plan_from_proto(b2_ctx* ctx, const uint8_t* buf, size_t len, b2_op* const* leaves, int n_leaves, b2_op** out);
/* The same decoder without a device: writes the pre-order operator list (one node per line, indented) into `out`
 * (NUL-terminated, truncated to `cap`) and the number of leaf nodes into *n_leaves.  For diagnostics and for
 * sizing the `leaves` array. */
int b2_plan_proto_describe(const uint8_t* buf, size_t len, char* out, size_t cap, int* n_leaves);
/* Run-time compiled scan kernels.  compile_expr accepts any expression (from_proto.rs:348-364); a HashAggregateExec whose
 * (predicate, group keys, aggregate arguments) shape is not among the pre-built fused kernels is compiled for sm_100a on
 * first use (NVRTC) and cached.  This entry point compiles the shape named by `signature` (the text b2_op_last_path shows
 * after "fused:", e.g. "p:lt(i0,LI0)|k:s0|a:f0,mul(f0,f1)") WITHOUT loading it - it needs no device - and returns the
 * size of the cubin in *cubin_bytes; B2_ERR_NOT_IMPLEMENTED when the shape is outside the fused vocabulary. */
int b2_jit_compile_check(const char* signature, size_t* cubin_bytes);
/* SortExec::try_new(Vec<PhysicalSortExpr{expr, SortOptions{descending, nulls_first}}>, input)  from_proto.rs:291-330.
 * Single input partition, one sorted output batch, stable.  `descending[i]` != 0 sorts expression i descending;
 * nulls_first is accepted for interface parity (columns on this path are non-null). */
int b2_sort_create(b2_ctx* ctx, b2_op* input, const b2_expr* exprs, const int* descending, const int* nulls_first,
                   int n_exprs, b2_op** out);
/* GlobalLimitExec::new(input, limit) (global != 0) / LocalLimitExec::new(input, limit)   from_proto.rs:165-171 */
int b2_limit_create(b2_ctx* ctx, b2_op* input, int64_t limit, int global, b2_op** out);
/* HashAggregateExec::try_new(mode, group_expr, aggr_expr, input, input_schema)  from_proto.rs:173-252.
 * Partial emits accumulator-state columns "<name>[sum]" / "<name>[count]" (SURVEY Appendix A.2);
 * Final reads them positionally after the group columns and needs `input_schema` = the schema of the
 * Partial's input (HashAggregateExecNode.input_schema, ballista.proto:382-383) to type the aggregates. */
int b2_hash_aggregate_create(b2_ctx* ctx, b2_op* input, int mode, const b2_expr* group_exprs, int n_groups,
                             const b2_agg* aggs, int n_aggs, const struct ArrowSchema* input_schema,
                             b2_op** out);
/* HashJoinExec::try_new(left, right, &on, &join_type)         from_proto.rs:253-276 (collect-left build) */
int b2_hash_join_create(b2_ctx* ctx, b2_op* left, b2_op* right, const char* const* on_left,
                        const char* const* on_right, int n_on, int join_type, b2_op** out);
/* RepartitionExec::try_new(input, Partitioning::{Hash(exprs,n)|RoundRobinBatch(n)})  from_proto.rs:133-164.
 * With a communicator attached (b2_comm_init) and n_out == nranks, Hash repartitioning is the in-box
 * shuffle: execute(p) scatters local rows into per-destination segments and exchanges them with an
 * all-to-all over NVLink (replaces utils.rs:49-84 + shuffle_reader.rs:77-99 + flight_service.rs:193-228). */
int b2_repartition_create(b2_ctx* ctx, b2_op* input, int scheme, const b2_expr* keys, int n_keys, int n_out,
                          b2_op** out);

/* ---- the ExecutionPlan trait ---- */
int b2_op_schema(b2_op* op, struct ArrowSchema* out);                   /* ExecutionPlan::schema              */
int b2_op_output_partitioning(b2_op* op, int* scheme, int* n_partitions); /* ::output_partitioning            */
int b2_op_n_children(b2_op* op);
b2_op* b2_op_child(b2_op* op, int i); /* ::children — a new handle sharing the node; b2_op_destroy it */
const char* b2_op_name(b2_op* op);                                       /* the operator's Debug name          */
/* ::execute(partition)  (called at rust/executor/src/flight_service.rs:117-121) */
int b2_op_execute(b2_op* op, int partition, b2_stream** out);
void b2_op_destroy(b2_op* op); /* drops this node's reference; children stay alive while referenced */
/* How the last execute() of this node ran: "fused:<kernel signature>" or "unfused" (diagnostics). */
const char* b2_op_last_path(b2_op* op);

/* ---- RecordBatchStream ---- */
/* next host RecordBatch (device->host copy inside); *eos=1 and out untouched at end of stream */
int b2_stream_next(b2_stream* s, struct ArrowArray* out, int* eos);
/* next HBM-resident batch (caller owns the returned handle) */
int b2_stream_next_device(b2_stream* s, b2_batch** out, int* eos);
void b2_stream_destroy(b2_stream* s);

/* ---- HBM-resident batches ---- */
int b2_batch_from_arrow(b2_ctx* ctx, const struct ArrowSchema* schema, struct ArrowArray* batch /*moved*/,
                        b2_batch** out);
int b2_batch_to_arrow(b2_batch* b, struct ArrowArray* out);
int b2_batch_schema(b2_batch* b, struct ArrowSchema* out);
int64_t b2_batch_num_rows(b2_batch* b);
int64_t b2_batch_num_bytes(b2_batch* b); /* PartitionStats.num_bytes analogue, utils.rs:79-83 */
void b2_batch_destroy(b2_batch* b);

/* ---- synthetic TPC-H-shaped tables generated directly in HBM (workload of rust/benchmarks/tpch;
 *      schemas rust/benchmarks/tpch/src/main.rs:267-360; generator spec in DESIGN.md) ----
 * table: "lineitem" | "orders" | "customer"; rows are the table's rows [row_begin,row_end) at `sf`;
 * columns: NULL = all generated columns, else a NULL-terminated list of column names. */
int b2_tpch_generate(b2_ctx* ctx, const char* table, double sf, int64_t row_begin, int64_t row_end,
                     const char* const* columns, b2_batch** out);
int64_t b2_tpch_num_rows(const char* table, double sf);

/* ---- multi-GPU: one process (ctx) per GPU ---- */
/* size of the opaque communicator id; rank 0 fills it with b2_comm_unique_id and the host plumbing
 * (torch.distributed in this repo) broadcasts it */
int b2_comm_unique_id(void* id128 /* 128 bytes */);
int b2_comm_init(b2_ctx* ctx, int rank, int nranks, const void* id128);
int b2_comm_destroy(b2_ctx* ctx);
/* All-gather of a small HBM batch over all ranks (merge step of Partial->Final, replaces the MergeExec
 * stage hop, rust/scheduler/src/planner.rs:136-148): out = concatenation in rank order. */
int b2_comm_allgather_batch(b2_ctx* ctx, b2_batch* local, b2_batch** out);
/* bytes this ctx sent over NVLink so far (shuffle roofline accounting) */
int64_t b2_comm_bytes_sent(b2_ctx* ctx);

/* ---- allocation-free entry point ----
 * No counterpart in the reference: added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * The hash split of RepartitionExec(Partitioning::Hash) over ONE Int32 / Date32 key column, with every intermediate
 * (destination byte per row, per-tile histogram) in a file-scope device array of the library instead of an
 * allocation: the call makes no cudaMalloc / cudaFree / allocating library call, creates no context object and does
 * not synchronise; it enqueues a memset and four kernels on `cuda_stream` of `device`.
 *   keys   [n_rows]      device, in : the key column's values
 *   perm   [n_rows]      device, out: row ids, destination-major; inside a destination the input order is kept
 *                                     (row perm[i] for starts[d] <= i < starts[d+1] goes to output partition d —
 *                                     the same rows in the same order as b2_repartition_create's partition d)
 *   starts [n_out + 1]   device, out: slice starts; starts[n_out] = n_rows
 * Limits: 0 <= n_rows <= 4,194,304 (2^22) and 1 <= n_out <= 32; beyond either the call returns B2_ERR_INVALID_ARG
 * and launches nothing.  The scratch array (about 4.5 MiB) is held by every GPU that loads the library until the
 * process exits.  All calls on one device share it: calls that overlap on the GPU (different streams, or replays of
 * two captured graphs at once) race and produce garbage; calls ordered on one stream are safe.  The first call on a
 * device looks the array's device address up (cudaGetSymbolAddress): make that call outside a graph capture. */
int b2_hash_split_i32_static(int device, void* cuda_stream, const int32_t* keys, int64_t n_rows, int n_out,
                             uint32_t* perm, int64_t* starts);

#ifdef __cplusplus
}
#endif
#endif /* BALLISTA_B200_H */
