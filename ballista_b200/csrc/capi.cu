// capi.cu — the extern "C" surface declared in include/ballista_b200.h.  Every entry point converts
// internal exceptions into status codes; nothing throws across the ABI.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>

#include "jit.h"
#include "ops.h"
#include "parquet.h"

using namespace b2;

static thread_local std::string g_last_error;

template <class F> static int guard(F&& f) {
  try {
    f();
    return B2_OK;
  } catch (const Error& e) {
    g_last_error = e.what();
    return e.code;
  } catch (const std::bad_alloc&) {
    g_last_error = "host allocation failed";
    return B2_ERR_OOM;
  } catch (const std::exception& e) {
    g_last_error = e.what();
    return B2_ERR_EXECUTION;
  } catch (...) {
    g_last_error = "unknown error";
    return B2_ERR_EXECUTION;
  }
}
#define REQUIRE(cond, msg) \
  if (!(cond)) fail(B2_ERR_INVALID_ARG, msg)

extern "C" {

const char* b2_last_error(void) { return g_last_error.c_str(); }
const char* b2_version(void) { return "ballista_b200 0.1 (sm_100a)"; }

int b2_ctx_create(int device, b2_ctx** out) {
  return guard([&] {
    REQUIRE(out, "out is null");
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0) {
      cudaGetLastError();
      fail(B2_ERR_CUDA, "no CUDA device available: this library has no CPU fallback");
    }
    REQUIRE(device >= 0 && device < n, "device index out of range");
    auto* h = new b2_ctx;
    h->c.device = device;
    try {
      B2_CUDA(cudaSetDevice(device));
      cudaDeviceProp prop;
      B2_CUDA(cudaGetDeviceProperties(&prop, device));
      if (prop.major < 10) fail(B2_ERR_CUDA, std::string("device '") + prop.name + "' is not sm_100-class");
      h->c.sm_count = prop.multiProcessorCount;
      h->c.trace = getenv("B2_TRACE") != nullptr;
      h->c.cache_off = getenv("B2_DISABLE_ALLOC_CACHE") != nullptr;
      B2_CUDA(cudaStreamCreateWithFlags(&h->c.stream, cudaStreamNonBlocking));
      B2_CUDA(cudaStreamCreateWithFlags(&h->c.copy_stream, cudaStreamNonBlocking));
      B2_CUDA(cudaDeviceGetDefaultMemPool(&h->c.pool, device));
      uint64_t threshold = UINT64_MAX;  // keep freed HBM in the pool: no cudaFree on the hot path
      B2_CUDA(cudaMemPoolSetAttribute(h->c.pool, cudaMemPoolAttrReleaseThreshold, &threshold));
    } catch (...) {
      delete h;
      throw;
    }
    *out = h;
  });
}
void b2_ctx_destroy(b2_ctx* ctx) {
  if (!ctx) return;
  cudaSetDevice(ctx->c.device);
  if (ctx->c.comm) {
    try { comm_destroy(&ctx->c); } catch (...) {}
  }
  cudaStreamSynchronize(ctx->c.stream);
  if (ctx->c.trace) {
    fprintf(stderr, "[b2 trace] %-60s %10s %8s\n", "site", "total ms", "calls");
    for (auto& kv : ctx->c.trace_acc)
      fprintf(stderr, "[b2 trace] %-60s %10.3f %8lld\n", kv.first.c_str(), kv.second.first * 1e3,
              (long long)kv.second.second);
  }
  ctx_trim_cache(&ctx->c);
  if (ctx->c.pinned_stage) cudaFreeHost(ctx->c.pinned_stage);
  cudaStreamDestroy(ctx->c.stream);
  cudaStreamDestroy(ctx->c.copy_stream);
  delete ctx;
}
void* b2_ctx_cuda_stream(b2_ctx* ctx) { return ctx ? (void*)ctx->c.stream : nullptr; }
int b2_ctx_synchronize(b2_ctx* ctx) {
  return guard([&] {
    REQUIRE(ctx, "ctx is null");
    B2_CUDA(cudaStreamSynchronize(ctx->c.stream));
  });
}
int64_t b2_ctx_kernel_launches(b2_ctx* ctx) { return ctx ? ctx->c.launches : 0; }
int b2_ctx_set_profiling(b2_ctx* ctx, int on) {
  return guard([&] {
    REQUIRE(ctx, "ctx is null");
    if (on == 2) {  // diagnostics: restart the B2_TRACE accumulators (after warm-up)
      ctx->c.trace_acc.clear();
      return;
    }
    ctx->c.profiling = on != 0;
  });
}
int b2_ctx_profile(b2_ctx* ctx, double* ms_total, int64_t* launches) {
  return guard([&] {
    REQUIRE(ctx && ms_total && launches, "null argument");
    B2_CUDA(cudaStreamSynchronize(ctx->c.stream));
    double total = 0;
    for (auto& ev : ctx->c.prof_events) {
      float ms = 0;
      if (cudaEventElapsedTime(&ms, ev.first, ev.second) == cudaSuccess) total += ms;
      cudaEventDestroy(ev.first);
      cudaEventDestroy(ev.second);
    }
    cudaGetLastError();
    *ms_total = total;
    *launches = (int64_t)ctx->c.prof_events.size();
    ctx->c.prof_events.clear();
  });
}
int b2_host_alloc(b2_ctx* ctx, size_t nbytes, void** out) {
  return guard([&] {
    REQUIRE(ctx && out, "null argument");
    ctx->c.activate();
    B2_CUDA(cudaHostAlloc(out, nbytes ? nbytes : 1, cudaHostAllocDefault));
  });
}
int b2_host_free(b2_ctx* ctx, void* p) {
  return guard([&] {
    REQUIRE(ctx, "ctx is null");
    if (p) B2_CUDA(cudaFreeHost(p));
  });
}

// ---- plan construction ----
static OpPtr op_of(b2_op* h) {
  REQUIRE(h && h->op, "operator handle is null");
  return h->op;
}
static b2_op* wrap(OpPtr op) {
  auto* h = new b2_op;
  h->op = std::move(op);
  return h;
}

int b2_memory_exec_create(b2_ctx* ctx, const struct ArrowSchema* schema, int n_partitions, b2_op** out) {
  return guard([&] {
    REQUIRE(ctx && schema && out, "null argument");
    *out = wrap(make_memory_exec(&ctx->c, schema_from_arrow(schema), n_partitions));
  });
}
int b2_memory_exec_add_batch(b2_op* op, int partition, struct ArrowArray* batch) {
  return guard([&] {
    REQUIRE(batch && batch->release, "batch is null or released");
    memory_exec_add_host(op_of(op).get(), partition, batch);
  });
}
int b2_memory_exec_add_device_batch(b2_op* op, int partition, b2_batch* batch) {
  return guard([&] {
    REQUIRE(batch && batch->b, "batch is null");
    memory_exec_add_device(op_of(op).get(), partition, batch->b);
  });
}
int b2_memory_exec_clear(b2_op* op) {
  return guard([&] { memory_exec_clear(op_of(op).get()); });
}

int b2_filter_create(b2_ctx* ctx, b2_op* input, const b2_expr* predicate, b2_op** out) {
  return guard([&] {
    REQUIRE(ctx && out && predicate, "null argument");
    OpPtr in = op_of(input);
    *out = wrap(make_filter(&ctx->c, in, decode_expr(predicate, in->schema())));
  });
}
int b2_coalesce_batches_create(b2_ctx* ctx, b2_op* input, int64_t target, b2_op** out) {
  return guard([&] {
    REQUIRE(ctx && out, "null argument");
    *out = wrap(make_coalesce(&ctx->c, op_of(input), target));
  });
}
int b2_merge_create(b2_ctx* ctx, b2_op* input, b2_op** out) {
  return guard([&] {
    REQUIRE(ctx && out, "null argument");
    *out = wrap(make_merge(&ctx->c, op_of(input)));
  });
}
int b2_projection_create(b2_ctx* ctx, b2_op* input, const b2_expr* exprs, int n_exprs, b2_op** out) {
  return guard([&] {
    REQUIRE(ctx && out && (exprs || n_exprs == 0), "null argument");
    OpPtr in = op_of(input);
    Schema s = in->schema();
    std::vector<ExprPtr> ev;
    for (int i = 0; i < n_exprs; ++i) ev.push_back(decode_expr(&exprs[i], s));
    *out = wrap(make_projection(&ctx->c, in, std::move(ev)));
  });
}
int b2_csv_decode(b2_ctx* ctx, const struct ArrowSchema* file_schema, const int* projection, int n_projection,
                  char delimiter, int has_header, const char* text, size_t len, b2_batch** out) {
  return guard([&] {
    REQUIRE(ctx && file_schema && out && (text || len == 0) && n_projection >= 0 && (projection || n_projection == 0),
            "null argument");
    ctx->c.activate();
    Schema fs = schema_from_arrow(file_schema);
    std::vector<int> proj(projection, projection + n_projection);
    if (proj.empty())
      for (size_t i = 0; i < fs.fields.size(); ++i) proj.push_back((int)i);
    auto* h = new b2_batch;
    h->b = csv_decode(&ctx->c, fs, proj, delimiter, has_header != 0, text, len);
    *out = h;
  });
}
int b2_parquet_describe(const uint8_t* file, size_t len, int64_t* num_rows, int* n_row_groups, int* n_columns,
                        char* text, size_t cap) {
  return guard([&] {
    REQUIRE(file, "null argument");
    pq::FileMeta m = pq::parse_footer(file, len);
    if (num_rows) *num_rows = m.num_rows;
    if (n_row_groups) *n_row_groups = (int)m.row_groups.size();
    if (n_columns) *n_columns = (int)m.columns.size();
    if (text && cap > 0) {
      std::string t = pq::describe(m, file, len);
      size_t n = t.size() < cap - 1 ? t.size() : cap - 1;
      memcpy(text, t.data(), n);
      text[n] = 0;
    }
  });
}
static std::vector<int> parquet_projection(const pq::FileMeta& m, const int* projection, int n_projection) {
  std::vector<int> proj;
  if (projection) proj.assign(projection, projection + n_projection);
  else
    for (size_t i = 0; i < m.columns.size(); ++i) proj.push_back((int)i);
  return proj;
}
int b2_parquet_schema(const uint8_t* file, size_t len, const int* projection, int n_projection, struct ArrowSchema* out) {
  return guard([&] {
    REQUIRE(file && out && n_projection >= 0, "null argument");
    pq::FileMeta m = pq::parse_footer(file, len);
    schema_to_arrow(pq::arrow_schema(m, parquet_projection(m, projection, n_projection)), out);
  });
}
int b2_parquet_decode(b2_ctx* ctx, const uint8_t* file, size_t len, const int* projection, int n_projection,
                      int row_group_begin, int row_group_end, b2_batch** out) {
  return guard([&] {
    REQUIRE(ctx && file && out && n_projection >= 0, "null argument");
    ctx->c.activate();
    std::vector<int> proj;
    if (projection) proj.assign(projection, projection + n_projection);
    else proj = parquet_projection(pq::parse_footer(file, len), nullptr, 0);
    auto* h = new b2_batch;
    try {
      h->b = parquet_decode(&ctx->c, file, len, proj, row_group_begin, row_group_end);
    } catch (...) {
      delete h;
      throw;
    }
    *out = h;
  });
}
int b2_plan_from_proto(b2_ctx* ctx, const uint8_t* buf, size_t len, b2_op* const* leaves, int n_leaves, b2_op** out) {
  return guard([&] {
    REQUIRE(ctx && buf && out && n_leaves >= 0 && (leaves || n_leaves == 0), "null argument");
    std::vector<OpPtr> ops;
    for (int i = 0; i < n_leaves; ++i) ops.push_back(op_of(leaves[i]));
    *out = wrap(plan_from_proto(&ctx->c, buf, len, ops));
  });
}
int b2_plan_proto_describe(const uint8_t* buf, size_t len, char* out, size_t cap, int* n_leaves) {
  return guard([&] {
    REQUIRE(buf && out && cap > 0, "null argument");
    std::string t = plan_proto_describe(buf, len, n_leaves);
    size_t n = t.size() < cap - 1 ? t.size() : cap - 1;
    memcpy(out, t.data(), n);
    out[n] = 0;
  });
}
int b2_jit_compile_check(const char* signature, size_t* cubin_bytes) {
  return guard([&] {
    REQUIRE(signature && cubin_bytes, "null argument");
    std::string why;
    *cubin_bytes = jit_compile_check(signature, &why);
    if (!*cubin_bytes) fail(B2_ERR_NOT_IMPLEMENTED, why);
  });
}
int b2_sort_create(b2_ctx* ctx, b2_op* input, const b2_expr* exprs, const int* descending, const int* nulls_first,
                   int n_exprs, b2_op** out) {
  return guard([&] {
    REQUIRE(ctx && exprs && descending && out && n_exprs > 0, "null argument");
    (void)nulls_first;
    OpPtr in = op_of(input);
    Schema s = in->schema();
    std::vector<ExprPtr> es;
    std::vector<char> desc;
    for (int i = 0; i < n_exprs; ++i) {
      es.push_back(decode_expr(&exprs[i], s));
      desc.push_back(descending[i] != 0);
    }
    *out = wrap(make_sort(&ctx->c, in, std::move(es), std::move(desc)));
  });
}
int b2_limit_create(b2_ctx* ctx, b2_op* input, int64_t limit, int global, b2_op** out) {
  return guard([&] {
    REQUIRE(ctx && out, "null argument");
    *out = wrap(make_limit(&ctx->c, op_of(input), limit, global != 0));
  });
}
int b2_hash_aggregate_create(b2_ctx* ctx, b2_op* input, int mode, const b2_expr* group_exprs, int n_groups,
                             const b2_agg* aggs, int n_aggs, const struct ArrowSchema* input_schema, b2_op** out) {
  return guard([&] {
    REQUIRE(ctx && out, "null argument");
    REQUIRE(mode == B2_AGG_PARTIAL || mode == B2_AGG_FINAL, "unknown AggregateMode");
    REQUIRE(n_groups >= 0 && n_aggs >= 0 && (n_groups + n_aggs) > 0, "aggregate needs groups or aggregates");
    OpPtr in = op_of(input);
    Schema s = in->schema();
    Schema arg_schema = input_schema ? schema_from_arrow(input_schema) : s;
    if (mode == B2_AGG_FINAL && !input_schema)
      fail(B2_ERR_INVALID_ARG, "input_schema in HashAggregateNode is missing.");  // from_proto.rs:214-219
    std::vector<ExprPtr> g;
    for (int i = 0; i < n_groups; ++i) g.push_back(decode_expr(&group_exprs[i], s));
    std::vector<AggSpec> a;
    for (int i = 0; i < n_aggs; ++i) {
      AggSpec sp;
      sp.fn = aggs[i].fn;
      sp.arg = decode_expr(&aggs[i].arg, arg_schema);
      sp.name = aggs[i].arg.name ? aggs[i].arg.name : "";
      a.push_back(sp);
    }
    *out = wrap(make_hash_aggregate(&ctx->c, in, mode, std::move(g), std::move(a), &arg_schema));
  });
}
int b2_hash_join_create(b2_ctx* ctx, b2_op* left, b2_op* right, const char* const* on_left,
                        const char* const* on_right, int n_on, int join_type, b2_op** out) {
  return guard([&] {
    REQUIRE(ctx && out && on_left && on_right && n_on > 0, "null argument");
    std::vector<std::string> l, r;
    for (int i = 0; i < n_on; ++i) {
      l.push_back(on_left[i]);
      r.push_back(on_right[i]);
    }
    *out = wrap(make_hash_join(&ctx->c, op_of(left), op_of(right), l, r, join_type));
  });
}
int b2_repartition_create(b2_ctx* ctx, b2_op* input, int scheme, const b2_expr* keys, int n_keys, int n_out,
                          b2_op** out) {
  return guard([&] {
    REQUIRE(ctx && out, "null argument");
    OpPtr in = op_of(input);
    Schema s = in->schema();
    std::vector<ExprPtr> k;
    for (int i = 0; i < n_keys; ++i) k.push_back(decode_expr(&keys[i], s));
    *out = wrap(make_repartition(&ctx->c, in, scheme, std::move(k), n_out));
  });
}

// ---- ExecutionPlan trait ----
int b2_op_schema(b2_op* op, struct ArrowSchema* out) {
  return guard([&] {
    REQUIRE(out, "out is null");
    schema_to_arrow(op_of(op)->schema(), out);
  });
}
int b2_op_output_partitioning(b2_op* op, int* scheme, int* n) {
  return guard([&] {
    REQUIRE(scheme && n, "null argument");
    op_of(op)->partitioning(scheme, n);
  });
}
int b2_op_n_children(b2_op* op) { return (op && op->op) ? (int)op->op->children.size() : 0; }
b2_op* b2_op_child(b2_op* op, int i) {
  if (!op || !op->op || i < 0 || i >= (int)op->op->children.size()) return nullptr;
  return wrap(op->op->children[i]);  // a new handle sharing the node; destroy it like any other
}
const char* b2_op_name(b2_op* op) { return (op && op->op) ? op->op->name() : ""; }
int b2_op_execute(b2_op* op, int partition, b2_stream** out) {
  return guard([&] {
    REQUIRE(out, "out is null");
    OpPtr o = op_of(op);
    o->ctx->activate();
    auto* s = new b2_stream;
    try {
      s->s = o->execute(partition);
    } catch (...) {
      delete s;
      throw;
    }
    *out = s;
  });
}
void b2_op_destroy(b2_op* op) { delete op; }
const char* b2_op_last_path(b2_op* op) { return (op && op->op) ? op->op->last_path.c_str() : ""; }

// ---- streams ----
int b2_stream_next_device(b2_stream* s, b2_batch** out, int* eos) {
  return guard([&] {
    REQUIRE(s && s->s && out && eos, "null argument");
    double t0 = wall_now();
    BatchPtr b = s->s->next_lazy();  // the handle may stand for a batch whose row count is still on the device
    if (b && b->ctx && b->ctx->trace) b->ctx->trace_add("b2_stream_next_device (total host time)", wall_now() - t0);
    if (!b) {
      *eos = 1;
      *out = nullptr;
      return;
    }
    *eos = 0;
    auto* h = new b2_batch;
    h->b = b;
    *out = h;
  });
}
int b2_stream_next(b2_stream* s, struct ArrowArray* out, int* eos) {
  return guard([&] {
    REQUIRE(s && s->s && out && eos, "null argument");
    BatchPtr b = s->s->next_lazy();  // exported with its header: one synchronisation learns the count and copies
    if (!b) {
      *eos = 1;
      return;
    }
    *eos = 0;
    batch_to_arrow(*b, out);
  });
}
void b2_stream_destroy(b2_stream* s) { delete s; }

// ---- batches ----
int b2_batch_from_arrow(b2_ctx* ctx, const struct ArrowSchema* schema, struct ArrowArray* batch, b2_batch** out) {
  return guard([&] {
    REQUIRE(ctx && schema && batch && batch->release && out, "null argument");
    ctx->c.activate();
    BatchPtr b = batch_from_arrow(&ctx->c, schema_from_arrow(schema), batch);
    B2_CUDA(cudaStreamSynchronize(ctx->c.stream));
    batch->release(batch);
    auto* h = new b2_batch;
    h->b = b;
    *out = h;
  });
}
int b2_batch_to_arrow(b2_batch* b, struct ArrowArray* out) {
  return guard([&] {
    REQUIRE(b && b->b && out, "null argument");
    b->b->ctx->activate();
    batch_to_arrow(*b->b, out);
  });
}
int b2_batch_schema(b2_batch* b, struct ArrowSchema* out) {
  return guard([&] {
    REQUIRE(b && b->b && out, "null argument");
    schema_to_arrow(b->b->schema, out);
  });
}
static bool resolve_handle(b2_batch* b) {
  if (!b || !b->b) return false;
  if (b->b->lazy && guard([&] {
        b->b->ctx->activate();
        resolve_batch(*b->b);
      }) != B2_OK)
    return false;
  return true;
}
int64_t b2_batch_num_rows(b2_batch* b) { return resolve_handle(b) ? b->b->num_rows : -1; }
int64_t b2_batch_num_bytes(b2_batch* b) { return resolve_handle(b) ? b->b->num_bytes() : -1; }
void b2_batch_destroy(b2_batch* b) { delete b; }

// ---- workload ----
int b2_tpch_generate(b2_ctx* ctx, const char* table, double sf, int64_t row_begin, int64_t row_end,
                     const char* const* columns, b2_batch** out) {
  return guard([&] {
    REQUIRE(ctx && table && out, "null argument");
    ctx->c.activate();
    BatchPtr b = tpch_generate(&ctx->c, table, sf, row_begin, row_end, columns);
    auto* h = new b2_batch;
    h->b = b;
    *out = h;
  });
}
int64_t b2_tpch_num_rows(const char* table, double sf) { return table ? tpch_num_rows(table, sf) : -1; }

// ---- multi-GPU ----
int b2_comm_unique_id(void* id128) {
  return guard([&] {
    REQUIRE(id128, "null argument");
    comm_unique_id(id128);
  });
}
int b2_comm_init(b2_ctx* ctx, int rank, int nranks, const void* id128) {
  return guard([&] {
    REQUIRE(ctx && id128 && nranks >= 1 && rank >= 0 && rank < nranks, "bad argument");
    ctx->c.activate();
    comm_init(&ctx->c, rank, nranks, id128);
  });
}
int b2_comm_destroy(b2_ctx* ctx) {
  return guard([&] {
    REQUIRE(ctx, "ctx is null");
    comm_destroy(&ctx->c);
  });
}
int b2_comm_allgather_batch(b2_ctx* ctx, b2_batch* local, b2_batch** out) {
  return guard([&] {
    REQUIRE(ctx && local && local->b && out, "null argument");
    ctx->c.activate();
    BatchPtr b = comm_allgather_batch(&ctx->c, local->b, /*allow_lazy=*/true);
    auto* h = new b2_batch;
    h->b = b;
    *out = h;
  });
}
int64_t b2_comm_bytes_sent(b2_ctx* ctx) { return ctx ? ctx->c.bytes_sent : 0; }

int b2_hash_split_i32_static(int device, void* cuda_stream, const int32_t* keys, int64_t n_rows, int n_out,
                             uint32_t* perm, int64_t* starts) {
  return guard([&] {
    hash_split_i32_static(device, static_cast<cudaStream_t>(cuda_stream), keys, n_rows, n_out, perm, starts);
  });
}

}  // extern "C"
