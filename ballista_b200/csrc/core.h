// core.h — internal types of libballista_b200: context, HBM buffers, columns, batches, schemas.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <chrono>
#include <functional>
#include <map>
#include <memory>
#include <mutex>
#include <stdexcept>
#include <string>
#include <vector>

#include "../../include/ballista_b200.h"

namespace b2 {

// ---- errors: thrown inside the library, converted to status codes at the C ABI ----
struct Error : std::runtime_error {
  int code;
  Error(int c, const std::string& m) : std::runtime_error(m), code(c) {}
};
[[noreturn]] inline void fail(int code, const std::string& m) { throw Error(code, m); }

#define B2_CUDA(expr)                                                                              \
  do {                                                                                             \
    cudaError_t _e = (expr);                                                                       \
    if (_e != cudaSuccess)                                                                         \
      ::b2::fail(_e == cudaErrorMemoryAllocation ? B2_ERR_OOM : B2_ERR_CUDA,                       \
                 std::string(#expr) + ": " + cudaGetErrorString(_e));                              \
  } while (0)

struct Comm;  // comm.cu

struct Ctx {
  int device = 0;
  int sm_count = 148;
  cudaStream_t stream = nullptr;
  cudaStream_t copy_stream = nullptr;  // H2D prefetch of the next host batch
  cudaMemPool_t pool = nullptr;
  int64_t launches = 0;
  Comm* comm = nullptr;
  int rank = 0, nranks = 1;
  int64_t bytes_sent = 0;
  // size-class cache of released HBM blocks (core.cu); B2_DISABLE_ALLOC_CACHE=1 turns it off
  std::mutex cache_mu;
  std::map<size_t, std::vector<void*>> cache;
  size_t cache_bytes = 0, cache_limit = (size_t)48 << 30;
  bool cache_off = false;
  void* pinned_stage = nullptr;  // 4 MiB pinned host buffer for the D2H of small result batches (core.cu)
  // optional per-launch timing of the dominant (aggregate scan) kernel with CUDA events on `stream`
  bool profiling = false;
  std::vector<std::pair<cudaEvent_t, cudaEvent_t>> prof_events;
  void prof_begin() {
    if (!profiling) return;
    cudaEvent_t a, b;
    cudaEventCreate(&a);
    cudaEventCreate(&b);
    cudaEventRecord(a, stream);
    prof_events.emplace_back(a, b);
  }
  void prof_end() {
    if (profiling && !prof_events.empty()) cudaEventRecord(prof_events.back().second, stream);
  }
  void activate() const { B2_CUDA(cudaSetDevice(device)); }
  // B2_TRACE=1: host wall time per stream synchronisation site and in the allocator, printed at ctx destroy
  bool trace = false;
  std::map<std::string, std::pair<double, int64_t>> trace_acc;
  void trace_add(const std::string& label, double seconds) {
    auto& e = trace_acc[label];
    e.first += seconds;
    e.second++;
  }
};
inline double wall_now() {
  return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}
inline void sync_stream(Ctx* ctx, const char* where) {
  if (!ctx->trace) {
    B2_CUDA(cudaStreamSynchronize(ctx->stream));
    return;
  }
  double t0 = wall_now();
  B2_CUDA(cudaStreamSynchronize(ctx->stream));
  ctx->trace_add(where, wall_now() - t0);
}
// device -> host read of a few bytes followed by a stream sync (the memcpy to pageable memory blocks)
inline void read_back(Ctx* ctx, void* dst, const void* src, size_t bytes, const char* where) {
  double t0 = ctx->trace ? wall_now() : 0;
  B2_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, ctx->stream));
  B2_CUDA(cudaStreamSynchronize(ctx->stream));
  if (ctx->trace) ctx->trace_add(where, wall_now() - t0);
}
// B2_TRACE only: synchronise and charge the wall time since the previous mark to `label` (phase timing)
struct PhaseTimer {
  Ctx* ctx;
  double t0;
  explicit PhaseTimer(Ctx* c) : ctx(c), t0(0) {
    if (ctx->trace) {
      cudaStreamSynchronize(ctx->stream);
      t0 = wall_now();
    }
  }
  void mark(const char* label) {
    if (!ctx->trace) return;
    cudaStreamSynchronize(ctx->stream);
    double t = wall_now();
    ctx->trace_add(label, t - t0);
    t0 = t;
  }
};
#define B2_STR2(x) #x
#define B2_STR(x) B2_STR2(x)
#define B2_SYNC(ctx) ::b2::sync_stream((ctx), __FILE__ ":" B2_STR(__LINE__))
#define B2_READBACK(ctx, dst, src, bytes) ::b2::read_back((ctx), (dst), (src), (bytes), __FILE__ ":" B2_STR(__LINE__))

// count + check a kernel launch
#define B2_LAUNCHED(ctx)                      \
  do {                                        \
    (ctx)->launches++;                        \
    B2_CUDA(cudaGetLastError());              \
  } while (0)

// stream-ordered HBM allocation
struct DevBuf {
  Ctx* ctx;
  void* ptr = nullptr;
  size_t bytes = 0;
  std::shared_ptr<DevBuf> parent;  // set: this is a window of `parent`'s allocation (dev_view) and owns nothing
  DevBuf(Ctx* c, size_t n);
  DevBuf(std::shared_ptr<DevBuf> p, size_t off, size_t n) : ctx(p->ctx), ptr(static_cast<char*>(p->ptr) + off), bytes(n), parent(std::move(p)) {}
  ~DevBuf();
  DevBuf(const DevBuf&) = delete;
  DevBuf& operator=(const DevBuf&) = delete;
  template <class T> T* as() const { return reinterpret_cast<T*>(ptr); }
};
using BufPtr = std::shared_ptr<DevBuf>;
void ctx_trim_cache(Ctx* ctx);  // returns every cached block to the stream-ordered pool
BufPtr dev_alloc(Ctx* ctx, size_t bytes);
inline BufPtr dev_view(const BufPtr& parent, size_t off, size_t bytes) { return std::make_shared<DevBuf>(parent, off, bytes); }
BufPtr dev_alloc_zero(Ctx* ctx, size_t bytes);

inline size_t dtype_width(int dt) {
  switch (dt) {
    case B2_BOOL: case B2_UINT8: return 1;  // device BOOL columns are one byte per row
    case B2_INT32: case B2_DATE32: return 4;
    case B2_INT64: case B2_UINT64: case B2_FLOAT64: return 8;
    default: return 0;
  }
}
const char* dtype_name(int dt);

struct Field {
  std::string name;
  int dtype = 0;
  bool nullable = false;
};
struct Schema {
  std::vector<Field> fields;
  int index_of(const std::string& name) const {
    for (size_t i = 0; i < fields.size(); ++i)
      if (fields[i].name == name) return (int)i;
    return -1;
  }
};

// One column resident in HBM.  Fixed width: `data` = values.  Utf8: `data` = int32 offsets[length+1]
// (starting at 0), `chars` = bytes.  `valid` (optional) = one byte per row, 1 = value present, 0 = NULL; absent
// means every row is valid (the TPC-H schema is non-nullable, main.rs:267-360, so the hot kernels never see it).
// Arrow's bit-packed validity bitmaps are expanded at import and packed again at export (validity.cu).
struct Column {
  int dtype = 0;
  int64_t length = 0;
  BufPtr data;
  BufPtr chars;
  int64_t chars_len = 0;
  BufPtr valid;
  const uint8_t* validity() const { return valid ? valid->as<uint8_t>() : nullptr; }
  template <class T> const T* values() const { return data ? data->as<T>() : nullptr; }
  int64_t num_bytes() const {
    if (dtype == B2_UTF8) return (length + 1) * 4 + chars_len;
    return length * (int64_t)dtype_width(dtype);
  }
};

struct Batch;
using BatchPtr = std::shared_ptr<Batch>;
// A small aggregate result whose row count is still on the device: the producing kernels ran speculatively and nothing
// was read back.  Columns are allocated for `bound` rows; `hdr` (int64, in HBM) = [0] status of the producing launches
// (0 = fine), [1] rows, [2 + c] character total of Utf8 column c.  The consumers that can work from the device-side
// count (the small Final aggregate, the Partial -> Final all-gather, the host export) do so, and a whole
// Partial -> Final -> export step synchronises once; everyone else calls resolve_batch() first.
struct LazyRows {
  BufPtr arena;     // when set: header and every column buffer are windows of this one allocation, [0, arena_used),
  size_t arena_used = 0;  // so the host export is ONE device -> host copy
  BufPtr hdr;
  int64_t bound = 0;
  bool redo_if_empty = false;        // an ungrouped aggregate over no rows is one NULL row, made by the checked path
  // recompute through the checked path when the speculation failed; gets the header's status word (0 when the
  // failure is the row count itself)
  std::function<BatchPtr(int)> redo;
};
struct Batch {
  Ctx* ctx = nullptr;
  Schema schema;
  std::vector<Column> cols;
  int64_t num_rows = 0;  // -1 while `lazy` is set
  std::shared_ptr<LazyRows> lazy;
  // >= 0: this batch is the complete output of one Partial aggregate whose first `unique_key_cols` columns are its
  // group keys, so no key occurs twice in it.  Any operator that builds a new batch leaves the default.
  int unique_key_cols = -1;
  int64_t num_bytes() const {
    int64_t b = 0;
    for (auto& c : cols) b += c.num_bytes();
    return b;
  }
};
// one readback: afterwards b.lazy is null and num_rows / column lengths are exact (core.cu)
void resolve_batch(Batch& b);
inline BatchPtr resolved(BatchPtr b) {
  if (b && b->lazy) resolve_batch(*b);
  return b;
}
// paths that move raw value buffers only (the exchange) refuse nullable data loudly instead of dropping the validity
inline void require_no_nulls(const Batch& b, const char* where) {
  for (size_t c = 0; c < b.cols.size(); ++c)
    if (b.cols[c].valid)
      fail(B2_ERR_NOT_IMPLEMENTED, std::string(where) + ": column '" + b.schema.fields[c].name + "' holds NULLs");
}

// Arrow C Data import / export (core.cu)
Schema schema_from_arrow(const ArrowSchema* s);
void schema_to_arrow(const Schema& s, ArrowSchema* out);
// host Arrow struct array -> HBM (async on ctx->stream unless `stream` given); releases `arr`
BatchPtr batch_from_arrow(Ctx* ctx, const Schema& schema, ArrowArray* arr, cudaStream_t stream = nullptr);
void batch_to_arrow(const Batch& b, ArrowArray* out);

// validity.cu — null handling shared by the operators (a null predicate is false, SUM / AVG /
// COUNT skip nulls, null join keys never match)
BufPtr valid_and(Ctx* ctx, const BufPtr& a, const BufPtr& b, int64_t n);             // null when both are absent
BufPtr valid_from_bitmap(Ctx* ctx, const uint8_t* host_bits, int64_t bit_off, int64_t n, cudaStream_t st);
// packs byte validity into an Arrow bitmap in HBM; *nulls_dev (int64 in HBM) receives the null count
// (n_dev: the row count lives on the device; rows at and beyond it are packed but not counted as NULLs)
BufPtr valid_to_bitmap(Ctx* ctx, const uint8_t* valid, int64_t n, int64_t* nulls_dev, const int64_t* n_dev = nullptr);
BufPtr valid_gather(Ctx* ctx, const uint8_t* valid, const uint32_t* idx, int64_t m);  // idx 0xFFFFFFFF -> NULL
BufPtr valid_not(Ctx* ctx, const uint8_t* valid_or_null, int64_t n);                 // IS NULL as a Boolean column
BufPtr valid_copy(Ctx* ctx, const uint8_t* valid_or_null, int64_t n);                // IS NOT NULL as a Boolean column

inline int grid_for(const Ctx* ctx, int ctas_per_sm, int64_t work_items, int per_cta) {
  int64_t need = (work_items + per_cta - 1) / per_cta;
  int64_t cap = (int64_t)ctx->sm_count * ctas_per_sm;
  if (need < 1) need = 1;
  return (int)(need < cap ? need : cap);
}

}  // namespace b2

// opaque C handles
struct b2_ctx { b2::Ctx c; };
struct b2_batch { b2::BatchPtr b; };
