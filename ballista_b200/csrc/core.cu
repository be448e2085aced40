// core.cu — context, HBM allocation, Arrow C Data Interface import/export.
#include <algorithm>
#include <cstdlib>
#include <cstring>

#include "core.h"

namespace b2 {

// Size-class cache in front of the stream-ordered pool.  Every buffer of a context is produced and consumed on
// the context's stream (the H2D prefetch stream is joined by an event before its buffers are read), so a block
// released by the host can be handed to the next request at once: work already enqueued on it precedes, in stream
// order, anything the new owner will enqueue.  cudaMallocAsync costs ~4 us a call and an operator pipeline makes
// ~70 allocations per Q3 step; a cache hit costs a map lookup.
static size_t size_class(size_t n) {
  // every buffer is padded so that 16-byte vector loads / TMA bulk copies of the last rows stay in bounds
  n = (n + 32 + 255) & ~(size_t)255;
  if (n <= (1u << 20)) {
    size_t p = 256;
    while (p < n) p <<= 1;
    return p;
  }
  return (n + (1u << 20) - 1) & ~(size_t)((1u << 20) - 1);
}
DevBuf::DevBuf(Ctx* c, size_t n) : ctx(c), bytes(n) {
  if (n == 0) return;
  n = size_class(n);
  {
    std::lock_guard<std::mutex> g(c->cache_mu);
    auto it = c->cache.find(n);
    if (it != c->cache.end() && !it->second.empty()) {
      ptr = it->second.back();
      it->second.pop_back();
      c->cache_bytes -= n;
      return;
    }
  }
  double t0 = c->trace ? wall_now() : 0;
  cudaError_t e = cudaMallocAsync(&ptr, n, c->pool, c->stream);
  if (e == cudaErrorMemoryAllocation) {  // give the cached blocks back and retry once
    cudaGetLastError();
    ctx_trim_cache(c);
    e = cudaMallocAsync(&ptr, n, c->pool, c->stream);
  }
  if (c->trace) c->trace_add("cudaMallocAsync", wall_now() - t0);
  if (e != cudaSuccess) {
    cudaGetLastError();
    fail(e == cudaErrorMemoryAllocation ? B2_ERR_OOM : B2_ERR_CUDA,
         "cudaMallocAsync(" + std::to_string(n) + " B): " + cudaGetErrorString(e));
  }
}
DevBuf::~DevBuf() {
  if (!ptr || parent) return;
  size_t n = size_class(bytes);
  {
    std::lock_guard<std::mutex> g(ctx->cache_mu);
    if (!ctx->cache_off && ctx->cache_bytes + n <= ctx->cache_limit) {
      ctx->cache[n].push_back(ptr);
      ctx->cache_bytes += n;
      return;
    }
  }
  double t0 = ctx->trace ? wall_now() : 0;
  cudaFreeAsync(ptr, ctx->stream);
  if (ctx->trace) ctx->trace_add("cudaFreeAsync", wall_now() - t0);
}
void ctx_trim_cache(Ctx* c) {
  std::lock_guard<std::mutex> g(c->cache_mu);
  for (auto& kv : c->cache)
    for (void* p : kv.second) cudaFreeAsync(p, c->stream);
  c->cache.clear();
  c->cache_bytes = 0;
}
BufPtr dev_alloc(Ctx* ctx, size_t bytes) { return std::make_shared<DevBuf>(ctx, bytes); }
BufPtr dev_alloc_zero(Ctx* ctx, size_t bytes) {
  auto b = dev_alloc(ctx, bytes);
  if (bytes) B2_CUDA(cudaMemsetAsync(b->ptr, 0, bytes, ctx->stream));
  return b;
}

void resolve_batch(Batch& b) {
  auto lz = b.lazy;
  if (!lz) return;
  int64_t h[2 + 24] = {0};
  size_t words = 2 + std::min<size_t>(b.cols.size(), 24);
  B2_READBACK(b.ctx, h, lz->hdr->ptr, words * 8);
  b.lazy = nullptr;
  const bool ok = h[0] == 0 && h[1] >= 0 && h[1] <= lz->bound && !(lz->redo_if_empty && h[1] == 0);
  if (ok) {
    b.num_rows = h[1];
    for (size_t c = 0; c < b.cols.size(); ++c) {
      b.cols[c].length = h[1];
      if (b.cols[c].dtype == B2_UTF8) b.cols[c].chars_len = h[2 + c];
    }
    return;
  }
  if (!lz->redo) fail(B2_ERR_EXECUTION, "internal: a speculative aggregate failed and cannot be redone");
  BatchPtr r = lz->redo((int)h[0]);
  if (r->lazy) resolve_batch(*r);
  b.cols = r->cols;
  b.num_rows = r->num_rows;
  b.unique_key_cols = r->unique_key_cols;
}

const char* dtype_name(int dt) {
  switch (dt) {
    case B2_BOOL: return "Boolean";
    case B2_INT32: return "Int32";
    case B2_INT64: return "Int64";
    case B2_UINT64: return "UInt64";
    case B2_FLOAT64: return "Float64";
    case B2_DATE32: return "Date32";
    case B2_UTF8: return "Utf8";
    case B2_UINT8: return "UInt8";
    default: return "?";
  }
}

static int dtype_from_format(const char* f) {
  if (!strcmp(f, "b")) return B2_BOOL;
  if (!strcmp(f, "i")) return B2_INT32;
  if (!strcmp(f, "l")) return B2_INT64;
  if (!strcmp(f, "L")) return B2_UINT64;
  if (!strcmp(f, "g")) return B2_FLOAT64;
  if (!strcmp(f, "tdD")) return B2_DATE32;
  if (!strcmp(f, "u")) return B2_UTF8;
  if (!strcmp(f, "C")) return B2_UINT8;
  fail(B2_ERR_NOT_IMPLEMENTED, std::string("Arrow format '") + f + "' is not supported on the GPU path");
}
static const char* format_from_dtype(int dt) {
  switch (dt) {
    case B2_BOOL: return "b";
    case B2_INT32: return "i";
    case B2_INT64: return "l";
    case B2_UINT64: return "L";
    case B2_FLOAT64: return "g";
    case B2_DATE32: return "tdD";
    case B2_UTF8: return "u";
    case B2_UINT8: return "C";
  }
  fail(B2_ERR_INVALID_ARG, "bad dtype");
}

Schema schema_from_arrow(const ArrowSchema* s) {
  if (!s || !s->format) fail(B2_ERR_INVALID_ARG, "null ArrowSchema");
  if (strcmp(s->format, "+s") != 0) fail(B2_ERR_ARROW, "schema must be a struct ('+s')");
  Schema out;
  for (int64_t i = 0; i < s->n_children; ++i) {
    const ArrowSchema* c = s->children[i];
    Field f;
    f.name = c->name ? c->name : "";
    f.dtype = dtype_from_format(c->format);
    f.nullable = (c->flags & ARROW_FLAG_NULLABLE) != 0;
    out.fields.push_back(f);
  }
  return out;
}

// ---- schema export ----
struct SchemaPriv {
  std::string format, name;
  std::vector<ArrowSchema> child_store;
  std::vector<ArrowSchema*> child_ptrs;
};
static void release_schema(ArrowSchema* s) {
  if (!s || !s->release) return;
  for (int64_t i = 0; i < s->n_children; ++i)
    if (s->children[i]->release) s->children[i]->release(s->children[i]);
  delete static_cast<SchemaPriv*>(s->private_data);
  s->release = nullptr;
}
static void fill_schema(ArrowSchema* out, const std::string& fmt, const std::string& name, bool nullable,
                        size_t n_children) {
  auto* p = new SchemaPriv;
  p->format = fmt;
  p->name = name;
  p->child_store.resize(n_children);
  for (auto& c : p->child_store) p->child_ptrs.push_back(&c);
  memset(out, 0, sizeof(*out));
  out->format = p->format.c_str();
  out->name = p->name.c_str();
  out->flags = nullable ? ARROW_FLAG_NULLABLE : 0;
  out->n_children = (int64_t)n_children;
  out->children = n_children ? p->child_ptrs.data() : nullptr;
  out->release = release_schema;
  out->private_data = p;
}
void schema_to_arrow(const Schema& s, ArrowSchema* out) {
  fill_schema(out, "+s", "", false, s.fields.size());
  for (size_t i = 0; i < s.fields.size(); ++i)
    fill_schema(out->children[i], format_from_dtype(s.fields[i].dtype), s.fields[i].name,
                s.fields[i].nullable, 0);
}

// ---- array import ----
__global__ void rebase_offsets_kernel(int32_t* off, int64_t n, int32_t base) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  for (; i < n; i += (int64_t)gridDim.x * blockDim.x) off[i] -= base;
}

BatchPtr batch_from_arrow(Ctx* ctx, const Schema& schema, ArrowArray* arr, cudaStream_t stream) {
  if (!arr) fail(B2_ERR_INVALID_ARG, "null ArrowArray");
  if ((size_t)arr->n_children != schema.fields.size())
    fail(B2_ERR_ARROW, "RecordBatch has " + std::to_string(arr->n_children) + " columns, schema has " +
                           std::to_string(schema.fields.size()));
  // null_count == -1 means "not computed" (Arrow C Data Interface): a struct-level bitmap is then looked at
  if (arr->null_count != 0 && arr->n_buffers > 0 && arr->buffers && arr->buffers[0]) {
    const uint8_t* sb = static_cast<const uint8_t*>(arr->buffers[0]);
    for (int64_t i = 0; i < arr->length; ++i) {
      int64_t bit = arr->offset + i;
      if (!((sb[bit >> 3] >> (bit & 7)) & 1)) fail(B2_ERR_NOT_IMPLEMENTED, "null struct rows (a RecordBatch has none)");
    }
  }
  cudaStream_t st = stream ? stream : ctx->stream;
  auto b = std::make_shared<Batch>();
  b->ctx = ctx;
  b->schema = schema;
  b->num_rows = arr->length;
  for (int64_t ci = 0; ci < arr->n_children; ++ci) {
    const ArrowArray* c = arr->children[ci];
    const Field& f = schema.fields[ci];
    int64_t off = c->offset + arr->offset;
    int64_t len = arr->length;
    if (c->length < arr->offset + arr->length) fail(B2_ERR_ARROW, "child array shorter than batch");
    Column col;
    col.dtype = f.dtype;
    col.length = len;
    // validity bitmap -> one byte per row in HBM.  null_count == -1 ("not computed") is treated as "may hold nulls".
    if (c->null_count != 0 && c->n_buffers > 0 && c->buffers[0] && len > 0)
      col.valid = valid_from_bitmap(ctx, static_cast<const uint8_t*>(c->buffers[0]), off, len, st);
    if (f.dtype == B2_UTF8) {
      const int32_t* ho = static_cast<const int32_t*>(c->buffers[1]);
      const char* hd = static_cast<const char*>(c->buffers[2]);
      col.data = dev_alloc(ctx, (size_t)(len + 1) * 4);
      int32_t base = 0, end = 0;
      if (len > 0 || ho) {
        base = ho ? ho[off] : 0;
        end = ho ? ho[off + len] : 0;
      }
      if (ho)
        B2_CUDA(cudaMemcpyAsync(col.data->ptr, ho + off, (size_t)(len + 1) * 4, cudaMemcpyHostToDevice, st));
      else
        B2_CUDA(cudaMemsetAsync(col.data->ptr, 0, 4, st));
      col.chars_len = end - base;
      col.chars = dev_alloc(ctx, (size_t)col.chars_len + 16);
      if (col.chars_len)
        B2_CUDA(cudaMemcpyAsync(col.chars->ptr, hd + base, (size_t)col.chars_len, cudaMemcpyHostToDevice, st));
      if (base != 0) {
        rebase_offsets_kernel<<<grid_for(ctx, 4, len + 1, 256), 256, 0, st>>>(col.data->as<int32_t>(), len + 1,
                                                                             base);
        B2_LAUNCHED(ctx);
      }
    } else if (f.dtype == B2_BOOL) {
      // Arrow booleans are bit-packed; device BOOL columns are one byte per row
      std::vector<uint8_t> tmp((size_t)len);
      const uint8_t* bits = static_cast<const uint8_t*>(c->buffers[1]);
      for (int64_t i = 0; i < len; ++i) tmp[(size_t)i] = (bits[(off + i) >> 3] >> ((off + i) & 7)) & 1;
      col.data = dev_alloc(ctx, (size_t)len + 16);
      if (len) {
        B2_CUDA(cudaMemcpyAsync(col.data->ptr, tmp.data(), (size_t)len, cudaMemcpyHostToDevice, st));
        B2_CUDA(cudaStreamSynchronize(st));
      }
    } else {
      size_t w = dtype_width(f.dtype);
      col.data = dev_alloc(ctx, (size_t)len * w + 16);
      if (len)
        B2_CUDA(cudaMemcpyAsync(col.data->ptr, static_cast<const char*>(c->buffers[1]) + (size_t)off * w,
                                (size_t)len * w, cudaMemcpyHostToDevice, st));
    }
    b->cols.push_back(std::move(col));
  }
  return b;
}

// ---- array export ----
struct ArrayPriv {
  std::vector<void*> owned;  // malloc'ed host buffers
  std::vector<const void*> buffers;
  std::vector<ArrowArray> child_store;
  std::vector<ArrowArray*> child_ptrs;
};
static void release_array(ArrowArray* a) {
  if (!a || !a->release) return;
  for (int64_t i = 0; i < a->n_children; ++i)
    if (a->children[i]->release) a->children[i]->release(a->children[i]);
  auto* p = static_cast<ArrayPriv*>(a->private_data);
  for (void* q : p->owned) free(q);
  delete p;
  a->release = nullptr;
}
static void* host_buf(ArrayPriv* p, size_t n) {
  void* q = malloc(n ? n : 1);
  if (!q) fail(B2_ERR_OOM, "host malloc failed");
  p->owned.push_back(q);
  return q;
}

// Device -> host copies of one exported batch.  Small results (the usual case: aggregate outputs) go through
// the context's pinned staging buffer: every copy is truly asynchronous and ONE synchronisation serves them all,
// instead of one blocking pageable copy per buffer.  Large batches are copied directly.
struct D2HStager {
  struct Item {
    void* dst;
    const void* src;
    size_t bytes;
  };
  Ctx* ctx;
  std::vector<Item> items;
  explicit D2HStager(Ctx* c) : ctx(c) {}
  void add(void* dst, const void* src, size_t bytes) { items.push_back(Item{dst, src, bytes}); }
  void flush() {
    constexpr size_t CAP = 4u << 20;
    size_t total = 0;
    for (auto& it : items) total += (it.bytes + 15) & ~(size_t)15;
    if (total <= CAP && !ctx->pinned_stage) {
      if (cudaHostAlloc(&ctx->pinned_stage, CAP, cudaHostAllocDefault) != cudaSuccess) {
        cudaGetLastError();
        ctx->pinned_stage = nullptr;
      }
    }
    if (total <= CAP && ctx->pinned_stage) {
      char* base = static_cast<char*>(ctx->pinned_stage);
      size_t off = 0;
      for (auto& it : items) {
        B2_CUDA(cudaMemcpyAsync(base + off, it.src, it.bytes, cudaMemcpyDeviceToHost, ctx->stream));
        off += (it.bytes + 15) & ~(size_t)15;
      }
      B2_SYNC(ctx);
      off = 0;
      for (auto& it : items) {
        memcpy(it.dst, base + off, it.bytes);
        off += (it.bytes + 15) & ~(size_t)15;
      }
    } else {
      for (auto& it : items) B2_CUDA(cudaMemcpyAsync(it.dst, it.src, it.bytes, cudaMemcpyDeviceToHost, ctx->stream));
      B2_SYNC(ctx);
    }
    items.clear();
  }
};

// Lazy batch whose header and columns are windows of one arena: ONE device -> host copy of the used prefix into one
// host block; the Arrow buffers point into that block, validity bytes are packed into bitmaps on the host (a lazy
// result holds at most a few hundred rows).  Returns false when the header says the speculation failed.
static bool export_arena(const Batch& b, const LazyRows& lz, ArrowArray* out, ArrayPriv* top, int64_t* status) {
  Ctx* ctx = b.ctx;
  char* host = static_cast<char*>(host_buf(top, lz.arena_used + 64));
  const char* dev0 = static_cast<const char*>(lz.arena->ptr);
  D2HStager stage(ctx);
  stage.add(host, dev0, lz.arena_used);
  stage.flush();
  auto at = [&](const BufPtr& buf) { return host + (static_cast<const char*>(buf->ptr) - dev0); };
  const int64_t* hdr = reinterpret_cast<const int64_t*>(at(lz.hdr));
  *status = hdr[0];
  if (!(hdr[0] == 0 && hdr[1] >= 0 && hdr[1] <= lz.bound && !(lz.redo_if_empty && hdr[1] == 0))) return false;
  const int64_t n = hdr[1];
  out->length = n;
  for (size_t ci = 0; ci < b.cols.size(); ++ci) {
    const Column& col = b.cols[ci];
    auto* p = new ArrayPriv;
    ArrowArray* a = &top->child_store[ci];
    a->length = n;
    a->release = release_array;
    a->private_data = p;
    if (col.dtype == B2_UTF8) p->buffers = {nullptr, at(col.data), at(col.chars)};
    else if (col.dtype == B2_BOOL) fail(B2_ERR_EXECUTION, "internal: Boolean column in an aggregate result");
    else p->buffers = {nullptr, at(col.data)};
    if (col.valid) {
      const uint8_t* v = reinterpret_cast<const uint8_t*>(at(col.valid));
      uint8_t* bits = static_cast<uint8_t*>(host_buf(p, (size_t)((n + 7) >> 3) + 1));
      memset(bits, 0, (size_t)((n + 7) >> 3) + 1);
      int64_t nulls = 0;
      for (int64_t i = 0; i < n; ++i) {
        if (v[i]) bits[i >> 3] |= (uint8_t)(1u << (i & 7));
        else ++nulls;
      }
      p->buffers[0] = bits;
      a->null_count = nulls;
    }
    a->n_buffers = (int64_t)p->buffers.size();
    a->buffers = p->buffers.data();
  }
  return true;
}

// A lazy batch (Batch::lazy) is exported without resolving it first: the header travels with the columns (copied at
// their allocated bound; aggregate results are a few rows), so the export's one synchronisation is also the one that
// learns the row count.  If the header then says the speculation failed, the batch is redone and exported again.
void batch_to_arrow(const Batch& b, ArrowArray* out) {
  Ctx* ctx = b.ctx;
  const std::shared_ptr<LazyRows> lz = b.lazy;
  auto* top = new ArrayPriv;
  top->child_store.resize(b.cols.size());
  for (auto& c : top->child_store) {
    memset(&c, 0, sizeof(c));
    top->child_ptrs.push_back(&c);
  }
  top->buffers.push_back(nullptr);
  memset(out, 0, sizeof(*out));
  out->length = b.num_rows;
  out->n_buffers = 1;
  out->buffers = top->buffers.data();
  out->n_children = (int64_t)b.cols.size();
  out->children = top->child_ptrs.data();
  out->release = release_array;
  out->private_data = top;
  int64_t hdr[2 + 24] = {0};
  bool redo = false;
  try {
    if (lz && b.cols.size() > 24) fail(B2_ERR_EXECUTION, "internal: lazy batch with more than 24 columns");
    if (lz && lz->arena) {
      for (size_t ci = 0; ci < b.cols.size(); ++ci)
        if (!b.cols[ci].data) fail(B2_ERR_EXECUTION, "internal: pruned column in an aggregate result");
      redo = !export_arena(b, *lz, out, top, &hdr[0]);
      if (!redo) return;
    } else {
    std::vector<std::pair<size_t, std::vector<uint8_t>>> bools;  // bit-pack after the sync
    bools.reserve(b.cols.size());  // element addresses are handed to the stager
    D2HStager stage(ctx);
    std::vector<BufPtr> keep;  // packed bitmaps / null counters live until the flush
    std::vector<int64_t> null_counts(b.cols.size(), 0);
    if (lz) stage.add(hdr, lz->hdr->ptr, (2 + b.cols.size()) * 8);
    for (size_t ci = 0; ci < b.cols.size(); ++ci) {
      const Column& col = b.cols[ci];
      if (!col.data) fail(B2_ERR_EXECUTION, "internal: column '" + b.schema.fields[ci].name +
                                                "' was pruned upstream but is being exported");
      const int64_t len = lz ? lz->bound : col.length;  // rows copied; the exported length is set after the sync
      auto* p = new ArrayPriv;
      ArrowArray* a = &top->child_store[ci];
      a->length = len;
      a->release = release_array;
      a->private_data = p;
      if (col.dtype == B2_UTF8) {
        const size_t nchars = lz ? (col.chars ? col.chars->bytes : 0) : (size_t)col.chars_len;
        void* ho = host_buf(p, (size_t)(len + 1) * 4);
        void* hd = host_buf(p, nchars);
        stage.add(ho, col.data->ptr, (size_t)(len + 1) * 4);
        if (nchars) stage.add(hd, col.chars->ptr, nchars);
        p->buffers = {nullptr, ho, hd};
      } else if (col.dtype == B2_BOOL) {
        bools.emplace_back(ci, std::vector<uint8_t>((size_t)len));
        if (len) stage.add(bools.back().second.data(), col.data->ptr, (size_t)len);
        void* hb = host_buf(p, (size_t)(len + 7) / 8);
        memset(hb, 0, (size_t)(len + 7) / 8);
        p->buffers = {nullptr, hb};
      } else {
        size_t nb = (size_t)len * dtype_width(col.dtype);
        void* hv = host_buf(p, nb);
        if (nb) stage.add(hv, col.data->ptr, nb);
        p->buffers = {nullptr, hv};
      }
      if (col.valid && len) {
        auto counter = dev_alloc(ctx, 16);
        auto bits = valid_to_bitmap(ctx, col.valid->as<uint8_t>(), len, counter->as<int64_t>(),
                                    lz ? lz->hdr->as<int64_t>() + 1 : nullptr);
        void* hb = host_buf(p, (size_t)((len + 7) >> 3));
        stage.add(hb, bits->ptr, (size_t)((len + 7) >> 3));
        stage.add(&null_counts[ci], counter->ptr, 8);
        p->buffers[0] = hb;
        keep.push_back(bits);
        keep.push_back(counter);
      }
      a->n_buffers = (int64_t)p->buffers.size();
      a->buffers = p->buffers.data();
    }
    stage.flush();
    if (lz) {
      redo = !(hdr[0] == 0 && hdr[1] >= 0 && hdr[1] <= lz->bound && !(lz->redo_if_empty && hdr[1] == 0));
      if (!redo) {
        out->length = hdr[1];
        for (size_t ci = 0; ci < b.cols.size(); ++ci) top->child_store[ci].length = hdr[1];
      }
    }
    for (size_t ci = 0; ci < b.cols.size(); ++ci) top->child_store[ci].null_count = null_counts[ci];
    for (auto& pr : bools) {
      auto* p = static_cast<ArrayPriv*>(top->child_store[pr.first].private_data);
      uint8_t* bits = static_cast<uint8_t*>(const_cast<void*>(p->buffers[1]));
      for (size_t i = 0; i < pr.second.size(); ++i)
        if (pr.second[i]) bits[i >> 3] |= (uint8_t)(1u << (i & 7));
    }
    }
  } catch (...) {
    release_array(out);
    throw;
  }
  if (redo) {
    release_array(out);
    if (!lz->redo) fail(B2_ERR_EXECUTION, "internal: a speculative aggregate failed and cannot be redone");
    BatchPtr r = lz->redo((int)hdr[0]);
    if (r->lazy) resolve_batch(*r);
    batch_to_arrow(*r, out);
  }
}

}  // namespace b2
