// comm.cu — the in-box exchange over NVLink (NCCL): replaces the reference's shuffle write + Flight fetch
// (rust/core/src/utils.rs:49-84, rust/core/src/execution_plans/shuffle_reader.rs:77-99,
// rust/executor/src/flight_service.rs:193-228) between GPUs of one box.  One process per GPU; the
// communicator id is created by rank 0 and broadcast by the host plumbing (torch.distributed).
#include <nccl.h>

#include <cstring>

#include "comm.h"
#include "ops.h"
#include "scan.cuh"

namespace b2 {

static_assert(sizeof(ncclUniqueId) == 128, "b2_comm_unique_id hands out 128 bytes");

void comm_unique_id(void* id128) {
  ncclUniqueId id;
  B2_NCCL(ncclGetUniqueId(&id));
  memcpy(id128, &id, sizeof id);
}
void comm_init(Ctx* ctx, int rank, int nranks, const void* id128) {
  if (ctx->comm) fail(B2_ERR_INVALID_ARG, "communicator already initialised");
  ncclUniqueId id;
  memcpy(&id, id128, sizeof id);
  auto* c = new Comm;
  ncclResult_t r = ncclCommInitRank(&c->nccl, nranks, id, rank);
  if (r != ncclSuccess) {
    delete c;
    fail(B2_ERR_NCCL, std::string("ncclCommInitRank: ") + ncclGetErrorString(r));
  }
  ctx->comm = c;
  ctx->rank = rank;
  ctx->nranks = nranks;
}
void comm_destroy(Ctx* ctx) {
  if (!ctx->comm) return;
  shuffle_arena_release(ctx);
  ncclCommDestroy(ctx->comm->nccl);
  delete ctx->comm;
  ctx->comm = nullptr;
  ctx->rank = 0;
  ctx->nranks = 1;
}

__global__ void lens_kernel(const int32_t* off, int64_t n, int32_t* out) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  for (; i < n; i += (int64_t)gridDim.x * blockDim.x) out[i] = off[i + 1] - off[i];
}
struct OffOut {
  int32_t* off;
  int64_t n;
  __device__ __forceinline__ void operator()(int64_t i, int v, int64_t pre) const {
    off[i] = (int32_t)pre;
    if (i == n - 1) off[n] = (int32_t)(pre + v);
  }
};

// exchange of per-rank int64 vectors (row / byte counts); returns the [nranks][len] matrix on the host
std::vector<int64_t> allgather_counts(Ctx* ctx, const std::vector<int64_t>& mine) {
  size_t len = mine.size();
  auto dsend = dev_alloc(ctx, len * 8);
  auto drecv = dev_alloc(ctx, len * 8 * ctx->nranks);
  B2_CUDA(cudaMemcpyAsync(dsend->ptr, mine.data(), len * 8, cudaMemcpyHostToDevice, ctx->stream));
  B2_NCCL(ncclAllGather(dsend->ptr, drecv->ptr, len, ncclInt64, ctx->comm->nccl, ctx->stream));
  std::vector<int64_t> all(len * ctx->nranks);
  B2_READBACK(ctx, all.data(), drecv->ptr, all.size() * 8);
  return all;
}


static size_t align16(size_t x) { return (x + 15) & ~(size_t)15; }

// ---- small-batch all-gather: the Partial -> Final hop of an aggregate (a handful of group rows per rank) ----
// Every rank packs header + columns into a slot whose size depends only on the schema, so ONE ncclAllGather moves
// everything and ONE readback (the headers) sizes the output; one kernel packs, one unpacks.  A rank whose batch
// does not fit its slot says so in the header and everybody falls back to the general, two-round protocol below.
constexpr int SMALL_ROWS = 256, SMALL_CHARS = 4096, SMALL_MAX_COLS = 24, SMALL_MAX_RANKS = 64, SMALL_MAX_STR = 4;
struct SmallCol {
  const void* src;       // pack: values (Utf8: int32 offsets); unpack: unused
  const uint8_t* valid;  // pack: validity bytes, or null (all rows valid)
  uint8_t* dst_valid;    // unpack: validity bytes of the gathered column, or null (no rank sent a NULL)
  int64_t valid_off;     // SMALL_ROWS validity bytes inside a slot
  const uint8_t* chars;  // pack: Utf8 characters
  void* dst;             // unpack: values (Utf8: int32 offsets)
  uint8_t* dst_chars;    // unpack: Utf8 characters
  int width;             // bytes per value (Utf8: 4, the staged lengths)
  int utf8;              // index into the char-prefix table, -1 for fixed width
  int64_t val_off, chr_off;  // offsets inside a slot
};
struct SmallArgs {
  int ncol, nr;
  const int64_t* lazy_hdr;  // pack: the batch's device-side header (status, rows, character totals), or null
  int64_t lazy_bound;
  int lazy_redo_if_empty;
  int64_t stride, rows, header_bytes;
  SmallCol c[SMALL_MAX_COLS];
  int64_t chars[SMALL_MAX_COLS];                         // pack: my character counts
  int64_t row_prefix[SMALL_MAX_RANKS + 1];               // unpack
  int64_t chr_prefix[SMALL_MAX_STR][SMALL_MAX_RANKS + 1];
};
// header word 0: rows, or -1 = does not fit the slot (everybody takes the general protocol), or -2 = this rank's lazy
// batch has to be redone first (everybody retries)
__global__ void pack_small_kernel(SmallArgs a, char* slot, int fits) {
  const SmallCol& c = a.c[blockIdx.y];
  int64_t rows = a.rows;
  int64_t my_chars = a.chars[blockIdx.y];
  int64_t verdict = fits ? a.rows : -1;
  if (a.lazy_hdr) {  // row count and character totals live on the device
    rows = a.lazy_hdr[1];
    my_chars = a.lazy_hdr[2 + blockIdx.y];
    if (a.lazy_hdr[0] != 0 || rows < 0 || rows > a.lazy_bound || (a.lazy_redo_if_empty && rows == 0)) {
      verdict = -2;
    } else {
      verdict = rows <= SMALL_ROWS ? rows : -1;
      for (int k = 0; k < a.ncol; ++k)
        if (a.c[k].utf8 >= 0 && a.lazy_hdr[2 + k] > SMALL_CHARS) verdict = -1;
    }
  }
  if (blockIdx.y == 0 && threadIdx.x == 0) {
    int64_t* h = reinterpret_cast<int64_t*>(slot);
    h[0] = verdict;
    for (int k = 0; k < a.ncol; ++k) {
      h[1 + k] = a.lazy_hdr ? (a.c[k].utf8 >= 0 ? a.lazy_hdr[2 + k] : 0) : a.chars[k];
      h[1 + a.ncol + k] = a.c[k].valid != nullptr;  // this rank's column may hold NULLs
    }
  }
  if (verdict < 0) return;
  char* base = slot + c.val_off;
  for (int64_t i = threadIdx.x; i < rows; i += blockDim.x) {
    slot[c.valid_off + i] = c.valid ? (char)c.valid[i] : (char)1;
    if (c.utf8 >= 0) {
      const int32_t* off = static_cast<const int32_t*>(c.src);
      reinterpret_cast<int32_t*>(base)[i] = off[i + 1] - off[i];
    } else if (c.width == 8) {
      reinterpret_cast<int64_t*>(base)[i] = static_cast<const int64_t*>(c.src)[i];
    } else if (c.width == 4) {
      reinterpret_cast<int32_t*>(base)[i] = static_cast<const int32_t*>(c.src)[i];
    } else {
      base[i] = static_cast<const char*>(c.src)[i];
    }
  }
  if (c.utf8 >= 0)
    for (int64_t k = threadIdx.x; k < my_chars; k += blockDim.x) slot[c.chr_off + k] = (char)c.chars[k];
}
__global__ void unpack_small_kernel(SmallArgs a, const char* recv) {
  const SmallCol& c = a.c[blockIdx.y];
  const int64_t total = a.row_prefix[a.nr];
  for (int64_t i = threadIdx.x; i < total; i += blockDim.x) {
    int r = 0;
    while (a.row_prefix[r + 1] <= i) ++r;
    int64_t j = i - a.row_prefix[r];
    const char* base = recv + (int64_t)r * a.stride + c.val_off;
    if (c.dst_valid) c.dst_valid[i] = (uint8_t)recv[(int64_t)r * a.stride + c.valid_off + j];
    if (c.utf8 >= 0) {
      const int32_t* lens = reinterpret_cast<const int32_t*>(base);
      int64_t o = a.chr_prefix[c.utf8][r];
      for (int64_t t = 0; t < j; ++t) o += lens[t];
      static_cast<int32_t*>(c.dst)[i] = (int32_t)o;
      if (i == total - 1) static_cast<int32_t*>(c.dst)[total] = (int32_t)a.chr_prefix[c.utf8][a.nr];
    } else if (c.width == 8) {
      static_cast<int64_t*>(c.dst)[i] = reinterpret_cast<const int64_t*>(base)[j];
    } else if (c.width == 4) {
      static_cast<int32_t*>(c.dst)[i] = reinterpret_cast<const int32_t*>(base)[j];
    } else {
      static_cast<char*>(c.dst)[i] = base[j];
    }
  }
  if (c.utf8 >= 0) {
    const int64_t tc = a.chr_prefix[c.utf8][a.nr];
    for (int64_t k = threadIdx.x; k < tc; k += blockDim.x) {
      int r = 0;
      while (a.chr_prefix[c.utf8][r + 1] <= k) ++r;
      c.dst_chars[k] = (uint8_t)recv[(int64_t)r * a.stride + c.chr_off + (k - a.chr_prefix[c.utf8][r])];
    }
  }
}

// The same unpack when nothing was read back (the local batch was lazy): row / character prefixes are derived from the
// gathered slot headers on the device, and the result's own header (Batch::lazy: status, rows, character totals) is
// written for whoever consumes the gathered batch.  A slot that did not fit (-1) or whose sender must redo (-2) makes
// the status non-zero: every rank sees the same headers, so every rank redoes the gather through the checked path.
__global__ void unpack_small_lazy_kernel(SmallArgs a, const char* recv, int64_t* hdr_out) {
  __shared__ int64_t row_prefix[SMALL_MAX_RANKS + 1];
  __shared__ int64_t chr_prefix[SMALL_MAX_RANKS + 1];
  __shared__ int bad;
  const SmallCol& c = a.c[blockIdx.y];
  if (threadIdx.x == 0) {
    bad = 0;
    row_prefix[0] = chr_prefix[0] = 0;
    for (int r = 0; r < a.nr; ++r) {
      const int64_t* h = reinterpret_cast<const int64_t*>(recv + (int64_t)r * a.stride);
      if (h[0] < 0) bad = h[0] == -1 ? AGG_TOO_BIG : AGG_INPUT_FAILED;
      row_prefix[r + 1] = row_prefix[r] + (h[0] > 0 ? h[0] : 0);
      chr_prefix[r + 1] = chr_prefix[r] + (c.utf8 >= 0 ? h[1 + blockIdx.y] : 0);
    }
    if (blockIdx.y == 0) {
      hdr_out[0] = bad;
      hdr_out[1] = row_prefix[a.nr];
    }
    if (c.utf8 >= 0) hdr_out[2 + blockIdx.y] = chr_prefix[a.nr];
  }
  __syncthreads();
  if (bad) return;
  const int64_t total = row_prefix[a.nr];
  for (int64_t i = threadIdx.x; i < total; i += blockDim.x) {
    int r = 0;
    while (row_prefix[r + 1] <= i) ++r;
    int64_t j = i - row_prefix[r];
    const char* base = recv + (int64_t)r * a.stride + c.val_off;
    if (c.dst_valid) c.dst_valid[i] = (uint8_t)recv[(int64_t)r * a.stride + c.valid_off + j];
    if (c.utf8 >= 0) {
      const int32_t* lens = reinterpret_cast<const int32_t*>(base);
      int64_t o = chr_prefix[r];
      for (int64_t t = 0; t < j; ++t) o += lens[t];
      static_cast<int32_t*>(c.dst)[i] = (int32_t)o;
      if (i == total - 1) static_cast<int32_t*>(c.dst)[total] = (int32_t)chr_prefix[a.nr];
    } else if (c.width == 8) {
      static_cast<int64_t*>(c.dst)[i] = reinterpret_cast<const int64_t*>(base)[j];
    } else if (c.width == 4) {
      static_cast<int32_t*>(c.dst)[i] = reinterpret_cast<const int32_t*>(base)[j];
    } else {
      static_cast<char*>(c.dst)[i] = base[j];
    }
  }
  if (c.utf8 >= 0) {
    if (total == 0 && threadIdx.x == 0) static_cast<int32_t*>(c.dst)[0] = 0;
    const int64_t tc = chr_prefix[a.nr];
    for (int64_t k = threadIdx.x; k < tc; k += blockDim.x) {
      int r = 0;
      while (chr_prefix[r + 1] <= k) ++r;
      c.dst_chars[k] = (uint8_t)recv[(int64_t)r * a.stride + c.chr_off + (k - chr_prefix[r])];
    }
  }
}

static BatchPtr allgather_general(Ctx* ctx, const Batch& local);

// Returns nullptr when the schema is outside the small protocol or some rank's batch did not fit its slot; *retry is
// set when some rank's lazy batch must be redone first (then every rank resolves and calls again).
// `lazy_src`: the local batch is lazy (a speculative Partial result) and the caller can take a lazy result: nothing is
// read back here either, the gathered batch carries its own device-side header.
static BatchPtr allgather_small(Ctx* ctx, const Batch& local, bool* retry, const BatchPtr& lazy_src = nullptr) {
  *retry = false;
  const int nr = ctx->nranks;
  const size_t ncol = local.cols.size();
  if (ncol > SMALL_MAX_COLS || nr > SMALL_MAX_RANKS || ncol == 0) return nullptr;
  SmallArgs a;
  memset(&a, 0, sizeof a);
  a.ncol = (int)ncol;
  a.nr = nr;
  a.rows = local.lazy ? 0 : local.num_rows;
  if (local.lazy) {
    a.lazy_hdr = local.lazy->hdr->as<int64_t>();
    a.lazy_bound = local.lazy->bound;
    a.lazy_redo_if_empty = local.lazy->redo_if_empty;
  }
  const size_t hdr_words = 1 + 2 * ncol;  // rows, character count per column, may-hold-NULLs flag per column
  a.header_bytes = (int64_t)align16(hdr_words * 8);
  int64_t off = a.header_bytes;
  int n_str = 0;
  bool fits = local.lazy || local.num_rows <= SMALL_ROWS;
  for (size_t c = 0; c < ncol; ++c) {
    const Column& col = local.cols[c];
    if (!col.data) return nullptr;
    SmallCol& sc = a.c[c];
    sc.src = col.data->ptr;
    sc.valid = col.validity();
    sc.valid_off = off;
    off += SMALL_ROWS;
    sc.utf8 = -1;
    sc.width = col.dtype == B2_UTF8 ? 4 : (int)dtype_width(col.dtype);
    sc.val_off = off;
    off += (int64_t)align16((size_t)SMALL_ROWS * sc.width);
    if (col.dtype == B2_UTF8) {
      if (n_str == SMALL_MAX_STR) return nullptr;
      sc.utf8 = n_str++;
      sc.chars = col.chars ? col.chars->as<uint8_t>() : nullptr;
      sc.chr_off = off;
      off += SMALL_CHARS;
      a.chars[c] = local.lazy ? 0 : col.chars_len;
      fits = fits && (local.lazy || col.chars_len <= SMALL_CHARS);
    }
  }
  a.stride = off;
  auto send = dev_alloc(ctx, (size_t)a.stride);
  auto recv = dev_alloc(ctx, (size_t)a.stride * nr);
  pack_small_kernel<<<dim3(1, (unsigned)ncol), 256, 0, ctx->stream>>>(a, send->as<char>(), fits ? 1 : 0);
  B2_LAUNCHED(ctx);
  B2_NCCL(ncclAllGather(send->ptr, recv->ptr, (size_t)a.stride, ncclChar, ctx->comm->nccl, ctx->stream));
  ctx->bytes_sent += a.stride * (nr - 1);
  if (lazy_src && local.lazy && ncol <= 24) {
    const int64_t bound = (int64_t)nr * SMALL_ROWS;
    auto out = std::make_shared<Batch>();
    out->ctx = ctx;
    out->schema = local.schema;
    out->num_rows = -1;
    for (size_t c = 0; c < ncol; ++c) {
      Column o;
      o.dtype = local.cols[c].dtype;
      o.length = bound;
      if (o.dtype == B2_UTF8) {
        o.data = dev_alloc(ctx, (size_t)(bound + 1) * 4 + 16);
        o.chars = dev_alloc(ctx, (size_t)nr * SMALL_CHARS + 16);
        a.c[c].dst_chars = o.chars->as<uint8_t>();
      } else {
        o.data = dev_alloc(ctx, (size_t)bound * dtype_width(o.dtype) + 16);
      }
      a.c[c].dst = o.data->ptr;
      o.valid = dev_alloc(ctx, (size_t)bound + 16);  // which ranks sent NULLs is not known on the host: always carried
      a.c[c].dst_valid = o.valid->as<uint8_t>();
      out->cols.push_back(std::move(o));
    }
    auto lz = std::make_shared<LazyRows>();
    lz->hdr = dev_alloc_zero(ctx, (2 + 24) * 8);
    lz->bound = bound;
    unpack_small_lazy_kernel<<<dim3(1, (unsigned)ncol), 256, 0, ctx->stream>>>(a, recv->as<char>(), lz->hdr->as<int64_t>());
    B2_LAUNCHED(ctx);
    BatchPtr src = lazy_src;
    lz->redo = [ctx, src](int) {  // collective: every rank saw the same slot headers and lands here together
      resolve_batch(*src);
      return comm_allgather_batch(ctx, src);
    };
    out->lazy = lz;
    return out;
  }
  // one strided readback: the header of every slot
  std::vector<int64_t> heads((size_t)nr * hdr_words);
  {
    double t0 = ctx->trace ? wall_now() : 0;
    B2_CUDA(cudaMemcpy2DAsync(heads.data(), hdr_words * 8, recv->ptr, (size_t)a.stride, hdr_words * 8, (size_t)nr,
                              cudaMemcpyDeviceToHost, ctx->stream));
    B2_CUDA(cudaStreamSynchronize(ctx->stream));
    if (ctx->trace) ctx->trace_add("comm.cu:allgather_small headers", wall_now() - t0);
  }
  bool general = false;
  for (int r = 0; r < nr; ++r) {
    if (heads[(size_t)r * hdr_words] == -1) general = true;  // somebody did not fit: general protocol (all ranks agree)
    if (heads[(size_t)r * hdr_words] == -2) *retry = true;   // somebody has to redo a speculative batch first
  }
  if (general) *retry = false;
  if (general || *retry) return nullptr;
  std::vector<char> any_null(ncol, 0);
  for (int r = 0; r < nr; ++r) {
    a.row_prefix[r + 1] = a.row_prefix[r] + heads[(size_t)r * hdr_words];
    for (size_t c = 0; c < ncol; ++c) {
      if (a.c[c].utf8 >= 0)
        a.chr_prefix[a.c[c].utf8][r + 1] = a.chr_prefix[a.c[c].utf8][r] + heads[(size_t)r * hdr_words + 1 + c];
      any_null[c] |= heads[(size_t)r * hdr_words + 1 + ncol + c] != 0;
    }
  }
  int64_t total = a.row_prefix[nr];
  auto out = std::make_shared<Batch>();
  out->ctx = ctx;
  out->schema = local.schema;
  out->num_rows = total;
  for (size_t c = 0; c < ncol; ++c) {
    Column o;
    o.dtype = local.cols[c].dtype;
    o.length = total;
    if (o.dtype == B2_UTF8) {
      o.data = dev_alloc_zero(ctx, (size_t)(total + 1) * 4);
      o.chars_len = a.chr_prefix[a.c[c].utf8][nr];
      o.chars = dev_alloc(ctx, (size_t)o.chars_len + 16);
      a.c[c].dst_chars = o.chars->as<uint8_t>();
    } else {
      o.data = dev_alloc(ctx, (size_t)total * dtype_width(o.dtype) + 16);
    }
    a.c[c].dst = o.data->ptr;
    a.c[c].dst_valid = nullptr;
    if (any_null[c]) {  // some rank's column carried validity (e.g. the NULL sum state of an empty partition)
      o.valid = dev_alloc(ctx, (size_t)total + 16);
      a.c[c].dst_valid = o.valid->as<uint8_t>();
    }
    out->cols.push_back(std::move(o));
  }
  if (total) {
    unpack_small_kernel<<<dim3(1, (unsigned)ncol), 256, 0, ctx->stream>>>(a, recv->as<char>());
    B2_LAUNCHED(ctx);
  }
  return out;
}

BatchPtr comm_allgather_batch(Ctx* ctx, const BatchPtr& local_ptr, bool allow_lazy) {
  if (!ctx->comm) fail(B2_ERR_INVALID_ARG, "no communicator (call b2_comm_init)");
  Batch& local = *local_ptr;
  bool retry = false;
  // a lazy batch is packed from its device-side header and (allow_lazy) unpacked without a readback: the
  // Partial -> Final hop adds no synchronisation of its own
  if (BatchPtr b = allgather_small(ctx, local, &retry, allow_lazy && local.lazy ? local_ptr : nullptr)) return b;  // carries validity
  if (local.lazy) resolve_batch(local);
  if (retry)
    if (BatchPtr b = allgather_small(ctx, local, &retry)) return b;
  require_no_nulls(local, "all-gather of a large batch");
  return allgather_general(ctx, local);
}

// General all-gather of a batch: counts first, then one NCCL all-gather of a packed staging

// buffer: per column a region of max_rows values (Utf8: lengths + a chars region of max_chars bytes).
static BatchPtr allgather_general(Ctx* ctx, const Batch& local) {
  int nr = ctx->nranks;
  size_t ncol = local.cols.size();
  std::vector<int64_t> mine(1 + ncol, 0);
  mine[0] = local.num_rows;
  for (size_t c = 0; c < ncol; ++c) mine[1 + c] = local.cols[c].dtype == B2_UTF8 ? local.cols[c].chars_len : 0;
  std::vector<int64_t> all = allgather_counts(ctx, mine);
  auto at = [&](int r, size_t k) { return all[(size_t)r * (1 + ncol) + k]; };
  int64_t max_rows = 0, total_rows = 0;
  std::vector<int64_t> max_chars(ncol, 0), total_chars(ncol, 0);
  for (int r = 0; r < nr; ++r) {
    max_rows = std::max(max_rows, at(r, 0));
    total_rows += at(r, 0);
    for (size_t c = 0; c < ncol; ++c) {
      max_chars[c] = std::max(max_chars[c], at(r, 1 + c));
      total_chars[c] += at(r, 1 + c);
    }
  }
  // staging layout
  std::vector<size_t> val_off(ncol), chr_off(ncol);
  size_t stride = 0;
  for (size_t c = 0; c < ncol; ++c) {
    size_t w = local.cols[c].dtype == B2_UTF8 ? 4 : dtype_width(local.cols[c].dtype);
    val_off[c] = stride;
    stride += align16((size_t)max_rows * w);
    if (local.cols[c].dtype == B2_UTF8) {
      chr_off[c] = stride;
      stride += align16((size_t)max_chars[c]);
    }
  }
  if (stride == 0) stride = 16;
  auto send = dev_alloc_zero(ctx, stride);
  auto recv = dev_alloc(ctx, stride * nr);
  int64_t n = local.num_rows;
  for (size_t c = 0; c < ncol; ++c) {
    const Column& col = local.cols[c];
    if (col.dtype == B2_UTF8) {
      if (n) {
        lens_kernel<<<grid_for(ctx, 4, n, 256), 256, 0, ctx->stream>>>(col.values<int32_t>(), n,
                                                                      (int32_t*)(send->as<char>() + val_off[c]));
        B2_LAUNCHED(ctx);
        if (col.chars_len)
          B2_CUDA(cudaMemcpyAsync(send->as<char>() + chr_off[c], col.chars->ptr, (size_t)col.chars_len,
                                  cudaMemcpyDeviceToDevice, ctx->stream));
      }
    } else if (n) {
      B2_CUDA(cudaMemcpyAsync(send->as<char>() + val_off[c], col.data->ptr, (size_t)n * dtype_width(col.dtype),
                              cudaMemcpyDeviceToDevice, ctx->stream));
    }
  }
  B2_NCCL(ncclAllGather(send->ptr, recv->ptr, stride, ncclChar, ctx->comm->nccl, ctx->stream));
  ctx->bytes_sent += (int64_t)stride * (nr - 1);
  // unpack in rank order
  auto out = std::make_shared<Batch>();
  out->ctx = ctx;
  out->schema = local.schema;
  out->num_rows = total_rows;
  for (size_t c = 0; c < ncol; ++c) {
    Column o;
    o.dtype = local.cols[c].dtype;
    o.length = total_rows;
    size_t w = o.dtype == B2_UTF8 ? 4 : dtype_width(o.dtype);
    BufPtr vals = dev_alloc_zero(ctx, (size_t)(total_rows + 1) * w + 16);
    int64_t row = 0;
    for (int r = 0; r < nr; ++r) {
      int64_t cnt = at(r, 0);
      if (cnt)
        B2_CUDA(cudaMemcpyAsync(vals->as<char>() + (size_t)row * w, recv->as<char>() + (size_t)r * stride + val_off[c],
                                (size_t)cnt * w, cudaMemcpyDeviceToDevice, ctx->stream));
      row += cnt;
    }
    if (o.dtype == B2_UTF8) {
      o.data = dev_alloc_zero(ctx, (size_t)(total_rows + 1) * 4);
      device_scan(ctx, I32In{vals->as<int32_t>()}, total_rows, OffOut{o.data->as<int32_t>(), total_rows},
                  (int64_t*)nullptr);
      o.chars_len = total_chars[c];
      o.chars = dev_alloc(ctx, (size_t)o.chars_len + 16);
      int64_t ch = 0;
      for (int r = 0; r < nr; ++r) {
        int64_t cnt = at(r, 1 + c);
        if (cnt)
          B2_CUDA(cudaMemcpyAsync(o.chars->as<char>() + ch, recv->as<char>() + (size_t)r * stride + chr_off[c],
                                  (size_t)cnt, cudaMemcpyDeviceToDevice, ctx->stream));
        ch += cnt;
      }
    } else {
      o.data = vals;
    }
    out->cols.push_back(std::move(o));
  }
  return out;
}

// All-to-all of a destination-major permuted batch: rows of slice d go to rank d.  Fixed-width columns
// are sent as they are; Utf8 columns travel as a lengths column plus a chars byte stream.
BatchPtr comm_all_to_all(Ctx* ctx, const Batch& b, const std::vector<int64_t>& counts) {
  require_no_nulls(b, "exchange over NVLink");
  if (!ctx->comm) fail(B2_ERR_INVALID_ARG, "no communicator (call b2_comm_init)");
  int nr = ctx->nranks, me = ctx->rank;
  if ((int)counts.size() != nr) fail(B2_ERR_INVALID_ARG, "all-to-all needs one slice per rank");
  size_t ncol = b.cols.size();
  // per destination: rows, then chars per Utf8 column
  std::vector<int64_t> send_off(nr + 1, 0);
  for (int d = 0; d < nr; ++d) send_off[d + 1] = send_off[d] + counts[d];
  std::vector<std::vector<int64_t>> chr_bounds(ncol);  // char offset at each slice boundary
  for (size_t c = 0; c < ncol; ++c) {
    if (b.cols[c].dtype != B2_UTF8 || !b.cols[c].data) continue;
    chr_bounds[c].assign(nr + 1, 0);
    std::vector<int32_t> tmp(nr + 1, 0);
    if (b.num_rows)
      for (int d = 0; d <= nr; ++d)
        B2_READBACK(ctx, &tmp[d], b.cols[c].values<int32_t>() + send_off[d], 4);
    for (int d = 0; d <= nr; ++d) chr_bounds[c][d] = tmp[d];
  }
  std::vector<int64_t> mine((size_t)nr * (1 + ncol), 0);
  for (int d = 0; d < nr; ++d) {
    mine[(size_t)d * (1 + ncol)] = counts[d];
    for (size_t c = 0; c < ncol; ++c)
      if (b.cols[c].dtype == B2_UTF8 && b.cols[c].data)
        mine[(size_t)d * (1 + ncol) + 1 + c] = chr_bounds[c][d + 1] - chr_bounds[c][d];
  }
  std::vector<int64_t> all = allgather_counts(ctx, mine);
  // what rank s sends to me
  auto from = [&](int s, size_t k) { return all[((size_t)s * nr + me) * (1 + ncol) + k]; };
  std::vector<int64_t> recv_off(nr + 1, 0);
  for (int s = 0; s < nr; ++s) recv_off[s + 1] = recv_off[s] + from(s, 0);
  int64_t total = recv_off[nr];
  if (total >= (1ll << 31) - 1) fail(B2_ERR_NOT_IMPLEMENTED, "received partition beyond 2^31-2 rows");
  auto out = std::make_shared<Batch>();
  out->ctx = ctx;
  out->schema = b.schema;
  out->num_rows = total;
  std::vector<BufPtr> keep;
  auto exchange = [&](const char* sendbuf, const std::vector<int64_t>& soff, char* recvbuf,
                      const std::vector<int64_t>& roff, size_t w) {
    B2_NCCL(ncclGroupStart());
    for (int p = 0; p < nr; ++p) {
      size_t sb = (size_t)(soff[p + 1] - soff[p]) * w, rb = (size_t)(roff[p + 1] - roff[p]) * w;
      if (sb) B2_NCCL(ncclSend(sendbuf + (size_t)soff[p] * w, sb, ncclChar, p, ctx->comm->nccl, ctx->stream));
      if (rb) B2_NCCL(ncclRecv(recvbuf + (size_t)roff[p] * w, rb, ncclChar, p, ctx->comm->nccl, ctx->stream));
      if (p != me) ctx->bytes_sent += (int64_t)sb;
    }
    B2_NCCL(ncclGroupEnd());
  };
  for (size_t c = 0; c < ncol; ++c) {
    const Column& col = b.cols[c];
    Column o;
    o.dtype = col.dtype;
    o.length = total;
    if (!col.data) {  // elided by column pruning: nothing to move
      out->cols.push_back(std::move(o));
      continue;
    }
    if (col.dtype == B2_UTF8) {
      auto lens = dev_alloc(ctx, (size_t)(b.num_rows + 1) * 4);
      if (b.num_rows) {
        lens_kernel<<<grid_for(ctx, 8, b.num_rows, 256), 256, 0, ctx->stream>>>(col.values<int32_t>(), b.num_rows,
                                                                               lens->as<int32_t>());
        B2_LAUNCHED(ctx);
      }
      auto rlens = dev_alloc(ctx, (size_t)(total + 1) * 4);
      exchange(lens->as<char>(), send_off, rlens->as<char>(), recv_off, 4);
      std::vector<int64_t> croff(nr + 1, 0);
      for (int s = 0; s < nr; ++s) croff[s + 1] = croff[s] + from(s, 1 + c);
      o.chars_len = croff[nr];
      o.chars = dev_alloc(ctx, (size_t)o.chars_len + 16);
      exchange(col.chars->as<char>(), chr_bounds[c], o.chars->as<char>(), croff, 1);
      o.data = dev_alloc_zero(ctx, (size_t)(total + 1) * 4);
      device_scan(ctx, I32In{rlens->as<int32_t>()}, total, OffOut{o.data->as<int32_t>(), total}, (int64_t*)nullptr);
      keep.push_back(lens);
      keep.push_back(rlens);
    } else {
      size_t w = dtype_width(col.dtype);
      o.data = dev_alloc(ctx, (size_t)total * w + 16);
      exchange(col.data->as<char>(), send_off, o.data->as<char>(), recv_off, w);
    }
    out->cols.push_back(std::move(o));
  }
  B2_SYNC(ctx);
  return out;
}

}  // namespace b2
