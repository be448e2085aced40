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

// All-gather of a small batch (the Partial -> Final merge hop).  One NCCL all-gather of a packed staging
// buffer: per column a region of max_rows values (Utf8: lengths + a chars region of max_chars bytes).
BatchPtr comm_allgather_batch(Ctx* ctx, const Batch& local) {
  if (!ctx->comm) fail(B2_ERR_INVALID_ARG, "no communicator (call b2_comm_init)");
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
