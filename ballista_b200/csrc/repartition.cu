// repartition.cu — hash repartition kernels (RepartitionExec(Partitioning::Hash)):
// destination = hash(packed key) % n_out; a stable, destination-major permutation is built with one
// compaction scan per destination, so every column is then moved by a single gather and each
// destination's rows are one contiguous slice (the send segment of the all-to-all).
#include "ops.h"
#include "scan.cuh"

namespace b2 {

__global__ void dest_kernel(const Key128* keys, int64_t n, int n_out, uint8_t* dest) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  for (; i < n; i += (int64_t)gridDim.x * blockDim.x) dest[i] = (uint8_t)(key_hash(keys[i]) % (u64)n_out);
}
struct DestIn {
  const uint8_t* dest;
  int d;
  __device__ __forceinline__ int operator()(int64_t i) const { return dest[i] == d; }
};
struct PermOut {
  uint32_t* idx;  // already offset to this destination's segment
  __device__ __forceinline__ void operator()(int64_t i, int v, int64_t pre) const {
    if (v) idx[pre] = (uint32_t)i;
  }
};

BufPtr hash_partition_indices(Ctx* ctx, const Key128* keys, int64_t n, int n_out, std::vector<int64_t>* counts) {
  if (n_out < 1 || n_out > 255) fail(B2_ERR_INVALID_ARG, "hash repartition supports 1..255 output partitions");
  if (n >= (1ll << 32) - 1) fail(B2_ERR_NOT_IMPLEMENTED, "batches beyond 2^32-2 rows");
  auto dest = dev_alloc(ctx, (size_t)n + 16);
  auto perm = dev_alloc(ctx, (size_t)(n + 1) * 4);
  counts->assign(n_out, 0);
  if (n == 0) return perm;
  dest_kernel<<<grid_for(ctx, 8, n, 256), 256, 0, ctx->stream>>>(keys, n, n_out, dest->as<uint8_t>());
  B2_LAUNCHED(ctx);
  int64_t base = 0;
  for (int d = 0; d < n_out; ++d) {
    int64_t c = 0;
    device_scan(ctx, DestIn{dest->as<uint8_t>(), d}, n, PermOut{perm->as<uint32_t>() + base}, &c);
    (*counts)[d] = c;
    base += c;
  }
  return perm;
}

// rows [begin, end) of a batch as a new batch (device-to-device copies; Utf8 offsets rebased)
__global__ void rebase_kernel(const int32_t* in, int64_t n, int32_t base, int32_t* out) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  for (; i < n; i += (int64_t)gridDim.x * blockDim.x) out[i] = in[i] - base;
}
BatchPtr slice_batch(Ctx* ctx, const Batch& b, int64_t begin, int64_t end) {
  auto out = std::make_shared<Batch>();
  out->ctx = ctx;
  out->schema = b.schema;
  int64_t m = end - begin;
  out->num_rows = m;
  for (auto& c : b.cols) {
    Column o;
    o.dtype = c.dtype;
    o.length = m;
    if (c.dtype == B2_UTF8) {
      int32_t ends[2] = {0, 0};
      if (c.length) {
        B2_CUDA(cudaMemcpyAsync(&ends[0], c.values<int32_t>() + begin, 4, cudaMemcpyDeviceToHost, ctx->stream));
        B2_READBACK(ctx, &ends[1], c.values<int32_t>() + end, 4);
      }
      o.data = dev_alloc_zero(ctx, (size_t)(m + 1) * 4);
      o.chars_len = ends[1] - ends[0];
      o.chars = dev_alloc(ctx, (size_t)o.chars_len + 16);
      if (m) {
        rebase_kernel<<<grid_for(ctx, 8, m + 1, 256), 256, 0, ctx->stream>>>(c.values<int32_t>() + begin, m + 1,
                                                                            ends[0], o.data->as<int32_t>());
        B2_LAUNCHED(ctx);
        if (o.chars_len)
          B2_CUDA(cudaMemcpyAsync(o.chars->ptr, c.chars->as<uint8_t>() + ends[0], (size_t)o.chars_len,
                                  cudaMemcpyDeviceToDevice, ctx->stream));
      }
    } else {
      size_t w = dtype_width(c.dtype);
      o.data = dev_alloc(ctx, (size_t)m * w + 16);
      if (m)
        B2_CUDA(cudaMemcpyAsync(o.data->ptr, c.data->as<char>() + (size_t)begin * w, (size_t)m * w,
                                cudaMemcpyDeviceToDevice, ctx->stream));
    }
    out->cols.push_back(std::move(o));
  }
  return out;
}

}  // namespace b2
