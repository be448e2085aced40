// validity.cu — NULL handling.  A column's validity lives in HBM as one byte per row (Column::valid; absent = no
// nulls), which lets every operator treat it like a mask column: gathers move it with the values, predicates AND it
// into their mask (a null predicate is false), aggregates turn it into per-argument counts (SUM / AVG /
// COUNT skip nulls, A.2), joins fold it into their row filter (null keys never match, A.3).  Arrow's bit-packed
// bitmaps (rust/core/proto/ballista.proto:14-45 IsNull / IsNotNull operate on them) exist only at the C Data
// Interface: expanded at import, packed at export.
#include "core.h"

namespace b2 {

__global__ void valid_and_kernel(const uint8_t* a, const uint8_t* b, int64_t n, uint8_t* out) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  for (; i < n; i += (int64_t)gridDim.x * blockDim.x) out[i] = (a[i] != 0) & (b[i] != 0);
}
BufPtr valid_and(Ctx* ctx, const BufPtr& a, const BufPtr& b, int64_t n) {
  if (!a) return b;
  if (!b) return a;
  auto out = dev_alloc(ctx, (size_t)n + 16);
  if (n) {
    valid_and_kernel<<<grid_for(ctx, 8, n, 256), 256, 0, ctx->stream>>>(a->as<uint8_t>(), b->as<uint8_t>(), n,
                                                                       out->as<uint8_t>());
    B2_LAUNCHED(ctx);
  }
  return out;
}

__global__ void bitmap_expand_kernel(const uint8_t* bits, int64_t bit_off, int64_t n, uint8_t* out) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  for (; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    int64_t b = bit_off + i;
    out[i] = (bits[b >> 3] >> (b & 7)) & 1;
  }
}
BufPtr valid_from_bitmap(Ctx* ctx, const uint8_t* host_bits, int64_t bit_off, int64_t n, cudaStream_t st) {
  int64_t first = bit_off >> 3, last = (bit_off + n + 7) >> 3;
  auto bits = dev_alloc(ctx, (size_t)(last - first) + 16);
  auto out = dev_alloc(ctx, (size_t)n + 16);
  if (n) {
    B2_CUDA(cudaMemcpyAsync(bits->ptr, host_bits + first, (size_t)(last - first), cudaMemcpyHostToDevice, st));
    bitmap_expand_kernel<<<grid_for(ctx, 8, n, 256), 256, 0, st>>>(bits->as<uint8_t>(), bit_off & 7, n,
                                                                  out->as<uint8_t>());
    B2_LAUNCHED(ctx);
  }
  return out;
}

// one thread packs 8 rows into one bitmap byte
__global__ void bitmap_pack_kernel(const uint8_t* valid, int64_t n, uint8_t* bits, unsigned long long* nulls,
                                   const int64_t* n_dev) {
  int64_t nb = (n + 7) >> 3;
  const int64_t counted = n_dev ? (*n_dev < n ? *n_dev : n) : n;
  int64_t j = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  unsigned long long missing = 0;
  for (; j < nb; j += (int64_t)gridDim.x * blockDim.x) {
    unsigned v = 0;
    for (int q = 0; q < 8; ++q) {
      int64_t i = j * 8 + q;
      if (i < n) {
        if (valid[i]) v |= 1u << q;
        else if (i < counted) missing++;
      }
    }
    bits[j] = (uint8_t)v;
  }
  if (missing) atomicAdd(nulls, missing);
}
BufPtr valid_to_bitmap(Ctx* ctx, const uint8_t* valid, int64_t n, int64_t* nulls_dev, const int64_t* n_dev) {
  auto bits = dev_alloc(ctx, (size_t)((n + 7) >> 3) + 16);
  B2_CUDA(cudaMemsetAsync(nulls_dev, 0, 8, ctx->stream));
  if (n) {
    bitmap_pack_kernel<<<grid_for(ctx, 8, (n + 7) >> 3, 256), 256, 0, ctx->stream>>>(
        valid, n, bits->as<uint8_t>(), reinterpret_cast<unsigned long long*>(nulls_dev), n_dev);
    B2_LAUNCHED(ctx);
  }
  return bits;
}

__global__ void valid_gather_kernel(const uint8_t* valid, const uint32_t* idx, int64_t m, uint8_t* out) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  for (; i < m; i += (int64_t)gridDim.x * blockDim.x) {
    uint32_t r = idx[i];
    out[i] = r == 0xFFFFFFFFu ? 0 : (valid ? valid[r] : 1);
  }
}
BufPtr valid_gather(Ctx* ctx, const uint8_t* valid, const uint32_t* idx, int64_t m) {
  auto out = dev_alloc(ctx, (size_t)m + 16);
  if (m) {
    valid_gather_kernel<<<grid_for(ctx, 8, m, 256), 256, 0, ctx->stream>>>(valid, idx, m, out->as<uint8_t>());
    B2_LAUNCHED(ctx);
  }
  return out;
}

__global__ void valid_not_kernel(const uint8_t* valid, int64_t n, uint8_t* out) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  for (; i < n; i += (int64_t)gridDim.x * blockDim.x) out[i] = valid[i] == 0;
}
BufPtr valid_not(Ctx* ctx, const uint8_t* valid, int64_t n) {
  auto out = dev_alloc(ctx, (size_t)n + 16);
  if (!n) return out;
  if (!valid) {
    B2_CUDA(cudaMemsetAsync(out->ptr, 0, (size_t)n, ctx->stream));
  } else {
    valid_not_kernel<<<grid_for(ctx, 8, n, 256), 256, 0, ctx->stream>>>(valid, n, out->as<uint8_t>());
    B2_LAUNCHED(ctx);
  }
  return out;
}
BufPtr valid_copy(Ctx* ctx, const uint8_t* valid, int64_t n) {
  auto out = dev_alloc(ctx, (size_t)n + 16);
  if (!n) return out;
  if (!valid) B2_CUDA(cudaMemsetAsync(out->ptr, 1, (size_t)n, ctx->stream));
  else B2_CUDA(cudaMemcpyAsync(out->ptr, valid, (size_t)n, cudaMemcpyDeviceToDevice, ctx->stream));
  return out;
}

}  // namespace b2
