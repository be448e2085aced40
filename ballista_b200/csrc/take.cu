// take.cu — selection vectors and gathers: mask -> ordered row indices (stream compaction), and
// index-driven column gathers (`take`), the building blocks of FilterExec (arrow::compute::filter),
// HashJoinExec output assembly and RepartitionExec (arrow::compute::take).  SURVEY §8a rows A, F, G.
#include <cstring>

#include "ops.h"
#include "scan.cuh"

namespace b2 {

struct IndexOut {
  uint32_t* idx;
  __device__ __forceinline__ void operator()(int64_t i, int v, int64_t pre) const {
    if (v) idx[pre] = (uint32_t)i;
  }
};

// order-preserving compaction of the rows where mask != 0
BufPtr mask_to_indices(Ctx* ctx, const uint8_t* mask, int64_t n, int64_t* count) {
  if (n >= (1ll << 32) - 1) fail(B2_ERR_NOT_IMPLEMENTED, "batches beyond 2^32-2 rows");
  auto idx = dev_alloc(ctx, (size_t)(n + 1) * 4);
  device_scan(ctx, MaskIn{mask}, n, IndexOut{idx->as<uint32_t>()}, count);
  return idx;
}

// idx == 0xFFFFFFFF marks a null-extended row (outer joins): its value slot is zero and its validity byte 0
template <class T> __global__ void gather_kernel(const T* in, const uint32_t* idx, int64_t m, T* out) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  for (; i < m; i += (int64_t)gridDim.x * blockDim.x) {
    uint32_t r = idx[i];
    out[i] = r == 0xFFFFFFFFu ? T() : in[r];
  }
}
struct StrOffOut {
  int32_t* off;
  int64_t n;
  __device__ __forceinline__ void operator()(int64_t i, int v, int64_t pre) const {
    off[i] = (int32_t)pre;
    if (i == n - 1) off[n] = (int32_t)(pre + v);
  }
};
__global__ void gather_chars_kernel(const int32_t* in_off, const uint8_t* in_chr, const uint32_t* idx, int64_t m,
                                    const int32_t* out_off, uint8_t* out_chr) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  for (; i < m; i += (int64_t)gridDim.x * blockDim.x) {
    int64_t r = idx[i];
    if (r == 0xFFFFFFFFll) continue;
    int32_t b = in_off[r], len = in_off[r + 1] - b, o = out_off[i];
    for (int q = 0; q < len; ++q) out_chr[o + q] = in_chr[b + q];
  }
}

Column take_column(Ctx* ctx, const Column& c, const uint32_t* idx, int64_t m, bool null_extended) {
  Column out;
  out.dtype = c.dtype;
  out.length = m;
  if (!c.data) return out;  // elided by column pruning upstream: stays elided
  if (c.valid || null_extended) out.valid = valid_gather(ctx, c.validity(), idx, m);
  int grid = grid_for(ctx, 8, m, 256);
  if (c.dtype == B2_UTF8) {
    out.data = dev_alloc_zero(ctx, (size_t)(m + 1) * 4);
    int64_t total = 0;
    device_scan(ctx, StrLenIn{c.values<int32_t>(), idx}, m, StrOffOut{out.data->as<int32_t>(), m}, &total);
    if (total >= (1ll << 31)) fail(B2_ERR_EXECUTION, "Utf8 column beyond 2 GiB after take");
    out.chars_len = total;
    out.chars = dev_alloc(ctx, (size_t)total + 16);
    if (m) {
      gather_chars_kernel<<<grid, 256, 0, ctx->stream>>>(c.values<int32_t>(), c.chars->as<uint8_t>(), idx, m,
                                                        out.data->as<int32_t>(), out.chars->as<uint8_t>());
      B2_LAUNCHED(ctx);
    }
    return out;
  }
  size_t w = dtype_width(c.dtype);
  out.data = dev_alloc(ctx, (size_t)m * w + 16);
  if (m) {
    switch (w) {
      case 1: gather_kernel<uint8_t><<<grid, 256, 0, ctx->stream>>>(c.values<uint8_t>(), idx, m, out.data->as<uint8_t>()); break;
      case 4: gather_kernel<int32_t><<<grid, 256, 0, ctx->stream>>>(c.values<int32_t>(), idx, m, out.data->as<int32_t>()); break;
      case 8: gather_kernel<int64_t><<<grid, 256, 0, ctx->stream>>>(c.values<int64_t>(), idx, m, out.data->as<int64_t>()); break;
      default: fail(B2_ERR_NOT_IMPLEMENTED, "take on this type");
    }
    B2_LAUNCHED(ctx);
  }
  return out;
}

// Several fixed-width columns gathered through ONE index vector in one launch (a join's output side): the index is read
// once and every thread has one independent load per column in flight.
struct GatherCols {
  int n;
  const char* in[8];
  char* out[8];
  int width[8];
};
__global__ void gather_multi_kernel(const __grid_constant__ GatherCols g, const uint32_t* __restrict__ idx, int64_t m) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  for (; i < m; i += (int64_t)gridDim.x * blockDim.x) {
    const uint32_t r = idx[i];
    int64_t v[8];
#pragma unroll
    for (int c = 0; c < 8; ++c)
      if (c < g.n) {
        if (g.width[c] == 8) v[c] = reinterpret_cast<const int64_t*>(g.in[c])[r];
        else if (g.width[c] == 4) v[c] = reinterpret_cast<const int32_t*>(g.in[c])[r];
        else v[c] = reinterpret_cast<const uint8_t*>(g.in[c])[r];
      }
#pragma unroll
    for (int c = 0; c < 8; ++c)
      if (c < g.n) {
        if (g.width[c] == 8) reinterpret_cast<int64_t*>(g.out[c])[i] = v[c];
        else if (g.width[c] == 4) reinterpret_cast<int32_t*>(g.out[c])[i] = (int32_t)v[c];
        else reinterpret_cast<uint8_t*>(g.out[c])[i] = (uint8_t)v[c];
      }
  }
}
std::vector<Column> take_columns(Ctx* ctx, const std::vector<const Column*>& cols, const uint32_t* idx, int64_t m,
                                 bool null_extended) {
  std::vector<Column> out(cols.size());
  std::vector<size_t> fused;
  for (size_t k = 0; k < cols.size(); ++k) {
    const Column& c = *cols[k];
    const size_t w = c.dtype == B2_UTF8 ? 0 : dtype_width(c.dtype);
    if (c.data && !c.valid && !null_extended && m && (w == 1 || w == 4 || w == 8)) fused.push_back(k);
    else out[k] = take_column(ctx, c, idx, m, null_extended);
  }
  if (fused.size() == 1) {
    out[fused[0]] = take_column(ctx, *cols[fused[0]], idx, m, null_extended);
    fused.clear();
  }
  for (size_t g0 = 0; g0 < fused.size(); g0 += 8) {
    GatherCols g;
    memset(&g, 0, sizeof g);
    g.n = (int)std::min<size_t>(8, fused.size() - g0);
    for (int q = 0; q < g.n; ++q) {
      const Column& c = *cols[fused[g0 + q]];
      Column& o = out[fused[g0 + q]];
      o.dtype = c.dtype;
      o.length = m;
      const size_t w = dtype_width(c.dtype);
      o.data = dev_alloc(ctx, (size_t)m * w + 16);
      g.in[q] = c.data->as<char>();
      g.out[q] = o.data->as<char>();
      g.width[q] = (int)w;
    }
    gather_multi_kernel<<<grid_for(ctx, 8, m, 256), 256, 0, ctx->stream>>>(g, idx, m);
    B2_LAUNCHED(ctx);
  }
  return out;
}

BatchPtr take_batch(Ctx* ctx, const Batch& b, const uint32_t* idx, int64_t m) {
  auto out = std::make_shared<Batch>();
  out->ctx = ctx;
  out->schema = b.schema;
  out->num_rows = m;
  for (auto& c : b.cols) out->cols.push_back(take_column(ctx, c, idx, m));
  return out;
}

// a column of `n` NULLs (the absent side of an outer join's unmatched rows)
Column null_column(Ctx* ctx, int dtype, int64_t n) {
  Column c;
  c.dtype = dtype;
  c.length = n;
  size_t w = dtype == B2_UTF8 ? 4 : dtype_width(dtype);
  c.data = dev_alloc_zero(ctx, (size_t)(n + 1) * w + 16);
  if (dtype == B2_UTF8) c.chars = dev_alloc(ctx, 16);
  c.valid = dev_alloc_zero(ctx, (size_t)n + 16);
  return c;
}

BatchPtr empty_batch(Ctx* ctx, const Schema& schema) {
  auto b = std::make_shared<Batch>();
  b->ctx = ctx;
  b->schema = schema;
  for (auto& f : schema.fields) {
    Column c;
    c.dtype = f.dtype;
    c.data = dev_alloc_zero(ctx, 16);
    if (f.dtype == B2_UTF8) c.chars = dev_alloc(ctx, 16);
    b->cols.push_back(c);
  }
  return b;
}

// concatenate batches of one schema (CoalesceBatchesExec / MergeExec collection / join build side)
__global__ void shift_offsets_kernel(const int32_t* in, int64_t n, int32_t delta, int32_t* out) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  for (; i < n; i += (int64_t)gridDim.x * blockDim.x) out[i] = in[i] + delta;
}

BatchPtr concat_batches(Ctx* ctx, const Schema& schema, const std::vector<BatchPtr>& batches) {
  if (batches.size() == 1) return batches[0];
  auto out = std::make_shared<Batch>();
  out->ctx = ctx;
  out->schema = schema;
  int64_t n = 0;
  for (auto& b : batches) n += b->num_rows;
  out->num_rows = n;
  for (size_t ci = 0; ci < schema.fields.size(); ++ci) {
    Column c;
    c.dtype = schema.fields[ci].dtype;
    c.length = n;
    if (!batches[0]->cols[ci].data) {  // elided by column pruning
      out->cols.push_back(std::move(c));
      continue;
    }
    bool any_valid = false;
    for (auto& b : batches) any_valid |= (bool)b->cols[ci].valid;
    if (any_valid) {
      c.valid = dev_alloc(ctx, (size_t)n + 16);
      int64_t row = 0;
      for (auto& b : batches) {
        const Column& s = b->cols[ci];
        if (s.length) {
          if (s.valid)
            B2_CUDA(cudaMemcpyAsync(c.valid->as<uint8_t>() + row, s.valid->ptr, (size_t)s.length, cudaMemcpyDeviceToDevice,
                                    ctx->stream));
          else
            B2_CUDA(cudaMemsetAsync(c.valid->as<uint8_t>() + row, 1, (size_t)s.length, ctx->stream));
        }
        row += s.length;
      }
    }
    if (c.dtype == B2_UTF8) {
      int64_t chars = 0;
      for (auto& b : batches) chars += b->cols[ci].chars_len;
      if (chars >= (1ll << 31)) fail(B2_ERR_EXECUTION, "Utf8 column beyond 2 GiB after concat");
      c.data = dev_alloc_zero(ctx, (size_t)(n + 1) * 4);
      c.chars = dev_alloc(ctx, (size_t)chars + 16);
      c.chars_len = chars;
      int64_t row = 0, ch = 0;
      for (auto& b : batches) {
        const Column& s = b->cols[ci];
        if (s.length) {
          shift_offsets_kernel<<<grid_for(ctx, 8, s.length + 1, 256), 256, 0, ctx->stream>>>(
              s.values<int32_t>(), s.length + 1, (int32_t)ch, c.data->as<int32_t>() + row);
          B2_LAUNCHED(ctx);
          if (s.chars_len)
            B2_CUDA(cudaMemcpyAsync(c.chars->as<uint8_t>() + ch, s.chars->ptr, (size_t)s.chars_len,
                                    cudaMemcpyDeviceToDevice, ctx->stream));
        }
        row += s.length;
        ch += s.chars_len;
      }
    } else {
      size_t w = dtype_width(c.dtype);
      c.data = dev_alloc(ctx, (size_t)n * w + 16);
      int64_t row = 0;
      for (auto& b : batches) {
        const Column& s = b->cols[ci];
        if (s.length)
          B2_CUDA(cudaMemcpyAsync(c.data->as<char>() + (size_t)row * w, s.data->ptr, (size_t)s.length * w,
                                  cudaMemcpyDeviceToDevice, ctx->stream));
        row += s.length;
      }
    }
    out->cols.push_back(std::move(c));
  }
  return out;
}

}  // namespace b2
