// scan.cuh — device-wide exclusive prefix sum of int32 values produced by a functor (three launches:
// per-CTA reduce, single-CTA scan of CTA sums, per-CTA scan with base).  Used by stream compaction,
// Utf8 gathers, join output sizing and the hash repartition.
#pragma once
#include "core.h"

namespace b2 {

constexpr int SCAN_THREADS = 256;
constexpr int SCAN_ITEMS = 8;
constexpr int SCAN_CHUNK = SCAN_THREADS * SCAN_ITEMS;

struct MaskIn {  // uint8 mask -> 0/1
  const uint8_t* m;
  __device__ __forceinline__ int operator()(int64_t i) const { return m[i] != 0; }
};
struct I32In {
  const int32_t* v;
  __device__ __forceinline__ int operator()(int64_t i) const { return v[i]; }
};
struct StrLenIn {  // length of string idx[i] (or i when idx is null)
  const int32_t* off;
  const uint32_t* idx;
  __device__ __forceinline__ int operator()(int64_t i) const {
    int64_t r = idx ? (int64_t)idx[i] : i;
    return r == 0xFFFFFFFFll ? 0 : off[r + 1] - off[r];
  }
};

__device__ __forceinline__ int block_exclusive_scan(int v, int* total) {
  __shared__ int warp_sums[SCAN_THREADS / 32];
  int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  int incl = v;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    int t = __shfl_up_sync(0xffffffffu, incl, d);
    if (lane >= d) incl += t;
  }
  if (lane == 31) warp_sums[warp] = incl;
  __syncthreads();
  int woff = 0, tot = 0;
#pragma unroll
  for (int w = 0; w < SCAN_THREADS / 32; ++w) {
    int s = warp_sums[w];
    if (w < warp) woff += s;
    tot += s;
  }
  __syncthreads();
  *total = tot;
  return woff + incl - v;
}

template <class In> __global__ void scan_reduce_kernel(In in, int64_t n, int64_t* block_sums) {
  int64_t base = (int64_t)blockIdx.x * SCAN_CHUNK + (int64_t)threadIdx.x * SCAN_ITEMS;
  int s = 0;
#pragma unroll
  for (int k = 0; k < SCAN_ITEMS; ++k)
    if (base + k < n) s += in(base + k);
  int total;
  block_exclusive_scan(s, &total);
  if (threadIdx.x == 0) block_sums[blockIdx.x] = total;
}
// single CTA: exclusive scan of the CTA sums in place; block_sums[nb] = grand total
static __global__ void scan_sums_kernel(int64_t* block_sums, int64_t nb) {
  __shared__ int64_t carry_s;
  __shared__ int64_t wsum[SCAN_THREADS / 32];
  if (threadIdx.x == 0) carry_s = 0;
  __syncthreads();
  int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int64_t b0 = 0; b0 < nb; b0 += SCAN_THREADS) {
    int64_t i = b0 + threadIdx.x;
    int64_t v = i < nb ? block_sums[i] : 0;
    int64_t incl = v;
    for (int d = 1; d < 32; d <<= 1) {
      int64_t t = __shfl_up_sync(0xffffffffu, incl, d);
      if (lane >= d) incl += t;
    }
    if (lane == 31) wsum[warp] = incl;
    __syncthreads();
    int64_t woff = 0, tot = 0;
    for (int w = 0; w < SCAN_THREADS / 32; ++w) {
      if (w < warp) woff += wsum[w];
      tot += wsum[w];
    }
    int64_t carry = carry_s;
    if (i < nb) block_sums[i] = carry + woff + incl - v;
    __syncthreads();
    if (threadIdx.x == 0) carry_s = carry + tot;
    __syncthreads();
  }
  if (threadIdx.x == 0) block_sums[nb] = carry_s;
}
// single CTA: in-place exclusive scan of v[0..n); v[n] = total.  Each thread owns 8 consecutive values per round
// (all loads in flight at once), so ~30 k tile counts take two rounds.
constexpr int SC_THREADS = 1024, SC_ITEMS = 8;
static __global__ void __launch_bounds__(SC_THREADS) scan_i64_kernel(int64_t* v, int64_t n) {
  __shared__ int64_t wsum[SC_THREADS / 32];
  __shared__ int64_t carry_s;
  if (threadIdx.x == 0) carry_s = 0;
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int64_t c0 = 0; c0 < n; c0 += SC_THREADS * SC_ITEMS) {
    int64_t i0 = c0 + (int64_t)threadIdx.x * SC_ITEMS;
    int64_t x[SC_ITEMS], s = 0;
#pragma unroll
    for (int j = 0; j < SC_ITEMS; ++j) {
      x[j] = i0 + j < n ? v[i0 + j] : 0;
      s += x[j];
    }
    int64_t incl = s;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      int64_t t = __shfl_up_sync(0xffffffffu, incl, d);
      if (lane >= d) incl += t;
    }
    if (lane == 31) wsum[warp] = incl;
    __syncthreads();
    int64_t woff = 0, tot = 0;
    for (int w = 0; w < SC_THREADS / 32; ++w) {
      int64_t t = wsum[w];
      if (w < warp) woff += t;
      tot += t;
    }
    int64_t pre = carry_s + woff + incl - s;
#pragma unroll
    for (int j = 0; j < SC_ITEMS; ++j) {
      if (i0 + j < n) v[i0 + j] = pre;
      pre += x[j];
    }
    __syncthreads();
    if (threadIdx.x == 0) carry_s += tot;
    __syncthreads();
  }
  if (threadIdx.x == 0) v[n] = carry_s;
}

// Multi-CTA form for long inputs (one count per 256-row warp tile of a 600 M-row probe side = 2.3 M values, which the
// single CTA above would walk in ~290 dependent rounds): CTA b scans chunk b of SC_CHUNK values in place, exclusive and
// chunk-local, and stores the chunk total in sums[b]; `scan_i64_kernel` then scans the totals, and the add kernel folds
// each chunk's offset back in and writes v[n] = grand total.  See `scan_i64` (repartition.cu) for the host side.
constexpr int64_t SC_CHUNK = (int64_t)SC_THREADS * SC_ITEMS;
static __global__ void __launch_bounds__(SC_THREADS) scan_i64_chunk_kernel(int64_t* v, int64_t n, int64_t* sums) {
  __shared__ int64_t wsum[SC_THREADS / 32];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  int64_t i0 = blockIdx.x * SC_CHUNK + (int64_t)threadIdx.x * SC_ITEMS;
  int64_t x[SC_ITEMS], s = 0;
#pragma unroll
  for (int j = 0; j < SC_ITEMS; ++j) {
    x[j] = i0 + j < n ? v[i0 + j] : 0;
    s += x[j];
  }
  int64_t incl = s;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    int64_t t = __shfl_up_sync(0xffffffffu, incl, d);
    if (lane >= d) incl += t;
  }
  if (lane == 31) wsum[warp] = incl;
  __syncthreads();
  int64_t woff = 0, tot = 0;
  for (int w = 0; w < SC_THREADS / 32; ++w) {
    int64_t t = wsum[w];
    if (w < warp) woff += t;
    tot += t;
  }
  int64_t pre = woff + incl - s;
#pragma unroll
  for (int j = 0; j < SC_ITEMS; ++j) {
    if (i0 + j < n) v[i0 + j] = pre;
    pre += x[j];
  }
  if (threadIdx.x == 0) sums[blockIdx.x] = tot;
}
static __global__ void __launch_bounds__(SC_THREADS) scan_i64_add_kernel(int64_t* v, int64_t n, const int64_t* __restrict__ sums,
                                                                        int64_t n_chunks) {
  const int64_t add = sums[blockIdx.x];
  int64_t i0 = blockIdx.x * SC_CHUNK + (int64_t)threadIdx.x * SC_ITEMS;
#pragma unroll
  for (int j = 0; j < SC_ITEMS; ++j)
    if (i0 + j < n) v[i0 + j] += add;
  if (blockIdx.x == 0 && threadIdx.x == 0) v[n] = sums[n_chunks];
}

// Out functor receives (i, value, exclusive_prefix)
template <class In, class Out>
__global__ void scan_apply_kernel(In in, int64_t n, const int64_t* block_sums, Out out) {
  int64_t base = (int64_t)blockIdx.x * SCAN_CHUNK + (int64_t)threadIdx.x * SCAN_ITEMS;
  int v[SCAN_ITEMS];
  int s = 0;
#pragma unroll
  for (int k = 0; k < SCAN_ITEMS; ++k) {
    v[k] = base + k < n ? in(base + k) : 0;
    s += v[k];
  }
  int total;
  int64_t pre = block_sums[blockIdx.x] + block_exclusive_scan(s, &total);
#pragma unroll
  for (int k = 0; k < SCAN_ITEMS; ++k) {
    if (base + k < n) out(base + k, v[k], pre);
    pre += v[k];
  }
}

void scan_i64(Ctx* ctx, int64_t* v, int64_t n);  // repartition.cu: in-place exclusive scan, v[n] = total (chunked when large)

// Runs the three launches.  Returns the device buffer of CTA sums; sums[nb] holds the grand total
// (int64).  `total_host` (optional) receives it after a stream sync.
template <class In, class Out>
BufPtr device_scan(Ctx* ctx, In in, int64_t n, Out out, int64_t* total_host) {
  int64_t nb = (n + SCAN_CHUNK - 1) / SCAN_CHUNK;
  auto sums = dev_alloc_zero(ctx, (size_t)(nb + 2) * 8);
  if (n > 0) {
    scan_reduce_kernel<In><<<(unsigned)nb, SCAN_THREADS, 0, ctx->stream>>>(in, n, sums->as<int64_t>());
    B2_LAUNCHED(ctx);
    if (nb > 4 * SCAN_THREADS) {  // tens of thousands of CTA sums: the single small CTA walks them in ~nb / 256 rounds
      scan_i64(ctx, sums->as<int64_t>(), nb);
    } else {
      scan_sums_kernel<<<1, SCAN_THREADS, 0, ctx->stream>>>(sums->as<int64_t>(), nb);
      B2_LAUNCHED(ctx);
    }
    scan_apply_kernel<In, Out><<<(unsigned)nb, SCAN_THREADS, 0, ctx->stream>>>(in, n, sums->as<int64_t>(), out);
    B2_LAUNCHED(ctx);
  }
  if (total_host) {
    *total_host = 0;
    if (n > 0) {
      B2_READBACK(ctx, total_host, sums->as<int64_t>() + nb, 8);
    }
  }
  return sums;
}

}  // namespace b2
