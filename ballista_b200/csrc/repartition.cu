// repartition.cu — hash repartition kernels (RepartitionExec(Partitioning::Hash)):
// destination = part_dest(part_hash(packed key), n_out) (table.cuh); a stable, destination-major permutation is built in two passes over
// the keys (per-tile histogram -> one scan -> stable scatter of row ids), so every column is then moved by a
// single gather and each destination's rows are one contiguous slice (the send segment of the all-to-all).
// Beyond 32 output partitions the permutation falls back to one compaction scan per destination.
#include <cstring>
#include <mutex>
#include <string>

#include "ops.h"
#include "rowpred.cuh"
#include "scan.cuh"

namespace b2 {

void scan_i64(Ctx* ctx, int64_t* v, int64_t n) {
  if (n <= 4 * SC_CHUNK) {
    scan_i64_kernel<<<1, SC_THREADS, 0, ctx->stream>>>(v, n);
    B2_LAUNCHED(ctx);
    return;
  }
  const int64_t n_chunks = (n + SC_CHUNK - 1) / SC_CHUNK;
  auto sums = dev_alloc(ctx, (size_t)(n_chunks + 1) * 8);
  scan_i64_chunk_kernel<<<(unsigned)n_chunks, SC_THREADS, 0, ctx->stream>>>(v, n, sums->as<int64_t>());
  B2_LAUNCHED(ctx);
  scan_i64(ctx, sums->as<int64_t>(), n_chunks);
  scan_i64_add_kernel<<<(unsigned)n_chunks, SC_THREADS, 0, ctx->stream>>>(v, n, sums->as<int64_t>(), n_chunks);
  B2_LAUNCHED(ctx);
}


__global__ void dest_kernel(const Key128* keys, int64_t n, int n_out, uint8_t* dest) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  for (; i < n; i += (int64_t)gridDim.x * blockDim.x) dest[i] = (uint8_t)part_dest(part_hash(keys[i]), n_out);
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

// ---- n_out <= 32: two passes over the keys (histogram, then stable scatter of row ids) ----
// A tile is RP_ITEMS sub-rows of RP_THREADS consecutive rows (thread t owns rows base + j * RP_THREADS + t, so
// every load is coalesced); the stable order inside a tile is (j, warp, lane).
constexpr int RP_THREADS = 256, RP_ITEMS = 8, RP_TILE = RP_THREADS * RP_ITEMS, RP_MAX_OUT = 32;
constexpr int RP_WARPS = RP_THREADS / 32;
constexpr uint8_t RP_DROPPED = 255;  // destination byte of a row the fused filter rejected

struct KeySrc {  // packed 16-byte keys, or a single Int32 / Date32 key column hashed as its packed form
  const Key128* k128;
  const int32_t* k32;
  bool table_order;  // split by the join table's bucket order (join.cu's partitioned build), not for an exchange
  __device__ __forceinline__ uint32_t hash(int64_t i) const {
    const uint32_t h = k128 ? part_hash(k128[i]) : part_hash((uint32_t)k32[i]);
    return table_order ? table_hash(h) : h;
  }
};

// pass 1: destination byte per row + per-tile histogram, stored destination-major (hist[d * n_tiles + tile]) so
// that ONE exclusive scan of the whole array yields every (destination, tile) output offset
struct SemiArgs {  // device view of a SemiFilter (ops.h); words == nullptr: no filter
  const uint32_t* words;
  uint32_t stride;
  uint32_t wmask[RP_MAX_OUT];
};
__global__ void __launch_bounds__(RP_THREADS)
    rp_hist_kernel(KeySrc keys, RowPred pred, const uint8_t* __restrict__ mask, SemiArgs semi, int64_t n, int n_out,
                   uint8_t* __restrict__ dest, int64_t* __restrict__ hist, int64_t n_tiles) {
  __shared__ int h[RP_MAX_OUT];
  for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    if (threadIdx.x < RP_MAX_OUT) h[threadIdx.x] = 0;
    __syncthreads();
#pragma unroll
    for (int j = 0; j < RP_ITEMS; ++j) {
      int64_t r = tile * RP_TILE + j * RP_THREADS + threadIdx.x;
      // rows rejected by the fused FilterExec (byte mask and / or inline predicate) get no destination
      bool live = r < n && (!mask || mask[r]) && row_pred(pred, r);
      int d = -1;
      if (live) {
        const uint32_t h = keys.hash(r);
        d = part_dest(h, n_out);
        if (semi.words) {  // semi-join reduction: the destination's build side cannot hold this key
          const uint32_t bits = bloom_bits(h);
          if ((__ldg(semi.words + (size_t)d * semi.stride + (h & semi.wmask[d])) & bits) != bits) {
            d = -1;
            live = false;
          }
        }
      }
      if (r < n) dest[r] = live ? (uint8_t)d : RP_DROPPED;
      unsigned peers = __match_any_sync(0xffffffffu, d);
      if (d >= 0 && (peers & ((1u << (threadIdx.x & 31)) - 1)) == 0) atomicAdd(&h[d], __popc(peers));
    }
    __syncthreads();
    if (threadIdx.x < n_out) hist[(int64_t)threadIdx.x * n_tiles + tile] = h[threadIdx.x];
    __syncthreads();
  }
}

// pass 2: stable rank of every row inside its (destination, tile) run -> perm[offset + rank] = row
__global__ void __launch_bounds__(RP_THREADS)
    rp_scatter_kernel(const uint8_t* __restrict__ dest, int64_t n, int n_out, const int64_t* __restrict__ off,
                      int64_t n_tiles, uint32_t* __restrict__ perm, const uint32_t* __restrict__ remap = nullptr) {
  __shared__ uint32_t cnt[RP_ITEMS][RP_WARPS][RP_MAX_OUT];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    for (int i = threadIdx.x; i < RP_ITEMS * RP_WARPS * RP_MAX_OUT; i += RP_THREADS) (&cnt[0][0][0])[i] = 0;
    __syncthreads();
    int d[RP_ITEMS], rank[RP_ITEMS];
#pragma unroll
    for (int j = 0; j < RP_ITEMS; ++j) {
      int64_t r = tile * RP_TILE + j * RP_THREADS + threadIdx.x;
      d[j] = r < n ? (int)dest[r] : -1;
      if (d[j] == RP_DROPPED) d[j] = -1;
      unsigned peers = __match_any_sync(0xffffffffu, d[j]);
      rank[j] = __popc(peers & ((1u << lane) - 1));
      if (d[j] >= 0 && rank[j] == 0) cnt[j][warp][d[j]] = __popc(peers);
    }
    __syncthreads();
    if (threadIdx.x < n_out) {  // per destination: exclusive scan over (j, warp) from the run's global offset
      uint32_t run = (uint32_t)off[(int64_t)threadIdx.x * n_tiles + tile];
      for (int j = 0; j < RP_ITEMS; ++j)
        for (int w = 0; w < RP_WARPS; ++w) {
          uint32_t c = cnt[j][w][threadIdx.x];
          cnt[j][w][threadIdx.x] = run;
          run += c;
        }
    }
    __syncthreads();
#pragma unroll
    for (int j = 0; j < RP_ITEMS; ++j)
      if (d[j] >= 0) {
        int64_t r = tile * RP_TILE + j * RP_THREADS + threadIdx.x;
        perm[cnt[j][warp][d[j]] + rank[j]] = remap ? remap[r] : (uint32_t)r;  // remap: the radix sort's current order
      }
    __syncthreads();
  }
}

static BufPtr partition_indices_scan_per_destination(Ctx* ctx, const Key128* keys, int64_t n, int n_out,
                                                     std::vector<int64_t>* counts) {
  auto dest = dev_alloc(ctx, (size_t)n + 16);
  auto perm = dev_alloc(ctx, (size_t)(n + 1) * 4);
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

// pass 2': instead of a permutation, move up to 8 fixed-width columns straight to their destination-major
// position (the shuffle's staging arena): element k of destination d's slice of column c lands at
// dst[c] + base[c * n_out + d] + k * width[c].  Reads are coalesced; a warp's stores fall into one run per destination.
struct MoveCols {
  int n;
  const char* src[8];
  char* dst[8];
  int width[8];
  const int64_t* base;  // device: [n][n_out] byte offset of slice d inside dst[c]
};
__global__ void __launch_bounds__(RP_THREADS)
    rp_move_kernel(const uint8_t* __restrict__ dest, int64_t n, int n_out, const int64_t* __restrict__ off,
                   int64_t n_tiles, MoveCols mc, uint32_t* __restrict__ perm) {
  __shared__ uint32_t cnt[RP_ITEMS][RP_WARPS][RP_MAX_OUT];
  __shared__ uint32_t start_s[RP_MAX_OUT];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (threadIdx.x < n_out) start_s[threadIdx.x] = (uint32_t)off[(int64_t)threadIdx.x * n_tiles];
  for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    for (int i = threadIdx.x; i < RP_ITEMS * RP_WARPS * RP_MAX_OUT; i += RP_THREADS) (&cnt[0][0][0])[i] = 0;
    __syncthreads();
    int d[RP_ITEMS], rank[RP_ITEMS];
#pragma unroll
    for (int j = 0; j < RP_ITEMS; ++j) {
      int64_t r = tile * RP_TILE + j * RP_THREADS + threadIdx.x;
      d[j] = r < n ? (int)dest[r] : -1;
      if (d[j] == RP_DROPPED) d[j] = -1;
      unsigned peers = __match_any_sync(0xffffffffu, d[j]);
      rank[j] = __popc(peers & ((1u << lane) - 1));
      if (d[j] >= 0 && rank[j] == 0) cnt[j][warp][d[j]] = __popc(peers);
    }
    __syncthreads();
    if (threadIdx.x < n_out) {
      uint32_t run = (uint32_t)off[(int64_t)threadIdx.x * n_tiles + tile];
      for (int j = 0; j < RP_ITEMS; ++j)
        for (int w = 0; w < RP_WARPS; ++w) {
          uint32_t c = cnt[j][w][threadIdx.x];
          cnt[j][w][threadIdx.x] = run;
          run += c;
        }
    }
    __syncthreads();
    uint32_t pos[RP_ITEMS];  // destination-major global position
#pragma unroll
    for (int j = 0; j < RP_ITEMS; ++j) pos[j] = d[j] >= 0 ? cnt[j][warp][d[j]] + rank[j] : 0u;
    if (perm) {
#pragma unroll
      for (int j = 0; j < RP_ITEMS; ++j)
        if (d[j] >= 0) perm[pos[j]] = (uint32_t)(tile * RP_TILE + j * RP_THREADS + threadIdx.x);
    }
    for (int c = 0; c < mc.n; ++c) {
      const int w = mc.width[c];
      if (w == 8) {
        int64_t v[RP_ITEMS];
#pragma unroll
        for (int j = 0; j < RP_ITEMS; ++j)
          if (d[j] >= 0) v[j] = reinterpret_cast<const int64_t*>(mc.src[c])[tile * RP_TILE + j * RP_THREADS + threadIdx.x];
#pragma unroll
        for (int j = 0; j < RP_ITEMS; ++j)
          if (d[j] >= 0)
            *reinterpret_cast<int64_t*>(mc.dst[c] + mc.base[c * n_out + d[j]] + (int64_t)(pos[j] - start_s[d[j]]) * 8) = v[j];
      } else if (w == 4) {
        int32_t v[RP_ITEMS];
#pragma unroll
        for (int j = 0; j < RP_ITEMS; ++j)
          if (d[j] >= 0) v[j] = reinterpret_cast<const int32_t*>(mc.src[c])[tile * RP_TILE + j * RP_THREADS + threadIdx.x];
#pragma unroll
        for (int j = 0; j < RP_ITEMS; ++j)
          if (d[j] >= 0)
            *reinterpret_cast<int32_t*>(mc.dst[c] + mc.base[c * n_out + d[j]] + (int64_t)(pos[j] - start_s[d[j]]) * 4) = v[j];
      } else {
#pragma unroll
        for (int j = 0; j < RP_ITEMS; ++j)
          if (d[j] >= 0)
            mc.dst[c][mc.base[c * n_out + d[j]] + (int64_t)(pos[j] - start_s[d[j]])] =
                mc.src[c][tile * RP_TILE + j * RP_THREADS + threadIdx.x];
      }
    }
    __syncthreads();
  }
}

// pass 2'' for a SPARSE split (a selective fused filter or semi-join filter dropped most rows): same result as
// `rp_move_kernel`, an order of magnitude fewer instructions per dropped row.  Each thread owns 8 CONSECUTIVE rows
// (one 8-byte load of destination bytes, one SIMD byte compare for "any survivor"); the tile's survivors are
// compacted in row order into shared memory by one block scan; warp 0 ranks them per destination with n_out ballots
// per 32 survivors; then one thread per survivor moves that row's columns.  Positions are the same stable
// destination-major ones: off[d][tile] + rank within (tile, d) in row order.
__global__ void __launch_bounds__(RP_THREADS)
    rp_move_sparse_kernel(const uint8_t* __restrict__ dest, int64_t n, int n_out, const int64_t* __restrict__ off,
                          int64_t n_tiles, MoveCols mc, uint32_t* __restrict__ perm) {
  __shared__ uint32_t ent[RP_TILE];   // survivor: row within the tile | destination << 16
  __shared__ uint32_t rnk[RP_TILE];   // its rank within (tile, destination)
  __shared__ uint32_t start_s[RP_MAX_OUT];
  const int lane = threadIdx.x & 31;
  if (threadIdx.x < n_out) start_s[threadIdx.x] = (uint32_t)off[(int64_t)threadIdx.x * n_tiles];
  __syncthreads();
  for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    const int64_t r0 = tile * RP_TILE + (int64_t)threadIdx.x * RP_ITEMS;
    unsigned long long bytes = ~0ull;  // RP_DROPPED in every lane of the word
    if (r0 + RP_ITEMS <= n) {
      bytes = __ldg(reinterpret_cast<const unsigned long long*>(dest + r0));  // dest is 256-byte aligned, r0 % 8 == 0
    } else {
      for (int j = 0; j < RP_ITEMS; ++j)
        if (r0 + j < n) bytes = (bytes & ~(0xFFull << (8 * j))) | ((unsigned long long)dest[r0 + j] << (8 * j));
    }
    // per byte: 0xFF where the row survives (destination byte != RP_DROPPED)
    const unsigned lo = __vcmpne4((unsigned)bytes, 0xFFFFFFFFu), hi = __vcmpne4((unsigned)(bytes >> 32), 0xFFFFFFFFu);
    const int mine = (__popc(lo) + __popc(hi)) >> 3;
    int total;
    int at = block_exclusive_scan(mine, &total);
    if (total == 0) continue;  // uniform
    if (mine) {
#pragma unroll
      for (int j = 0; j < RP_ITEMS; ++j) {
        const unsigned d = (unsigned)(bytes >> (8 * j)) & 0xFFu;
        if (d != RP_DROPPED) ent[at++] = (unsigned)(threadIdx.x * RP_ITEMS + j) | (d << 16);
      }
    }
    __syncthreads();
    if (threadIdx.x < 32) {  // stable rank per destination: n_out ballots per 32 survivors
      uint32_t seen = 0;     // lane d (< n_out) carries the running count of destination d
      for (int c0 = 0; c0 < total; c0 += 32) {
        const int c = c0 + lane;
        const int d = c < total ? (int)(ent[c] >> 16) : -1;
        uint32_t my_rank = 0;
        for (int q = 0; q < n_out; ++q) {
          const unsigned m = __ballot_sync(0xffffffffu, d == q);
          const uint32_t before = __shfl_sync(0xffffffffu, seen, q);
          if (d == q) my_rank = before + __popc(m & ((1u << lane) - 1u));
          if (lane == q) seen += __popc(m);
        }
        if (c < total) rnk[c] = my_rank;
      }
    }
    __syncthreads();
    for (int c = threadIdx.x; c < total; c += RP_THREADS) {
      const uint32_t e = ent[c];
      const int d = (int)(e >> 16);
      const int64_t row = tile * RP_TILE + (e & 0xFFFFu);
      const uint32_t pos = (uint32_t)off[(int64_t)d * n_tiles + tile] + rnk[c];  // destination-major global position
      if (perm) perm[pos] = (uint32_t)row;
      for (int k = 0; k < mc.n; ++k) {
        const int w = mc.width[k];
        char* to = mc.dst[k] + mc.base[k * n_out + d] + (int64_t)(pos - start_s[d]) * w;
        if (w == 8) {
          *reinterpret_cast<int64_t*>(to) = reinterpret_cast<const int64_t*>(mc.src[k])[row];
        } else if (w == 4) {
          *reinterpret_cast<int32_t*>(to) = reinterpret_cast<const int32_t*>(mc.src[k])[row];
        } else {
          for (int q = 0; q < w; ++q) to[q] = mc.src[k][row * w + q];
        }
      }
    }
    __syncthreads();
  }
}

// pass 2-push for the PUSH exchange (shuffle.cu): the same stable destination-major positions, but every destination's
// rows live in another GPU's memory, reached over NVLink.  A warp of `rp_move_kernel` stores 32 / n_out elements per
// destination per instruction — fine for local HBM behind the L2, 32-byte packets on the wire.  Here a tile's values of
// one column are first put in destination-major order in shared memory, then written out run by run with consecutive
// lanes on consecutive addresses, so a destination receives whole 128-byte lines.  addr[c * n_out + d] is the absolute
// address of this source's first element of column c in destination d's receive block.
struct PushCols {
  int n;
  const char* src[8];
  int width[8];
  const int64_t* addr;  // device: [n][n_out]
};
__global__ void __launch_bounds__(RP_THREADS, 4)
    rp_push_kernel(const uint8_t* __restrict__ dest, int64_t n, int n_out, const int64_t* __restrict__ off,
                   int64_t n_tiles, PushCols pc) {
  __shared__ uint32_t cnt[RP_ITEMS][RP_WARPS][RP_MAX_OUT];
  __shared__ uint32_t tcount[RP_MAX_OUT], tstart[RP_MAX_OUT + 1];
  __shared__ int64_t goff[RP_MAX_OUT];   // row offset of this tile's run inside my slice at destination d
  __shared__ int64_t addr_s[8 * RP_MAX_OUT];
  __shared__ uint8_t dst_of[RP_TILE];    // destination of the tile-local sorted position
  __shared__ int64_t stage[RP_TILE];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int i = threadIdx.x; i < pc.n * n_out; i += RP_THREADS) addr_s[i] = pc.addr[i];
  for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    for (int i = threadIdx.x; i < RP_ITEMS * RP_WARPS * RP_MAX_OUT; i += RP_THREADS) (&cnt[0][0][0])[i] = 0;
    __syncthreads();
    int d[RP_ITEMS], rank[RP_ITEMS];
#pragma unroll
    for (int j = 0; j < RP_ITEMS; ++j) {
      int64_t r = tile * RP_TILE + j * RP_THREADS + threadIdx.x;
      d[j] = r < n ? (int)dest[r] : -1;
      if (d[j] == RP_DROPPED) d[j] = -1;
      unsigned peers = __match_any_sync(0xffffffffu, d[j]);
      rank[j] = __popc(peers & ((1u << lane) - 1));
      if (d[j] >= 0 && rank[j] == 0) cnt[j][warp][d[j]] = __popc(peers);
    }
    __syncthreads();
    if (threadIdx.x < n_out) {  // per destination: exclusive scan over (j, warp), tile-local
      uint32_t run = 0;
      for (int j = 0; j < RP_ITEMS; ++j)
        for (int w = 0; w < RP_WARPS; ++w) {
          uint32_t c = cnt[j][w][threadIdx.x];
          cnt[j][w][threadIdx.x] = run;
          run += c;
        }
      tcount[threadIdx.x] = run;
      goff[threadIdx.x] = off[(int64_t)threadIdx.x * n_tiles + tile] - off[(int64_t)threadIdx.x * n_tiles];
    }
    __syncthreads();
    if (threadIdx.x == 0) {
      uint32_t run = 0;
      for (int q = 0; q < n_out; ++q) {
        tstart[q] = run;
        run += tcount[q];
      }
      tstart[n_out] = run;
    }
    __syncthreads();
    const uint32_t total = tstart[n_out];
    if (total == 0) continue;  // uniform: every row of the tile was dropped by the fused filter
    uint32_t lpos[RP_ITEMS];
#pragma unroll
    for (int j = 0; j < RP_ITEMS; ++j) {
      lpos[j] = d[j] >= 0 ? tstart[d[j]] + cnt[j][warp][d[j]] + rank[j] : 0u;
      if (d[j] >= 0) dst_of[lpos[j]] = (uint8_t)d[j];
    }
    for (int c = 0; c < pc.n; ++c) {
      const int w = pc.width[c];
      __syncthreads();  // dst_of written / the previous column's write-out has finished with `stage`
      if (w == 8) {
        int64_t v[RP_ITEMS];
#pragma unroll
        for (int j = 0; j < RP_ITEMS; ++j)
          if (d[j] >= 0) v[j] = reinterpret_cast<const int64_t*>(pc.src[c])[tile * RP_TILE + j * RP_THREADS + threadIdx.x];
#pragma unroll
        for (int j = 0; j < RP_ITEMS; ++j)
          if (d[j] >= 0) stage[lpos[j]] = v[j];
        __syncthreads();
        for (int q = 0; q < n_out; ++q) {  // run by run; a warp's 32 stores cover one aligned 256-byte span
          const int64_t a0 = addr_s[c * n_out + q] + goff[q] * 8;
          const int lead = (int)((a0 >> 3) & 31), cnt_q = (int)tcount[q];
          for (int k = (int)threadIdx.x - lead; k < cnt_q; k += RP_THREADS)
            if (k >= 0) *reinterpret_cast<int64_t*>(a0 + (int64_t)k * 8) = stage[tstart[q] + k];
        }
      } else if (w == 4) {
        int32_t* stage4 = reinterpret_cast<int32_t*>(stage);
        int32_t v[RP_ITEMS];
#pragma unroll
        for (int j = 0; j < RP_ITEMS; ++j)
          if (d[j] >= 0) v[j] = reinterpret_cast<const int32_t*>(pc.src[c])[tile * RP_TILE + j * RP_THREADS + threadIdx.x];
#pragma unroll
        for (int j = 0; j < RP_ITEMS; ++j)
          if (d[j] >= 0) stage4[lpos[j]] = v[j];
        __syncthreads();
        for (int q = 0; q < n_out; ++q) {  // a warp's 32 stores cover one aligned 128-byte line
          const int64_t a0 = addr_s[c * n_out + q] + goff[q] * 4;
          const int lead = (int)((a0 >> 2) & 31), cnt_q = (int)tcount[q];
          for (int k = (int)threadIdx.x - lead; k < cnt_q; k += RP_THREADS)
            if (k >= 0) *reinterpret_cast<int32_t*>(a0 + (int64_t)k * 4) = stage4[tstart[q] + k];
        }
      } else {
        uint8_t* stage1 = reinterpret_cast<uint8_t*>(stage);
#pragma unroll
        for (int j = 0; j < RP_ITEMS; ++j)
          if (d[j] >= 0)
            for (int b = 0; b < w; ++b)
              stage1[lpos[j] * w + b] = (uint8_t)pc.src[c][(tile * RP_TILE + j * RP_THREADS + threadIdx.x) * w + b];
        __syncthreads();
        for (uint32_t i = threadIdx.x; i < total; i += RP_THREADS) {
          const int q = dst_of[i];
          uint8_t* to = reinterpret_cast<uint8_t*>(addr_s[c * n_out + q] + (goff[q] + (int64_t)(i - tstart[q])) * w);
          for (int b = 0; b < w; ++b) to[b] = stage1[i * w + b];
        }
      }
    }
    __syncthreads();
  }
  __threadfence_system();  // the stores above cross NVLink: performed before the kernel (and the barrier behind it) ends
}

// ---- host side: phase 1 (destinations, histogram, scan: no host sync), then counts / permutation / column move ----
SplitState hash_split_begin(Ctx* ctx, const Key128* keys, const int32_t* key32, int64_t n, int n_out,
                            const RowPred* pred, const uint8_t* mask, const SemiFilter* semi, bool table_order) {
  if (n_out < 1 || n_out > RP_MAX_OUT) fail(B2_ERR_INVALID_ARG, "hash_split_begin supports 1..32 output partitions");
  if (n >= (1ll << 32) - 1) fail(B2_ERR_NOT_IMPLEMENTED, "batches beyond 2^32-2 rows");
  SplitState st;
  st.n = n;
  st.n_out = n_out;
  st.n_tiles = (n + RP_TILE - 1) / RP_TILE;
  st.dest = dev_alloc(ctx, (size_t)n + 16);
  st.hist = dev_alloc_zero(ctx, (size_t)(st.n_tiles * n_out + 2) * 8);
  if (n == 0) return st;
  int grid = grid_for(ctx, 8, n, RP_TILE);
  SemiArgs sa{nullptr, 0, {0}};
  if (semi) {
    if (semi->n != n_out) fail(B2_ERR_EXECUTION, "internal: semi-join filter does not match the exchange width");
    sa.words = semi->words->as<uint32_t>();
    sa.stride = semi->stride;
    for (int d = 0; d < n_out; ++d) sa.wmask[d] = semi->wmask[d];
  }
  rp_hist_kernel<<<grid, RP_THREADS, 0, ctx->stream>>>(KeySrc{keys, key32, table_order}, pred ? *pred : RowPred(), mask, sa, n, n_out,
                                                       st.dest->as<uint8_t>(), st.hist->as<int64_t>(), st.n_tiles);
  B2_LAUNCHED(ctx);
  scan_i64(ctx, st.hist->as<int64_t>(), st.n_tiles * n_out);
  return st;
}
// start offset of every destination = scanned hist[d * n_tiles]: gathers the n_out + 1 starts into `out` (device)
__global__ void rp_starts_kernel(const int64_t* __restrict__ off, int64_t n_tiles, int n_out, int64_t n,
                                 int64_t* __restrict__ out) {
  int d = threadIdx.x;
  if (d <= n_out) out[d] = n ? off[(int64_t)d * n_tiles] : 0;  // off[n_out * n_tiles] = rows that passed the filter
}
void hash_split_starts_device(Ctx* ctx, const SplitState& st, int64_t* out_dev) {
  rp_starts_kernel<<<1, 64, 0, ctx->stream>>>(st.hist->as<int64_t>(), st.n_tiles, st.n_out, st.n, out_dev);
  B2_LAUNCHED(ctx);
}
void hash_split_counts(Ctx* ctx, const SplitState& st, std::vector<int64_t>* counts) {
  counts->assign(st.n_out, 0);
  if (st.n == 0) return;
  auto tmp = dev_alloc(ctx, (size_t)(st.n_out + 1) * 8);
  hash_split_starts_device(ctx, st, tmp->as<int64_t>());
  std::vector<int64_t> starts(st.n_out + 1, 0);
  B2_READBACK(ctx, starts.data(), tmp->ptr, starts.size() * 8);
  for (int d = 0; d < st.n_out; ++d) (*counts)[d] = starts[d + 1] - starts[d];
}
BufPtr hash_split_perm(Ctx* ctx, const SplitState& st) {
  auto perm = dev_alloc(ctx, (size_t)(st.n + 1) * 4);
  if (st.n == 0) return perm;
  rp_scatter_kernel<<<grid_for(ctx, 8, st.n, RP_TILE), RP_THREADS, 0, ctx->stream>>>(
      st.dest->as<uint8_t>(), st.n, st.n_out, st.hist->as<int64_t>(), st.n_tiles, perm->as<uint32_t>());
  B2_LAUNCHED(ctx);
  return perm;
}
void hash_split_move(Ctx* ctx, const SplitState& st, int n_cols, const void* const* src, void* const* dst,
                     const int* width, const int64_t* base_dev, uint32_t* perm_or_null, int64_t n_live) {
  if (st.n == 0) return;
  if (n_cols > 8) fail(B2_ERR_EXECUTION, "internal: hash_split_move handles 8 columns per launch");
  MoveCols mc;
  memset(&mc, 0, sizeof mc);
  mc.n = n_cols;
  for (int c = 0; c < n_cols; ++c) {
    mc.src[c] = static_cast<const char*>(src[c]);
    mc.dst[c] = static_cast<char*>(dst[c]);
    mc.width[c] = width[c];
  }
  mc.base = base_dev;
  if (n_live >= 0 && n_live * 8 < st.n) {  // the caller knows that fewer than 1 row in 8 survived the fused filters
    rp_move_sparse_kernel<<<grid_for(ctx, 6, st.n, RP_TILE), RP_THREADS, 0, ctx->stream>>>(
        st.dest->as<uint8_t>(), st.n, st.n_out, st.hist->as<int64_t>(), st.n_tiles, mc, perm_or_null);
    B2_LAUNCHED(ctx);
    return;
  }
  rp_move_kernel<<<grid_for(ctx, 8, st.n, RP_TILE), RP_THREADS, 0, ctx->stream>>>(
      st.dest->as<uint8_t>(), st.n, st.n_out, st.hist->as<int64_t>(), st.n_tiles, mc, perm_or_null);
  B2_LAUNCHED(ctx);
}

// the push form of hash_split_move: addr_dev[c * n_out + d] = absolute (possibly peer-mapped) address of this source's
// first element of column c at destination d
void hash_split_push(Ctx* ctx, const SplitState& st, int n_cols, const void* const* src, const int* width,
                     const int64_t* addr_dev, int64_t n_live) {
  if (st.n == 0) return;
  if (n_cols > 8) fail(B2_ERR_EXECUTION, "internal: hash_split_push handles 8 columns per launch");
  if (n_live >= 0 && n_live * 8 < st.n) {  // sparse survivors: one thread per surviving row, absolute addresses as bases
    MoveCols mc;
    memset(&mc, 0, sizeof mc);
    mc.n = n_cols;
    for (int c = 0; c < n_cols; ++c) {
      mc.src[c] = static_cast<const char*>(src[c]);
      mc.dst[c] = nullptr;
      mc.width[c] = width[c];
    }
    mc.base = addr_dev;
    rp_move_sparse_kernel<<<grid_for(ctx, 6, st.n, RP_TILE), RP_THREADS, 0, ctx->stream>>>(
        st.dest->as<uint8_t>(), st.n, st.n_out, st.hist->as<int64_t>(), st.n_tiles, mc, nullptr);
    B2_LAUNCHED(ctx);
    return;
  }
  PushCols pc;
  memset(&pc, 0, sizeof pc);
  pc.n = n_cols;
  for (int c = 0; c < n_cols; ++c) {
    pc.src[c] = static_cast<const char*>(src[c]);
    pc.width[c] = width[c];
    if (width[c] > 8) fail(B2_ERR_EXECUTION, "internal: hash_split_push moves columns of at most 8 bytes per value");
  }
  pc.addr = addr_dev;
  rp_push_kernel<<<grid_for(ctx, 4, st.n, RP_TILE), RP_THREADS, 0, ctx->stream>>>(
      st.dest->as<uint8_t>(), st.n, st.n_out, st.hist->as<int64_t>(), st.n_tiles, pc);
  B2_LAUNCHED(ctx);
}

BufPtr hash_partition_indices(Ctx* ctx, const Key128* keys, const int32_t* key32, int64_t n, int n_out,
                              std::vector<int64_t>* counts, const RowPred* pred, const uint8_t* mask) {
  if (n_out < 1 || n_out > 255) fail(B2_ERR_INVALID_ARG, "hash repartition supports 1..255 output partitions");
  if (n >= (1ll << 32) - 1) fail(B2_ERR_NOT_IMPLEMENTED, "batches beyond 2^32-2 rows");
  counts->assign(n_out, 0);
  if (n == 0) return dev_alloc(ctx, 4);
  if (n_out > RP_MAX_OUT) {
    if (!keys || pred || mask) fail(B2_ERR_EXECUTION, "internal: packed keys / no fused filter beyond 32 output partitions");
    return partition_indices_scan_per_destination(ctx, keys, n, n_out, counts);
  }
  SplitState st = hash_split_begin(ctx, keys, key32, n, n_out, pred, mask);
  BufPtr perm = hash_split_perm(ctx, st);
  hash_split_counts(ctx, st, counts);
  return perm;
}

// ------------------------------------------------------------------------------------------------------
// SortExec support: stable LSD radix sort of row ids over order-preserving 64-bit key words, 5 bits per pass,
// reusing the split machinery above (a pass IS a stable 32-way split of the current order by one digit).
// ------------------------------------------------------------------------------------------------------
// key words: unsigned integers whose order is the column's sort order (descending = complement)
__global__ void sort_word_i32_kernel(const int32_t* __restrict__ v, int64_t n, bool desc, u64* __restrict__ w) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  for (; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    u64 x = (u64)((uint32_t)v[i] ^ 0x80000000u);
    w[i] = desc ? (~x & 0xFFFFFFFFull) : x;
  }
}
__global__ void sort_word_i64_kernel(const int64_t* __restrict__ v, int64_t n, bool desc, bool is_unsigned,
                                     u64* __restrict__ w) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  for (; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    u64 x = (u64)v[i] ^ (is_unsigned ? 0ull : 0x8000000000000000ull);
    w[i] = desc ? ~x : x;
  }
}
// IEEE total order with NaN last in ascending order (arrow's sort places NaN after all numbers)
__global__ void sort_word_f64_kernel(const double* __restrict__ v, int64_t n, bool desc, u64* __restrict__ w) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  for (; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    double d = v[i];
    u64 b = (u64)__double_as_longlong(d);
    u64 x = (d != d) ? ~0ull : ((b >> 63) ? ~b : (b | 0x8000000000000000ull));
    w[i] = desc ? ~x : x;
  }
}
// Utf8: bytes [8 * word, 8 * word + 8) big-endian, zero padded (byte strings of at most 16 bytes; longer ones raise
// status so the host can refuse instead of mis-sorting)
__global__ void sort_word_str_kernel(const int32_t* __restrict__ off, const uint8_t* __restrict__ chr, int64_t n, int word,
                                     bool desc, u64* __restrict__ w, int* status) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  for (; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    int32_t b = off[i], len = off[i + 1] - b;
    if (len > 16) *status = 1;
    u64 x = 0;
    for (int q = 0; q < 8; ++q) {
      int p = word * 8 + q;
      x = (x << 8) | (p < len ? (u64)chr[b + p] : 0ull);
    }
    w[i] = desc ? ~x : x;
  }
}
// digit of the current order's key + per-tile histogram (destination-major, like rp_hist_kernel)
__global__ void __launch_bounds__(RP_THREADS)
    sort_digit_kernel(const u64* __restrict__ word, const uint32_t* __restrict__ order, int64_t n, int shift,
                      uint8_t* __restrict__ dest, int64_t* __restrict__ hist, int64_t n_tiles) {
  __shared__ int h[RP_MAX_OUT];
  for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    if (threadIdx.x < RP_MAX_OUT) h[threadIdx.x] = 0;
    __syncthreads();
#pragma unroll
    for (int j = 0; j < RP_ITEMS; ++j) {
      int64_t r = tile * RP_TILE + j * RP_THREADS + threadIdx.x;
      int d = r < n ? (int)((word[order[r]] >> shift) & 31ull) : -1;
      if (d >= 0) dest[r] = (uint8_t)d;
      unsigned peers = __match_any_sync(0xffffffffu, d);
      if (d >= 0 && (peers & ((1u << (threadIdx.x & 31)) - 1)) == 0) atomicAdd(&h[d], __popc(peers));
    }
    __syncthreads();
    if (threadIdx.x < RP_MAX_OUT) hist[(int64_t)threadIdx.x * n_tiles + tile] = h[threadIdx.x];
    __syncthreads();
  }
}
__global__ void iota_kernel(uint32_t* v, int64_t n) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  for (; i < n; i += (int64_t)gridDim.x * blockDim.x) v[i] = (uint32_t)i;
}

// words[k] / bits[k]: key words from the MOST significant (first sort column) to the least; returns the sorted
// order of row ids (stable: equal keys keep their input order)
BufPtr radix_sort_order(Ctx* ctx, const std::vector<const u64*>& words, const std::vector<int>& bits, int64_t n) {
  if (n >= (1ll << 32) - 1) fail(B2_ERR_NOT_IMPLEMENTED, "sort beyond 2^32-2 rows");
  auto a = dev_alloc(ctx, (size_t)(n + 1) * 4), b = dev_alloc(ctx, (size_t)(n + 1) * 4);
  if (n == 0) return a;
  int grid = grid_for(ctx, 8, n, RP_TILE);
  iota_kernel<<<grid_for(ctx, 8, n, 256), 256, 0, ctx->stream>>>(a->as<uint32_t>(), n);
  B2_LAUNCHED(ctx);
  int64_t n_tiles = (n + RP_TILE - 1) / RP_TILE;
  auto dest = dev_alloc(ctx, (size_t)n + 16);
  auto hist = dev_alloc(ctx, (size_t)(n_tiles * RP_MAX_OUT + 2) * 8);
  for (int k = (int)words.size() - 1; k >= 0; --k) {
    for (int shift = 0; shift < bits[k]; shift += 5) {
      sort_digit_kernel<<<grid, RP_THREADS, 0, ctx->stream>>>(words[k], a->as<uint32_t>(), n, shift, dest->as<uint8_t>(),
                                                              hist->as<int64_t>(), n_tiles);
      B2_LAUNCHED(ctx);
      scan_i64(ctx, hist->as<int64_t>(), n_tiles * RP_MAX_OUT);
      rp_scatter_kernel<<<grid, RP_THREADS, 0, ctx->stream>>>(dest->as<uint8_t>(), n, RP_MAX_OUT, hist->as<int64_t>(),
                                                              n_tiles, b->as<uint32_t>(), a->as<uint32_t>());
      B2_LAUNCHED(ctx);
      std::swap(a, b);
    }
  }
  return a;
}

// order-preserving key words of one sort column (1 word; Utf8: 2)
void sort_key_words(Ctx* ctx, const Column& c, bool desc, std::vector<BufPtr>* words, std::vector<int>* bits, int* status_dev) {
  int64_t n = c.length;
  int grid = grid_for(ctx, 8, n, 256);
  auto alloc = [&]() { return dev_alloc(ctx, (size_t)(n + 1) * 8); };
  switch (c.dtype) {
    case B2_INT32: case B2_DATE32: {
      auto w = alloc();
      if (n) { sort_word_i32_kernel<<<grid, 256, 0, ctx->stream>>>(c.values<int32_t>(), n, desc, w->as<u64>()); B2_LAUNCHED(ctx); }
      words->push_back(w); bits->push_back(32);
      break;
    }
    case B2_INT64: case B2_UINT64: {
      auto w = alloc();
      if (n) { sort_word_i64_kernel<<<grid, 256, 0, ctx->stream>>>(c.values<int64_t>(), n, desc, c.dtype == B2_UINT64, w->as<u64>()); B2_LAUNCHED(ctx); }
      words->push_back(w); bits->push_back(64);
      break;
    }
    case B2_FLOAT64: {
      auto w = alloc();
      if (n) { sort_word_f64_kernel<<<grid, 256, 0, ctx->stream>>>(c.values<double>(), n, desc, w->as<u64>()); B2_LAUNCHED(ctx); }
      words->push_back(w); bits->push_back(64);
      break;
    }
    case B2_UTF8: {
      for (int word = 0; word < 2; ++word) {
        auto w = alloc();
        if (n) {
          sort_word_str_kernel<<<grid, 256, 0, ctx->stream>>>(c.values<int32_t>(), c.chars->as<uint8_t>(), n, word, desc,
                                                              w->as<u64>(), status_dev);
          B2_LAUNCHED(ctx);
        }
        words->push_back(w); bits->push_back(64);
      }
      break;
    }
    default: fail(B2_ERR_NOT_IMPLEMENTED, std::string("sort key of type ") + dtype_name(c.dtype));
  }
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
    if (!c.data) {  // elided by column pruning
      out->cols.push_back(std::move(o));
      continue;
    }
    if (c.valid) {
      o.valid = dev_alloc(ctx, (size_t)m + 16);
      if (m)
        B2_CUDA(cudaMemcpyAsync(o.valid->ptr, c.valid->as<uint8_t>() + begin, (size_t)m, cudaMemcpyDeviceToDevice,
                                ctx->stream));
    }
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


// ---- allocation-free hash split of one Int32 / Date32 key column (b2_hash_split_i32_static) ----
// Same two passes as hash_split_begin + hash_split_perm, but the destination bytes and the per-tile histogram live in a
// file-scope device array instead of stream-ordered allocations, so the call can sit inside a CUDA graph capture or a
// serving loop that may not allocate.  Every GPU that loads the library holds the array (~4.5 MiB) until exit.  All calls
// on one device share it: calls whose kernels overlap on the GPU (different streams) race; calls on one stream are fine.
constexpr int64_t SPLIT_STATIC_MAX_ROWS = int64_t(1) << 22;
constexpr int64_t SPLIT_STATIC_MAX_TILES = SPLIT_STATIC_MAX_ROWS / RP_TILE;
struct SplitStaticScratch {
  int64_t hist[RP_MAX_OUT * SPLIT_STATIC_MAX_TILES + 2];  // destination-major tile counts, scanned in place (+ total)
  uint8_t dest[SPLIT_STATIC_MAX_ROWS + 16];               // destination byte per row
};
static __device__ __align__(256) unsigned char g_split_static[sizeof(SplitStaticScratch)];

// The array's name is a host-side stand-in in host code: its device address is looked up once per device, with that
// device current, and kept here.
constexpr int SPLIT_STATIC_MAX_DEVICES = 64;
static std::mutex g_split_static_mu;
static SplitStaticScratch* g_split_static_ptr[SPLIT_STATIC_MAX_DEVICES];
static int g_split_static_sms[SPLIT_STATIC_MAX_DEVICES];

struct DeviceGuard {  // makes `device` current for the scope and restores the caller's device
  int prev = -1;
  explicit DeviceGuard(int device) {
    B2_CUDA(cudaGetDevice(&prev));
    if (prev != device) B2_CUDA(cudaSetDevice(device));
    else prev = -1;
  }
  ~DeviceGuard() {
    if (prev >= 0) cudaSetDevice(prev);
  }
};

int64_t hash_split_i32_static_max_rows() { return SPLIT_STATIC_MAX_ROWS; }

void hash_split_i32_static(int device, cudaStream_t stream, const int32_t* keys, int64_t n, int n_out, uint32_t* perm,
                           int64_t* starts) {
  if (device < 0 || device >= SPLIT_STATIC_MAX_DEVICES) fail(B2_ERR_INVALID_ARG, "b2_hash_split_i32_static: device out of range");
  if (n_out < 1 || n_out > RP_MAX_OUT) fail(B2_ERR_INVALID_ARG, "b2_hash_split_i32_static supports 1..32 output partitions");
  if (n < 0) fail(B2_ERR_INVALID_ARG, "b2_hash_split_i32_static: negative row count");
  if (n > SPLIT_STATIC_MAX_ROWS)
    fail(B2_ERR_INVALID_ARG, "b2_hash_split_i32_static: " + std::to_string(n) + " rows exceed the static scratch's maximum of " +
                                 std::to_string(SPLIT_STATIC_MAX_ROWS) + " rows per call (split the batch, or use RepartitionExec)");
  if (!starts || (n && (!keys || !perm))) fail(B2_ERR_INVALID_ARG, "b2_hash_split_i32_static: null pointer");
  DeviceGuard on(device);
  SplitStaticScratch* sc;
  int sms;
  {
    std::lock_guard<std::mutex> lock(g_split_static_mu);
    if (!g_split_static_ptr[device]) {
      void* p = nullptr;
      B2_CUDA(cudaGetSymbolAddress(&p, g_split_static));
      int v = 0;
      B2_CUDA(cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, device));
      g_split_static_sms[device] = v;
      g_split_static_ptr[device] = static_cast<SplitStaticScratch*>(p);
    }
    sc = g_split_static_ptr[device];
    sms = g_split_static_sms[device];
  }
  const int64_t n_tiles = (n + RP_TILE - 1) / RP_TILE;
  // the array keeps the previous call's contents: the counts this call scans start at zero
  B2_CUDA(cudaMemsetAsync(sc->hist, 0, (size_t)(n_tiles * n_out + 2) * 8, stream));
  if (n) {
    int64_t grid = std::min<int64_t>(n_tiles, (int64_t)sms * 8);
    rp_hist_kernel<<<(unsigned)grid, RP_THREADS, 0, stream>>>(KeySrc{nullptr, keys, false}, RowPred(), nullptr,
                                                             SemiArgs{nullptr, 0, {0}}, n, n_out, sc->dest, sc->hist, n_tiles);
    B2_CUDA(cudaGetLastError());
    scan_i64_kernel<<<1, SC_THREADS, 0, stream>>>(sc->hist, n_tiles * n_out);
    B2_CUDA(cudaGetLastError());
    rp_scatter_kernel<<<(unsigned)grid, RP_THREADS, 0, stream>>>(sc->dest, n, n_out, sc->hist, n_tiles, perm);
    B2_CUDA(cudaGetLastError());
  }
  rp_starts_kernel<<<1, 64, 0, stream>>>(sc->hist, n_tiles, n_out, n, starts);
  B2_CUDA(cudaGetLastError());
}

}  // namespace b2
