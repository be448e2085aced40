// join.cu — hash-join build and probe kernels (HashJoinExec).
//
// Build: open-addressing table of distinct packed keys in HBM, claimed with one 128-bit CAS; every build
// row is pushed onto its key's chain (head[slot] / next[row]) and the key's match count is bumped, so the
// probe sizes its output without walking chains.  Probe is two-pass (count -> scan -> write) and emits
// (build_row, probe_row) pairs in probe order, which is the reference's output order.
#include "ops.h"
#include "scan.cuh"

namespace b2 {

constexpr uint32_t NIL = 0xFFFFFFFFu;

__global__ void join_build_kernel(const Key128* keys, int64_t n, Key128* tkeys, u64 cap, uint32_t* head,
                                  uint32_t* count, uint32_t* next, int* n_keys, int* status) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  TableRef t{tkeys, nullptr, nullptr, cap, n_keys};
  for (; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    u64 slot = table_upsert(t, keys[i]);
    if (slot == ~0ull) {
      atomicExch(status, 1);
      continue;
    }
    uint32_t old = atomicExch(&head[slot], (uint32_t)i);
    next[i] = old;
    atomicAdd(&count[slot], 1u);
  }
}

std::shared_ptr<JoinTable> join_build(Ctx* ctx, const Key128* keys, int64_t n) {
  if (n >= (int64_t)NIL) fail(B2_ERR_NOT_IMPLEMENTED, "join build side beyond 2^32-2 rows");
  auto t = std::make_shared<JoinTable>();
  t->ctx = ctx;
  t->n_build = n;
  u64 cap = 16;
  while (cap < 2 * (u64)n + 16) cap <<= 1;
  t->cap = cap;
  t->keys = dev_alloc(ctx, cap * sizeof(Key128));
  t->head = dev_alloc(ctx, cap * 4);
  t->count = dev_alloc_zero(ctx, cap * 4);
  t->next = dev_alloc(ctx, (size_t)(n + 1) * 4);
  B2_CUDA(cudaMemsetAsync(t->keys->ptr, 0xFF, t->keys->bytes, ctx->stream));
  B2_CUDA(cudaMemsetAsync(t->head->ptr, 0xFF, t->head->bytes, ctx->stream));
  auto st = dev_alloc_zero(ctx, 16);
  if (n) {
    join_build_kernel<<<grid_for(ctx, 8, n, 256), 256, 0, ctx->stream>>>(
        keys, n, t->keys->as<Key128>(), cap, t->head->as<uint32_t>(), t->count->as<uint32_t>(),
        t->next->as<uint32_t>(), st->as<int>(), st->as<int>() + 1);
    B2_LAUNCHED(ctx);
  }
  return t;
}

struct ProbeCountIn {  // matches of probe row i
  const Key128* tkeys;
  u64 cap;
  const uint32_t* count;
  const Key128* pkeys;
  __device__ __forceinline__ int operator()(int64_t i) const {
    u64 slot = table_find(tkeys, cap, pkeys[i]);
    return slot == ~0ull ? 0 : (int)count[slot];
  }
};
struct ProbeWriteOut {
  const Key128* tkeys;
  u64 cap;
  const uint32_t* head;
  const uint32_t* next;
  const Key128* pkeys;
  uint32_t* left_idx;
  uint32_t* right_idx;
  __device__ __forceinline__ void operator()(int64_t i, int v, int64_t pre) const {
    if (!v) return;
    u64 slot = table_find(tkeys, cap, pkeys[i]);
    uint32_t r = head[slot];
    int64_t o = pre;
    while (r != NIL) {
      left_idx[o] = r;
      right_idx[o] = (uint32_t)i;
      ++o;
      r = next[r];
    }
  }
};

int64_t join_probe(Ctx* ctx, const JoinTable& t, const Key128* pkeys, int64_t n, BufPtr* left_idx, BufPtr* right_idx) {
  ProbeCountIn in{t.keys->as<Key128>(), t.cap, t.count->as<uint32_t>(), pkeys};
  // pass 1: count + scan (offsets are recomputed inside the apply pass, so only the CTA sums persist)
  int64_t nb = (n + SCAN_CHUNK - 1) / SCAN_CHUNK;
  auto sums = dev_alloc_zero(ctx, (size_t)(nb + 2) * 8);
  int64_t total = 0;
  if (n > 0) {
    scan_reduce_kernel<ProbeCountIn><<<(unsigned)nb, SCAN_THREADS, 0, ctx->stream>>>(in, n, sums->as<int64_t>());
    B2_LAUNCHED(ctx);
    scan_sums_kernel<<<1, SCAN_THREADS, 0, ctx->stream>>>(sums->as<int64_t>(), nb);
    B2_LAUNCHED(ctx);
    B2_CUDA(cudaMemcpyAsync(&total, sums->as<int64_t>() + nb, 8, cudaMemcpyDeviceToHost, ctx->stream));
    B2_CUDA(cudaStreamSynchronize(ctx->stream));
  }
  if (total >= (int64_t)NIL) fail(B2_ERR_NOT_IMPLEMENTED, "join output beyond 2^32-2 rows in one batch");
  *left_idx = dev_alloc(ctx, (size_t)(total + 1) * 4);
  *right_idx = dev_alloc(ctx, (size_t)(total + 1) * 4);
  if (total > 0) {
    ProbeWriteOut out{t.keys->as<Key128>(), t.cap, t.head->as<uint32_t>(), t.next->as<uint32_t>(), pkeys,
                      (*left_idx)->as<uint32_t>(), (*right_idx)->as<uint32_t>()};
    scan_apply_kernel<ProbeCountIn, ProbeWriteOut><<<(unsigned)nb, SCAN_THREADS, 0, ctx->stream>>>(
        in, n, sums->as<int64_t>(), out);
    B2_LAUNCHED(ctx);
  }
  return total;
}


// ------------------------------------------------------------------------------------------------------
// 32-bit-key join (every TPC-H join: one Int32 / Date32 key column per side).
//
// The table is ONE array of 8-byte slots, (key << 32) | head_row, so a probe costs a single 8-byte load and
// a build insert a single 64-bit CAS; EMPTY is all ones (a claimed slot always has head_row != NIL).  Build
// rows with equal keys are chained through next[] (head swap by CAS) and raise flags[0]; for primary-key
// builds (flags[0] == 0) next[] is never read.  The slot index keeps the low 3 key bits and mixes the rest:
// eight neighbouring keys share one 64-byte line, so a probe side clustered on the key (lineitem by
// l_orderkey) walks the table with L1/L2 locality instead of one DRAM sector per row.
//
// Both kernels take the FilterExec mask of their input (or null) and the ORIGINAL, uncompacted key column:
// the filter is never materialised, row ids are those of the unfiltered batch, and output columns are
// gathered once, for matching rows only (late materialisation).
// ------------------------------------------------------------------------------------------------------
__host__ __device__ __forceinline__ uint32_t fmix32(uint32_t h) {
  h ^= h >> 16;
  h *= 0x85ebca6bu;
  h ^= h >> 13;
  h *= 0xc2b2ae35u;
  h ^= h >> 16;
  return h;
}
__device__ __forceinline__ u64 slot_of(uint32_t k, u64 capmask) {
  return (((u64)fmix32(k >> 3) << 3) + (k & 7u)) & capmask;
}
__device__ __forceinline__ u64 ld_slot_volatile(const u64* p) {
  u64 v;
  asm volatile("ld.volatile.global.u64 %0, [%1];" : "=l"(v) : "l"(p));
  return v;
}

__global__ void join_build32_kernel(const int32_t* __restrict__ key, const uint8_t* __restrict__ mask, int64_t n,
                                    u64* slots, u64 capmask, uint32_t* next, int* flags) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  for (; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    if (mask && !mask[i]) continue;
    uint32_t k = (uint32_t)key[i];
    u64 h = slot_of(k, capmask);
    u64 mine = ((u64)k << 32) | (uint32_t)i;
    for (u64 probe = 0; probe <= capmask; ++probe) {
      u64 cur = ld_slot_volatile(&slots[h]);
      if (cur == KEY_EMPTY) {
        u64 old = atomicCAS((unsigned long long*)&slots[h], (unsigned long long)KEY_EMPTY, (unsigned long long)mine);
        if (old == KEY_EMPTY) {
          next[i] = NIL;
          break;
        }
        cur = old;
      }
      bool placed = false;
      while ((uint32_t)(cur >> 32) == k) {  // equal key: push this row onto the chain
        u64 old = atomicCAS((unsigned long long*)&slots[h], (unsigned long long)cur, (unsigned long long)mine);
        if (old == cur) {
          next[i] = (uint32_t)cur;
          flags[0] = 1;
          placed = true;
          break;
        }
        cur = old;
      }
      if (placed) break;
      h = (h + 1) & capmask;
    }
  }
}

__global__ void mask_count_kernel(const uint8_t* __restrict__ mask, int64_t n, unsigned long long* out) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  unsigned c = 0;
  for (; i < n; i += (int64_t)gridDim.x * blockDim.x) c += mask[i] != 0;
  for (int d = 16; d > 0; d >>= 1) c += __shfl_xor_sync(0xffffffffu, c, d);
  if ((threadIdx.x & 31) == 0 && c) atomicAdd(out, (unsigned long long)c);
}

std::shared_ptr<JoinTable> join_build32(Ctx* ctx, const int32_t* key, const uint8_t* mask, int64_t n) {
  if (n >= (int64_t)NIL) fail(B2_ERR_NOT_IMPLEMENTED, "join build side beyond 2^32-2 rows");
  auto t = std::make_shared<JoinTable>();
  t->ctx = ctx;
  t->n_build = n;
  t->k32 = true;
  t->flags = dev_alloc_zero(ctx, 16);
  // size the table for the rows that pass the fused filter, so that a selective build side stays L2-resident
  int64_t n_in = n;
  if (mask && n > (1 << 18)) {
    mask_count_kernel<<<grid_for(ctx, 8, n, 1024), 256, 0, ctx->stream>>>(
        mask, n, reinterpret_cast<unsigned long long*>(t->flags->as<char>() + 8));
    B2_LAUNCHED(ctx);
    unsigned long long c = 0;
    B2_CUDA(cudaMemcpyAsync(&c, t->flags->as<char>() + 8, 8, cudaMemcpyDeviceToHost, ctx->stream));
    B2_CUDA(cudaStreamSynchronize(ctx->stream));
    n_in = (int64_t)c;
  }
  u64 cap = 64;
  while (cap < 2 * (u64)n_in + 16) cap <<= 1;
  t->cap = cap;
  t->slots = dev_alloc(ctx, cap * 8);
  t->next = dev_alloc(ctx, (size_t)(n + 1) * 4);
  B2_CUDA(cudaMemsetAsync(t->slots->ptr, 0xFF, cap * 8, ctx->stream));
  if (n) {
    join_build32_kernel<<<grid_for(ctx, 8, n, 256), 256, 0, ctx->stream>>>(key, mask, n, t->slots->as<u64>(), cap - 1,
                                                                         t->next->as<uint32_t>(), t->flags->as<int>());
    B2_LAUNCHED(ctx);
  }
  t->dups_known = false;
  return t;
}

constexpr int JP_THREADS = SCAN_THREADS;  // block_exclusive_scan is written for this CTA size
constexpr int JP_ITEMS = 4;
constexpr int JP_TILE = JP_THREADS * JP_ITEMS;

// tile status word of the single-pass ordered compaction: flag (2 bits) | value (62 bits)
constexpr u64 ST_AGG = 1ull << 62, ST_PREFIX = 2ull << 62, ST_VAL = (1ull << 62) - 1;
__device__ __forceinline__ void st_status(u64* p, u64 v) {
  asm volatile("st.volatile.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

// One pass over the probe side: filter mask -> table lookup -> (build_row, probe_row) pairs in PROBE ORDER
// (the reference's output order).  Tiles are handed out by an atomic ticket, so tile t-1 has always started
// before tile t; each tile publishes its match count, then resolves its output offset with a warp-wide
// decoupled look-back over its predecessors' status words (no second probe pass, no global scan kernels).
// COUNT_ONLY sizes the output of a join whose build side has duplicate keys (chains are walked to count).
template <bool COUNT_ONLY, bool DUPS>
__global__ void __launch_bounds__(JP_THREADS)
    join_probe32_kernel(const int32_t* __restrict__ key, const uint8_t* __restrict__ mask, int64_t n, bool vec,
                        const u64* __restrict__ slots, u64 capmask, const uint32_t* __restrict__ next,
                        uint32_t* __restrict__ left_idx, uint32_t* __restrict__ right_idx, u64* status,
                        unsigned long long* ticket, unsigned long long* total_out) {
  __shared__ unsigned long long base_s, tile_s;
  const int64_t n_tiles = (n + JP_TILE - 1) / JP_TILE;
  const int lane = threadIdx.x & 31;
  while (true) {
    if (threadIdx.x == 0) tile_s = atomicAdd(ticket, 1ull);
    __syncthreads();
    const int64_t tile = (int64_t)tile_s;
    if (tile >= n_tiles) break;
    int64_t r0 = tile * JP_TILE + (int64_t)threadIdx.x * JP_ITEMS;
    uint32_t k[JP_ITEMS];
    bool pass[JP_ITEMS];
    if (vec && r0 + JP_ITEMS <= n) {
      int4 v = ld_stream_i32x4(key + r0);
      k[0] = v.x; k[1] = v.y; k[2] = v.z; k[3] = v.w;
      uint32_t m = mask ? __ldg(reinterpret_cast<const uint32_t*>(mask + r0)) : 0x01010101u;
#pragma unroll
      for (int j = 0; j < JP_ITEMS; ++j) pass[j] = ((m >> (8 * j)) & 0xFFu) != 0;
    } else {
#pragma unroll
      for (int j = 0; j < JP_ITEMS; ++j) {
        bool in = r0 + j < n;
        k[j] = in ? (uint32_t)key[r0 + j] : 0u;
        pass[j] = in && (!mask || mask[r0 + j]);
      }
    }
    u64 h[JP_ITEMS], cur[JP_ITEMS];
#pragma unroll
    for (int j = 0; j < JP_ITEMS; ++j) {  // all first-probe loads in flight before any is used
      h[j] = slot_of(k[j], capmask);
      cur[j] = pass[j] ? __ldg(&slots[h[j]]) : KEY_EMPTY;
    }
    uint32_t head[JP_ITEMS];
    int cnt = 0;
#pragma unroll
    for (int j = 0; j < JP_ITEMS; ++j) {
      head[j] = NIL;
      u64 c = cur[j], hh = h[j];
      while (c != KEY_EMPTY) {
        if ((uint32_t)(c >> 32) == k[j]) {
          head[j] = (uint32_t)c;
          break;
        }
        hh = (hh + 1) & capmask;
        c = __ldg(&slots[hh]);
      }
      if (head[j] != NIL) {
        if (DUPS) {
          for (uint32_t r = head[j]; r != NIL; r = next[r]) ++cnt;
        } else {
          ++cnt;
        }
      }
    }
    int total;
    int off = block_exclusive_scan(cnt, &total);
    if (COUNT_ONLY) {
      if (threadIdx.x == 0 && total) atomicAdd(total_out, (unsigned long long)total);
      continue;
    }
    if (threadIdx.x < 32) {
      u64 excl = 0;
      if (tile == 0) {
        if (lane == 0) st_status(&status[0], ST_PREFIX | (u64)total);
      } else {
        if (lane == 0) st_status(&status[tile], ST_AGG | (u64)total);
        int64_t look = tile - 1;
        while (true) {
          int64_t idx = look - lane;
          u64 sv = idx >= 0 ? ld_slot_volatile(&status[idx]) : ST_PREFIX;
          while (__any_sync(0xffffffffu, (sv >> 62) == 0)) {
            if ((sv >> 62) == 0) sv = ld_slot_volatile(&status[idx]);
          }
          unsigned pm = __ballot_sync(0xffffffffu, (sv >> 62) == 2);
          int first = pm ? __ffs(pm) - 1 : 32;
          u64 v = lane <= first ? (sv & ST_VAL) : 0ull;
          excl += warp_sum(v);
          if (pm) break;
          look -= 32;
        }
        if (lane == 0) st_status(&status[tile], ST_PREFIX | (excl + (u64)total));
      }
      if (lane == 0) {
        base_s = excl;
        if (tile == n_tiles - 1) *total_out = excl + (u64)total;
      }
    }
    __syncthreads();
    int64_t o = (int64_t)base_s + off;
#pragma unroll
    for (int j = 0; j < JP_ITEMS; ++j) {
      uint32_t r = head[j];
      while (r != NIL) {
        left_idx[o] = r;
        right_idx[o] = (uint32_t)(r0 + j);
        ++o;
        if (!DUPS) break;
        r = next[r];
      }
    }
  }
}

int64_t join_probe32(Ctx* ctx, JoinTable& t, const int32_t* key, const uint8_t* mask, int64_t n, BufPtr* left_idx,
                     BufPtr* right_idx) {
  if (n >= (int64_t)NIL) fail(B2_ERR_NOT_IMPLEMENTED, "batches beyond 2^32-2 rows");
  if (!t.dups_known) {  // one readback per build: does any key occur twice on the build side?
    int f = 0;
    B2_CUDA(cudaMemcpyAsync(&f, t.flags->ptr, 4, cudaMemcpyDeviceToHost, ctx->stream));
    B2_CUDA(cudaStreamSynchronize(ctx->stream));
    t.has_dups = f != 0;
    t.dups_known = true;
  }
  int64_t n_tiles = (n + JP_TILE - 1) / JP_TILE;
  // [0] ticket, [1] total, [2..] tile status words
  auto work = dev_alloc_zero(ctx, (size_t)(n_tiles + 4) * 8);
  auto* ticket = work->as<unsigned long long>();
  auto* total_dev = ticket + 1;
  u64* status = work->as<u64>() + 2;
  bool vec = ((uintptr_t)key & 15) == 0 && ((uintptr_t)mask & 3) == 0;
  int grid = grid_for(ctx, 8, n, JP_TILE);
  const u64* slots = t.slots->as<u64>();
  const uint32_t* next = t.next->as<uint32_t>();
  unsigned long long total = 0;
  int64_t capacity = n;  // primary-key build: at most one match per probe row
  if (n > 0 && t.has_dups) {
    join_probe32_kernel<true, true><<<grid, JP_THREADS, 0, ctx->stream>>>(key, mask, n, vec, slots, t.cap - 1, next,
                                                                         nullptr, nullptr, status, ticket, total_dev);
    B2_LAUNCHED(ctx);
    B2_CUDA(cudaMemcpyAsync(&total, total_dev, 8, cudaMemcpyDeviceToHost, ctx->stream));
    B2_CUDA(cudaStreamSynchronize(ctx->stream));
    if (total >= (unsigned long long)NIL) fail(B2_ERR_NOT_IMPLEMENTED, "join output beyond 2^32-2 rows in one batch");
    capacity = (int64_t)total;
    B2_CUDA(cudaMemsetAsync(ticket, 0, 16, ctx->stream));
  }
  *left_idx = dev_alloc(ctx, (size_t)(capacity + 1) * 4);
  *right_idx = dev_alloc(ctx, (size_t)(capacity + 1) * 4);
  if (n > 0 && capacity > 0) {
    if (t.has_dups)
      join_probe32_kernel<false, true><<<grid, JP_THREADS, 0, ctx->stream>>>(
          key, mask, n, vec, slots, t.cap - 1, next, (*left_idx)->as<uint32_t>(), (*right_idx)->as<uint32_t>(), status,
          ticket, total_dev);
    else
      join_probe32_kernel<false, false><<<grid, JP_THREADS, 0, ctx->stream>>>(
          key, mask, n, vec, slots, t.cap - 1, next, (*left_idx)->as<uint32_t>(), (*right_idx)->as<uint32_t>(), status,
          ticket, total_dev);
    B2_LAUNCHED(ctx);
    B2_CUDA(cudaMemcpyAsync(&total, total_dev, 8, cudaMemcpyDeviceToHost, ctx->stream));
    B2_CUDA(cudaStreamSynchronize(ctx->stream));
  }
  return (int64_t)total;
}

}  // namespace b2
