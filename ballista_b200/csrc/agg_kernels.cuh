// agg_kernels.cuh — the two aggregate kernels over a row source (see agg.cuh).
#pragma once
#include "agg.cuh"

namespace b2 {

template <int NA, int NC> struct SmallShared {
  unsigned int state[SMEM_SLOTS];  // 0 empty, 1 being written, 2 ready
  unsigned short gid[SMEM_SLOTS];
  Key128 key[SMEM_SLOTS];
  Key128 gkey[SMEM_GROUPS];
  double acc[SMEM_GROUPS][NA ? NA : 1];
  u64 cnt[SMEM_GROUPS][NC ? NC : 1];
  int n_groups;
  int abort;
  int any_valid;
};

// find-or-insert `key` in the CTA table; returns dense group id, or -1 when the CTA ran out of group ids
template <int NA, int NC>
__device__ __forceinline__ int smem_group_id(SmallShared<NA, NC>& sh, const Key128& key) {
  unsigned h = (unsigned)key_hash(key) & (SMEM_SLOTS - 1);
  for (int probe = 0; probe < 4 * SMEM_SLOTS; ++probe) {
    if (*(volatile int*)&sh.abort) return -1;  // the CTA already ran out of ids: stop probing a full table
    unsigned st = *(volatile unsigned*)&sh.state[h];
    if (st == 2u) {
      Key128 k = sh.key[h];
      if (key_eq(k, key)) return sh.gid[h] == 0xFFFF ? -1 : (int)sh.gid[h];  // 0xFFFF: seen after the CTA ran out of ids
      h = (h + 1) & (SMEM_SLOTS - 1);
    } else if (st == 0u) {
      if (atomicCAS(&sh.state[h], 0u, 1u) == 0u) {
        int g = atomicAdd(&sh.n_groups, 1);
        if (g >= SMEM_GROUPS) {
          sh.abort = 1;
          g = 0xFFFF;
        } else {
          sh.gkey[g] = key;
        }
        sh.key[h] = key;
        sh.gid[h] = (unsigned short)g;
        __threadfence_block();
        *(volatile unsigned*)&sh.state[h] = 2u;
        return g == 0xFFFF ? -1 : g;
      }
    }
    // st == 1 (or lost the claim): another thread is publishing this slot; look again
  }
  sh.abort = 1;
  return -1;
}

template <class S>
__global__ void __launch_bounds__(AGG_THREADS)
agg_small_kernel(const typename S::Params prm, int64_t n, TableRef tbl, int* status) {
  constexpr int NA = S::NACC, NC = S::NCNT;
  constexpr int G = S::KEYED ? REG_GROUPS : 1;
  __shared__ SmallShared<NA, NC> sh;
  for (int t = threadIdx.x; t < SMEM_SLOTS; t += AGG_THREADS) sh.state[t] = 0;
  for (int t = threadIdx.x; t < SMEM_GROUPS; t += AGG_THREADS) {
#pragma unroll
    for (int a = 0; a < NA; ++a) sh.acc[t][a] = 0.0;
#pragma unroll
    for (int a = 0; a < NC; ++a) sh.cnt[t][a] = 0;
  }
  if (threadIdx.x == 0) {
    sh.n_groups = S::KEYED ? 0 : 1;
    sh.abort = 0;
    sh.any_valid = 0;
    if (!S::KEYED) sh.gkey[0] = Key128{0, 0};
  }
  __syncthreads();

  double acc[G][NA ? NA : 1];
  u64 cnt[G][NC ? NC : 1];
#pragma unroll
  for (int g = 0; g < G; ++g) {
#pragma unroll
    for (int a = 0; a < NA; ++a) acc[g][a] = 0.0;
#pragma unroll
    for (int a = 0; a < NC; ++a) cnt[g][a] = 0;
  }
  bool key_overflow = false, any_valid = false;

  int64_t n_tiles = (n + TILE_ROWS - 1) / TILE_ROWS;
  for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    int64_t r0 = tile * TILE_ROWS + (int64_t)threadIdx.x * ROWS_PER_THREAD;
    typename S::Rows rows;
    if (r0 < n) S::load(prm, r0, n, rows);
#pragma unroll
    for (int k = 0; k < ROWS_PER_THREAD; ++k) {
      bool valid = (r0 + k < n) && S::pred(prm, rows, k);
      int gid = -1;
      double v[NA ? NA : 1];
      u64 c[NC ? NC : 1];
      if (valid) {
        any_valid = true;
        S::vals(prm, rows, k, v, c);
        if (S::KEYED) {
          Key128 key;
          S::key(prm, rows, k, key, key_overflow);
          gid = smem_group_id(sh, key);
        } else {
          gid = 0;
        }
      }
#pragma unroll
      for (int g = 0; g < G; ++g) {
        if (gid == g) {
#pragma unroll
          for (int a = 0; a < NA; ++a) acc[g][a] += v[a];
#pragma unroll
          for (int a = 0; a < NC; ++a) cnt[g][a] += c[a];
        }
      }
      if (gid >= G) {
#pragma unroll
        for (int a = 0; a < NA; ++a) smem_add(&sh.acc[gid][a], v[a]);
#pragma unroll
        for (int a = 0; a < NC; ++a) smem_add(&sh.cnt[gid][a], c[a]);
      }
    }
    if (S::KEYED) {
      // a CTA that ran out of group ids aborts the whole launch: the host redoes the batch on the
      // high-cardinality kernel (this launch wrote only into a scratch table)
      if (*(volatile int*)&sh.abort) atomicExch(status, AGG_ABORT_CARDINALITY);
      if (*(volatile int*)status == AGG_ABORT_CARDINALITY) return;
    }
  }
  if (key_overflow) atomicExch(status, AGG_KEY_TOO_WIDE);
  if (any_valid) sh.any_valid = 1;

  // registers -> warp -> shared
  int lane = threadIdx.x & 31;
#pragma unroll
  for (int g = 0; g < G; ++g) {
#pragma unroll
    for (int a = 0; a < NA; ++a) {
      double s = warp_sum(acc[g][a]);
      if (lane == 0 && s != 0.0) smem_add(&sh.acc[g][a], s);
    }
#pragma unroll
    for (int a = 0; a < NC; ++a) {
      u64 s = warp_sum(cnt[g][a]);
      if (lane == 0 && s != 0) smem_add(&sh.cnt[g][a], s);
    }
  }
  __syncthreads();
  // one flush per CTA into the HBM table
  int ng = sh.n_groups < SMEM_GROUPS ? sh.n_groups : SMEM_GROUPS;
  for (int g = threadIdx.x; g < ng; g += AGG_THREADS) {
    if (!S::KEYED && !sh.any_valid) continue;  // ungrouped CTA that saw no valid row
    u64 slot = table_upsert(tbl, sh.gkey[g]);
    if (slot == ~0ull) {
      atomicExch(status, AGG_TABLE_FULL);
      continue;
    }
#pragma unroll
    for (int a = 0; a < NA; ++a) red_add_f64(&tbl.acc[(u64)a * tbl.cap + slot], sh.acc[g][a]);
#pragma unroll
    for (int a = 0; a < NC; ++a) red_add_u64(&tbl.cnt[(u64)a * tbl.cap + slot], sh.cnt[g][a]);
  }
}

template <class S>
__global__ void __launch_bounds__(AGG_THREADS)
agg_global_kernel(const typename S::Params prm, int64_t row_begin, int64_t row_end, int64_t n, TableRef tbl,
                  int* status) {
  constexpr int NA = S::NACC, NC = S::NCNT;
  // slots this CTA claimed in the current tile: listed in tbl.slot_list behind ONE reservation per tile
  __shared__ uint32_t s_claim[TILE_ROWS];
  __shared__ int s_nclaim, s_base;
  bool key_overflow = false;
  int64_t first_tile = row_begin / TILE_ROWS;
  int64_t end_tile = (row_end + TILE_ROWS - 1) / TILE_ROWS;
  for (int64_t tile = first_tile + blockIdx.x; tile < end_tile; tile += gridDim.x) {
    if (threadIdx.x == 0) s_nclaim = 0;
    __syncthreads();
    int64_t r0 = tile * TILE_ROWS + (int64_t)threadIdx.x * ROWS_PER_THREAD;
    typename S::Rows rows;
    if (r0 < n) S::load(prm, r0, n, rows);
    // Consecutive rows with one key (a join output in probe order: Q3 has ~2.6 lineitem rows per order) are summed
    // in registers and go to the table once: each table visit is a dependent chain of a key load, possibly a
    // 128-bit CAS, and the reductions, and the kernel is bound by the latency of those chains.
    Key128 run_key{0, 0};
    bool run_open = false;
    double run_v[NA ? NA : 1];
    u64 run_c[NC ? NC : 1];
    auto flush = [&]() {
      bool claimed;
      u64 slot = table_upsert_quiet(tbl, run_key, &claimed);
      if (slot == ~0ull) {
        atomicExch(status, AGG_TABLE_FULL);
        return;
      }
      if (claimed) s_claim[atomicAdd(&s_nclaim, 1)] = (uint32_t)slot;
#pragma unroll
      for (int a = 0; a < NA; ++a) red_add_f64(&tbl.acc[(u64)a * tbl.cap + slot], run_v[a]);
#pragma unroll
      for (int a = 0; a < NC; ++a) red_add_u64(&tbl.cnt[(u64)a * tbl.cap + slot], run_c[a]);
    };
#pragma unroll
    for (int k = 0; k < ROWS_PER_THREAD; ++k) {
      bool valid = (r0 + k < row_end) && S::pred(prm, rows, k);
      if (!valid) continue;
      double v[NA ? NA : 1];
      u64 c[NC ? NC : 1];
      S::vals(prm, rows, k, v, c);
      Key128 key{0, 0};
      if (S::KEYED) S::key(prm, rows, k, key, key_overflow);
      if (run_open && key_eq(key, run_key)) {
#pragma unroll
        for (int a = 0; a < NA; ++a) run_v[a] += v[a];
#pragma unroll
        for (int a = 0; a < NC; ++a) run_c[a] += c[a];
        continue;
      }
      if (run_open) flush();
      run_key = key;
      run_open = true;
#pragma unroll
      for (int a = 0; a < NA; ++a) run_v[a] = v[a];
#pragma unroll
      for (int a = 0; a < NC; ++a) run_c[a] = c[a];
    }
    if (run_open) flush();
    __syncthreads();
    const int nclaim = s_nclaim;
    if (nclaim) {  // uniform
      if (threadIdx.x == 0) s_base = atomicAdd(tbl.n_groups, nclaim);
      __syncthreads();
      if (tbl.slot_list)
        for (int i = threadIdx.x; i < nclaim; i += AGG_THREADS) tbl.slot_list[s_base + i] = s_claim[i];
    }
    __syncthreads();
  }
  if (key_overflow) atomicExch(status, AGG_KEY_TOO_WIDE);
}

}  // namespace b2
