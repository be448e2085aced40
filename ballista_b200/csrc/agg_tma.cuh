// agg_tma.cuh — the bandwidth path of the fused scan -> filter -> expression -> partial-aggregate kernel.
//
// One persistent CTA per SM.  A producer warp streams 960-row tiles of every bound column into a ring of
// shared-memory stages with 1-D TMA bulk copies (cp.async.bulk ... mbarrier::complete_tx), so the bytes in
// flight are decoupled from the register file: a ~48 KB stage x 4 stages per SM keeps ~28 MB outstanding
// chip-wide, far beyond the ~5 MB that HBM latency x bandwidth needs.  15 consumer warps wait on the stage's
// "full" mbarrier, evaluate predicate / key / aggregate arguments from shared memory (conflict-free: lane ==
// row), accumulate the first 4 group ids in registers, and hand the stage back through its "empty" mbarrier.
// Group lookup / overflow groups / the once-per-CTA flush are those of agg_small_kernel (agg_kernels.cuh).
//
// Utf8 columns: the int32 offsets of a tile are copied like any fixed-width column; the tile's character
// range [off[r0], off[r0+rows]) is data dependent, so the producer warp prefetches the boundary offsets of
// its next 32 tiles with one lane-parallel load, and copies the characters when they fit the stage's
// character window (sized on the host from the column's average string length); otherwise the consumers
// read that tile's characters straight from global memory.
#pragma once
#include "agg_kernels.cuh"

namespace b2 {

// 15 consumer warps + 1 producer warp = 16 warps, so the register file gives every thread 128 registers
constexpr int TMA_CONSUMERS = 480;
constexpr int TMA_THREADS = TMA_CONSUMERS + 32;
constexpr int TMA_TILE = 2 * TMA_CONSUMERS;  // rows per stage (960: every column tile starts 16-byte aligned)
constexpr int TMA_RPT = TMA_TILE / TMA_CONSUMERS;
constexpr int TMA_MAX_STAGES = 8;

struct TmaLayout {
  int n_stages, stage_bytes;
  int f_off[fx::MAX_F], i_off[fx::MAX_I], so_off[fx::MAX_S], sc_off[fx::MAX_S], sc_cap[fx::MAX_S];
};

// ---- mbarrier / bulk-copy PTX ----
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_fence_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "B2_WAIT:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra B2_DONE;\n"
      "bra B2_WAIT;\n"
      "B2_DONE:\n"
      "}\n" ::"r"(smem_u32(bar)), "r"(parity)
      : "memory");
}
// global -> shared bulk copy; bytes and both addresses are multiples of 16
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   smem_u32(dst)),
               "l"(src), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}

__device__ __forceinline__ uint32_t key_hash32(const Key128& k) {
  uint32_t x = (uint32_t)k.lo ^ (uint32_t)(k.lo >> 32) * 0x85EBCA6Bu ^ (uint32_t)k.hi * 0xC2B2AE35u ^
               (uint32_t)(k.hi >> 32) * 0x27D4EB2Fu;
  x *= 0x9E3779B1u;
  return x >> (32 - 9);  // SMEM_SLOTS == 512
}
static_assert(SMEM_SLOTS == 512, "key_hash32 returns 9 bits");

// find-or-insert with the cheap hash (same protocol as smem_group_id)
template <int NA, int NC>
__device__ __forceinline__ int smem_group_id_fast(SmallShared<NA, NC>& sh, const Key128& key) {
  unsigned h = key_hash32(key);
  for (int probe = 0; probe < 4 * SMEM_SLOTS; ++probe) {
    unsigned st = *(volatile unsigned*)&sh.state[h];
    if (st == 2u) {
      Key128 k = sh.key[h];
      if (key_eq(k, key)) return (int)sh.gid[h];
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
  }
  sh.abort = 1;
  return -1;
}

template <class S>
__global__ void __launch_bounds__(TMA_THREADS, 1)
agg_tma_kernel(const typename S::Params prm, const TmaLayout lay, int64_t n, TableRef tbl, int* status) {
  constexpr int NA = S::NACC, NC = S::NCNT;
  constexpr int NF = S::NF, NI = S::NI, NS = S::NS;
  constexpr int G = S::KEYED ? REG_GROUPS : 1;
  extern __shared__ __align__(128) unsigned char ring[];
  __shared__ SmallShared<NA, NC> sh;
  __shared__ __align__(8) uint64_t full_bar[TMA_MAX_STAGES];
  __shared__ __align__(8) uint64_t empty_bar[TMA_MAX_STAGES];
  __shared__ int cbase[TMA_MAX_STAGES][NS ? NS : 1];  // first character held by the stage, or -1: read global

  const int tid = threadIdx.x;
  for (int t = tid; t < SMEM_SLOTS; t += TMA_THREADS) sh.state[t] = 0;
  for (int t = tid; t < SMEM_GROUPS; t += TMA_THREADS) {
#pragma unroll
    for (int a = 0; a < NA; ++a) sh.acc[t][a] = 0.0;
#pragma unroll
    for (int a = 0; a < NC; ++a) sh.cnt[t][a] = 0;
  }
  if (tid == 0) {
    sh.n_groups = S::KEYED ? 0 : 1;
    sh.abort = 0;
    sh.any_valid = 0;
    if (!S::KEYED) sh.gkey[0] = Key128{0, 0};
    for (int s = 0; s < lay.n_stages; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], TMA_CONSUMERS / 32);
    }
    mbar_fence_init();
  }
  __syncthreads();

  const int64_t n_tiles = (n + TMA_TILE - 1) / TMA_TILE;
  const int64_t my_tiles = blockIdx.x < n_tiles ? (n_tiles - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
  const int nst = lay.n_stages;

  if (tid >= TMA_CONSUMERS) {
    // ------------------------------ producer warp ------------------------------
    const int lane = tid & 31;
    int stage = 0;
    uint32_t phase = 0;
    for (int64_t it0 = 0; it0 < my_tiles; it0 += 32) {
      int32_t o0[NS ? NS : 1], o1[NS ? NS : 1];
      if (NS > 0) {
        int64_t it = it0 + lane;
        bool ok = it < my_tiles;
        int64_t r0 = (blockIdx.x + it * gridDim.x) * (int64_t)TMA_TILE;
        int64_t r1 = r0 + TMA_TILE < n ? r0 + TMA_TILE : n;
#pragma unroll
        for (int c = 0; c < NS; ++c) {
          o0[c] = ok ? __ldg(prm.soff[c] + r0) : 0;
          o1[c] = ok ? __ldg(prm.soff[c] + r1) : 0;
        }
      }
      const int kmax = my_tiles - it0 < 32 ? (int)(my_tiles - it0) : 32;
      for (int k = 0; k < kmax; ++k) {
        int32_t b0[NS ? NS : 1], b1[NS ? NS : 1];
#pragma unroll
        for (int c = 0; c < NS; ++c) {
          b0[c] = __shfl_sync(0xffffffffu, o0[c], k);
          b1[c] = __shfl_sync(0xffffffffu, o1[c], k);
        }
        if (lane == 0) {
          const int64_t r0 = (blockIdx.x + (it0 + k) * gridDim.x) * (int64_t)TMA_TILE;
          const int rows = (int)(n - r0 < TMA_TILE ? n - r0 : TMA_TILE);
          mbar_wait(&empty_bar[stage], phase ^ 1u);
          unsigned char* st = ring + (size_t)stage * lay.stage_bytes;
          if (*(volatile int*)&sh.abort) {
            mbar_arrive(&full_bar[stage]);  // drain: consumers skip the tile
          } else {
            const uint32_t bf = (uint32_t)(rows * 8 + 15) & ~15u, bi = (uint32_t)(rows * 4 + 15) & ~15u;
            const uint32_t bo = (uint32_t)((rows + 1) * 4 + 15) & ~15u;
            uint32_t total = NF * bf + NI * bi + NS * bo;
            uint32_t cb[NS ? NS : 1], csrc[NS ? NS : 1];
#pragma unroll
            for (int c = 0; c < NS; ++c) {
              const uint32_t a0 = (uint32_t)b0[c] & ~15u;
              const uint32_t len = ((uint32_t)b1[c] - a0 + 15u) & ~15u;
              csrc[c] = a0;
              cb[c] = len <= (uint32_t)lay.sc_cap[c] ? len : 0u;
              cbase[stage][c] = cb[c] || b1[c] == b0[c] ? (int)a0 : -1;
              total += cb[c];
            }
            mbar_arrive_expect_tx(&full_bar[stage], total);
#pragma unroll
            for (int c = 0; c < NF; ++c) bulk_g2s(st + lay.f_off[c], prm.f[c] + r0, bf, &full_bar[stage]);
#pragma unroll
            for (int c = 0; c < NI; ++c) bulk_g2s(st + lay.i_off[c], prm.i[c] + r0, bi, &full_bar[stage]);
#pragma unroll
            for (int c = 0; c < NS; ++c) {
              bulk_g2s(st + lay.so_off[c], prm.soff[c] + r0, bo, &full_bar[stage]);
              if (cb[c]) bulk_g2s(st + lay.sc_off[c], prm.schr[c] + csrc[c], cb[c], &full_bar[stage]);
            }
          }
        }
        __syncwarp();
        if (++stage == nst) {
          stage = 0;
          phase ^= 1u;
        }
      }
    }
  } else {
    // ------------------------------ consumer warps ------------------------------
    double acc[G][NA ? NA : 1];
    unsigned cnt[G];
#pragma unroll
    for (int g = 0; g < G; ++g) {
      cnt[g] = 0;
#pragma unroll
      for (int a = 0; a < NA; ++a) acc[g][a] = 0.0;
    }
    bool key_overflow = false, any_valid = false;
    int stage = 0;
    uint32_t phase = 0;
    for (int64_t it = 0; it < my_tiles; ++it) {
      const int64_t r0 = (blockIdx.x + it * gridDim.x) * (int64_t)TMA_TILE;
      const int rows = (int)(n - r0 < TMA_TILE ? n - r0 : TMA_TILE);
      mbar_wait(&full_bar[stage], phase);
      const unsigned char* st = ring + (size_t)stage * lay.stage_bytes;
      if (!*(volatile int*)&sh.abort) {
        fx::Row row[TMA_RPT];
        bool in[TMA_RPT];
#pragma unroll
        for (int k = 0; k < TMA_RPT; ++k) {
          const int j = tid + k * TMA_CONSUMERS;
          in[k] = j < rows;
          const int jj = in[k] ? j : 0;
#pragma unroll
          for (int c = 0; c < NF; ++c) row[k].f[c] = reinterpret_cast<const double*>(st + lay.f_off[c])[jj];
#pragma unroll
          for (int c = 0; c < NI; ++c) row[k].i[c] = reinterpret_cast<const int32_t*>(st + lay.i_off[c])[jj];
#pragma unroll
          for (int c = 0; c < NS; ++c) {
            const int32_t* so = reinterpret_cast<const int32_t*>(st + lay.so_off[c]);
            const int32_t b = so[jj], e = so[jj + 1];
            const int cb0 = cbase[stage][c];
            StrVal s;
            s.len = e - b;
            s.bytes = 0;
            const int m = s.len < 15 ? s.len : 15;
            if (cb0 >= 0) {
              const unsigned char* ch = st + lay.sc_off[c] + (b - cb0);
              if (m == 1) {
                s.bytes = (u128)ch[0];
              } else {
                for (int q = 0; q < m; ++q) s.bytes |= (u128)ch[q] << (8 * q);
              }
            } else {
              for (int q = 0; q < m; ++q) s.bytes |= (u128)__ldg(prm.schr[c] + b + q) << (8 * q);
            }
            row[k].s[c] = s;
          }
        }
#pragma unroll
        for (int k = 0; k < TMA_RPT; ++k) {
          const bool valid = in[k] && S::Pred_::eval(row[k], prm.lits);
          int gid = -1;
          double v[NA ? NA : 1];
          if (valid) {
            any_valid = true;
            fx::EvalF64<typename S::Accs_>::run(row[k], prm.lits, v);
            if (S::KEYED) {
              u128 kv = 0;
              fx::PackKey<typename S::Keys_, S::SB>::run(row[k], prm.lits, kv, 0, key_overflow);
              gid = smem_group_id_fast(sh, Key128{(u64)kv, (u64)(kv >> 64)});
            } else {
              gid = 0;
            }
          }
#pragma unroll
          for (int g = 0; g < G; ++g) {
            const bool hit = gid == g;
            cnt[g] += hit ? 1u : 0u;
#pragma unroll
            for (int a = 0; a < NA; ++a)
              if (hit) acc[g][a] += v[a];
          }
          if (gid >= G) {
#pragma unroll
            for (int a = 0; a < NA; ++a) smem_add(&sh.acc[gid][a], v[a]);
            smem_add(&sh.cnt[gid][0], (u64)1);
          }
        }
      }
      __syncwarp();
      if ((tid & 31) == 0) mbar_arrive(&empty_bar[stage]);
      if (++stage == nst) {
        stage = 0;
        phase ^= 1u;
      }
    }
    if (key_overflow) atomicExch(status, AGG_KEY_TOO_WIDE);
    if (any_valid) sh.any_valid = 1;
    const int lane = tid & 31;
#pragma unroll
    for (int g = 0; g < G; ++g) {
#pragma unroll
      for (int a = 0; a < NA; ++a) {
        double s = warp_sum(acc[g][a]);
        if (lane == 0 && s != 0.0) smem_add(&sh.acc[g][a], s);
      }
      u64 c = warp_sum((u64)cnt[g]);
      if (lane == 0 && c != 0) smem_add(&sh.cnt[g][0], c);
    }
  }
  __syncthreads();
  if (S::KEYED && sh.abort) {
    if (tid == 0) atomicExch(status, AGG_ABORT_CARDINALITY);
    return;
  }
  const int ng = sh.n_groups < SMEM_GROUPS ? sh.n_groups : SMEM_GROUPS;
  for (int g = tid; g < ng; g += TMA_THREADS) {
    if (!S::KEYED && !sh.any_valid) continue;
    u64 slot = table_upsert(tbl, sh.gkey[g]);
    if (slot == ~0ull) {
      atomicExch(status, AGG_TABLE_FULL);
      continue;
    }
#pragma unroll
    for (int a = 0; a < NA; ++a) red_add_f64(&tbl.acc[(u64)a * tbl.cap + slot], sh.acc[g][a]);
    red_add_u64(&tbl.cnt[slot], sh.cnt[g][0]);
  }
}

}  // namespace b2
