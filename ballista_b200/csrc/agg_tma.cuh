// agg_tma.cuh — the bandwidth path of the fused scan -> filter -> expression -> partial-aggregate kernel.
//
// One persistent CTA per SM.  A producer warp streams 960-row tiles of every bound column into a ring of
// shared-memory stages with 1-D TMA bulk copies (cp.async.bulk ... mbarrier::complete_tx), so the bytes in
// flight are decoupled from the register file: a ~45 KB stage x 4 stages per SM keeps ~27 MB outstanding
// chip-wide, far beyond the ~5 MB that HBM latency x bandwidth needs.  15 consumer warps wait on the stage's
// "full" mbarrier, evaluate predicate / key / aggregate arguments from shared memory (conflict-free: lane ==
// row), accumulate the first 4 group ids in registers, and hand the stage back through its "empty" mbarrier.
//
// The fixed-width part of a stage has a compile-time layout (shared-memory loads use immediate offsets from
// one per-tile base register); only the character windows of Utf8 columns are sized at run time.
//
// Utf8 columns: the int32 offsets of a tile are copied like any fixed-width column; the tile's character
// range [off[r0], off[r0+rows]) is data dependent, so the producer warp prefetches the boundary offsets of
// its next 32 tiles with one lane-parallel load, and copies the characters when they fit the stage's
// character window (sized on the host from the column's average string length); otherwise the consumers
// read that tile's characters straight from global memory.
#pragma once
#ifndef __CUDACC_RTC__
#include <type_traits>
#endif

#include "agg.cuh"

namespace b2 {

// 15 consumer warps + 1 producer warp = 16 warps, so the register file gives every thread 128 registers
constexpr int TMA_CONSUMERS = 480;
constexpr int TMA_THREADS = TMA_CONSUMERS + 32;
constexpr int TMA_RPT = 2;                         // rows per consumer thread per tile
constexpr int TMA_TILE = TMA_RPT * TMA_CONSUMERS;  // rows per stage (960: every column tile starts 16-byte aligned)
constexpr int TMA_MAX_STAGES = 8;
constexpr int TMA_SO_BYTES = ((TMA_TILE + 4) * 4 + 127) / 128 * 128;  // offsets tile: TILE+1 int32, padded

// compile-time stage layout of the fixed-width columns of a source
template <class S> struct TmaStage {
  static constexpr int F_OFF = 0;
  static constexpr int I_OFF = F_OFF + S::NF * TMA_TILE * 8;
  static constexpr int SO_OFF = I_OFF + S::NI * TMA_TILE * 4;
  static constexpr int SC_OFF = SO_OFF + S::NS * TMA_SO_BYTES;  // character windows start here
};
struct TmaLayout {  // run-time part
  int n_stages, stage_bytes;
  int sc_off[fx::MAX_S], sc_cap[fx::MAX_S];  // byte offset (from the stage base) and capacity of each window
};

// ---- mbarrier / bulk-copy PTX (all addresses are 32-bit shared-window addresses) ----
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_fence_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void mbar_arrive_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "B2_WAIT:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1, 0x989680;\n"  // suspend-time hint: sleep, don't spin
      "@p bra B2_DONE;\n"
      "bra B2_WAIT;\n"
      "B2_DONE:\n"
      "}\n" ::"r"(bar), "r"(parity)
      : "memory");
}
// global -> shared bulk copy; bytes and both addresses are multiples of 16
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
               "l"(src), "r"(bytes), "r"(bar)
               : "memory");
}

// ---- per-CTA group directory: open-addressing table key -> dense group id ----
// KW = u64 (narrow kernels: packed key of <= 8 bytes) or u128.
// meta word: 0 = empty, 1 = being published, otherwise 2 | gid << 8
template <int NA, class KW> struct TmaShared {
  unsigned meta[SMEM_SLOTS];
  KW key[SMEM_SLOTS];
  KW gkey[SMEM_GROUPS];
  double acc[SMEM_GROUPS][NA ? NA : 1];
  u64 cnt[SMEM_GROUPS];
  int n_groups;
  int abort;
  int any_valid;
};
static_assert(SMEM_SLOTS == 512, "key_hash32 returns 9 bits");
__device__ __forceinline__ uint32_t key_hash32(u64 k) {
  uint32_t x = ((uint32_t)k ^ (uint32_t)(k >> 32) * 0x85EBCA6Bu) * 0x9E3779B1u;
  return x >> (32 - 9);
}
__device__ __forceinline__ uint32_t key_hash32(u128 k) {
  const u64 lo = (u64)k, hi = (u64)(k >> 64);
  uint32_t x = (uint32_t)lo ^ (uint32_t)(lo >> 32) * 0x85EBCA6Bu ^ (uint32_t)hi * 0xC2B2AE35u ^
               (uint32_t)(hi >> 32) * 0x27D4EB2Fu;
  x *= 0x9E3779B1u;
  return x >> (32 - 9);
}
// find-or-insert; returns the dense group id, or -1 when the CTA ran out of group ids (launch aborts)
template <int NA, class KW> __device__ __forceinline__ int tma_group_id(TmaShared<NA, KW>& sh, const KW key) {
  unsigned h = key_hash32(key);
  for (int probe = 0; probe < 4 * SMEM_SLOTS; ++probe) {
    if (*(volatile int*)&sh.abort) return -1;  // the CTA already ran out of ids: stop probing a full table
    const unsigned m = *(volatile unsigned*)&sh.meta[h];
    if (m > 1u) {
      if (sh.key[h] == key) return (m >> 8) == 0xFFFFu ? -1 : (int)(m >> 8);  // 0xFFFF: CTA ran out of ids
      h = (h + 1) & (SMEM_SLOTS - 1);
    } else if (m == 0u) {
      if (atomicCAS(&sh.meta[h], 0u, 1u) == 0u) {
        int g = atomicAdd(&sh.n_groups, 1);
        if (g >= SMEM_GROUPS) {
          sh.abort = 1;
          g = 0xFFFF;
        } else {
          sh.gkey[g] = key;
        }
        sh.key[h] = key;
        __threadfence_block();
        *(volatile unsigned*)&sh.meta[h] = 2u | ((unsigned)g << 8);
        return g == 0xFFFF ? -1 : g;
      }
    }
    // m == 1 (or lost the claim): another thread is publishing this slot; look again
  }
  sh.abort = 1;
  return -1;
}
// slow path of the directory lookup (first probe missed): out of line, the hot loop stays straight-line
template <int NA, class KW> __device__ __noinline__ int tma_group_id_slow(TmaShared<NA, KW>* sh, KW key) {
  return tma_group_id(*sh, key);
}
// hot path: one probe with both loads issued up front
template <int NA, class KW> __device__ __forceinline__ int tma_group_id_fast(TmaShared<NA, KW>& sh, const KW key) {
  const unsigned h = key_hash32(key);
  const unsigned m = *(volatile unsigned*)&sh.meta[h];
  const KW k = sh.key[h];
  if (m > 1u && k == key && (m >> 8) != 0xFFFFu) return (int)(m >> 8);
  return tma_group_id_slow(&sh, key);
}

// ---- one Utf8 value of row j of a stage ----
// narrow (CH = 3 or 7 characters kept): straight-line — aligned word loads + funnel shifts + length mask.
// Longer values raise `overflow` (the host reruns the batch with the wide kernel).
template <int CH>
__device__ __forceinline__ StrValT<u64> tma_load_str_narrow(const unsigned char* st, int so_off, int sc_off, int cb0,
                                                            const uint8_t* gchars, int j, bool in, bool& overflow) {
  const int32_t* so = reinterpret_cast<const int32_t*>(st + so_off);
  const int32_t b = in ? so[j] : (cb0 >= 0 ? cb0 : 0);
  StrValT<u64> s;
  s.len = in ? so[j + 1] - b : 0;
  if (s.len > CH) overflow = true;
  const int m = s.len < CH ? s.len : CH;
  if (cb0 >= 0) {  // uniform per stage
    const int p = sc_off + (b - cb0);
    const uint32_t* w = reinterpret_cast<const uint32_t*>(st + (p & ~3));
    const unsigned shift = 8u * (unsigned)(p & 3);
    if (CH <= 3) {
      const uint32_t v = __funnelshift_r(w[0], w[1], shift);
      s.bytes = v & ((1u << (8 * m)) - 1u);
    } else {
      const uint32_t w0 = w[0], w1 = w[1], w2 = w[2];
      const u64 v = ((u64)__funnelshift_r(w1, w2, shift) << 32) | __funnelshift_r(w0, w1, shift);
      s.bytes = v & ((1ull << (8 * m)) - 1ull);
    }
  } else {
    s.bytes = 0;
    for (int q = 0; q < m; ++q) s.bytes |= (u64)__ldg(gchars + b + q) << (8 * q);
  }
  return s;
}
// wide: first <= 15 bytes
__device__ __forceinline__ StrValT<u128> tma_load_str_wide(const unsigned char* st, int so_off, int sc_off, int cb0,
                                                           const uint8_t* gchars, int j, bool in) {
  const int32_t* so = reinterpret_cast<const int32_t*>(st + so_off);
  const int32_t b = in ? so[j] : (cb0 >= 0 ? cb0 : 0);
  StrValT<u128> s;
  s.len = in ? so[j + 1] - b : 0;
  s.bytes = 0;
  const int m = s.len < 15 ? s.len : 15;
  if (cb0 >= 0) {
    const unsigned char* ch = st + sc_off + (b - cb0);
    for (int q = 0; q < m; ++q) s.bytes |= (u128)ch[q] << (8 * q);
  } else {
    for (int q = 0; q < m; ++q) s.bytes |= (u128)__ldg(gchars + b + q) << (8 * q);
  }
  return s;
}

template <class S, bool NARROW> struct TmaKeyTraits {
  using KS = fx::CountStr<typename S::Keys_>;
  static constexpr int SB = NARROW ? fx::str_budget_narrow(KS::INTS, KS::N) : S::SB;
  // characters a narrow kernel keeps per Utf8 value: 3 when every string is a short key and nothing is
  // compared with a literal, else 7
  static constexpr int CH = (S::NLS == 0 && S::KEYED && KS::N >= S::NS && SB <= 4) ? 3 : 7;
};

template <class S, bool NARROW>
__global__ void __launch_bounds__(TMA_THREADS, 1)
agg_tma_kernel(const typename S::Params prm, const TmaLayout lay, int64_t n, TableRef tbl, int* status) {
  constexpr int NA = S::NACC;
  constexpr int NF = S::NF, NI = S::NI, NS = S::NS;
  constexpr int G = S::KEYED ? REG_GROUPS : 1;
  constexpr int SB = TmaKeyTraits<S, NARROW>::SB;
  using W = typename std::conditional<NARROW, u64, u128>::type;  // string word == key word
  using L = TmaStage<S>;
  extern __shared__ __align__(128) unsigned char ring[];
  __shared__ TmaShared<NA, W> sh;
  __shared__ __align__(8) uint64_t full_bar[TMA_MAX_STAGES];
  __shared__ __align__(8) uint64_t empty_bar[TMA_MAX_STAGES];
  __shared__ int cbase[TMA_MAX_STAGES][NS ? NS : 1];  // first character held by the stage, or -1: read global

  const int tid = threadIdx.x;
  const int nst = lay.n_stages;
  for (int t = tid; t < SMEM_SLOTS; t += TMA_THREADS) sh.meta[t] = 0;
  for (int t = tid; t < SMEM_GROUPS; t += TMA_THREADS) {
#pragma unroll
    for (int a = 0; a < NA; ++a) sh.acc[t][a] = 0.0;
    sh.cnt[t] = 0;
  }
  if (tid == 0) {
    sh.n_groups = S::KEYED ? 0 : 1;
    sh.abort = 0;
    sh.any_valid = 0;
    if (!S::KEYED) sh.gkey[0] = 0;
    for (int s = 0; s < nst; ++s) {
      mbar_init(smem_u32(&full_bar[s]), 1);
      mbar_init(smem_u32(&empty_bar[s]), TMA_CONSUMERS / 32);
    }
    mbar_fence_init();
  }
  __syncthreads();

  const int64_t n_tiles = (n + TMA_TILE - 1) / TMA_TILE;
  const int64_t my_tiles = blockIdx.x < n_tiles ? (n_tiles - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
  const uint32_t full0 = smem_u32(&full_bar[0]), empty0 = smem_u32(&empty_bar[0]);
  const uint32_t ring0 = smem_u32(ring);

  if (tid >= TMA_CONSUMERS) {
    // ------------------------------ producer warp ------------------------------
    const int lane = tid & 31;
    int stage = 0;
    uint32_t phase = 0;
    for (int64_t it0 = 0; it0 < my_tiles; it0 += 32) {
      int32_t o0[NS ? NS : 1], o1[NS ? NS : 1];
      if (NS > 0) {
        const int64_t it = it0 + lane;
        const bool ok = it < my_tiles;
        const int64_t r0 = (blockIdx.x + it * gridDim.x) * (int64_t)TMA_TILE;
        const int64_t r1 = r0 + TMA_TILE < n ? r0 + TMA_TILE : n;
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
          const uint32_t fb = full0 + 8u * stage;
          mbar_wait(empty0 + 8u * stage, phase ^ 1u);
          const uint32_t st = ring0 + (uint32_t)stage * (uint32_t)lay.stage_bytes;
          if (*(volatile int*)&sh.abort) {
            mbar_arrive(fb);  // drain: consumers skip the tile
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
              // + 16: consumers may read up to 11 bytes past the last character (aligned-word loads)
              cb[c] = len + 16u <= (uint32_t)lay.sc_cap[c] ? len : 0u;
              cbase[stage][c] = (cb[c] || b1[c] == b0[c]) ? (int)a0 : -1;
              total += cb[c];
            }
            mbar_arrive_expect_tx(fb, total);
#pragma unroll
            for (int c = 0; c < NF; ++c) bulk_g2s(st + L::F_OFF + c * TMA_TILE * 8, prm.f[c] + r0, bf, fb);
#pragma unroll
            for (int c = 0; c < NI; ++c) bulk_g2s(st + L::I_OFF + c * TMA_TILE * 4, prm.i[c] + r0, bi, fb);
#pragma unroll
            for (int c = 0; c < NS; ++c) {
              bulk_g2s(st + L::SO_OFF + c * TMA_SO_BYTES, prm.soff[c] + r0, bo, fb);
              if (cb[c]) bulk_g2s(st + lay.sc_off[c], prm.schr[c] + csrc[c], cb[c], fb);
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
    unsigned cnt[G];  // rows per register-resident group (a thread sees < 2^32 rows)
#pragma unroll
    for (int g = 0; g < G; ++g) {
      cnt[g] = 0;
#pragma unroll
      for (int a = 0; a < NA; ++a) acc[g][a] = 0.0;
    }
    bool key_overflow = false, any_valid = false;
    int stage = 0;
    uint32_t phase = 0;
    int sc_off[NS ? NS : 1];
#pragma unroll
    for (int c = 0; c < NS; ++c) sc_off[c] = lay.sc_off[c];
    const int tail_rows = (int)(n - (n_tiles - 1) * (int64_t)TMA_TILE);  // rows of the table's last tile
    const bool own_tail = tail_rows < TMA_TILE && (int64_t)blockIdx.x == (n_tiles - 1) % gridDim.x;

    // one tile; FULL = all TMA_TILE rows present (no tail masking in the hot path)
    auto process = [&](auto full_tag, const unsigned char* st, int rows) {
      constexpr bool FULL = decltype(full_tag)::value;
      int cb0[NS ? NS : 1];
#pragma unroll
      for (int c = 0; c < NS; ++c) cb0[c] = cbase[stage][c];
      fx::RowT<W> row[TMA_RPT];
      bool in[TMA_RPT];
#pragma unroll
      for (int k = 0; k < TMA_RPT; ++k) {
        const int j = tid + k * TMA_CONSUMERS;  // rows past the tail read stale shared memory; masked
        in[k] = FULL || j < rows;
#pragma unroll
        for (int c = 0; c < NF; ++c)
          row[k].f[c] = reinterpret_cast<const double*>(st + L::F_OFF + c * TMA_TILE * 8)[j];
#pragma unroll
        for (int c = 0; c < NI; ++c)
          row[k].i[c] = reinterpret_cast<const int32_t*>(st + L::I_OFF + c * TMA_TILE * 4)[j];
#pragma unroll
        for (int c = 0; c < NS; ++c) {
          if constexpr (NARROW)
            row[k].s[c] = tma_load_str_narrow<TmaKeyTraits<S, NARROW>::CH>(st, L::SO_OFF + c * TMA_SO_BYTES, sc_off[c],
                                                                          cb0[c], prm.schr[c], j, in[k], key_overflow);
          else
            row[k].s[c] = tma_load_str_wide(st, L::SO_OFF + c * TMA_SO_BYTES, sc_off[c], cb0[c], prm.schr[c], j, in[k]);
        }
      }
#pragma unroll
      for (int k = 0; k < TMA_RPT; ++k) {
        const bool valid = in[k] && S::Pred_::eval(row[k], prm.lits);
        double v[NA ? NA : 1];
        fx::EvalF64<typename S::Accs_>::run(row[k], prm.lits, v);  // unconditional; unused when !valid
        int gid = 0;
        if (S::KEYED) {
          W kv = 0;
          bool ovf = false;
          fx::PackKey<typename S::Keys_, SB>::run(row[k], prm.lits, kv, 0, ovf);
          key_overflow |= ovf && valid;
          gid = -1;
          if (valid) gid = tma_group_id_fast(sh, kv);
        } else {
          any_valid |= valid;
          gid = valid ? 0 : -1;
        }
#pragma unroll
        for (int g = 0; g < G; ++g) {
          const bool hit = gid == g;
          if (hit) cnt[g] += 1u;
#pragma unroll
          for (int a = 0; a < NA; ++a)
            if (hit) acc[g][a] += v[a];
        }
        if (S::KEYED && __any_sync(0xffffffffu, gid >= G)) {  // warp-uniform: groups beyond the register-resident ones
          if (gid >= G) {
#pragma unroll
            for (int a = 0; a < NA; ++a) smem_add(&sh.acc[gid][a], v[a]);
            smem_add(&sh.cnt[gid], (u64)1);
          }
        }
      }
    };

    for (int64_t it = 0; it < my_tiles; ++it) {
      mbar_wait(full0 + 8u * stage, phase);
      const unsigned char* st = ring + (size_t)stage * lay.stage_bytes;
      if (!*(volatile int*)&sh.abort) {
        if (own_tail && it == my_tiles - 1)
          process(std::false_type{}, st, tail_rows);
        else
          process(std::true_type{}, st, TMA_TILE);
      }
      __syncwarp();
      if ((tid & 31) == 0) mbar_arrive(empty0 + 8u * stage);
      if (++stage == nst) {
        stage = 0;
        phase ^= 1u;
      }
    }
    if (key_overflow) atomicExch(status, AGG_KEY_TOO_WIDE);
    if (any_valid || S::KEYED) sh.any_valid = 1;
    const int lane = tid & 31;
#pragma unroll
    for (int g = 0; g < G; ++g) {
#pragma unroll
      for (int a = 0; a < NA; ++a) {
        double s = warp_sum(acc[g][a]);
        if (lane == 0 && s != 0.0) smem_add(&sh.acc[g][a], s);
      }
      u64 c = warp_sum((u64)cnt[g]);
      if (lane == 0 && c != 0) smem_add(&sh.cnt[g], c);
    }
  }
  __syncthreads();
  if (S::KEYED && sh.abort) {
    // a CTA that ran out of group ids aborts the whole launch: the host redoes the batch on the
    // high-cardinality kernel (this launch wrote only into a scratch table)
    if (tid == 0) atomicExch(status, AGG_ABORT_CARDINALITY);
    return;
  }
  const int ng = sh.n_groups < SMEM_GROUPS ? sh.n_groups : SMEM_GROUPS;
  for (int g = tid; g < ng; g += TMA_THREADS) {
    if (!S::KEYED && !sh.any_valid) continue;  // ungrouped CTA that saw no valid row
    u128 wide;
    if (NARROW) {
      wide = 0;
      fx::WidenKey<typename S::Keys_, SB, S::SB>::run((u64)sh.gkey[g], 0, wide, 0);
    } else {
      wide = sh.gkey[g];
    }
    u64 slot = table_upsert(tbl, Key128{(u64)wide, (u64)(wide >> 64)});
    if (slot == ~0ull) {
      atomicExch(status, AGG_TABLE_FULL);
      continue;
    }
#pragma unroll
    for (int a = 0; a < NA; ++a) red_add_f64(&tbl.acc[(u64)a * tbl.cap + slot], sh.acc[g][a]);
    red_add_u64(&tbl.cnt[slot], sh.cnt[g]);
  }
}

}  // namespace b2
