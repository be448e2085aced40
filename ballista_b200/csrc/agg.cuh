// agg.cuh — the hash-aggregate kernels (HashAggregateExec Partial/Final).
//
// Two kernels, both templated on a row *source*:
//   agg_small_kernel  : low-cardinality path.  Per-CTA shared-memory open-addressing table maps a packed key
//                       to a dense group id; groups 0..3 accumulate in registers (no atomics at all on the
//                       hot path), further groups in shared memory; each CTA flushes once into the HBM
//                       table with red.global.add.  Aborts (status=1) when a CTA sees > SMEM_GROUPS groups.
//   agg_global_kernel : high-cardinality path.  Every row claims/finds its slot in the HBM table with one
//                       128-bit CAS and accumulates with result-less red.global.add.f64/u64.
// A source is either fused (FusedSrc: scan + predicate + expression evaluation from raw columns, nothing
// materialised) or materialised (ColumnSrc: precomputed key/value/mask columns from the unfused path).
#pragma once
#include "fused.cuh"

namespace b2 {

constexpr int AGG_THREADS = 256;
constexpr int ROWS_PER_THREAD = 4;
constexpr int TILE_ROWS = AGG_THREADS * ROWS_PER_THREAD;
constexpr int SMEM_SLOTS = 512;   // per-CTA key table
constexpr int SMEM_GROUPS = 128;  // max dense group ids per CTA
constexpr int REG_GROUPS = 4;     // group ids accumulated in registers


// ------------------------------------------------------------------------------------------------------
// fused source: columns bound to slots, predicate / keys / aggregate arguments evaluated in registers
// ------------------------------------------------------------------------------------------------------
struct FusedParams {
  const double* f[fx::MAX_F];
  const int32_t* i[fx::MAX_I];
  const int32_t* soff[fx::MAX_S];
  const uint8_t* schr[fx::MAX_S];
  float savg[fx::MAX_S];  // average string bytes per row of each bound Utf8 column (sizes the TMA character window)
  int narrow_lits;        // every Utf8 literal has <= 7 bytes (the narrow kernels can represent them)
  fx::Lits lits;
};

template <class Pred, class Keys, class Accs> struct FusedSrc {
  using KT = fx::ListTraits<Keys>;
  using AT = fx::ListTraits<Accs>;
  static constexpr int NF = fx::cmax(Pred::NF, fx::cmax(KT::NF, AT::NF));
  static constexpr int NI = fx::cmax(Pred::NI, fx::cmax(KT::NI, AT::NI));
  static constexpr int NS = fx::cmax(Pred::NS, fx::cmax(KT::NS, AT::NS));
  static constexpr int NLS = fx::cmax(Pred::NLS, fx::cmax(KT::NLS, AT::NLS));  // Utf8 literal slots
  static constexpr int NACC = Accs::N;
  static constexpr int NCNT = 1;
  static constexpr bool KEYED = Keys::N > 0;
  static constexpr int SB = fx::str_budget(fx::CountStr<Keys>::INTS, fx::CountStr<Keys>::N);
  using Params = FusedParams;
  using Pred_ = Pred;
  using Keys_ = Keys;
  using Accs_ = Accs;
  struct Rows {
    fx::Row r[ROWS_PER_THREAD];
  };
  B2_SIG(return "p:" + Pred::sig() + "|k:" + KT::sig() + "|a:" + AT::sig();)

  // all global loads of a 4-row group are issued before any use (vector loads when the group is full)
  __device__ __forceinline__ static void load(const Params& p, int64_t r0, int64_t n, Rows& rows) {
    if (r0 + ROWS_PER_THREAD <= n) {
#pragma unroll
      for (int c = 0; c < NF; ++c) {
        double2 a = ld_stream_f64x2(p.f[c] + r0);
        double2 b = ld_stream_f64x2(p.f[c] + r0 + 2);
        rows.r[0].f[c] = a.x; rows.r[1].f[c] = a.y; rows.r[2].f[c] = b.x; rows.r[3].f[c] = b.y;
      }
#pragma unroll
      for (int c = 0; c < NI; ++c) {
        int4 a = ld_stream_i32x4(p.i[c] + r0);
        rows.r[0].i[c] = a.x; rows.r[1].i[c] = a.y; rows.r[2].i[c] = a.z; rows.r[3].i[c] = a.w;
      }
#pragma unroll
      for (int c = 0; c < NS; ++c) {
        int4 o = ld_stream_i32x4(p.soff[c] + r0);
        int32_t o4 = __ldg(p.soff[c] + r0 + 4);
        int32_t b[5] = {o.x, o.y, o.z, o.w, o4};
        if (o4 - o.x == 4 && o.y - o.x == 1 && o.z - o.y == 1 && o.w - o.z == 1) {
          // four 1-byte strings: consecutive bytes (TPC-H flag columns)
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            rows.r[k].s[c].len = 1;
            rows.r[k].s[c].bytes = (u128)__ldg(p.schr[c] + o.x + k);
          }
        } else {
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            StrVal s;
            s.len = b[k + 1] - b[k];
            s.bytes = 0;
            int m = s.len < 15 ? s.len : 15;
            for (int q = 0; q < m; ++q) s.bytes |= (u128)__ldg(p.schr[c] + b[k] + q) << (8 * q);
            rows.r[k].s[c] = s;
          }
        }
      }
    } else {
#pragma unroll
      for (int k = 0; k < ROWS_PER_THREAD; ++k) {
        int64_t r = r0 + k;
        if (r < n) {
#pragma unroll
          for (int c = 0; c < NF; ++c) rows.r[k].f[c] = p.f[c][r];
#pragma unroll
          for (int c = 0; c < NI; ++c) rows.r[k].i[c] = p.i[c][r];
#pragma unroll
          for (int c = 0; c < NS; ++c) rows.r[k].s[c] = load_str(p.soff[c], p.schr[c], r);
        } else {
#pragma unroll
          for (int c = 0; c < NF; ++c) rows.r[k].f[c] = 0.0;
#pragma unroll
          for (int c = 0; c < NI; ++c) rows.r[k].i[c] = 0;
#pragma unroll
          for (int c = 0; c < NS; ++c) { rows.r[k].s[c].len = 0; rows.r[k].s[c].bytes = 0; }
        }
      }
    }
  }
  __device__ __forceinline__ static bool pred(const Params& p, const Rows& rows, int k) {
    return Pred::eval(rows.r[k], p.lits);
  }
  __device__ __forceinline__ static void key(const Params& p, const Rows& rows, int k, Key128& key, bool& overflow) {
    u128 v = 0;
    fx::PackKey<Keys, SB>::run(rows.r[k], p.lits, v, 0, overflow);
    key.lo = (u64)v;
    key.hi = (u64)(v >> 64);
  }
  __device__ __forceinline__ static void vals(const Params& p, const Rows& rows, int k, double* v, u64* c) {
    fx::EvalF64<Accs>::run(rows.r[k], p.lits, v);
    c[0] = 1;
  }
};

// ------------------------------------------------------------------------------------------------------
// materialised source: key / value / count / mask columns produced by the unfused path (eval.cu)
// ------------------------------------------------------------------------------------------------------
template <int NA, int NC> struct ColumnParams {
  const Key128* keys;   // null => ungrouped
  const uint8_t* mask;  // null => all rows
  const double* val[NA ? NA : 1];
  const u64* cnt[NC ? NC : 1];
  u64 cnt_default[NC ? NC : 1];  // used when cnt[c] is null (1 = row count, 0 = padding)
};
template <int NA, int NC, bool K> struct ColumnSrc {
  static constexpr int NACC = NA, NCNT = NC;
  static constexpr bool KEYED = K;
  using Params = ColumnParams<NA, NC>;
  struct Rows {
    int64_t r0;
  };
  __device__ __forceinline__ static void load(const Params&, int64_t r0, int64_t, Rows& rows) { rows.r0 = r0; }
  __device__ __forceinline__ static bool pred(const Params& p, const Rows& rows, int k) {
    return p.mask ? p.mask[rows.r0 + k] != 0 : true;
  }
  __device__ __forceinline__ static void key(const Params& p, const Rows& rows, int k, Key128& key, bool&) {
    key = p.keys[rows.r0 + k];
  }
  __device__ __forceinline__ static void vals(const Params& p, const Rows& rows, int k, double* v, u64* c) {
#pragma unroll
    for (int a = 0; a < NA; ++a) v[a] = p.val[a] ? p.val[a][rows.r0 + k] : 0.0;
#pragma unroll
    for (int a = 0; a < NC; ++a) c[a] = p.cnt[a] ? p.cnt[a][rows.r0 + k] : p.cnt_default[a];
  }
};

// shared-memory f64 add (CAS loop in hardware terms; only reached for group ids >= REG_GROUPS and flushes)
__device__ __forceinline__ void smem_add(double* p, double v) { atomicAdd(p, v); }
__device__ __forceinline__ void smem_add(u64* p, u64 v) { atomicAdd(p, v); }

}  // namespace b2
