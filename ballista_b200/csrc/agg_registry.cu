// agg_registry.cu — instantiated fused scan->filter->aggregate kernels and their registry, the binding of
// runtime plans to them, and the materialised (unfused) source.
#include <cstring>
#include <mutex>

#include <cstdlib>
#include <type_traits>

#include "agg.h"
#include "agg_kernels.cuh"
#include "agg_tma.cuh"
#include "jit.h"

namespace b2 {

template <class S> struct is_fused : std::false_type {};
template <class P, class K, class A> struct is_fused<FusedSrc<P, K, A>> : std::true_type {};

static int align_up(int x, int a) { return (x + a - 1) / a * a; }

// shared-memory ring layout of agg_tma_kernel for one bound source; false when the TMA path cannot be used
// (a column pointer that is not 16-byte aligned, e.g. a sliced view, or no room for two stages)
static bool build_tma_layout_rt(const FusedParams& p, int nf, int ni, int ns, size_t static_smem, TmaLayout* lay) {
  memset(lay, 0, sizeof *lay);
  auto aligned = [](const void* q) { return ((uintptr_t)q & 15u) == 0; };
  for (int c = 0; c < nf; ++c)
    if (!aligned(p.f[c])) return false;
  for (int c = 0; c < ni; ++c)
    if (!aligned(p.i[c])) return false;
  int off = nf * TMA_TILE * 8 + ni * TMA_TILE * 4 + ns * TMA_SO_BYTES;  // == TmaStage<S>::SC_OFF
  for (int c = 0; c < ns; ++c) {
    if (!aligned(p.soff[c]) || !aligned(p.schr[c])) return false;
    int cap = (int)(p.savg[c] * 1.5f * TMA_TILE) + 64;
    if (cap < 256) cap = 256;
    if (cap > 16 * TMA_TILE) cap = 16 * TMA_TILE;
    cap = align_up(cap, 128);
    lay->sc_off[c] = off;
    lay->sc_cap[c] = cap;
    off += cap;
  }
  lay->stage_bytes = align_up(off > 0 ? off : 128, 128);
  int budget = 232448 - (int)static_smem - 256;
  int stages = budget / lay->stage_bytes;
  if (const char* e = getenv("B2_TMA_STAGES")) stages = std::min(stages, atoi(e));
  if (stages > TMA_MAX_STAGES) stages = TMA_MAX_STAGES;
  lay->n_stages = stages;
  return stages >= 2;
}
template <class S> static bool build_tma_layout(const FusedParams& p, size_t static_smem, TmaLayout* lay) {
  static_assert(TmaStage<S>::SC_OFF == S::NF * TMA_TILE * 8 + S::NI * TMA_TILE * 4 + S::NS * TMA_SO_BYTES, "stage layout");
  return build_tma_layout_rt(p, S::NF, S::NI, S::NS, static_smem, lay);
}

// launchers of a run-time compiled shape (jit.cu): the same four kernels make_launch<S> binds for a registry shape
static AggLaunch make_launch_jit(Ctx* ctx, const FusedParams& prm, std::shared_ptr<JitKernels> jk, int nf, int ni, int ns,
                                 int nacc, bool keyed) {
  AggLaunch l;
  l.na = nacc;
  l.nc = 1;
  l.keyed = keyed;
  l.small = [ctx, prm, jk](int64_t n, TableRef tbl, int* status) {
    int64_t tiles = (n + TILE_ROWS - 1) / TILE_ROWS;
    int64_t cap = (int64_t)jk->per_sm_small * ctx->sm_count;
    FusedParams p = prm;
    void* args[] = {&p, &n, &tbl, &status};
    ctx->prof_begin();
    jit_launch(jk->small, (int)(tiles < cap ? tiles : cap), AGG_THREADS, 0, ctx->stream, args);
    ctx->prof_end();
    ctx->launches++;
  };
  l.global = [ctx, prm, jk](int64_t rb, int64_t re, int64_t n, TableRef tbl, int* status) {
    int64_t tiles = (re - rb + TILE_ROWS - 1) / TILE_ROWS;
    int64_t cap = (int64_t)jk->per_sm_global * ctx->sm_count;
    FusedParams p = prm;
    void* args[] = {&p, &rb, &re, &n, &tbl, &status};
    ctx->prof_begin();
    jit_launch(jk->global, (int)(tiles < cap ? tiles : cap), AGG_THREADS, 0, ctx->stream, args);
    ctx->prof_end();
    ctx->launches++;
  };
  auto tma = [&](void* fn, int static_smem, std::function<void(int64_t, TableRef, int*)>* out) {
    TmaLayout lay;
    if (!fn || !build_tma_layout_rt(prm, nf, ni, ns, (size_t)static_smem, &lay)) return false;
    *out = [ctx, prm, lay, fn, jk](int64_t n, TableRef tbl, int* status) {
      int64_t tiles = (n + TMA_TILE - 1) / TMA_TILE;
      int grid = (int)(tiles < ctx->sm_count ? tiles : ctx->sm_count);
      FusedParams p = prm;
      TmaLayout ly = lay;
      void* args[] = {&p, &ly, &n, &tbl, &status};
      ctx->prof_begin();
      jit_launch(fn, grid, TMA_THREADS, (size_t)ly.n_stages * ly.stage_bytes, ctx->stream, args);
      ctx->prof_end();
      ctx->launches++;
    };
    return true;
  };
  if (!getenv("B2_DISABLE_TMA")) {
    std::function<void(int64_t, TableRef, int*)> narrow, wide;
    bool have_narrow = jk->narrow_ok && prm.narrow_lits && !getenv("B2_DISABLE_NARROW") &&
                       tma(jk->tma_narrow, jk->static_smem_narrow, &narrow);
    bool have_wide = (!jk->narrow_ok || ns > 0 || !have_narrow) && tma(jk->tma_wide, jk->static_smem_wide, &wide);
    if (have_narrow) {
      l.small_wide = have_wide ? wide : l.small;
      l.small = narrow;
      l.tma = true;
      l.narrow = true;
    } else if (have_wide) {
      l.small = wide;
      l.tma = true;
    }
  }
  return l;
}

template <class S> static AggLaunch make_launch(Ctx* ctx, const typename S::Params& prm) {
  AggLaunch l;
  l.na = S::NACC;
  l.nc = S::NCNT;
  l.keyed = S::KEYED;
  static int per_sm_small = 0, per_sm_global = 0;
  if (!per_sm_small) {
    B2_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm_small, agg_small_kernel<S>, AGG_THREADS, 0));
    B2_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm_global, agg_global_kernel<S>, AGG_THREADS, 0));
    if (per_sm_small < 1) per_sm_small = 1;
    if (per_sm_global < 1) per_sm_global = 1;
  }
  int ps = per_sm_small, pg = per_sm_global;
  l.small = [ctx, prm, ps](int64_t n, TableRef tbl, int* status) {
    int64_t tiles = (n + TILE_ROWS - 1) / TILE_ROWS;
    int64_t cap = (int64_t)ps * ctx->sm_count;
    int grid = (int)(tiles < cap ? tiles : cap);
    if (is_fused<S>::value) ctx->prof_begin();
    agg_small_kernel<S><<<grid, AGG_THREADS, 0, ctx->stream>>>(prm, n, tbl, status);
    if (is_fused<S>::value) ctx->prof_end();
    B2_LAUNCHED(ctx);
  };
  l.global = [ctx, prm, pg](int64_t rb, int64_t re, int64_t n, TableRef tbl, int* status) {
    int64_t tiles = (re - rb + TILE_ROWS - 1) / TILE_ROWS;
    int64_t cap = (int64_t)pg * ctx->sm_count;
    int grid = (int)(tiles < cap ? tiles : cap);
    if (is_fused<S>::value) ctx->prof_begin();
    agg_global_kernel<S><<<grid, AGG_THREADS, 0, ctx->stream>>>(prm, rb, re, n, tbl, status);
    if (is_fused<S>::value) ctx->prof_end();
    B2_LAUNCHED(ctx);
  };
  if constexpr (is_fused<S>::value) {
    // bandwidth path: persistent TMA-fed kernel, one CTA per SM (agg_tma.cuh).  Narrow variant (8-byte keys,
    // Utf8 values of <= 7 bytes) when the key shape allows it; the wide variant is its overflow fallback.
    using KS = fx::CountStr<typename S::Keys_>;
    constexpr bool NARROW_OK = fx::narrow_key_ok(KS::INTS, KS::N);
    constexpr bool NEED_WIDE = !NARROW_OK || S::NS > 0;
    auto tma_launcher = [ctx, prm](auto kernel, std::function<void(int64_t, TableRef, int*)>* out) {
      cudaFuncAttributes fa;
      B2_CUDA(cudaFuncGetAttributes(&fa, kernel));
      B2_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                   232448 - (int)fa.sharedSizeBytes));
      TmaLayout lay;
      if (!build_tma_layout<S>(prm, fa.sharedSizeBytes, &lay)) return false;
      *out = [ctx, prm, lay, kernel](int64_t n, TableRef tbl, int* status) {
        int64_t tiles = (n + TMA_TILE - 1) / TMA_TILE;
        int grid = (int)(tiles < ctx->sm_count ? tiles : ctx->sm_count);
        size_t dyn = (size_t)lay.n_stages * lay.stage_bytes;
        ctx->prof_begin();
        kernel<<<grid, TMA_THREADS, dyn, ctx->stream>>>(prm, lay, n, tbl, status);
        ctx->prof_end();
        B2_LAUNCHED(ctx);
      };
      return true;
    };
    if (!getenv("B2_DISABLE_TMA")) {
      std::function<void(int64_t, TableRef, int*)> narrow, wide;
      bool have_narrow = false, have_wide = false;
      if constexpr (NARROW_OK) {
        if (prm.narrow_lits && !getenv("B2_DISABLE_NARROW")) have_narrow = tma_launcher(agg_tma_kernel<S, true>, &narrow);
      }
      if constexpr (NEED_WIDE) have_wide = tma_launcher(agg_tma_kernel<S, false>, &wide);
      if (have_narrow) {
        l.small_wide = have_wide ? wide : l.small;
        l.small = narrow;
        l.tma = true;
        l.narrow = true;
      } else if (have_wide) {
        l.small = wide;
        l.tma = true;
      }
    }
  }
  return l;
}

// ------------------------------------------------------------------------------------------------------
// registry of fused shapes.  Slots are numbered in first-encounter order over (predicate, keys, args),
// exactly as expr_signature() numbers them.  Literals are runtime values.
// ------------------------------------------------------------------------------------------------------
using namespace fx;

template <class S> static void add_entry(std::map<std::string, FusedEntry>& m) {
  FusedEntry e;
  e.sig = S::sig();
  e.nf = S::NF;
  e.ni = S::NI;
  e.ns = S::NS;
  e.na = S::NACC;
  e.keyed = S::KEYED;
  e.bind = [](Ctx* ctx, const FusedParams& p) { return make_launch<S>(ctx, p); };
  m[e.sig] = e;
}

// TPC-H Q1 (queries/q1.sql): l_shipdate <= d; keys (l_returnflag, l_linestatus);
//   args qty, price, price*(1-disc), price*(1-disc)*(1+tax), disc
using Q1Pred = Le<I<0>, LI<0>>;
using Q1Keys = List<S<0>, S<1>>;
using Q1Args = List<F<0>, F<1>, Mul<F<1>, Sub<LF<0>, F<2>>>, Mul<Mul<F<1>, Sub<LF<0>, F<2>>>, Add<LF<0>, F<3>>>, F<2>>;
// TPC-H Q6 (queries/q6.sql): shipdate in [a,b) and discount between lo and hi and quantity < q; sum(price*disc)
using Q6Pred = And<And<And<And<Ge<I<0>, LI<0>>, Lt<I<0>, LI<1>>>, Ge<F<0>, LF<0>>>, Le<F<0>, LF<1>>>, Lt<F<1>, LF<2>>>;
using Q6Args = List<Mul<F<2>, F<0>>>;
// simple shapes used by tests / ad-hoc plans
using SumF0 = List<F<0>>;

static std::map<std::string, FusedEntry> build_registry() {
  std::map<std::string, FusedEntry> m;
  add_entry<FusedSrc<Q1Pred, Q1Keys, Q1Args>>(m);
  add_entry<FusedSrc<True, Q1Keys, Q1Args>>(m);
  add_entry<FusedSrc<Q6Pred, List<>, Q6Args>>(m);
  add_entry<FusedSrc<True, List<>, SumF0>>(m);
  add_entry<FusedSrc<True, List<I<0>>, SumF0>>(m);
  add_entry<FusedSrc<Le<I<0>, LI<0>>, List<I<1>>, SumF0>>(m);
  return m;
}
const std::map<std::string, FusedEntry>& fused_registry() {
  static std::map<std::string, FusedEntry> m = build_registry();
  return m;
}

static StrVal pack_literal(const std::string& s) {
  StrVal v;
  v.len = (int)s.size();
  v.bytes = 0;
  for (size_t q = 0; q < s.size() && q < 15; ++q) v.bytes |= (u128)(uint8_t)s[q] << (8 * q);
  return v;
}

bool bind_fused_agg(Ctx* ctx, const ExprPtr& pred, const std::vector<ExprPtr>& groups,
                    const std::vector<ExprPtr>& args, const Batch& batch, AggLaunch* out, std::string* sig_out) {
  SlotBinder b;
  std::string sig = "p:" + (pred ? expr_signature(pred, b) : std::string("true")) + "|k:";
  for (size_t i = 0; i < groups.size(); ++i) sig += (i ? "," : "") + expr_signature(groups[i], b);
  sig += "|a:";
  for (size_t i = 0; i < args.size(); ++i) sig += (i ? "," : "") + expr_signature(args[i], b);
  if (sig_out) *sig_out = sig;
  auto& reg = fused_registry();
  auto it = reg.find(sig);
  std::shared_ptr<JitKernels> jk;
  if (it == reg.end()) {
    // not a pre-built shape: compile it (once per shape and process; the cubin is also cached on disk)
    if (getenv("B2_DISABLE_JIT")) return false;
    std::string why;
    jk = jit_fused_kernels(sig, &why);
    if (!jk) {
      if (sig_out) *sig_out = sig + " [no fused kernel: " + why.substr(0, 200) + "]";
      return false;
    }
  }
  if ((int)b.f64_cols.size() > MAX_F || (int)b.i32_cols.size() > MAX_I || (int)b.str_cols.size() > MAX_S ||
      (int)b.f64_lits.size() > MAX_LF || (int)b.i64_lits.size() > MAX_LI || (int)b.str_lits.size() > MAX_LS)
    return false;
  for (auto& s : b.str_lits)
    if (s.size() > 15) return false;
  FusedParams p;
  memset(&p, 0, sizeof p);
  for (size_t k = 0; k < b.f64_cols.size(); ++k) p.f[k] = batch.cols[b.f64_cols[k]].values<double>();
  for (size_t k = 0; k < b.i32_cols.size(); ++k) p.i[k] = batch.cols[b.i32_cols[k]].values<int32_t>();
  for (size_t k = 0; k < b.str_cols.size(); ++k) {
    p.soff[k] = batch.cols[b.str_cols[k]].values<int32_t>();
    p.schr[k] = batch.cols[b.str_cols[k]].chars->as<uint8_t>();
    const Column& sc = batch.cols[b.str_cols[k]];
    p.savg[k] = sc.length ? (float)sc.chars_len / (float)sc.length : 0.f;
  }
  p.narrow_lits = 1;
  for (auto& lit : b.str_lits)
    if (lit.size() > 7) p.narrow_lits = 0;
  for (size_t k = 0; k < b.f64_lits.size(); ++k) p.lits.f[k] = b.f64_lits[k];
  for (size_t k = 0; k < b.i64_lits.size(); ++k) p.lits.i[k] = b.i64_lits[k];
  for (size_t k = 0; k < b.str_lits.size(); ++k) p.lits.s[k] = pack_literal(b.str_lits[k]);
  if (jk) {
    *out = make_launch_jit(ctx, p, jk, (int)b.f64_cols.size(), (int)b.i32_cols.size(), (int)b.str_cols.size(),
                           (int)args.size(), !groups.empty());
    if (sig_out) *sig_out = sig + (jk->compiled_now ? " {jit}" : " {jit,cached}");
  } else {
    *out = it->second.bind(ctx, p);
  }
  return true;
}

// ------------------------------------------------------------------------------------------------------
// materialised source (unfused path).  Instantiated for padded (NA, NC) in {1,2,4,8} x {1,3}.
// ------------------------------------------------------------------------------------------------------
template <int NA, int NC, bool K>
static AggLaunch bind_cols(Ctx* ctx, const Key128* keys, const uint8_t* mask, const std::vector<const double*>& vals,
                           const std::vector<const u64*>& cnts) {
  ColumnParams<NA, NC> p;
  memset(&p, 0, sizeof p);
  p.keys = keys;
  p.mask = mask;
  for (size_t a = 0; a < vals.size(); ++a) p.val[a] = vals[a];
  for (size_t a = 0; a < cnts.size(); ++a) {
    p.cnt[a] = cnts[a];
    p.cnt_default[a] = 1;  // a null entry counts rows; padding slots beyond cnts.size() keep 0
  }
  return make_launch<ColumnSrc<NA, NC, K>>(ctx, p);
}
template <int NA, int NC>
static AggLaunch bind_cols_k(Ctx* ctx, const Key128* keys, const uint8_t* mask, const std::vector<const double*>& vals,
                             const std::vector<const u64*>& cnts) {
  return keys ? bind_cols<NA, NC, true>(ctx, keys, mask, vals, cnts) : bind_cols<NA, NC, false>(ctx, keys, mask, vals, cnts);
}
template <int NA>
static AggLaunch bind_cols_c(Ctx* ctx, const Key128* keys, const uint8_t* mask, const std::vector<const double*>& vals,
                             const std::vector<const u64*>& cnts) {
  size_t nc = cnts.size();
  if (nc <= 1) return bind_cols_k<NA, 1>(ctx, keys, mask, vals, cnts);
  if (nc <= 4) return bind_cols_k<NA, 4>(ctx, keys, mask, vals, cnts);
  if (nc <= 12) return bind_cols_k<NA, 12>(ctx, keys, mask, vals, cnts);
  fail(B2_ERR_NOT_IMPLEMENTED, "more than 12 count / integer-sum accumulators in one aggregate");
}
AggLaunch bind_column_agg(Ctx* ctx, const Key128* keys, const uint8_t* mask, const std::vector<const double*>& vals,
                          const std::vector<const u64*>& cnts) {
  size_t na = vals.size();
  if (na <= 1) return bind_cols_c<1>(ctx, keys, mask, vals, cnts);
  if (na <= 2) return bind_cols_c<2>(ctx, keys, mask, vals, cnts);
  if (na <= 4) return bind_cols_c<4>(ctx, keys, mask, vals, cnts);
  if (na <= 8) return bind_cols_c<8>(ctx, keys, mask, vals, cnts);
  fail(B2_ERR_NOT_IMPLEMENTED, "more than 8 distinct Float64 accumulators in one aggregate");
}

// ------------------------------------------------------------------------------------------------------
// key packing for the unfused path
// ------------------------------------------------------------------------------------------------------
struct PackPart {
  int dtype, byte_off, bytes;
  const void* data;
  const uint8_t* chars;
  const uint8_t* valid;
};
struct PackArgs {
  PackPart part[8];
  int n;
};
__global__ void pack_keys_kernel(PackArgs a, int64_t n, Key128* out, int* status, int null_tag) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  bool overflow = false;
  for (; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    u128 key = 0;
    bool is_null = false;
    for (int p = 0; p < a.n; ++p) {
      const PackPart& pp = a.part[p];
      if (pp.valid && !pp.valid[i]) {
        is_null = true;
        continue;
      }
      if (pp.dtype == B2_UTF8) {
        StrVal s = load_str((const int32_t*)pp.data, pp.chars, i);
        if (s.len > pp.bytes - 1) overflow = true;
        u128 m = (((u128)1 << (8 * (pp.bytes - 1))) - 1);
        key |= ((u128)(uint8_t)s.len | ((s.bytes & m) << 8)) << (8 * pp.byte_off);
      } else if (pp.bytes == 8) {
        key |= (u128)(u64)((const int64_t*)pp.data)[i] << (8 * pp.byte_off);
      } else {
        key |= (u128)(uint32_t)((const int32_t*)pp.data)[i] << (8 * pp.byte_off);
      }
    }
    out[i] = is_null ? Key128{0, (u64)null_tag << 56} : Key128{(u64)key, (u64)(key >> 64)};
  }
  if (overflow) atomicExch(status, AGG_KEY_TOO_WIDE);
}

BufPtr pack_keys(Ctx* ctx, const KeyLayout& layout, const std::vector<Column>& cols, int64_t n, int* status_dev,
                 int null_tag) {
  if (layout.parts.size() > 8) fail(B2_ERR_NOT_IMPLEMENTED, "more than 8 key columns");
  if (!null_tag)
    for (auto& c : cols)
      if (c.valid) fail(B2_ERR_NOT_IMPLEMENTED, "NULLs in a group / partitioning key column");
  PackArgs a;
  memset(&a, 0, sizeof a);
  a.n = (int)layout.parts.size();
  for (int p = 0; p < a.n; ++p) {
    a.part[p].dtype = layout.parts[p].dtype;
    a.part[p].byte_off = layout.parts[p].byte_off;
    a.part[p].bytes = layout.parts[p].bytes;
    a.part[p].data = cols[p].data->ptr;
    a.part[p].chars = cols[p].chars ? cols[p].chars->as<uint8_t>() : nullptr;
    a.part[p].valid = cols[p].validity();
  }
  auto out = dev_alloc(ctx, (size_t)(n + 1) * sizeof(Key128));
  if (n) {
    pack_keys_kernel<<<grid_for(ctx, 8, n, 256), 256, 0, ctx->stream>>>(a, n, out->as<Key128>(), status_dev, null_tag);
    B2_LAUNCHED(ctx);
  }
  return out;
}

}  // namespace b2
