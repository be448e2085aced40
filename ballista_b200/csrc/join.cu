// join.cu — hash-join build and probe kernels (HashJoinExec).
//
// Build: open-addressing table of distinct packed keys in HBM, claimed with one 128-bit CAS; every build
// row is pushed onto its key's chain (head[slot] / next[row]) and the key's match count is bumped, so the
// probe sizes its output without walking chains.  Probe is two-pass (count -> scan -> write) and emits
// (build_row, probe_row) pairs in probe order, which is the reference's output order.
#include <algorithm>
#include <cmath>
#include <limits>

#include "ops.h"
#include "rowpred.cuh"
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
    B2_READBACK(ctx, &total, sums->as<int64_t>() + nb, 8);
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
// The table is ONE array of 8-byte slots, (key << 32) | head_row, grouped in buckets of 4 slots = one 32-byte
// sector.  A probe loads its key's whole bucket with two 16-byte loads of the same sector and is finished
// unless the bucket is full without a match (rare below 50 % load), so neither hits nor misses walk a
// divergent linear-probing chain; a build insert is one 64-bit CAS on the first empty slot of the bucket
// (slots of a bucket fill in order, so the first EMPTY slot ends a lookup).  EMPTY is all ones (a claimed slot
// always has head_row != NIL).  Build rows with equal keys are chained through next[] (head swap by CAS) and
// raise flags[0]; for primary-key builds (flags[0] == 0) next[] is never read.  The table is sized for the
// rows that pass the fused filter, so a selective build side stays L2-resident.
//
// Beside the slots the build keeps a blocked Bloom filter of its keys (table.cuh: `bloom_bits`; one 32-bit word
// per key, >= 16 bits per key): the probe tests it before touching the table, and a partitioned join across ranks
// publishes it so that the probe side's exchange drops non-matching rows before they move (ops_join.cu).
//
// Both kernels take the FilterExec mask of their input (or null) and the ORIGINAL, uncompacted key column:
// the filter is never materialised, row ids are those of the unfiltered batch, and output columns are
// gathered once, for matching rows only (late materialisation).
// ------------------------------------------------------------------------------------------------------
constexpr int BUCKET = 4;  // slots per bucket (one 32-byte sector)
__device__ __forceinline__ u64 ld_slot_volatile(const u64* p) {
  u64 v;
  asm volatile("ld.volatile.global.u64 %0, [%1];" : "=l"(v) : "l"(p));
  return v;
}

static void flatten_and(const ExprPtr& e, std::vector<ExprPtr>* terms) {
  if (e->kind == B2_EXPR_BINARY && e->op == B2_OP_AND) {
    flatten_and(e->ch[0], terms);
    flatten_and(e->ch[1], terms);
    return;
  }
  terms->push_back(e);
}
static int flip_cmp(int op) {
  switch (op) {
    case B2_OP_LT: return B2_OP_GT;
    case B2_OP_LTEQ: return B2_OP_GTEQ;
    case B2_OP_GT: return B2_OP_LT;
    case B2_OP_GTEQ: return B2_OP_LTEQ;
    default: return op;
  }
}
bool compile_row_pred(const ExprPtr& pred, const Batch& batch, RowPred* out) {
  std::vector<ExprPtr> terms;
  flatten_and(pred, &terms);
  if (terms.empty() || terms.size() > 4) return false;
  RowPred rp;
  for (auto& t : terms) {
    if (t->kind != B2_EXPR_BINARY) return false;
    int op = t->op;
    if (op != B2_OP_EQ && op != B2_OP_LT && op != B2_OP_LTEQ && op != B2_OP_GT && op != B2_OP_GTEQ) return false;
    ExprPtr c = t->ch[0], l = t->ch[1];
    if (c->kind == B2_EXPR_LITERAL && l->kind == B2_EXPR_COLUMN) {
      std::swap(c, l);
      op = flip_cmp(op);
    }
    if (c->kind != B2_EXPR_COLUMN || l->kind != B2_EXPR_LITERAL) return false;
    const Column& col = batch.cols[c->col];
    if (col.valid) return false;  // NULLs: the generic mask path folds them to false
    // a second term on a column already constrained narrows that term's range (d1 <= x AND x < d2 is ONE range test:
    // the one-term forms of the kernels, one load per row)
    int k = -1;
    for (int e = 0; e < rp.n; ++e)
      if (rp.col[e] == col.data->ptr) k = e;
    const bool merged = k >= 0;
    if (!merged) {
      k = rp.n++;
      rp.col[k] = col.data->ptr;
      rp.ilo[k] = 0;
      rp.iwidth[k] = 0;
      rp.flo[k] = rp.fhi[k] = 0;
    }
    if ((col.dtype == B2_INT32 || col.dtype == B2_DATE32) &&
        (l->dtype == B2_INT32 || l->dtype == B2_DATE32 || l->dtype == B2_INT64)) {
      const long long MIN = INT32_MIN, MAX = INT32_MAX;
      long long lo = MIN, hi = MAX, v = l->i;
      switch (op) {
        case B2_OP_EQ: lo = hi = v; break;
        case B2_OP_LT: hi = v - 1; break;
        case B2_OP_LTEQ: hi = v; break;
        case B2_OP_GT: lo = v + 1; break;
        default: lo = v; break;
      }
      lo = std::max(lo, MIN);
      hi = std::min(hi, MAX);
      if (merged) {
        if (rp.is_f64[k]) return false;
        lo = std::max<long long>(lo, rp.ilo[k]);
        hi = std::min<long long>(hi, (long long)rp.ilo[k] + (long long)rp.iwidth[k]);
      }
      rp.is_f64[k] = 0;
      if (lo > hi) return false;  // constant-false comparison: rare, the generic mask path handles it
      rp.ilo[k] = (int)lo;
      rp.iwidth[k] = (unsigned)(hi - lo);
    } else if (col.dtype == B2_FLOAT64 && l->dtype == B2_FLOAT64) {
      double v = l->f, inf = std::numeric_limits<double>::infinity();
      if (v != v) return false;  // NaN literal
      double lo = -inf, hi = inf;
      switch (op) {
        case B2_OP_EQ: lo = hi = v; break;
        case B2_OP_LT: hi = std::nextafter(v, -inf); break;
        case B2_OP_LTEQ: hi = v; break;
        case B2_OP_GT: lo = std::nextafter(v, inf); break;
        default: lo = v; break;
      }
      if (merged) {
        if (!rp.is_f64[k]) return false;
        lo = std::max(lo, rp.flo[k]);
        hi = std::min(hi, rp.fhi[k]);
        if (lo > hi) return false;  // empty range: the generic mask path
      }
      rp.is_f64[k] = 1;
      rp.flo[k] = lo;
      rp.fhi[k] = hi;
    } else {
      return false;
    }
  }
  *out = rp;
  return true;
}

// `perm` (optional): the rows to insert, already filtered and ordered by the top bits of their bucket (a 32-way split,
// repartition.cu): consecutive threads then fill one slice of the table at a time and the inserts hit L2, where rows
// in input order scatter 8-byte CAS operations over the whole table in HBM.
__global__ void join_build32_kernel(const int32_t* __restrict__ key, RowPred pred, const uint8_t* __restrict__ mask,
                                    const uint32_t* __restrict__ perm, int64_t n, u64* slots, u64 bmask, int bshift,
                                    uint32_t* next, int* flags, int ahead) {
  int64_t j = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (; j < n; j += stride) {
    // Experiment (B2_BUILD_AHEAD=k, off by default): request the bucket of the row this thread inserts k rounds from now
    // into L2.  Measured on Q3 SF100 (14.7 M inserts into 256 MB): 0 -> 4.38 ms per step, 2 -> 4.64, 4 -> 4.63: the build
    // is bound by the rate of random 32-byte sectors HBM sustains, not by the latency of one insert's chain.
    {
      const int64_t ja = j + ahead * stride;
      if (ahead && ja < n) {
        const int64_t ia = perm ? (int64_t)perm[ja] : ja;
        if (perm || ((!mask || mask[ia]) && row_pred(pred, ia))) {
          const u64 ba = (u64)(table_hash(fmix32((uint32_t)key[ia])) >> bshift);
          asm volatile("prefetch.global.L2 [%0];" ::"l"(slots + ba * BUCKET));
        }
      }
    }
    const int64_t i = perm ? (int64_t)perm[j] : j;
    if (!perm) {
      if (mask && !mask[i]) continue;
      if (!row_pred(pred, i)) continue;
    }
    uint32_t k = (uint32_t)key[i];
    const uint32_t h = fmix32(k);
    u64 b = (u64)(table_hash(h) >> bshift);
    u64 mine = ((u64)k << 32) | (uint32_t)i;
    bool placed = false;
    for (u64 probe = 0; probe <= bmask && !placed; ++probe) {
      u64* bs = slots + b * BUCKET;
      for (int sidx = 0; sidx < BUCKET && !placed; ++sidx) {
        u64 cur = ld_slot_volatile(bs + sidx);
        if (cur == KEY_EMPTY) {
          u64 old = atomicCAS((unsigned long long*)(bs + sidx), (unsigned long long)KEY_EMPTY, (unsigned long long)mine);
          if (old == KEY_EMPTY) {
            next[i] = NIL;
            placed = true;
            break;
          }
          cur = old;
        }
        while ((uint32_t)(cur >> 32) == k) {  // equal key: push this row onto the chain
          u64 old = atomicCAS((unsigned long long*)(bs + sidx), (unsigned long long)cur, (unsigned long long)mine);
          if (old == cur) {
            next[i] = (uint32_t)cur;
            flags[0] = 1;
            placed = true;
            break;
          }
          cur = old;
        }
      }
      b = (b + 1) & bmask;
    }
  }
}

// Bloom filter over the build keys, as its own pass before the table build: the filter (<= 4 B per key) stays in L2
// under the reductions, which it does not when they are interleaved with the table's random 32-byte sector traffic
__global__ void bloom_build_kernel(const int32_t* __restrict__ key, RowPred pred, const uint8_t* __restrict__ mask, int64_t n,
                                   uint32_t* bloom, uint32_t wmask) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  for (; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    if (mask && !mask[i]) continue;
    if (!row_pred(pred, i)) continue;
    const uint32_t h = fmix32((uint32_t)key[i]);
    atomicOr(bloom + (h & wmask), bloom_bits(h));
  }
}

__global__ void pass_count_kernel(RowPred pred, const uint8_t* __restrict__ mask, int64_t n, unsigned long long* out) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  unsigned c = 0;
  for (; i < n; i += (int64_t)gridDim.x * blockDim.x) c += (!mask || mask[i]) && row_pred(pred, i);
  for (int d = 16; d > 0; d >>= 1) c += __shfl_xor_sync(0xffffffffu, c, d);
  if ((threadIdx.x & 31) == 0 && c) atomicAdd(out, (unsigned long long)c);
}

std::shared_ptr<JoinTable> join_build32(Ctx* ctx, const int32_t* key, const RowPred& pred, const uint8_t* mask,
                                        int64_t n) {
  if (n >= (int64_t)NIL) fail(B2_ERR_NOT_IMPLEMENTED, "join build side beyond 2^32-2 rows");
  auto t = std::make_shared<JoinTable>();
  t->ctx = ctx;
  t->n_build = n;
  t->k32 = true;
  t->flags = dev_alloc_zero(ctx, 16);
  // size the table for the rows that pass the fused filter, so that a selective build side stays L2-resident
  int64_t n_in = n;
  // Large build sides are first split 32 ways by the top bits of their table bucket (destination bytes, per-tile
  // histogram, one scan, stable scatter of row ids: the hash-repartition kernels), which also counts the rows that pass
  // the fused filter; the insert kernel then walks the permutation.
  BufPtr perm;
  // (measured on Q3 SF100, 14.7 M build rows into a 256 MB table: the pre-sorted build takes 1.14 ms + 0.2 ms of split
  // against 0.72 ms for inserts in input order - the L2 atomics of one 8 MB slice serialise harder than CAS operations
  // spread over HBM - so it stays an experiment behind B2_PARTITIONED_BUILD=1)
  const bool big = n > (1 << 21) && getenv("B2_PARTITIONED_BUILD");
  if (big) {
    SplitState st = hash_split_begin(ctx, nullptr, key, n, 32, &pred, mask, nullptr, /*table_order=*/true);
    std::vector<int64_t> counts;
    hash_split_counts(ctx, st, &counts);
    n_in = 0;
    for (int64_t c : counts) n_in += c;
    if (n_in > (1 << 21)) perm = hash_split_perm(ctx, st);  // a small surviving set builds an L2-resident table anyway
  } else if ((mask || pred.n) && n > (1 << 18)) {
    pass_count_kernel<<<grid_for(ctx, 8, n, 1024), 256, 0, ctx->stream>>>(
        pred, mask, n, reinterpret_cast<unsigned long long*>(t->flags->as<char>() + 8));
    B2_LAUNCHED(ctx);
    unsigned long long c = 0;
    B2_READBACK(ctx, &c, t->flags->as<char>() + 8, 8);
    n_in = (int64_t)c;
  }
  u64 cap = 64;  // slots; load factor <= 0.5, BUCKET slots per bucket
  while (cap < 2 * (u64)n_in + 16) cap <<= 1;
  t->cap = cap;
  int lb = 0;
  while ((1ull << lb) < cap / BUCKET) ++lb;
  t->bshift = 32 - lb;
  t->slots = dev_alloc(ctx, cap * 8);
  t->next = dev_alloc(ctx, (size_t)(n + 1) * 4);
  B2_CUDA(cudaMemsetAsync(t->slots->ptr, 0xFF, cap * 8, ctx->stream));
  u64 words = 1024;  // Bloom filter: >= 16 bits per build key, 32-bit blocks
  while (words * 2 < (u64)n_in) words <<= 1;
  t->bloom_words = words;
  t->bloom = dev_alloc(ctx, words * 4);
  B2_CUDA(cudaMemsetAsync(t->bloom->ptr, 0, words * 4, ctx->stream));
  if (n) {
    bloom_build_kernel<<<grid_for(ctx, 8, n, 256), 256, 0, ctx->stream>>>(key, pred, mask, n, t->bloom->as<uint32_t>(),
                                                                        (uint32_t)(words - 1));
    B2_LAUNCHED(ctx);
    const int64_t work = perm ? n_in : n;
    if (work) {
      join_build32_kernel<<<grid_for(ctx, 8, work, 256), 256, 0, ctx->stream>>>(
          key, pred, mask, perm ? perm->as<uint32_t>() : nullptr, work, t->slots->as<u64>(), cap / BUCKET - 1, t->bshift,
          t->next->as<uint32_t>(), t->flags->as<int>(), getenv("B2_BUILD_AHEAD") ? atoi(getenv("B2_BUILD_AHEAD")) : 0);
      B2_LAUNCHED(ctx);
    }
  }
  t->dups_known = false;
  return t;
}

constexpr int JP_THREADS = SCAN_THREADS;  // block_exclusive_scan is written for this CTA size
constexpr int JP_ITEMS = 8;
constexpr int JP_TILE = JP_THREADS * JP_ITEMS;

// predicate over the JP_ITEMS consecutive rows of one thread as a BIT MASK (bit j = row r0 + j passes), 128-bit loads
// (r0 is a multiple of 8): three instructions per Int32 row (subtract, unsigned compare, select into the mask)
__device__ __forceinline__ unsigned row_pred_bits(const RowPred& p, int64_t r0, unsigned pm) {
  for (int t = 0; t < p.n; ++t) {
    if (p.is_f64[t]) {
      const double* c = static_cast<const double*>(p.col[t]) + r0;
      const double lo = p.flo[t], hi = p.fhi[t];
#pragma unroll
      for (int j = 0; j < JP_ITEMS; j += 2) {
        double2 v = ld_stream_f64x2(c + j);
        if (!in_range_f64(v.x, lo, hi)) pm &= ~(1u << j);
        if (!in_range_f64(v.y, lo, hi)) pm &= ~(2u << j);
      }
    } else {
      const int32_t* c = static_cast<const int32_t*>(p.col[t]) + r0;
      const int lo = p.ilo[t];
      const unsigned w = p.iwidth[t];
      int4 v0 = ld_stream_i32x4(c), v1 = ld_stream_i32x4(c + 4);
      if (!in_range_i32(v0.x, lo, w)) pm &= ~1u;
      if (!in_range_i32(v0.y, lo, w)) pm &= ~2u;
      if (!in_range_i32(v0.z, lo, w)) pm &= ~4u;
      if (!in_range_i32(v0.w, lo, w)) pm &= ~8u;
      if (!in_range_i32(v1.x, lo, w)) pm &= ~16u;
      if (!in_range_i32(v1.y, lo, w)) pm &= ~32u;
      if (!in_range_i32(v1.z, lo, w)) pm &= ~64u;
      if (!in_range_i32(v1.w, lo, w)) pm &= ~128u;
    }
  }
  return pm;
}

// Probe-side kernel: filter (inline predicate and / or mask) -> Bloom test -> table lookup -> (build_row,
// probe_row) pairs in PROBE ORDER (the reference's output order).  The unit of work is a WARP TILE of JP_WTILE = 256
// consecutive rows and warps never synchronise with each other:
//   (1) every lane tests its 8 consecutive rows against the predicate (128-bit loads) into an 8-bit pass mask.  A probe
//       side clustered on the key (lineitem: ~4 rows per order) repeats keys in consecutive rows, so a passing row
//       whose predecessor passed with the same key CONTINUES that row's run instead of being looked up: a second 8-bit
//       mask; run heads = pass & ~continues.  Both masks stay in shared memory for step (3b);
//   (2) the run heads are compacted, in row order, into the warp's slice of shared memory (one shuffle scan);
//   (3) the warp walks the compacted heads 32 at a time: Bloom word, and only on a Bloom hit the table bucket; the
//       hits of a chunk are compacted in place by ballot, and the output of the hits (run length from the owner lane's
//       masks: the zeros of `continues` after the head's bit) is placed by a second shuffle scan.
// Only the rows that survive a stage pay for the next one: the hash / Bloom arithmetic runs over the compacted
// heads (13% of lineitem's rows in Q3), the bucket walk over the Bloom hits (~1%), the run length is computed for
// hits only.  Three modes, none of which probes a row twice on the primary-key path:
//   PROBE_STAGE  : pairs of warp tile t are written, compacted, at [t * JP_WTILE, ...) of staging buffers (a tile has
//                  at most JP_WTILE matches when build keys are unique) and tile_cnt[t] = match count; a scan +
//                  `join_compact_kernel` then move each tile's run to its final offset.
//   PROBE_COUNT  : tile_cnt[t] only (build side with duplicate keys: chains are walked to size the output).
//   PROBE_WRITE  : pairs go straight to tile_cnt[t] (already scanned to offsets) + offset within tile.
enum { PROBE_STAGE = 0, PROBE_COUNT = 1, PROBE_WRITE = 2 };
constexpr int JP_WTILE = 32 * JP_ITEMS;
// PRED: 0 = no inline predicate, 1 = exactly one Int32 / Date32 range term (every TPC-H date filter: its column pointer
// and bounds live in registers across the tile loop), 2 = the general RowPred loop
template <int MODE, bool DUPS, int PRED>
// (register bound: the predicate-free form is held to 32 registers = 8 CTAs per SM; the one-term form reaches 32 on its own
// under the looser bound, and forcing it there costs 3 % - Q3 SF100 4.34 vs 4.47 ms per step, same box)
__global__ void __launch_bounds__(JP_THREADS, PRED == 0 ? 8 : 5)
    join_probe32_kernel(const int32_t* __restrict__ key, const __grid_constant__ RowPred pred,
                        const uint8_t* __restrict__ mask, int64_t n, bool vec, const u64* __restrict__ slots, u64 bmask,
                        int bshift, const uint32_t* __restrict__ bloom, uint32_t wmask, const uint32_t* __restrict__ next,
                        uint32_t* __restrict__ left_idx, uint32_t* __restrict__ right_idx, int64_t* tile_cnt) {
  // run heads of the warp's tile in row order: key | (row within the tile) << 32; rewritten in place by the hits of
  // step (3a): build row | the same position word
  __shared__ u64 s_ent[JP_TILE];
  __shared__ unsigned s_bits[JP_THREADS];  // per lane: pass mask | continues mask << 8
  const int lane = threadIdx.x & 31;
  u64* went = s_ent + (threadIdx.x >> 5) * JP_WTILE;
  unsigned* wbits = s_bits + (threadIdx.x & ~31);
  const int64_t n_tiles = (n + JP_WTILE - 1) / JP_WTILE;
  const int64_t warp0 = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
  const int64_t n_warps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  const int32_t* const p1_col = PRED == 1 ? static_cast<const int32_t*>(pred.col[0]) : nullptr;
  const int p1_lo = PRED == 1 ? pred.ilo[0] : 0;
  const unsigned p1_w = PRED == 1 ? pred.iwidth[0] : 0u;
  for (int64_t tile = warp0; tile < n_tiles; tile += n_warps) {
    const int64_t r0 = tile * JP_WTILE + (int64_t)lane * JP_ITEMS;
    uint32_t k[JP_ITEMS];
    unsigned pm;  // bit j: row r0 + j takes part in the join
    if (vec && r0 + JP_ITEMS <= n) {
      const int4 v0 = ld_stream_i32x4(key + r0), v1 = ld_stream_i32x4(key + r0 + 4);
      int4 d0 = make_int4(0, 0, 0, 0), d1 = d0;
      if (PRED == 1) {  // requested together with the keys: one round trip to HBM per tile
        d0 = ld_stream_i32x4(p1_col + r0);
        d1 = ld_stream_i32x4(p1_col + r0 + 4);
      }
      k[0] = v0.x; k[1] = v0.y; k[2] = v0.z; k[3] = v0.w;
      k[4] = v1.x; k[5] = v1.y; k[6] = v1.z; k[7] = v1.w;
      pm = 0xFFu;
      if (mask) {
        const u64 m = __ldg(reinterpret_cast<const u64*>(mask + r0));
        pm = 0;
#pragma unroll
        for (int j = 0; j < JP_ITEMS; ++j) pm |= (((m >> (8 * j)) & 0xFFull) != 0 ? 1u : 0u) << j;
      }
      if (PRED == 1) {
        if (!in_range_i32(d0.x, p1_lo, p1_w)) pm &= ~1u;
        if (!in_range_i32(d0.y, p1_lo, p1_w)) pm &= ~2u;
        if (!in_range_i32(d0.z, p1_lo, p1_w)) pm &= ~4u;
        if (!in_range_i32(d0.w, p1_lo, p1_w)) pm &= ~8u;
        if (!in_range_i32(d1.x, p1_lo, p1_w)) pm &= ~16u;
        if (!in_range_i32(d1.y, p1_lo, p1_w)) pm &= ~32u;
        if (!in_range_i32(d1.z, p1_lo, p1_w)) pm &= ~64u;
        if (!in_range_i32(d1.w, p1_lo, p1_w)) pm &= ~128u;
      } else if (PRED == 2) {
        pm = row_pred_bits(pred, r0, pm);
      }
    } else {
      pm = 0;
#pragma unroll
      for (int j = 0; j < JP_ITEMS; ++j) {
        bool in = r0 + j < n;
        k[j] = in ? (uint32_t)key[r0 + j] : 0u;
        if (in && (!mask || mask[r0 + j]) && row_pred(pred, r0 + j)) pm |= 1u << j;
      }
    }
    // bit j of `same`: row j repeats the key of row j - 1 (runs do not continue across lanes: the first row of a lane
    // is always a head, which costs one extra lookup per lane at most)
    unsigned same = 0;
#pragma unroll
    for (int j = 1; j < JP_ITEMS; ++j) same |= (k[j] == k[j - 1] ? 1u : 0u) << j;
    const unsigned cont = pm & (pm << 1) & same;
    const unsigned heads = pm & ~cont;
    wbits[lane] = pm | (cont << 8);
    const int nh = __popc(heads);
    int incl = nh;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      int t = __shfl_up_sync(0xffffffffu, incl, d);
      if (lane >= d) incl += t;
    }
    const int n_heads = __shfl_sync(0xffffffffu, incl, 31);
    int off = incl - nh;
#pragma unroll
    for (int j = 0; j < JP_ITEMS; ++j)
      if ((heads >> j) & 1u) went[off++] = (u64)k[j] | ((u64)(uint32_t)(lane * JP_ITEMS + j) << 32);
    __syncwarp();
    // (3a) Bloom word, then the bucket, for 32 heads at a time; the hits are compacted in place, in row order
    int n_hits = 0;
    for (int c0 = 0; c0 < n_heads; c0 += 32) {
      const u64 e = c0 + lane < n_heads ? went[c0 + lane] : ~0ull;
      const uint32_t kk = (uint32_t)e;
      const uint32_t h = fmix32(kk);
      const uint32_t bits = bloom_bits(h);
      const uint32_t word = c0 + lane < n_heads ? __ldg(bloom + (h & wmask)) : 0u;
      uint32_t hd = NIL;
      if ((word & bits) == bits) {  // lanes past the last head: word == 0, never true
        u64 bb = (u64)(table_hash(h) >> bshift);
        while (true) {
          const ulonglong2* bp = reinterpret_cast<const ulonglong2*>(slots + bb * BUCKET);
          const ulonglong2 x = __ldg(bp), y = __ldg(bp + 1);
          if ((uint32_t)(x.x >> 32) == kk && x.x != KEY_EMPTY) hd = (uint32_t)x.x;
          if ((uint32_t)(x.y >> 32) == kk && x.y != KEY_EMPTY) hd = (uint32_t)x.y;
          if ((uint32_t)(y.x >> 32) == kk && y.x != KEY_EMPTY) hd = (uint32_t)y.x;
          if ((uint32_t)(y.y >> 32) == kk && y.y != KEY_EMPTY) hd = (uint32_t)y.y;
          if (hd != NIL || y.y == KEY_EMPTY) break;  // found, or the bucket has a free slot: key absent
          bb = (bb + 1) & bmask;                      // full bucket without a match: overflow into the next one
        }
      }
      const unsigned found = __ballot_sync(0xffffffffu, hd != NIL);
      __syncwarp();  // every lane holds its entry: slots below c0 + 32 may be overwritten
      if (hd != NIL) went[n_hits + __popc(found & ((1u << lane) - 1u))] = (e & 0xFFFFFFFF00000000ull) | hd;
      n_hits += __popc(found);
    }
    __syncwarp();
    // (3b) output: 32 hits at a time, placed by a shuffle scan over (run length x chain length)
    const int64_t base = MODE == PROBE_WRITE ? tile_cnt[tile] : tile * JP_WTILE;
    int64_t done = 0;  // matches of the chunks before this one
    for (int c0 = 0; c0 < n_hits; c0 += 32) {
      const bool active = c0 + lane < n_hits;
      const u64 e = active ? went[c0 + lane] : 0ull;
      const uint32_t hd = (uint32_t)e, pos = (uint32_t)(e >> 32);
      // run length: 1 + the rows after the head that continue its run (consecutive ones of the owner lane's mask)
      const unsigned after = (wbits[pos >> 3] >> (8 + (pos & 7u) + 1));
      const int run = active ? __ffs((int)~after) : 0;
      int cnt = 0;
      if (active) {
        if (DUPS) {
          int chain = 0;
          for (uint32_t r = hd; r != NIL; r = next[r]) ++chain;
          cnt = chain * run;
        } else {
          cnt = run;
        }
      }
      int inc2 = cnt;
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        int t = __shfl_up_sync(0xffffffffu, inc2, d);
        if (lane >= d) inc2 += t;
      }
      const int chunk_total = __shfl_sync(0xffffffffu, inc2, 31);
      if (MODE != PROBE_COUNT && active) {
        int64_t o = base + done + (inc2 - cnt);
        const uint32_t row = (uint32_t)(tile * JP_WTILE + pos);
        for (int i = 0; i < run; ++i) {  // probe-row major, chain order within a row
          uint32_t r = hd;
          do {
            left_idx[o] = r;
            right_idx[o] = row + i;
            ++o;
            if (!DUPS) break;
            r = next[r];
          } while (r != NIL);
        }
      }
      done += chunk_total;
    }
    if (MODE != PROBE_WRITE && lane == 0) tile_cnt[tile] = done;
    __syncwarp();  // the warp's slices are rewritten by its next tile
  }
}

// moves each tile's staged run of pairs to its final offset (tile_off = scanned counts)
__global__ void join_compact_kernel(const uint32_t* __restrict__ stage_l, const uint32_t* __restrict__ stage_r,
                                    const int64_t* __restrict__ tile_off, int64_t n_tiles, uint32_t* __restrict__ out_l,
                                    uint32_t* __restrict__ out_r) {
  const int lane = threadIdx.x & 31;
  int64_t w = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
  int64_t nw = ((int64_t)gridDim.x * blockDim.x) >> 5;
  for (int64_t t = w; t < n_tiles; t += nw) {  // one warp per tile: sparse joins have a handful of pairs per tile
    int64_t o = tile_off[t];
    int c = (int)(tile_off[t + 1] - o);
    int64_t src = t * JP_WTILE;
    for (int i = lane; i < c; i += 32) {
      out_l[o + i] = stage_l[src + i];
      out_r[o + i] = stage_r[src + i];
    }
  }
}

// ---- several Int32 key columns: the table is keyed on the first pair, the others are checked on the candidate pairs ----
__device__ __forceinline__ bool verify_pair(const JoinVerify& v, uint32_t b, uint32_t p) {
  for (int k = 0; k < v.n; ++k) {
    if (v.bval[k] && !v.bval[k][b]) return false;
    if (v.pval[k] && !v.pval[k][p]) return false;
    if (__ldg(v.bcol[k] + b) != __ldg(v.pcol[k] + p)) return false;
  }
  return true;
}
// primary-key path: one warp per tile checks the tile's staged pairs and compacts the survivors IN PLACE, in order,
// before the tile counts are scanned - the scan and `join_compact_kernel` then only see verified pairs
__global__ void join_verify_stage_kernel(uint32_t* __restrict__ stage_l, uint32_t* __restrict__ stage_r,
                                         int64_t* __restrict__ tile_cnt, int64_t n_tiles,
                                         const __grid_constant__ JoinVerify v) {
  const int lane = threadIdx.x & 31;
  int64_t w = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
  const int64_t nw = ((int64_t)gridDim.x * blockDim.x) >> 5;
  for (int64_t t = w; t < n_tiles; t += nw) {
    const int c = (int)tile_cnt[t];
    if (c == 0) continue;
    const int64_t src = t * JP_WTILE;
    int kept = 0;
    for (int i0 = 0; i0 < c; i0 += 32) {
      const int i = i0 + lane;
      uint32_t b = 0, p = 0;
      bool ok = false;
      if (i < c) {
        b = stage_l[src + i];
        p = stage_r[src + i];
        ok = verify_pair(v, b, p);
      }
      const unsigned m = __ballot_sync(0xffffffffu, ok);
      if (ok) {  // destination index <= i: never ahead of a pair that is still to be read
        const int o = kept + __popc(m & ((1u << lane) - 1u));
        stage_l[src + o] = b;
        stage_r[src + o] = p;
      }
      kept += __popc(m);
    }
    if (lane == 0) tile_cnt[t] = kept;
  }
}
// chained (duplicate build keys) path: the written pairs are checked and compacted by one scan, in pair order
struct VerifyIn {
  const uint32_t* bi;
  const uint32_t* pi;
  JoinVerify v;
  __device__ __forceinline__ int operator()(int64_t i) const { return verify_pair(v, bi[i], pi[i]) ? 1 : 0; }
};
struct PairOut {
  const uint32_t* bi;
  const uint32_t* pi;
  uint32_t* bo;
  uint32_t* po;
  __device__ __forceinline__ void operator()(int64_t i, int v, int64_t pre) const {
    if (v) {
      bo[pre] = bi[i];
      po[pre] = pi[i];
    }
  }
};
static int64_t verify_pairs_scan(Ctx* ctx, const JoinVerify& v, int64_t m, BufPtr* build_idx, BufPtr* probe_idx) {
  if (m == 0) return 0;
  VerifyIn in{(*build_idx)->as<uint32_t>(), (*probe_idx)->as<uint32_t>(), v};
  BufPtr bo = dev_alloc(ctx, (size_t)(m + 1) * 4), po = dev_alloc(ctx, (size_t)(m + 1) * 4);
  int64_t kept = 0;
  device_scan(ctx, in, m, PairOut{in.bi, in.pi, bo->as<uint32_t>(), po->as<uint32_t>()}, &kept);
  *build_idx = bo;
  *probe_idx = po;
  return kept;
}

int64_t join_probe32(Ctx* ctx, JoinTable& t, const int32_t* key, const RowPred& pred, const uint8_t* mask, int64_t n,
                     BufPtr* left_idx, BufPtr* right_idx, const JoinVerify* verify) {
  if (n >= (int64_t)NIL) fail(B2_ERR_NOT_IMPLEMENTED, "batches beyond 2^32-2 rows");
  // Whether any key occurs twice on the build side is learnt with the first probe's own readback: that probe runs
  // the primary-key path speculatively and is redone through the chain-walking path if the flag turns out set.
  int64_t total = 0;
  if (n == 0) {
    *left_idx = dev_alloc(ctx, 4);
    *right_idx = dev_alloc(ctx, 4);
    return 0;
  }
  int64_t n_tiles = (n + JP_WTILE - 1) / JP_WTILE;
  auto tiles = dev_alloc(ctx, (size_t)(n_tiles + 2) * 8);
  int64_t* tile_cnt = tiles->as<int64_t>();
  bool vec = ((uintptr_t)key & 15) == 0 && ((uintptr_t)mask & 7) == 0;
  for (int k = 0; k < pred.n; ++k) vec = vec && ((uintptr_t)pred.col[k] & 15) == 0;
  // persistent grid = exactly the co-resident CTAs, so the static tile striding has no ragged second wave
  static int per_sm = 0;
  if (!per_sm) {
    B2_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, join_probe32_kernel<PROBE_STAGE, false, 2>, JP_THREADS, 0));
    if (per_sm < 1) per_sm = 1;
  }
  const int pk = pred.n == 0 ? 0 : (pred.n == 1 && !pred.is_f64[0] ? 1 : 2);
  auto launch = [&](auto mode, auto dups, uint32_t* li, uint32_t* ri, int grid_) {
    constexpr int M = decltype(mode)::value;
    constexpr bool D = decltype(dups)::value;
    auto go = [&](auto kernel) {
      // persistent grid = the CTAs of THIS variant that are co-resident (the one-term predicate form needs fewer
      // registers than the general one and fits more)
      int fit = 0;
      if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&fit, kernel, JP_THREADS, 0) != cudaSuccess) cudaGetLastError();
      if (getenv("B2_PROBE_CTAS")) fit = atoi(getenv("B2_PROBE_CTAS"));
      if (fit > 0) grid_ = grid_for(ctx, fit, n, JP_TILE);
      kernel<<<grid_, JP_THREADS, 0, ctx->stream>>>(key, pred, mask, n, vec, t.slots->as<u64>(), t.cap / BUCKET - 1,
                                                  t.bshift, t.bloom->as<uint32_t>(), (uint32_t)(t.bloom_words - 1),
                                                  t.next->as<uint32_t>(), li, ri, tile_cnt);
    };
    if (pk == 0) go(join_probe32_kernel<M, D, 0>);
    else if (pk == 1) go(join_probe32_kernel<M, D, 1>);
    else go(join_probe32_kernel<M, D, 2>);
    B2_LAUNCHED(ctx);
  };
  using std::integral_constant;
  int grid = grid_for(ctx, per_sm, n, JP_TILE);
  const u64* slots = t.slots->as<u64>();
  const uint32_t* next = t.next->as<uint32_t>();
  const uint32_t* bloom = t.bloom->as<uint32_t>();
  const uint32_t wmask = (uint32_t)(t.bloom_words - 1);
  if (t.dups_known && t.has_dups) {
    launch(integral_constant<int, PROBE_COUNT>{}, std::true_type{}, nullptr, nullptr, grid);
    scan_i64(ctx, tile_cnt, n_tiles);
    B2_READBACK(ctx, &total, tile_cnt + n_tiles, 8);
    if (total >= (int64_t)NIL) fail(B2_ERR_NOT_IMPLEMENTED, "join output beyond 2^32-2 rows in one batch");
    *left_idx = dev_alloc(ctx, (size_t)(total + 1) * 4);
    *right_idx = dev_alloc(ctx, (size_t)(total + 1) * 4);
    if (total)
      launch(integral_constant<int, PROBE_WRITE>{}, std::true_type{}, (*left_idx)->as<uint32_t>(),
             (*right_idx)->as<uint32_t>(), grid);
    return verify ? verify_pairs_scan(ctx, *verify, total, left_idx, right_idx) : total;
  }
  // primary-key build: one probe pass into tile-local staging runs, then scan + compaction of the runs
  auto stage_l = dev_alloc(ctx, (size_t)n_tiles * JP_WTILE * 4);
  auto stage_r = dev_alloc(ctx, (size_t)n_tiles * JP_WTILE * 4);
  launch(integral_constant<int, PROBE_STAGE>{}, std::false_type{}, stage_l->as<uint32_t>(), stage_r->as<uint32_t>(), grid);
  if (verify) {
    join_verify_stage_kernel<<<grid_for(ctx, 8, n_tiles * 32, 256), 256, 0, ctx->stream>>>(
        stage_l->as<uint32_t>(), stage_r->as<uint32_t>(), tile_cnt, n_tiles, *verify);
    B2_LAUNCHED(ctx);
  }
  scan_i64(ctx, tile_cnt, n_tiles);
  if (!t.dups_known) {
    int f = 0;
    B2_CUDA(cudaMemcpyAsync(&f, t.flags->ptr, 4, cudaMemcpyDeviceToHost, ctx->stream));
    B2_READBACK(ctx, &total, tile_cnt + n_tiles, 8);
    t.dups_known = true;
    t.has_dups = f != 0;
    if (t.has_dups) return join_probe32(ctx, t, key, pred, mask, n, left_idx, right_idx, verify);  // speculation failed
  } else {
    B2_READBACK(ctx, &total, tile_cnt + n_tiles, 8);
  }
  *left_idx = dev_alloc(ctx, (size_t)(total + 1) * 4);
  *right_idx = dev_alloc(ctx, (size_t)(total + 1) * 4);
  if (total) {
    join_compact_kernel<<<grid_for(ctx, 8, n_tiles * 32, 256), 256, 0, ctx->stream>>>(
        stage_l->as<uint32_t>(), stage_r->as<uint32_t>(), tile_cnt, n_tiles, (*left_idx)->as<uint32_t>(),
        (*right_idx)->as<uint32_t>());
    B2_LAUNCHED(ctx);
  }
  return total;
}

}  // namespace b2
