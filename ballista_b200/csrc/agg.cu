// agg.cu — group table management, kernel registry, strategy runner and emit kernels for
// HashAggregateExec (constructed at rust/core/src/serde/physical_plan/from_proto.rs:173-252).
#include "agg.h"

#include <cstring>

#include "agg_kernels.cuh"

namespace b2 {

// ------------------------------------------------------------------------------------------------------
// key layout
// ------------------------------------------------------------------------------------------------------
KeyLayout make_key_layout(const std::vector<int>& dtypes) {
  int int_bytes = 0, n_str = 0;
  for (int dt : dtypes) {
    if (dt == B2_INT32 || dt == B2_DATE32) int_bytes += 4;
    else if (dt == B2_INT64) int_bytes += 8;
    else if (dt == B2_UTF8) n_str++;
    else fail(B2_ERR_NOT_IMPLEMENTED, std::string("group/join key of type ") + dtype_name(dt));
  }
  int sb = 0;
  if (n_str) {
    sb = (15 - int_bytes) / n_str;
    if (sb > 15) sb = 15;
    if (sb < 2) fail(B2_ERR_NOT_IMPLEMENTED, "key wider than 15 packed bytes");
  }
  if (int_bytes > 15) fail(B2_ERR_NOT_IMPLEMENTED, "key wider than 15 packed bytes");
  KeyLayout l;
  int off = 0;
  for (int dt : dtypes) {
    int b = dt == B2_UTF8 ? sb : (dt == B2_INT64 ? 8 : 4);
    l.parts.push_back(KeyPart{dt, off, b});
    off += b;
  }
  return l;
}

// ------------------------------------------------------------------------------------------------------
// group table
// ------------------------------------------------------------------------------------------------------
std::shared_ptr<GroupTable> make_group_table(Ctx* ctx, u64 cap, int na, int nc, bool init) {
  if (cap > (1ull << 32)) fail(B2_ERR_NOT_IMPLEMENTED, "group table beyond 2^32 slots");
  auto t = std::make_shared<GroupTable>();
  t->ctx = ctx;
  t->cap = cap;
  t->na = na;
  t->nc = nc;
  t->keys = dev_alloc(ctx, cap * sizeof(Key128));
  t->acc = dev_alloc(ctx, cap * sizeof(double) * (size_t)(na ? na : 1));
  t->cnt = dev_alloc(ctx, cap * sizeof(u64) * (size_t)(nc ? nc : 1));
  t->n_groups = dev_alloc(ctx, 16);
  t->slot_list = dev_alloc(ctx, cap * 4 + 16);  // written before it is read: no initialisation
  if (init) reset_group_table(*t);
  return t;
}
void reset_group_table(GroupTable& t) {
  B2_CUDA(cudaMemsetAsync(t.keys->ptr, 0xFF, t.keys->bytes, t.ctx->stream));
  B2_CUDA(cudaMemsetAsync(t.acc->ptr, 0, t.acc->bytes, t.ctx->stream));
  B2_CUDA(cudaMemsetAsync(t.cnt->ptr, 0, t.cnt->bytes, t.ctx->stream));
  B2_CUDA(cudaMemsetAsync(t.n_groups->ptr, 0, 16, t.ctx->stream));
}

// walks the source's slot list (its claimed groups), not its whole capacity
// (256 threads per CTA; the groups a CTA claims in one round are listed behind ONE reservation of dst.n_groups, as in
// agg_global_kernel: a regrow moves up to millions of groups, and a counter update per claim queues on one L2 address)
__global__ void __launch_bounds__(256) table_merge_kernel(TableRef src, TableRef dst, int na, int nc, int* status) {
  __shared__ uint32_t s_claim[256];
  __shared__ int s_nclaim, s_base;
  const u64 ng = (u64)*src.n_groups;
  for (u64 g0 = blockIdx.x * (u64)blockDim.x; g0 < ng; g0 += (u64)gridDim.x * blockDim.x) {  // uniform per CTA
    if (threadIdx.x == 0) s_nclaim = 0;
    __syncthreads();
    const u64 g = g0 + threadIdx.x;
    if (g < ng) {
      const u64 i = src.slot_list[g];
      Key128 k = src.keys[i];
      bool claimed;
      u64 slot = table_upsert_quiet(dst, k, &claimed);
      if (slot == ~0ull) {
        atomicExch(status, AGG_TABLE_FULL);
      } else {
        if (claimed) s_claim[atomicAdd(&s_nclaim, 1)] = (uint32_t)slot;
        for (int a = 0; a < na; ++a) red_add_f64(&dst.acc[(u64)a * dst.cap + slot], src.acc[(u64)a * src.cap + i]);
        for (int a = 0; a < nc; ++a) red_add_u64(&dst.cnt[(u64)a * dst.cap + slot], src.cnt[(u64)a * src.cap + i]);
      }
    }
    __syncthreads();
    const int nclaim = s_nclaim;
    if (nclaim) {  // uniform
      if (threadIdx.x == 0) s_base = atomicAdd(dst.n_groups, nclaim);
      __syncthreads();
      if (dst.slot_list)
        for (int q = threadIdx.x; q < nclaim; q += blockDim.x) dst.slot_list[s_base + q] = s_claim[q];
    }
    __syncthreads();
  }
}

// ------------------------------------------------------------------------------------------------------
// runner
// ------------------------------------------------------------------------------------------------------
static u64 next_pow2(u64 x) {
  u64 p = 16;
  while (p < x) p <<= 1;
  return p;
}
constexpr u64 DELTA_CAP = 65536;
constexpr u64 SPEC_CAP = 4096;  // table of the speculative single-batch path (its emit bound is SPEC_BOUND groups)
constexpr int64_t GLOBAL_CHUNK_ROWS = 32ll << 20;  // multiple of TILE_ROWS

AggRunner::AggRunner(Ctx* c, int na_, int nc_, bool keyed_) : ctx(c), na(na_), nc(nc_), keyed(keyed_) {
  status = dev_alloc_zero(ctx, 16);
}
void AggRunner::reset() {
  main.reset();
  delta.reset();
  group_bound = 0;
  groups_exact = false;
  speculative = false;
  path.clear();
  B2_CUDA(cudaMemsetAsync(status->ptr, 0, 16, ctx->stream));
}

void AggRunner::check_status(int st) {
  if (st == AGG_KEY_TOO_WIDE)
    fail(B2_ERR_NOT_IMPLEMENTED, "a Utf8 group key value is longer than the packed-key budget (15 bytes per key)");
  if (st == AGG_TABLE_FULL) fail(B2_ERR_EXECUTION, "internal: group table full");
}

void AggRunner::ensure_main(int64_t extra) {
  u64 need = next_pow2(2 * (u64)(group_bound + extra) + 16);
  // The bound counts ROWS of the coming batch, not groups, and the high-cardinality kernel's time tracks the table's
  // footprint (each new group's first touch misses L2): beyond 2 M slots size for a worst-case load of 0.75 rather
  // than 0.5.  Linear probing stays short because real inputs repeat keys (Q3: 2.6 rows per group => load 0.27).
  if (need > (1ull << 21)) need = std::max<u64>(1ull << 21, next_pow2((u64)(group_bound + extra) * 4 / 3 + 16));
  if (!main) {
    main = make_group_table(ctx, std::max<u64>(need, keyed ? 1024 : 16), na, nc);
    return;
  }
  if (need <= main->cap) return;
  auto bigger = make_group_table(ctx, need, na, nc);
  int64_t walk = group_bound > 0 ? group_bound : 1;
  table_merge_kernel<<<grid_for(ctx, 8, walk, 256), 256, 0, ctx->stream>>>(main->ref(), bigger->ref(), na, nc,
                                                                         status->as<int>());
  B2_LAUNCHED(ctx);
  main = bigger;
}

bool AggRunner::consume_unchecked(const AggLaunch& l, int64_t n) {
  if (force_global || main || n <= 0) return false;
  main = make_group_table(ctx, keyed ? SPEC_CAP : 16, na, nc);
  // the status word was zeroed at construction / reset(); the key-packing kernel of the unfused path may already
  // have raised AGG_KEY_TOO_WIDE in it, which the emit's readback reports like a status of the scan itself
  l.small(n, main->ref(), status->as<int>());
  speculative = true;
  path = "small";
  return true;
}

void AggRunner::consume(const AggLaunch& l, int64_t n) {
  if (n <= 0) return;
  groups_exact = false;
  int host[4] = {0, 0, 0, 0};
  if (!force_global) {
    if (narrow_overflowed && l.small_wide) {
      AggLaunch w = l;
      w.small = l.small_wide;
      w.small_wide = nullptr;
      return consume(w, n);
    }
    if (!delta) delta = make_group_table(ctx, keyed ? DELTA_CAP : 16, na, nc);
    else reset_group_table(*delta);
    l.small(n, delta->ref(), status->as<int>());
    B2_CUDA(cudaMemcpyAsync(&host[0], status->ptr, 4, cudaMemcpyDeviceToHost, ctx->stream));
    B2_READBACK(ctx, &host[1], delta->n_groups->ptr, 4);
    if (host[0] == AGG_OK) {
      ensure_main(host[1]);
      if (host[1] > 0) {
        table_merge_kernel<<<grid_for(ctx, 8, host[1], 256), 256, 0, ctx->stream>>>(delta->ref(), main->ref(), na, nc,
                                                                                  status->as<int>());
        B2_LAUNCHED(ctx);
      }
      group_bound += host[1];
      if (path.empty() || path == "small") path = narrow_overflowed ? "small+widekeys" : "small";
      return;
    }
    B2_CUDA(cudaMemsetAsync(status->ptr, 0, 16, ctx->stream));
    if (host[0] == AGG_KEY_TOO_WIDE && l.small_wide) {
      // a narrow-key kernel met a Utf8 key value of more than its budget: redo the batch with wide keys
      AggLaunch w = l;
      w.small = l.small_wide;
      w.small_wide = nullptr;
      narrow_overflowed = true;
      return consume(w, n);
    }
    check_status(host[0]);
    // too many groups for the per-CTA tables: redo this batch (and all later ones) on the HBM table
    force_global = true;
    delta.reset();
  }
  path = "global";
  for (int64_t b = 0; b < n; b += GLOBAL_CHUNK_ROWS) {
    int64_t e = b + GLOBAL_CHUNK_ROWS < n ? b + GLOBAL_CHUNK_ROWS : n;
    ensure_main(keyed ? e - b : 1);
    l.global(b, e, n, main->ref(), status->as<int>());
    B2_CUDA(cudaMemcpyAsync(&host[0], status->ptr, 4, cudaMemcpyDeviceToHost, ctx->stream));
    B2_READBACK(ctx, &host[1], main->n_groups->ptr, 4);
    check_status(host[0]);
    group_bound = host[1];
    groups_exact = true;  // status checked and the count read after the last kernel on the table
  }
}

int64_t AggRunner::finish() {
  if (!main) {
    ensure_main(0);
    groups_exact = true;
    group_bound = 0;
  }
  if (groups_exact) return group_bound;
  int host[2] = {0, 0};
  B2_CUDA(cudaMemcpyAsync(&host[0], main->n_groups->ptr, 4, cudaMemcpyDeviceToHost, ctx->stream));
  B2_READBACK(ctx, &host[1], status->ptr, 4);
  check_status(host[1]);
  group_bound = host[0];
  groups_exact = true;
  return group_bound;
}

// ------------------------------------------------------------------------------------------------------
// emit kernels
// ------------------------------------------------------------------------------------------------------
__device__ __forceinline__ u128 key_bits(const Key128& k) { return ((u128)k.hi << 64) | k.lo; }
// byte `idx` (0..15) of a packed key, without 128-bit arithmetic
__device__ __forceinline__ uint8_t key_byte(const Key128& k, int idx) {
  return (uint8_t)(idx < 8 ? k.lo >> (8 * idx) : k.hi >> (8 * (idx - 8)));
}

constexpr size_t EMIT_ARENA_BYTES = 384 << 10;
BufPtr EmitPlan::alloc(size_t bytes) {
  if (arena) {
    size_t off = (arena_used + 15) & ~(size_t)15;
    if (off + bytes + 16 <= arena->bytes) {
      arena_used = off + bytes;
      return dev_view(arena, off, bytes);
    }
    arena = nullptr;  // does not fit: ordinary buffers, the export copies column by column
  }
  return dev_alloc(ctx, bytes);
}
Column EmitPlan::new_fixed(int dtype) {
  Column c;
  c.dtype = dtype;
  c.length = bound;
  c.data = alloc((size_t)bound * dtype_width(dtype) + 16);
  return c;
}

// ---- all output columns of one aggregate in ONE launch: blockIdx.y selects the column ----
enum { EMIT_KEY_I32 = 0, EMIT_KEY_I64, EMIT_KEY_STR, EMIT_F64, EMIT_U64, EMIT_AVG };
struct EmitCol {
  int kind;
  int byte_off;      // key columns: byte offset of the part inside the packed key
  const void* src;   // EMIT_F64 / EMIT_AVG: sums; EMIT_U64: counts
  const void* src2;  // EMIT_AVG: counts
  const u64* vsrc;   // validity source (count of non-null inputs), or null
  void* out;         // values (Utf8: int32 offsets)
  uint8_t* chr;      // Utf8 characters
  uint8_t* vout;     // validity bytes
};
constexpr int EMIT_MAX_COLS = 24;
struct EmitArgs {
  int n;
  EmitCol c[EMIT_MAX_COLS];
};
// hdr (int64): [0] = status word, [1] = group count, [2 + y] = character total of column y (Utf8 key columns).
// ng_dev == null: the host knows the group count (`bound`).
__global__ void __launch_bounds__(256) emit_all_kernel(const __grid_constant__ EmitArgs args, const Key128* __restrict__ keys,
                                                       const uint32_t* __restrict__ slots, int64_t bound,
                                                       const int* ng_dev, const int* status, int64_t* hdr) {
  const EmitCol& col = args.c[blockIdx.y];
  int64_t ng = ng_dev ? (int64_t)*ng_dev : bound;
  if (blockIdx.x == 0 && blockIdx.y == 0 && threadIdx.x == 0 && hdr) {
    hdr[0] = status ? (int64_t)*status : 0;
    hdr[1] = ng;
  }
  if (ng > bound) return;  // speculation failed: the host sees hdr[1] > bound and redoes the batch
  if (col.kind == EMIT_KEY_STR) {
    // one CTA scans the lengths chunk by chunk (group counts with Utf8 keys are small), then writes the bytes
    if (blockIdx.x != 0) return;
    __shared__ int32_t warp_sums[32];
    __shared__ int32_t carry_s;
    if (threadIdx.x == 0) carry_s = 0;
    __syncthreads();
    int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
    int32_t* off = static_cast<int32_t*>(col.out);
    if (ng == 0 && threadIdx.x == 0) off[0] = 0;
    for (int64_t base = 0; base < ng; base += blockDim.x) {
      int64_t i = base + threadIdx.x;
      Key128 kk{0, 0};
      int len = 0;
      if (i < ng) {
        kk = keys[slots[i]];
        len = (int)key_byte(kk, col.byte_off);
      }
      int v = len;
      for (int d = 1; d < 32; d <<= 1) {
        int t = __shfl_up_sync(0xffffffffu, v, d);
        if (lane >= d) v += t;
      }
      if (lane == 31) warp_sums[warp] = v;
      __syncthreads();
      int woff = 0, total = 0;
      for (int w = 0; w < nwarps; ++w) {
        if (w < warp) woff += warp_sums[w];
        total += warp_sums[w];
      }
      int carry = carry_s;
      int excl = carry + woff + v - len;
      if (i < ng) {
        off[i] = excl;
        for (int q = 0; q < len; ++q) col.chr[excl + q] = key_byte(kk, col.byte_off + 1 + q);
        if (i == ng - 1) off[ng] = excl + len;
      }
      __syncthreads();
      if (threadIdx.x == 0) carry_s = carry + total;
      __syncthreads();
    }
    if (threadIdx.x == 0 && hdr) hdr[2 + blockIdx.y] = carry_s;
    return;
  }
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  for (; i < ng; i += (int64_t)gridDim.x * blockDim.x) {
    uint32_t sl = slots[i];
    switch (col.kind) {
      case EMIT_KEY_I32:
        static_cast<int32_t*>(col.out)[i] = (int32_t)(uint32_t)(key_bits(keys[sl]) >> (8 * col.byte_off));
        break;
      case EMIT_KEY_I64:
        static_cast<int64_t*>(col.out)[i] = (int64_t)(u64)(key_bits(keys[sl]) >> (8 * col.byte_off));
        break;
      case EMIT_F64: static_cast<double*>(col.out)[i] = static_cast<const double*>(col.src)[sl]; break;
      case EMIT_U64: static_cast<u64*>(col.out)[i] = static_cast<const u64*>(col.src)[sl]; break;
      default: {  // AVG evaluate(): sum / count as f64; NULL when nothing was counted
        const u64 c = static_cast<const u64*>(col.src2)[sl];
        static_cast<double*>(col.out)[i] = c ? __ddiv_rn(static_cast<const double*>(col.src)[sl], (double)c) : 0.0;
        break;
      }
    }
    if (col.vout) col.vout[i] = col.vsrc[sl] > 0;
  }
}

EmitPlan::EmitPlan(Ctx* c, const GroupTable& t, int64_t b, bool use_arena) : ctx(c), table(t), bound(b) {
  if (use_arena) {
    arena = dev_alloc(ctx, EMIT_ARENA_BYTES);
    hdr = alloc((size_t)(2 + 24) * 8);
  }
}
void EmitPlan::push(int kind, int byte_off, const void* src, const void* src2, const u64* vsrc, Column col) {
  if (vsrc) col.valid = alloc((size_t)bound + 16);
  descs.push_back(EmitDesc{kind, byte_off, src, src2, vsrc});
  cols.push_back(std::move(col));
}
void EmitPlan::key(const KeyPart& part) {
  if (part.dtype == B2_UTF8) {
    Column c;
    c.dtype = B2_UTF8;
    c.length = bound;
    c.data = alloc((size_t)(bound + 1) * 4 + 16);
    c.chars = alloc((size_t)bound * 15 + 16);
    push(EMIT_KEY_STR, part.byte_off, nullptr, nullptr, nullptr, std::move(c));
    return;
  }
  push(part.bytes == 8 ? EMIT_KEY_I64 : EMIT_KEY_I32, part.byte_off, nullptr, nullptr, nullptr,
       new_fixed(part.dtype));
}
void EmitPlan::sum(int a, int v) {
  push(EMIT_F64, 0, table.acc->as<double>() + (u64)a * table.cap, nullptr,
       v >= 0 ? table.cnt->as<u64>() + (u64)v * table.cap : nullptr, new_fixed(B2_FLOAT64));
}
void EmitPlan::isum(int a, int dtype, int v) {
  push(EMIT_U64, 0, table.cnt->as<u64>() + (u64)a * table.cap, nullptr,
       v >= 0 ? table.cnt->as<u64>() + (u64)v * table.cap : nullptr, new_fixed(dtype));
}
void EmitPlan::count(int a) {
  push(EMIT_U64, 0, table.cnt->as<u64>() + (u64)a * table.cap, nullptr, nullptr, new_fixed(B2_UINT64));
}
void EmitPlan::avg(int a, int cidx, bool track_null) {
  const u64* counts = table.cnt->as<u64>() + (u64)cidx * table.cap;
  push(EMIT_AVG, 0, table.acc->as<double>() + (u64)a * table.cap, counts, track_null ? counts : nullptr,
       new_fixed(B2_FLOAT64));
}
bool EmitPlan::run(int64_t exact_ng, const int* status_dev, int64_t* ng_out, int* status_out) {
  if (cols.size() > EMIT_MAX_COLS) fail(B2_ERR_NOT_IMPLEMENTED, "aggregate with more than 24 output columns");
  const bool spec = exact_ng < 0;
  if (status_out) *status_out = 0;
  if (!spec && exact_ng == 0) {
    for (auto& c : cols) {
      c.length = 0;
      if (c.dtype == B2_UTF8) B2_CUDA(cudaMemsetAsync(c.data->ptr, 0, 4, ctx->stream));
    }
    *ng_out = 0;
    return true;
  }
  EmitArgs args;
  memset(&args, 0, sizeof args);
  args.n = (int)cols.size();
  bool any_str = false;
  for (int k = 0; k < args.n; ++k) {
    const EmitDesc& d = descs[k];
    Column& c = cols[k];
    args.c[k] = EmitCol{d.kind, d.byte_off, d.src, d.src2, d.vsrc, c.data->ptr, c.chars ? c.chars->as<uint8_t>() : nullptr,
                        c.valid ? c.valid->as<uint8_t>() : nullptr};
    any_str |= d.kind == EMIT_KEY_STR;
  }
  const int64_t rows = spec ? bound : exact_ng;
  BufPtr hdr;
  if (spec || any_str) hdr = dev_alloc_zero(ctx, (size_t)(2 + EMIT_MAX_COLS) * 8);
  dim3 grid((unsigned)grid_for(ctx, 8, rows, 256), (unsigned)args.n);
  emit_all_kernel<<<grid, 256, 0, ctx->stream>>>(args, table.keys->as<Key128>(), table.slot_list->as<uint32_t>(), rows,
                                                 spec ? table.n_groups->as<int>() : nullptr, status_dev,
                                                 hdr ? hdr->as<int64_t>() : nullptr);
  B2_LAUNCHED(ctx);
  int64_t ng = exact_ng;
  int64_t h[2 + EMIT_MAX_COLS] = {0};
  if (hdr) {  // ONE readback: status, group count, character totals
    B2_READBACK(ctx, h, hdr->ptr, sizeof h);
    if (spec) {
      ng = h[1];
      if (status_out) *status_out = (int)h[0];
      if (h[0] != AGG_OK || ng > bound) return false;
    }
  }
  for (int k = 0; k < args.n; ++k) {
    cols[k].length = ng;
    if (descs[k].kind == EMIT_KEY_STR) cols[k].chars_len = h[2 + k];
  }
  *ng_out = ng;
  return true;
}

BufPtr EmitPlan::run_lazy(const int* status_dev) {
  if (cols.size() > EMIT_MAX_COLS) fail(B2_ERR_NOT_IMPLEMENTED, "aggregate with more than 24 output columns");
  EmitArgs args;
  memset(&args, 0, sizeof args);
  args.n = (int)cols.size();
  for (int k = 0; k < args.n; ++k) {
    const EmitDesc& d = descs[k];
    Column& c = cols[k];
    args.c[k] = EmitCol{d.kind, d.byte_off, d.src, d.src2, d.vsrc, c.data->ptr, c.chars ? c.chars->as<uint8_t>() : nullptr,
                        c.valid ? c.valid->as<uint8_t>() : nullptr};
    c.length = bound;
  }
  if (!hdr) hdr = dev_alloc(ctx, (size_t)(2 + EMIT_MAX_COLS) * 8);
  B2_CUDA(cudaMemsetAsync(hdr->ptr, 0, (size_t)(2 + EMIT_MAX_COLS) * 8, ctx->stream));
  dim3 grid((unsigned)grid_for(ctx, 8, bound, 256), (unsigned)args.n);
  emit_all_kernel<<<grid, 256, 0, ctx->stream>>>(args, table.keys->as<Key128>(), table.slot_list->as<uint32_t>(), bound,
                                                 table.n_groups->as<int>(), status_dev, hdr->as<int64_t>());
  B2_LAUNCHED(ctx);
  return hdr;
}

// ------------------------------------------------------------------------------------------------------
// small Final aggregate (see agg.h)
// ------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(1024) final_tiny_kernel(const __grid_constant__ TinyArgs a, TableRef t, int na_tbl, int nc_tbl, int* status) {
  __shared__ int64_t s_rows[TINY_BATCHES];
  __shared__ int s_bad;
  __shared__ unsigned long long s_cap;
  if (threadIdx.x == 0) {
    int bad = 0;
    int64_t total = 0;
    for (int b = 0; b < a.nb; ++b) {
      const TinyBatch& tb = a.b[b];
      int64_t n = tb.rows;
      if (tb.hdr) {
        n = tb.hdr[1];
        if (tb.hdr[0] != 0 || n < 0 || n > tb.bound || (tb.redo_if_empty && n == 0)) bad = AGG_INPUT_FAILED;
      }
      s_rows[b] = n;
      total += n;
    }
    if (!bad && total > TINY_MAX_ROWS) bad = AGG_TOO_BIG;
    s_bad = bad;
    if (bad) atomicExch(status, bad);
    // The table arrives uninitialised and only the part that is used gets initialised: with a single launch the
    // hashed capacity is sized from the rows actually present (Q1: 4 rows -> 64 slots), the arrays keep their stride.
    u64 cap = t.cap;
    if (a.single) {
      cap = t.cap < 64 ? t.cap : 64;
      while (cap < t.cap && cap < 4 * (u64)(total > 0 ? total : 1)) cap <<= 1;
    }
    s_cap = cap;
  }
  __syncthreads();
  const u64 stride = t.cap;  // acc[x * stride + slot], cnt[x * stride + slot]
  t.cap = s_cap;
  if (a.init) {
    for (u64 i = threadIdx.x; i < t.cap; i += blockDim.x) {
      t.keys[i] = Key128{KEY_EMPTY, KEY_EMPTY};
      for (int x = 0; x < na_tbl; ++x) t.acc[(u64)x * stride + i] = 0.0;
      for (int x = 0; x < nc_tbl; ++x) t.cnt[(u64)x * stride + i] = 0;
    }
    if (threadIdx.x == 0) *t.n_groups = 0;
    __syncthreads();
  }
  if (s_bad) return;
  int err = 0;
  for (int b = 0; b < a.nb; ++b) {
    const TinyBatch& tb = a.b[b];
    const int64_t n = s_rows[b];
    for (int64_t r = threadIdx.x; r < n; r += blockDim.x) {
      u128 key = 0;
      for (int p = 0; p < a.nkeys; ++p) {
        const TinyCol& kc = tb.key[p];
        if (kc.valid && !kc.valid[r]) {
          err = AGG_NULL_KEY;
          continue;
        }
        if (a.part_dtype[p] == B2_UTF8) {
          StrVal sv = load_str((const int32_t*)kc.data, kc.chars, r);
          if (sv.len > a.part_bytes[p] - 1) err = AGG_KEY_TOO_WIDE;
          u128 m = (((u128)1 << (8 * (a.part_bytes[p] - 1))) - 1);
          key |= ((u128)(uint8_t)sv.len | ((sv.bytes & m) << 8)) << (8 * a.part_off[p]);
        } else if (a.part_bytes[p] == 8) {
          key |= (u128)(u64)((const int64_t*)kc.data)[r] << (8 * a.part_off[p]);
        } else {
          key |= (u128)(uint32_t)((const int32_t*)kc.data)[r] << (8 * a.part_off[p]);
        }
      }
      u64 slot = table_upsert(t, Key128{(u64)key, (u64)(key >> 64)});
      if (slot == ~0ull) {
        err = AGG_TABLE_FULL;
        continue;
      }
      for (int x = 0; x < a.na; ++x) {
        const TinyCol& vc = tb.val[x];
        if (!vc.data) continue;
        const bool ok = !vc.valid || vc.valid[r];
        if (ok) red_add_f64(&t.acc[(u64)x * stride + slot], ((const double*)vc.data)[r]);
        if (a.vslot[x] >= 0 && ok) red_add_u64(&t.cnt[(u64)a.vslot[x] * stride + slot], 1ull);
      }
      for (int x = 0; x < a.nc; ++x) {
        const TinyCol& cc = tb.cnt[x];
        if (!cc.data) continue;
        const bool ok = !cc.valid || cc.valid[r];
        if (ok) red_add_u64(&t.cnt[(u64)x * stride + slot], ((const u64*)cc.data)[r]);
        if (a.cvslot[x] >= 0 && ok) red_add_u64(&t.cnt[(u64)a.cvslot[x] * stride + slot], 1ull);
      }
    }
  }
  if (err) atomicExch(status, err);
}
void final_tiny_launch(Ctx* ctx, const TinyArgs& a, const GroupTable& t, int* status) {
  final_tiny_kernel<<<1, 1024, 0, ctx->stream>>>(a, t.ref(), t.na ? t.na : 1, t.nc ? t.nc : 1, status);
  B2_LAUNCHED(ctx);
}

}  // namespace b2
