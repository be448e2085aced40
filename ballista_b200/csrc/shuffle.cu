// shuffle.cu — the hash-repartition shuffle inside one box, over NVLink peer memory.  Replaces the reference's
// ShuffleWriter -> Arrow IPC file -> Flight do_get -> ShuffleReaderExec hop (rust/core/src/utils.rs:49-84,
// rust/executor/src/flight_service.rs:193-228, rust/core/src/execution_plans/shuffle_reader.rs:77-99).
//
// Every rank owns one exchange ARENA (cudaMalloc) that all peers of the box map through CUDA IPC.  One exchange:
//   1. split, phase 1 (repartition.cu): destination byte per row, per-tile histogram, one scan.  No host sync.
//   2. ONE host-synchronised all-gather carrying, per rank: its per-destination row (and Utf8 character) counts,
//      taken straight from the scanned histogram on the device, plus its arena generation and IPC handle.  Stream
//      order makes its completion the barrier "every peer has finished pulling my previous exchange", so the arena
//      may be overwritten; a rank that had to regrow has already published the new handle in the same message.
//   3. split, phase 2: ONE kernel moves every fixed-width column, rows read coalesced, straight to its
//      destination-major place in the arena.  Slice d of a column starts at an offset congruent (mod 16) to where
//      rank d will store it, so the pull below runs on 16-byte accesses end to end.
//   4. a stream-ordered barrier (1-word all-reduce, no host sync): all arenas are staged.
//   5. ONE pull kernel: every rank copies its slice of every column of every peer's arena over NVLink (peer
//      loads) directly into its final, contiguous output columns: no receive staging, no copy-out.
// NCCL carries only the metadata all-gather and the barrier; when peer mapping is unavailable the grouped
// ncclSend/ncclRecv path of comm.cu is used instead (B2_SHUFFLE=nccl forces it).
//
// PUSH form (the default when every moved column is fixed-width; B2_SHUFFLE=pull keeps the form above): steps 3-5
// become ONE kernel.  Every rank owns a RECEIVE arena (comm.h) mapped by its peers and offers, in the metadata
// message of step 2, the free span the exchange may land in.  After the all-gather every rank knows every
// destination's total and where its own rows start there, so phase 2 of the split (`rp_push_kernel`) stores each
// row's values straight into the destination's final column over NVLink, whole runs at a time; the barrier follows
// and the output columns are windows of the receive block: no staging copy, no copy-out.  An offer that turns out too
// small (all ranks see the same numbers) sends this one exchange down the pull form and grows the arena for the next.
#include <cstdio>
#include <cstdlib>
#include <cstring>

#include "comm.h"
#include "ops.h"
#include "scan.cuh"

namespace b2 {

static size_t align256(size_t x) { return (x + 255) & ~(size_t)255; }

void shuffle_arena_release(Ctx* ctx) {
  Comm* c = ctx->comm;
  if (!c) return;
  for (size_t r = 0; r < c->peer_arena.size(); ++r)
    if (c->peer_arena[r] && (int)r != ctx->rank) cudaIpcCloseMemHandle(c->peer_arena[r]);
  c->peer_arena.clear();
  c->peer_gen.clear();
  if (c->arena) cudaFree(c->arena);
  c->arena = nullptr;
  c->arena_cap = 0;
  for (size_t r = 0; r < c->peer_recv.size(); ++r)
    if ((int)r != ctx->rank)
      for (auto& m : c->peer_recv[r]) cudaIpcCloseMemHandle(m.second);
  c->peer_recv.clear();
  c->recv_arenas.clear();  // each arena is freed with the last output block that still windows it
}

struct ArenaMsg {  // what every rank publishes in step A
  int64_t need;    // bytes it will stage in this exchange
  int64_t gen;     // arena generation after a possible regrow
  int64_t ok;      // 1 = arena usable for peer mapping
  int64_t pad;
  cudaIpcMemHandle_t handle;  // 64 bytes
};
static_assert(sizeof(ArenaMsg) == 96, "ArenaMsg is exchanged as 12 int64 words");

// (Re)allocates this rank's arena for `need` bytes and fills its message.  A retired arena is returned through
// `retired` and must be freed only after the metadata all-gather (peers may still be pulling from it).
static void arena_local(Ctx* ctx, size_t need, ArenaMsg* mine, void** retired) {
  Comm* c = ctx->comm;
  memset(mine, 0, sizeof *mine);
  mine->need = (int64_t)need;
  mine->ok = 1;
  *retired = nullptr;
  if (c->peer_arena.empty()) {
    c->peer_arena.assign(ctx->nranks, nullptr);
    c->peer_gen.assign(ctx->nranks, -1);
  }
  if (need > c->arena_cap) {
    *retired = c->arena;
    c->arena = nullptr;
    size_t cap = align256(need + need / 4 + (64u << 20));
    cudaError_t e = cudaMalloc(&c->arena, cap);
    if (e != cudaSuccess) {
      cudaGetLastError();
      c->arena = nullptr;
      c->arena_cap = 0;
      mine->ok = 0;
    } else {
      c->arena_cap = cap;
      c->arena_gen++;
    }
  }
  if (mine->ok) {
    cudaError_t e = cudaIpcGetMemHandle(&mine->handle, c->arena);
    if (e != cudaSuccess) {
      cudaGetLastError();
      mine->ok = 0;
    }
  }
  mine->gen = c->arena_gen;
}
// Maps the arenas of peers whose generation changed; agrees on the outcome across ranks when anything changed.
static bool arena_map_peers(Ctx* ctx, const std::vector<ArenaMsg>& msgs, void* retired) {
  Comm* c = ctx->comm;
  int nr = ctx->nranks, me = ctx->rank;
  bool ok = true, remap = false;
  for (int r = 0; r < nr; ++r) {
    ok = ok && msgs[r].ok;
    remap = remap || msgs[r].gen != c->peer_gen[r];
  }
  if (ok && remap) {
    for (int r = 0; r < nr; ++r) {
      if (r == me) {
        c->peer_arena[r] = c->arena;
        c->peer_gen[r] = c->arena_gen;
        continue;
      }
      if (c->peer_gen[r] == msgs[r].gen && c->peer_arena[r]) continue;
      if (c->peer_arena[r]) cudaIpcCloseMemHandle(c->peer_arena[r]);
      c->peer_arena[r] = nullptr;
      cudaError_t e = cudaIpcOpenMemHandle(&c->peer_arena[r], msgs[r].handle, cudaIpcMemLazyEnablePeerAccess);
      if (e != cudaSuccess) {
        cudaGetLastError();
        c->peer_arena[r] = nullptr;
        ok = false;
      } else {
        c->peer_gen[r] = msgs[r].gen;
      }
    }
  }
  if (!ok || remap) {
    // a mapping can fail on one rank only: agree on the outcome.  This round trip also makes it safe to free a
    // retired arena: every peer has closed its mapping of it by now.
    std::vector<int64_t> votes = allgather_counts(ctx, std::vector<int64_t>{ok ? 1 : 0});
    for (int64_t v : votes) ok = ok && v;
  }
  if (retired) cudaFree(retired);
  c->p2p_state = ok ? 1 : -1;
  return ok;
}

// ---- push form: receive-arena offer, block layout, peer mappings ----
struct PushMsg {  // rides in the metadata message behind ArenaMsg
  int64_t ok;   // 1 = this rank offers [off, off + cap) of arena `gen` and wants the push form
  int64_t gen;
  int64_t off;
  int64_t cap;
  cudaIpcMemHandle_t handle;  // 64 bytes
};
static_assert(sizeof(PushMsg) == 96, "PushMsg is exchanged as 12 int64 words");
constexpr size_t PUSH_MAX_ARENAS = 6;

// layout of one destination's receive block holding `rows` rows: fixed-width columns in schema order (elided ones
// take no room), each 256-byte aligned with slack behind it for the vector loads of the consumers
static size_t push_layout(const Batch& like, int64_t rows, std::vector<size_t>* col_off) {
  size_t off = 0;
  col_off->assign(like.cols.size(), 0);
  for (size_t c = 0; c < like.cols.size(); ++c) {
    if (!like.cols[c].data) continue;
    (*col_off)[c] = off;
    off += align256((size_t)rows * dtype_width(like.cols[c].dtype) + 64);
  }
  return off ? off : 256;
}
static bool push_wanted(const Batch& local) {
  const char* e = getenv("B2_SHUFFLE");
  if (e && strcmp(e, "push") != 0) return false;  // "pull" / "nccl"
  for (auto& c : local.cols)
    if (c.data && (c.dtype == B2_UTF8 || dtype_width(c.dtype) == 0)) return false;
  return true;
}
// picks (or allocates) the arena whose largest free span the coming exchange may land in
static void push_offer(Ctx* ctx, size_t estimate, PushMsg* msg, std::shared_ptr<RecvArena>* arena) {
  Comm* c = ctx->comm;
  memset(msg, 0, sizeof *msg);
  arena->reset();
  if (c->push_state < 0) return;
  size_t best_off = 0, best_len = 0;
  for (auto& a : c->recv_arenas) {
    size_t o, l;
    a->largest(&o, &l);
    if (l > best_len) {
      best_off = o;
      best_len = l;
      *arena = a;
    }
  }
  // grow on what was learned from an offer that was too small; the very first offer is sized from the local rows
  const size_t want = c->recv_arenas.empty() ? std::max(estimate, c->recv_predict) : c->recv_predict;
  if (best_len < want || c->recv_arenas.empty()) {
    if (c->recv_arenas.size() >= PUSH_MAX_ARENAS) {
      if (!*arena) return;
    } else {
      const char* fl = getenv("B2_PUSH_ARENA_MIN_MB");  // tests shrink the floor to reach the grow / too-small paths
      size_t cap = align256(std::max<size_t>(want * 5 / 2, (size_t)(fl ? atoi(fl) : 256) << 20));
      size_t free_b = 0, total_b = 0;
      if (cudaMemGetInfo(&free_b, &total_b) != cudaSuccess) cudaGetLastError();
      if (cap > free_b / 3) cap = align256(free_b / 3);
      void* p = nullptr;
      if (cap >= want && cudaMalloc(&p, cap) == cudaSuccess) {
        auto a = std::make_shared<RecvArena>();
        a->base = p;
        a->cap = cap;
        a->gen = ((int64_t)(ctx->rank + 1) << 32) | ++c->recv_gen;
        a->spans.reset(cap);
        c->recv_arenas.push_back(a);
        *arena = a;
        best_off = 0;
        best_len = cap;
      } else {
        cudaGetLastError();
        if (!*arena) return;
      }
    }
  }
  if (cudaIpcGetMemHandle(&msg->handle, (*arena)->base) != cudaSuccess) {
    cudaGetLastError();
    arena->reset();
    return;
  }
  msg->ok = 1;
  msg->gen = (*arena)->gen;
  msg->off = (int64_t)best_off;
  msg->cap = (int64_t)best_len;
}
// maps the offered arenas not seen before.  Every rank sees the same offers, so "something new" is a global fact
// and the vote below is entered by all ranks or by none.
static bool push_map_peers(Ctx* ctx, const std::vector<PushMsg>& msgs, std::vector<char*>* base) {
  Comm* c = ctx->comm;
  const int nr = ctx->nranks, me = ctx->rank;
  if (c->peer_recv.empty()) c->peer_recv.assign(nr, {});
  base->assign(nr, nullptr);
  bool ok = true, fresh = false;
  for (int r = 0; r < nr; ++r) {
    if (r == me) continue;
    auto it = c->peer_recv[r].find(msgs[r].gen);
    if (it != c->peer_recv[r].end()) {
      (*base)[r] = static_cast<char*>(it->second);
      continue;
    }
    fresh = true;
    void* p = nullptr;
    cudaError_t e = cudaIpcOpenMemHandle(&p, msgs[r].handle, cudaIpcMemLazyEnablePeerAccess);
    if (e != cudaSuccess) {
      cudaGetLastError();
      ok = false;
    } else {
      c->peer_recv[r][msgs[r].gen] = p;
      (*base)[r] = static_cast<char*>(p);
    }
  }
  // has any rank offered an arena that was new to its peers?  (my own offer is new to them iff it is new to me)
  bool mine_fresh = !c->recv_published[msgs[me].gen];
  c->recv_published[msgs[me].gen] = true;
  if (nr > 1 && (fresh || mine_fresh)) {
    std::vector<int64_t> votes = allgather_counts(ctx, std::vector<int64_t>{ok ? 1 : 0});
    for (int64_t v : votes) ok = ok && v;
  }
  if (!ok) c->push_state = -1;
  return ok;
}

// ---- arena layout of one rank's staged batch: a pure function of its row / character totals, so every peer
// derives every other rank's layout from the exchanged counts.  Every column region leaves 16 bytes of slack per
// destination slice for the co-alignment padding. ----
struct ColRegion {
  size_t val_off;  // fixed width: values; Utf8: int32 lengths
  size_t chr_off;  // Utf8: character bytes
};
static size_t arena_layout(const Schema& schema, int64_t rows, const std::vector<int64_t>& chars, int nranks,
                           std::vector<ColRegion>* regions) {
  size_t off = 0;
  regions->assign(schema.fields.size(), ColRegion{0, 0});
  for (size_t c = 0; c < schema.fields.size(); ++c) {
    int dt = schema.fields[c].dtype;
    size_t w = dt == B2_UTF8 ? 4 : dtype_width(dt);
    (*regions)[c].val_off = off;
    off += align256((size_t)rows * w + 32 * (size_t)nranks + 16);
    if (dt == B2_UTF8) {
      (*regions)[c].chr_off = off;
      off += align256((size_t)chars[c] + 32 * (size_t)nranks + 16);
    }
  }
  return off;
}
// byte offset of slice d inside a column region of source s: slices are packed in destination order, each start
// rounded up to 16 and then shifted so that it is congruent (mod 16) to the receiver's store offset `dst_bytes[d]`
static void slice_bases(const std::vector<int64_t>& slice_bytes, const std::vector<int64_t>& dst_bytes,
                        std::vector<int64_t>* base) {
  size_t nr = slice_bytes.size();
  base->assign(nr, 0);
  int64_t end = 0;
  for (size_t d = 0; d < nr; ++d) {
    int64_t b = ((end + 15) & ~15ll) + (dst_bytes[d] & 15);
    (*base)[d] = b;
    end = b + slice_bytes[d];
  }
}

__global__ void stage_lens_kernel(const int32_t* __restrict__ off, int64_t n, int32_t* __restrict__ out) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  for (; i < n; i += (int64_t)gridDim.x * blockDim.x) out[i] = off[i + 1] - off[i];
}

// ---- the pull.  One descriptor per (source rank, column buffer); blockIdx.y selects the descriptor, the CTAs of
// a row stride over it.  Source and destination are congruent mod 16 by construction, so after a head of < 16
// bytes everything moves as 16-byte peer loads (NVLink) and 16-byte local stores. ----
struct PullSeg {
  const char* src;
  char* dst;
  int64_t bytes;
};
constexpr int PULL_UNROLL = 4;
template <int PULL_UNROLL>
__global__ void __launch_bounds__(256) pull_kernel(const PullSeg* __restrict__ segs, int seg0) {
  PullSeg s = segs[seg0 + blockIdx.y];
  if (s.bytes <= 0) return;
  const int64_t tid = blockIdx.x * (int64_t)blockDim.x + threadIdx.x, nth = (int64_t)gridDim.x * blockDim.x;
  if ((((uintptr_t)s.src ^ (uintptr_t)s.dst) & 15) != 0) {  // not congruent (never on the planned path): bytes
    for (int64_t i = tid; i < s.bytes; i += nth) s.dst[i] = s.src[i];
    return;
  }
  int64_t head = (16 - ((uintptr_t)s.dst & 15)) & 15;
  if (head > s.bytes) head = s.bytes;
  int64_t body = (s.bytes - head) >> 4, tail = s.bytes - head - (body << 4);
  if (tid < head) s.dst[tid] = s.src[tid];
  if (tid < tail) s.dst[head + (body << 4) + tid] = s.src[head + (body << 4) + tid];
  const int4* __restrict__ src = reinterpret_cast<const int4*>(s.src + head);
  int4* __restrict__ dst = reinterpret_cast<int4*>(s.dst + head);
  int64_t i = tid;
  for (; i + (PULL_UNROLL - 1) * nth < body; i += PULL_UNROLL * nth) {
    int4 v[PULL_UNROLL];
#pragma unroll
    for (int u = 0; u < PULL_UNROLL; ++u) v[u] = src[i + u * nth];
#pragma unroll
    for (int u = 0; u < PULL_UNROLL; ++u) dst[i + u * nth] = v[u];
  }
  for (; i < body; i += nth) dst[i] = src[i];
}

struct LensToOff {
  int32_t* off;
  int64_t n;
  __device__ __forceinline__ void operator()(int64_t i, int v, int64_t pre) const {
    off[i] = (int32_t)pre;
    if (i == n - 1) off[n] = (int32_t)(pre + v);
  }
};

// Shuffle of `local` given phase 1 of its hash split (`st`, n_out == nranks).  Returns this rank's output
// partition, or nullptr when peer mapping is unavailable (the caller then takes the NCCL send/recv path).
BatchPtr comm_shuffle_p2p(Ctx* ctx, const Batch& local, const SplitState& st, bool* pushed) {
  if (pushed) *pushed = false;
  require_no_nulls(local, "exchange over NVLink");
  if (!ctx->comm) fail(B2_ERR_INVALID_ARG, "no communicator (call b2_comm_init)");
  Comm* cm = ctx->comm;
  if (cm->p2p_state < 0) return nullptr;
  if (getenv("B2_SHUFFLE") && !strcmp(getenv("B2_SHUFFLE"), "nccl")) return nullptr;
  const int nr = ctx->nranks, me = ctx->rank;
  const size_t ncol = local.cols.size();
  const int64_t n = local.num_rows;
  if (st.n_out != nr || st.n != n) fail(B2_ERR_EXECUTION, "internal: split state does not match the exchange");
  PhaseTimer pt(ctx);

  // Utf8 columns (rare on shuffled sides) go through the permutation + generic take; their per-destination
  // character counts join the metadata message
  bool has_utf8 = false;
  for (size_t c = 0; c < ncol; ++c) has_utf8 = has_utf8 || (local.cols[c].data && local.cols[c].dtype == B2_UTF8);
  auto elided = [&](size_t c) { return !local.cols[c].data; };  // pruned upstream: not moved, stays elided
  std::vector<Column> taken(ncol);
  std::vector<std::vector<int64_t>> chr_bounds(ncol);
  std::vector<int64_t> my_chars(ncol, 0), my_send_off(nr + 1, 0);
  if (has_utf8) {
    std::vector<int64_t> counts;
    hash_split_counts(ctx, st, &counts);
    for (int d = 0; d < nr; ++d) my_send_off[d + 1] = my_send_off[d] + counts[d];
    BufPtr perm = hash_split_perm(ctx, st);
    for (size_t c = 0; c < ncol; ++c) {
      if (local.cols[c].dtype != B2_UTF8 || elided(c)) continue;
      taken[c] = take_column(ctx, local.cols[c], perm->as<uint32_t>(), my_send_off[nr]);
      my_chars[c] = taken[c].chars_len;
      chr_bounds[c].assign(nr + 1, 0);
      if (n) {
        std::vector<int32_t> tmp(nr + 1, 0);
        for (int d = 0; d <= nr; ++d)
          B2_CUDA(cudaMemcpyAsync(&tmp[d], taken[c].values<int32_t>() + my_send_off[d], 4, cudaMemcpyDeviceToHost,
                                  ctx->stream));
        B2_SYNC(ctx);
        for (int d = 0; d <= nr; ++d) chr_bounds[c][d] = tmp[d];
      }
    }
  }

  // ---- step 2: metadata all-gather (counts from the device histogram + arena handle) ----
  std::vector<ColRegion> my_regions;
  size_t need = arena_layout(local.schema, n, my_chars, nr, &my_regions);
  ArenaMsg my_msg;
  void* retired = nullptr;
  arena_local(ctx, need, &my_msg, &retired);
  // the push form's offer: the free span of my receive arena this exchange may land in
  PushMsg my_push;
  std::shared_ptr<RecvArena> my_recv;
  memset(&my_push, 0, sizeof my_push);
  if (push_wanted(local)) {
    std::vector<size_t> tmp;
    size_t est = push_layout(local, n, &tmp);
    push_offer(ctx, est + est / 4 + ((size_t)1 << 20), &my_push, &my_recv);
  }
  const size_t AW = (sizeof(ArenaMsg) + sizeof(PushMsg)) / 8, L = AW + (size_t)(nr + 1) + (size_t)nr * ncol;
  std::vector<int64_t> host_words(L, 0);
  memcpy(host_words.data(), &my_msg, sizeof my_msg);
  memcpy(host_words.data() + sizeof(ArenaMsg) / 8, &my_push, sizeof my_push);
  for (int d = 0; d < nr; ++d)
    for (size_t c = 0; c < ncol; ++c)
      if (!chr_bounds[c].empty()) host_words[AW + nr + 1 + (size_t)d * ncol + c] = chr_bounds[c][d + 1] - chr_bounds[c][d];
  auto dsend = dev_alloc(ctx, L * 8);
  auto drecv = dev_alloc(ctx, L * 8 * nr);
  B2_CUDA(cudaMemcpyAsync(dsend->ptr, host_words.data(), L * 8, cudaMemcpyHostToDevice, ctx->stream));
  hash_split_starts_device(ctx, st, dsend->as<int64_t>() + AW);
  B2_NCCL(ncclAllGather(dsend->ptr, drecv->ptr, L, ncclInt64, cm->nccl, ctx->stream));
  std::vector<int64_t> all(L * nr);
  B2_READBACK(ctx, all.data(), drecv->ptr, all.size() * 8);
  pt.mark("shuffle 1-2: hist + scan + metadata all-gather");
  std::vector<ArenaMsg> msgs(nr);
  for (int r = 0; r < nr; ++r) memcpy(&msgs[r], all.data() + (size_t)r * L, sizeof(ArenaMsg));
  if (!arena_map_peers(ctx, msgs, retired)) return nullptr;
  // my staged rows = the rows that passed the fused filter: the layout every peer derives for me
  arena_layout(local.schema, all[(size_t)me * L + AW + nr], my_chars, nr, &my_regions);
  auto cnt = [&](int s, int d) { return all[(size_t)s * L + AW + d + 1] - all[(size_t)s * L + AW + d]; };
  auto chars = [&](int s, int d, size_t c) { return all[(size_t)s * L + AW + nr + 1 + (size_t)d * ncol + c]; };
  // R[d][s]: row offset, inside destination d's output, of the rows coming from source s (CR: characters)
  std::vector<std::vector<int64_t>> R(nr, std::vector<int64_t>(nr + 1, 0));
  std::vector<std::vector<std::vector<int64_t>>> CR(ncol);
  for (int d = 0; d < nr; ++d)
    for (int s = 0; s < nr; ++s) R[d][s + 1] = R[d][s] + cnt(s, d);
  for (size_t c = 0; c < ncol; ++c) {
    if (local.cols[c].dtype != B2_UTF8 || elided(c)) continue;
    CR[c].assign(nr, std::vector<int64_t>(nr + 1, 0));
    for (int d = 0; d < nr; ++d)
      for (int s = 0; s < nr; ++s) CR[c][d][s + 1] = CR[c][d][s] + chars(s, d, c);
  }
  // slice bases of column c staged by source s (value region; `chr` = character region of a Utf8 column)
  auto bases_of = [&](int s, size_t c, bool chr, std::vector<int64_t>* base) {
    int dt = local.cols[c].dtype;
    int64_t w = chr ? 1 : (dt == B2_UTF8 ? 4 : (int64_t)dtype_width(dt));
    std::vector<int64_t> sb(nr), db(nr);
    for (int d = 0; d < nr; ++d) {
      sb[d] = (chr ? chars(s, d, c) : cnt(s, d)) * w;
      db[d] = (chr ? CR[c][d][s] : R[d][s]) * w;
    }
    slice_bases(sb, db, base);
  };
  int64_t total = R[me][nr];
  if (total >= (1ll << 31) - 1) fail(B2_ERR_NOT_IMPLEMENTED, "received partition beyond 2^31-2 rows");
  char* arena = static_cast<char*>(cm->arena);

  // ---- push form: every destination offered room for its total -> ONE kernel stores my rows into the final columns ----
  {
    std::vector<PushMsg> pm(nr);
    bool fits = true;
    for (int r = 0; r < nr; ++r) {
      memcpy(&pm[r], all.data() + (size_t)r * L + sizeof(ArenaMsg) / 8, sizeof(PushMsg));
      fits = fits && pm[r].ok;
    }
    std::vector<std::vector<size_t>> col_off(nr);
    std::vector<size_t> block_need(nr, 0);
    for (int d = 0; d < nr && fits; ++d) block_need[d] = push_layout(local, R[d][nr], &col_off[d]);
    for (int d = 0; d < nr && fits; ++d)
      if ((int64_t)block_need[d] > pm[d].cap) {
        fits = false;  // every rank computes this from the same numbers: all of them take the pull form below
        for (int q = 0; q < nr; ++q)
          if (q == me && (int64_t)block_need[q] > pm[q].cap)
            cm->recv_predict = std::max(cm->recv_predict, block_need[q] + block_need[q] / 4);
      }
    if (getenv("B2_PUSH_DEBUG"))
      fprintf(stderr, "[push r%d] ok=%lld offer gen=%llx off=%lld cap=%lld need=%zu fits=%d arenas=%zu predict=%zu\n", me,
              (long long)pm[me].ok, (long long)pm[me].gen, (long long)pm[me].off, (long long)pm[me].cap, block_need[me],
              (int)fits, cm->recv_arenas.size(), cm->recv_predict);
    std::vector<char*> pbase;
    if (fits && push_map_peers(ctx, pm, &pbase)) {
      const size_t boff = (size_t)pm[me].off, blen = align256(block_need[me]);
      if (!my_recv || !my_recv->take(boff, blen)) fail(B2_ERR_EXECUTION, "internal: the offered receive span is gone");
      pbase[me] = static_cast<char*>(my_recv->base);
      // the block: returns its span to the arena when the last window of it is dropped
      DevBuf* raw = new DevBuf(ctx, 0);
      raw->ptr = pbase[me] + boff;
      raw->bytes = blen;
      std::shared_ptr<RecvArena> keep = my_recv;
      BufPtr block(raw, [keep, boff, blen](DevBuf* b) {
        keep->give(boff, blen);
        b->ptr = nullptr;
        delete b;
      });
      std::vector<const void*> src;
      std::vector<int> width;
      std::vector<int64_t> addr;
      int64_t row_bytes = 0;
      for (size_t c = 0; c < ncol; ++c) {
        if (elided(c)) continue;
        const int64_t w = (int64_t)dtype_width(local.cols[c].dtype);
        row_bytes += w;
        src.push_back(local.cols[c].data->ptr);
        width.push_back((int)w);
        for (int d = 0; d < nr; ++d)
          addr.push_back((int64_t)(uintptr_t)(pbase[d] + pm[d].off + col_off[d][c]) + R[d][me] * w);
      }
      if (n && !src.empty()) {
        auto daddr = dev_alloc(ctx, addr.size() * 8);
        B2_CUDA(cudaMemcpyAsync(daddr->ptr, addr.data(), addr.size() * 8, cudaMemcpyHostToDevice, ctx->stream));
        ctx->prof_begin();  // b2_ctx_set_profiling: CUDA events around the exchange kernel (bench.py's NVLink roofline)
        for (size_t g = 0; g < src.size(); g += 8) {
          int k = (int)std::min<size_t>(8, src.size() - g);
          hash_split_push(ctx, st, k, src.data() + g, width.data() + g, daddr->as<int64_t>() + g * nr,
                          all[(size_t)me * L + AW + nr]);
        }
        ctx->prof_end();
      }
      for (int d = 0; d < nr; ++d)
        if (d != me) ctx->bytes_sent += cnt(me, d) * row_bytes;  // bytes crossing NVLink
      pt.mark("shuffle 3p: push over NVLink");
      {  // stream-ordered barrier: every rank's stores into my block have been performed
        auto flag = dev_alloc_zero(ctx, 16);
        B2_NCCL(ncclAllReduce(flag->ptr, flag->ptr, 1, ncclInt32, ncclSum, cm->nccl, ctx->stream));
      }
      pt.mark("shuffle 4p: barrier");
      auto out = std::make_shared<Batch>();
      out->ctx = ctx;
      out->schema = local.schema;
      out->num_rows = total;
      for (size_t c = 0; c < ncol; ++c) {
        Column o;
        o.dtype = local.cols[c].dtype;
        o.length = total;
        if (!elided(c)) o.data = dev_view(block, col_off[me][c], (size_t)total * dtype_width(o.dtype) + 64);
        out->cols.push_back(std::move(o));
      }
      if (pushed) *pushed = true;
      return out;
    }
  }

  // ---- step 3: move my columns into my arena, destination-major, slices co-aligned with their receivers ----
  {
    std::vector<const void*> src;
    std::vector<void*> dst;
    std::vector<int> width;
    std::vector<int64_t> base_table;
    for (size_t c = 0; c < ncol; ++c) {
      const Column& col = local.cols[c];
      std::vector<int64_t> base;
      if (elided(c)) continue;
      if (col.dtype == B2_UTF8) {
        if (!n) continue;
        bases_of(me, c, false, &base);
        for (int d = 0; d < nr; ++d) {
          int64_t k = cnt(me, d);
          if (!k) continue;
          stage_lens_kernel<<<grid_for(ctx, 4, k, 256), 256, 0, ctx->stream>>>(
              taken[c].values<int32_t>() + my_send_off[d], k, (int32_t*)(arena + my_regions[c].val_off + base[d]));
          B2_LAUNCHED(ctx);
        }
        bases_of(me, c, true, &base);
        for (int d = 0; d < nr; ++d)
          if (chars(me, d, c))
            B2_CUDA(cudaMemcpyAsync(arena + my_regions[c].chr_off + base[d],
                                    taken[c].chars->as<char>() + chr_bounds[c][d], (size_t)chars(me, d, c),
                                    cudaMemcpyDeviceToDevice, ctx->stream));
        continue;
      }
      bases_of(me, c, false, &base);
      src.push_back(col.data->ptr);
      dst.push_back(arena);
      width.push_back((int)dtype_width(col.dtype));
      for (int d = 0; d < nr; ++d) base_table.push_back((int64_t)my_regions[c].val_off + base[d]);
    }
    if (n && !src.empty()) {
      auto dbase = dev_alloc(ctx, base_table.size() * 8);
      B2_CUDA(cudaMemcpyAsync(dbase->ptr, base_table.data(), base_table.size() * 8, cudaMemcpyHostToDevice, ctx->stream));
      for (size_t g = 0; g < src.size(); g += 8) {
        int k = (int)std::min<size_t>(8, src.size() - g);
        hash_split_move(ctx, st, k, src.data() + g, dst.data() + g, width.data() + g, dbase->as<int64_t>() + g * nr,
                        nullptr, all[(size_t)me * L + AW + nr]);  // my surviving rows: picks the sparse move kernel
      }
    }
  }
  pt.mark("shuffle 3: move columns into the arena");
  // ---- step 4: stream-ordered barrier: every arena is staged ----
  {
    auto flag = dev_alloc_zero(ctx, 16);
    B2_NCCL(ncclAllReduce(flag->ptr, flag->ptr, 1, ncclInt32, ncclSum, cm->nccl, ctx->stream));
  }
  pt.mark("shuffle 4: barrier");

  // ---- step 5: pull my slice of every column of every source into the final output columns ----
  auto out = std::make_shared<Batch>();
  out->ctx = ctx;
  out->schema = local.schema;
  out->num_rows = total;
  std::vector<BufPtr> lens_bufs(ncol);
  for (size_t c = 0; c < ncol; ++c) {
    Column o;
    o.dtype = local.cols[c].dtype;
    o.length = total;
    if (elided(c)) {
      out->cols.push_back(std::move(o));
      continue;
    }
    if (o.dtype == B2_UTF8) {
      if (CR[c][me][nr] >= (1ll << 31)) fail(B2_ERR_EXECUTION, "Utf8 column beyond 2 GiB after the shuffle");
      o.chars_len = CR[c][me][nr];
      o.chars = dev_alloc(ctx, (size_t)o.chars_len + 16);
      o.data = dev_alloc_zero(ctx, (size_t)(total + 1) * 4);
      lens_bufs[c] = dev_alloc(ctx, (size_t)(total + 1) * 4);
    } else {
      o.data = dev_alloc(ctx, (size_t)total * dtype_width(o.dtype) + 16);
    }
    out->cols.push_back(std::move(o));
  }
  std::vector<PullSeg> segs;
  int64_t max_seg = 0;
  for (int k = 0; k < nr; ++k) {
    int s = (me + k) % nr;  // local slice first, then the peers in an order that spreads ranks over sources
    int64_t rows_s = all[(size_t)s * L + AW + nr];
    std::vector<int64_t> chars_s(ncol, 0);
    for (int d = 0; d < nr; ++d)
      for (size_t c = 0; c < ncol; ++c) chars_s[c] += chars(s, d, c);
    std::vector<ColRegion> reg;
    arena_layout(local.schema, rows_s, chars_s, nr, &reg);
    const char* base_ptr = static_cast<const char*>(cm->peer_arena[s]);
    size_t first = segs.size();
    for (size_t c = 0; c < ncol; ++c) {
      int dt = local.cols[c].dtype;
      std::vector<int64_t> base;
      if (elided(c)) continue;
      bases_of(s, c, false, &base);
      if (dt == B2_UTF8) {
        segs.push_back(PullSeg{base_ptr + reg[c].val_off + base[me], lens_bufs[c]->as<char>() + R[me][s] * 4,
                               cnt(s, me) * 4});
        bases_of(s, c, true, &base);
        segs.push_back(PullSeg{base_ptr + reg[c].chr_off + base[me], out->cols[c].chars->as<char>() + CR[c][me][s],
                               chars(s, me, c)});
      } else {
        int64_t w = (int64_t)dtype_width(dt);
        segs.push_back(PullSeg{base_ptr + reg[c].val_off + base[me], out->cols[c].data->as<char>() + R[me][s] * w,
                               cnt(s, me) * w});
      }
    }
    for (size_t q = first; q < segs.size(); ++q) {
      max_seg = std::max<int64_t>(max_seg, segs[q].bytes);
      if (s != me) ctx->bytes_sent += segs[q].bytes;  // bytes crossing NVLink (pulled here = sent by the peer)
    }
  }
  if (!segs.empty() && max_seg > 0) {
    auto dsegs = dev_alloc(ctx, segs.size() * sizeof(PullSeg));
    B2_CUDA(cudaMemcpyAsync(dsegs->ptr, segs.data(), segs.size() * sizeof(PullSeg), cudaMemcpyHostToDevice, ctx->stream));
    int64_t want = (max_seg / 16 + 256 * PULL_UNROLL - 1) / (256 * PULL_UNROLL);
    int64_t cap = std::max<int64_t>(1, (int64_t)ctx->sm_count * 8 / (int64_t)segs.size());
    dim3 g((unsigned)std::max<int64_t>(1, std::min(want, cap)), (unsigned)segs.size());
    // experiment knobs (defaults = the measured best): B2_PULL_UNROLL 4|8, B2_PULL_CTAS_PER_SM, B2_PULL_SPLIT=1
    // launches the local slice (the first ncol-ish descriptors) separately from the peer slices
    int unroll = getenv("B2_PULL_UNROLL") ? atoi(getenv("B2_PULL_UNROLL")) : 4;
    int per_sm = getenv("B2_PULL_CTAS_PER_SM") ? atoi(getenv("B2_PULL_CTAS_PER_SM")) : 8;
    cap = std::max<int64_t>(1, (int64_t)ctx->sm_count * per_sm / (int64_t)segs.size());
    g.x = (unsigned)std::max<int64_t>(1, std::min(want, cap));
    size_t n_local = segs.size() / (size_t)nr;  // descriptors of the local slice come first
    auto launch = [&](size_t first, size_t count) {
      if (!count) return;
      dim3 gg(g.x, (unsigned)count);
      if (unroll == 8)
        pull_kernel<8><<<gg, 256, 0, ctx->stream>>>(dsegs->as<PullSeg>(), (int)first);
      else
        pull_kernel<4><<<gg, 256, 0, ctx->stream>>>(dsegs->as<PullSeg>(), (int)first);
      B2_LAUNCHED(ctx);
    };
    ctx->prof_begin();  // b2_ctx_set_profiling: CUDA events around the exchange step (bench.py's NVLink roofline)
    if (getenv("B2_PULL_SPLIT")) {
      launch(0, n_local);
      pt.mark("shuffle 5a: local slice copy");
      launch(n_local, segs.size() - n_local);
    } else {
      launch(0, segs.size());
    }
    ctx->prof_end();
  }
  pt.mark("shuffle 5: pull over NVLink");
  for (size_t c = 0; c < ncol; ++c)
    if (local.cols[c].dtype == B2_UTF8 && total && !elided(c))
      device_scan(ctx, I32In{lens_bufs[c]->as<int32_t>()}, total, LensToOff{out->cols[c].data->as<int32_t>(), total},
                  (int64_t*)nullptr);
  return out;
}

}  // namespace b2
