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

}  // namespace b2
