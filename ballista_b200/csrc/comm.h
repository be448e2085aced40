// comm.h — the per-context communicator: NCCL for rendezvous-style collectives (counts, partial-state
// all-gather) and the peer-mapped exchange arena the shuffle kernels read over NVLink (shuffle.cu).
#pragma once
#include <nccl.h>

#include <map>
#include <memory>
#include <mutex>
#include <string>
#include <vector>

#include "core.h"

namespace b2 {

// Receive arena of the PUSH exchange (shuffle.cu): a cudaMalloc'ed region every peer maps through CUDA IPC and stores
// into.  One exchange takes one block of it; the exchange's output columns are windows of that block, which goes back
// to the free list when the last of them is dropped (any thread), so nothing is copied out of the arena.
struct RecvArena {
  void* base = nullptr;
  size_t cap = 0;
  int64_t gen = 0;  // unique per (rank, arena): what the peers key their mappings on
  std::mutex mu;
  std::map<size_t, size_t> free_spans;  // offset -> length, coalesced
  ~RecvArena() {
    if (base) cudaFree(base);
  }
  void largest(size_t* off, size_t* len) {
    std::lock_guard<std::mutex> g(mu);
    *off = 0;
    *len = 0;
    for (auto& s : free_spans)
      if (s.second > *len) {
        *off = s.first;
        *len = s.second;
      }
  }
  bool take(size_t off, size_t len) {  // [off, off + len) must lie at the start of a free span
    std::lock_guard<std::mutex> g(mu);
    auto it = free_spans.find(off);
    if (it == free_spans.end() || it->second < len) return false;
    size_t rest = it->second - len;
    free_spans.erase(it);
    if (rest) free_spans[off + len] = rest;
    return true;
  }
  void give(size_t off, size_t len) {
    std::lock_guard<std::mutex> g(mu);
    auto nx = free_spans.lower_bound(off);
    if (nx != free_spans.end() && off + len == nx->first) {
      len += nx->second;
      nx = free_spans.erase(nx);
    }
    if (nx != free_spans.begin()) {
      auto pv = std::prev(nx);
      if (pv->first + pv->second == off) {
        pv->second += len;
        return;
      }
    }
    free_spans[off] = len;
  }
  bool idle() {
    std::lock_guard<std::mutex> g(mu);
    return free_spans.size() == 1 && free_spans.begin()->second == cap;
  }
};

struct Comm {
  ncclComm_t nccl = nullptr;
  // push exchange: this rank's receive arenas and the peers' arenas mapped so far (by generation)
  std::vector<std::shared_ptr<RecvArena>> recv_arenas;
  int64_t recv_gen = 0;
  size_t recv_predict = 0;                          // bytes the next exchange is expected to receive
  std::vector<std::map<int64_t, void*>> peer_recv;  // [nranks]: generation -> mapped base
  std::map<int64_t, bool> recv_published;           // own arenas the peers have been asked to map
  int push_state = 0;                               // -1: a peer mapping failed once, the pull path is used from then on
  // exchange arena: one cudaMalloc'ed region per rank, CUDA-IPC-mapped by every peer of the box
  void* arena = nullptr;
  size_t arena_cap = 0;
  int64_t arena_gen = 0;                // bumped on every (re)allocation
  std::vector<void*> peer_arena;        // [nranks]; own entry = arena
  std::vector<int64_t> peer_gen;        // generation each mapping was opened at
  int p2p_state = 0;                    // 0 unknown, 1 usable, -1 unavailable (NCCL send/recv fallback)
};

#define B2_NCCL(expr)                                                                         \
  do {                                                                                        \
    ncclResult_t _r = (expr);                                                                 \
    if (_r != ncclSuccess) ::b2::fail(B2_ERR_NCCL, std::string(#expr) + ": " + ncclGetErrorString(_r)); \
  } while (0)

// exchange of per-rank int64 vectors (row / byte counts); returns the [nranks][len] matrix on the host
std::vector<int64_t> allgather_counts(Ctx* ctx, const std::vector<int64_t>& mine);
void shuffle_arena_release(Ctx* ctx);

}  // namespace b2
