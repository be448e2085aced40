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
#include "span_list.h"

namespace b2 {

// Receive arena of the PUSH exchange (shuffle.cu): a cudaMalloc'ed region every peer maps through CUDA IPC and stores
// into.  One exchange takes one block of it; the exchange's output columns are windows of that block, which goes back
// to the free list when the last of them is dropped (any thread), so nothing is copied out of the arena.
struct RecvArena {
  void* base = nullptr;
  size_t cap = 0;
  int64_t gen = 0;  // unique per (rank, arena): what the peers key their mappings on
  SpanList spans;   // free byte spans of [0, cap)
  ~RecvArena() {
    if (base) cudaFree(base);
  }
  void largest(size_t* off, size_t* len) { spans.largest(off, len); }
  bool take(size_t off, size_t len) { return spans.take(off, len); }
  void give(size_t off, size_t len) { spans.give(off, len); }
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
