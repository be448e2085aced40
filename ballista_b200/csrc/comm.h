// comm.h — the per-context communicator: NCCL for rendezvous-style collectives (counts, partial-state
// all-gather) and the peer-mapped exchange arena the shuffle kernels read over NVLink (shuffle.cu).
#pragma once
#include <nccl.h>

#include <string>
#include <vector>

#include "core.h"

namespace b2 {

struct Comm {
  ncclComm_t nccl = nullptr;
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
