#include <cstdio>
#include <cstdint>
#include <cstring>
#include "../../ballista_b200/csrc/table.cuh"
using namespace b2;
#define B2_UTF8 7
struct PackPart { int dtype, byte_off, bytes; const void* data; const uint8_t* chars; const uint8_t* valid; };
struct PackArgs { PackPart part[8]; int n; };
__global__ void pack_keys_kernel(PackArgs a, int64_t n, Key128* out, int* status, int null_tag) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  bool overflow = false;
  for (; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    u128 key = 0;
    bool is_null = false;
    for (int p = 0; p < a.n; ++p) {
      const PackPart& pp = a.part[p];
      if (pp.valid && !pp.valid[i]) { is_null = true; continue; }
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
  if (overflow) atomicExch(status, 2);
}
int main() {
  const char* s = "INDONESIAABCDEFGHIJKLMNXY";
  int32_t off[4] = {0, 9, 23, 25};
  uint8_t* dc; int32_t* dof; Key128* dk; int* st;
  cudaMalloc(&dc, 64); cudaMalloc(&dof, 64); cudaMalloc(&dk, 64); cudaMalloc(&st, 16); cudaMemset(st, 0, 16);
  cudaMemcpy(dc, s, 25, cudaMemcpyHostToDevice); cudaMemcpy(dof, off, 16, cudaMemcpyHostToDevice);
  PackArgs a; memset(&a, 0, sizeof a); a.n = 1; a.part[0] = PackPart{B2_UTF8, 0, 15, dof, dc, nullptr};
  pack_keys_kernel<<<1, 256>>>(a, 3, dk, st, 0);
  Key128 h[3]; cudaMemcpy(h, dk, 48, cudaMemcpyDeviceToHost);
  for (int i = 0; i < 3; ++i) printf("key %d lo=%016llx hi=%016llx\n", i, h[i].lo, h[i].hi);
  return 0;
}
