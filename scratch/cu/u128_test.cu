#include <cstdio>
#include <cstdint>
typedef unsigned __int128 u128;
typedef unsigned long long u64;
__global__ void k(const unsigned char* chr, int len, int off, u64* out) {
  u128 bytes = 0;
  int m = len < 15 ? len : 15;
  for (int q = 0; q < m; ++q) bytes |= (u128)chr[q] << (8 * q);
  out[0] = (u64)bytes; out[1] = (u64)(bytes >> 64);
  u128 mm = (((u128)1 << (8 * 14)) - 1);
  u128 key = 0;
  key |= ((u128)(uint8_t)len | ((bytes & mm) << 8)) << (8 * off);
  out[2] = (u64)key; out[3] = (u64)(key >> 64);
  u128 bits = key >> (8 * off);
  for (int q = 0; q < len; ++q) ((unsigned char*)(out + 4))[q] = (uint8_t)(bits >> (8 * (q + 1)));
}
int main() {
  unsigned char* d; u64* o;
  cudaMalloc(&d, 32); cudaMalloc(&o, 128); cudaMemset(o, 0, 128);
  const char* s = "ABCDEFGHIJKLMN";
  cudaMemcpy(d, s, 14, cudaMemcpyHostToDevice);
  k<<<1, 1>>>(d, 14, 0, o);
  u64 h[16]; cudaMemcpy(h, o, 128, cudaMemcpyDeviceToHost);
  printf("bytes lo=%016llx hi=%016llx\nkey   lo=%016llx hi=%016llx\nchars=%.14s\n", h[0], h[1], h[2], h[3], (char*)(h + 4));
  return 0;
}
