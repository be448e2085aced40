// table.cuh — device-side building blocks shared by the aggregate / join / repartition kernels:
// 128-bit packed keys, the global open-addressing group table (128-bit CAS claim + red.add accumulate),
// streaming loads, warp reductions.
#pragma once
#ifndef __CUDACC_RTC__
#include <cuda_runtime.h>
#include <stdint.h>

#include <type_traits>
#else
// compiled at run time by NVRTC (jit.cu): no host headers; libcu++ supplies the few type traits used here
#include <cuda/std/type_traits>
typedef signed char int8_t;
typedef unsigned char uint8_t;
typedef short int16_t;
typedef unsigned short uint16_t;
typedef int int32_t;
typedef unsigned int uint32_t;
typedef long long int64_t;
typedef unsigned long long uint64_t;
typedef unsigned long long uintptr_t;
namespace std {
using cuda::std::conditional;
using cuda::std::enable_if;
using cuda::std::false_type;
using cuda::std::integral_constant;
using cuda::std::is_integral;
using cuda::std::true_type;
}  // namespace std
#endif
// sig(): the signature text of an expression template (host only; the registry is keyed on it)
#ifdef __CUDACC_RTC__
#define B2_SIG(...)
#else
#define B2_SIG(...) \
  static std::string sig() { __VA_ARGS__ }
#endif

namespace b2 {

// status word of an aggregate launch (read back by the host)
// 4..6: raised by the small Final kernel (its host side redoes the aggregate through the checked path)
enum { AGG_OK = 0, AGG_ABORT_CARDINALITY = 1, AGG_KEY_TOO_WIDE = 2, AGG_TABLE_FULL = 3, AGG_INPUT_FAILED = 4,
       AGG_TOO_BIG = 5, AGG_NULL_KEY = 6 };

typedef unsigned long long u64;

// 128-bit unsigned integer as two explicit 64-bit words.  The compiler's own `unsigned __int128` is NOT used in device
// code: nvcc 12.9 drops the carry into the high word in `((x & m) << 8) << (8 * off)` of the key-packing kernel (a
// character at byte 7 of a string key came back as 0: scratch/cu/pack_test.cu reproduces it), so every shift here is
// spelled out.
struct u128 {
  u64 lo, hi;
  u128() = default;
  template <class T, class = typename std::enable_if<std::is_integral<T>::value>::type>
  __host__ __device__ constexpr u128(T v) : lo((u64)v), hi(0) {}
  template <class T, class = typename std::enable_if<std::is_integral<T>::value>::type>
  __host__ __device__ explicit constexpr operator T() const { return (T)lo; }
};
__host__ __device__ __forceinline__ constexpr u128 make_u128(u64 lo, u64 hi) {
  u128 r(lo);
  r.hi = hi;
  return r;
}
__host__ __device__ __forceinline__ constexpr u128 operator|(u128 a, u128 b) { return make_u128(a.lo | b.lo, a.hi | b.hi); }
__host__ __device__ __forceinline__ constexpr u128 operator&(u128 a, u128 b) { return make_u128(a.lo & b.lo, a.hi & b.hi); }
__host__ __device__ __forceinline__ constexpr u128 operator~(u128 a) { return make_u128(~a.lo, ~a.hi); }
__host__ __device__ __forceinline__ u128& operator|=(u128& a, u128 b) {
  a.lo |= b.lo;
  a.hi |= b.hi;
  return a;
}
__host__ __device__ __forceinline__ constexpr bool operator==(u128 a, u128 b) { return a.lo == b.lo && a.hi == b.hi; }
__host__ __device__ __forceinline__ constexpr bool operator!=(u128 a, u128 b) { return !(a == b); }
__host__ __device__ __forceinline__ constexpr u128 operator<<(u128 a, int s) {
  return s <= 0 ? a : s < 64 ? make_u128(a.lo << s, (a.hi << s) | (a.lo >> (64 - s))) : s < 128 ? make_u128(0, a.lo << (s - 64)) : make_u128(0, 0);
}
__host__ __device__ __forceinline__ constexpr u128 operator>>(u128 a, int s) {
  return s <= 0 ? a : s < 64 ? make_u128((a.lo >> s) | (a.hi << (64 - s)), a.hi >> s) : s < 128 ? make_u128(a.hi >> (s - 64), 0) : make_u128(0, 0);
}
__host__ __device__ __forceinline__ constexpr u128 operator-(u128 a, u128 b) {
  return make_u128(a.lo - b.lo, a.hi - b.hi - (a.lo < b.lo ? 1ull : 0ull));
}

// ---- packed group / join keys: 15 payload bytes, byte 15 always 0 so no key equals EMPTY ----
struct Key128 {
  u64 lo, hi;
};
__host__ __device__ __forceinline__ bool key_eq(const Key128& a, const Key128& b) { return a.lo == b.lo && a.hi == b.hi; }
constexpr u64 KEY_EMPTY = ~0ull;

__host__ __device__ __forceinline__ u64 mix64(u64 x) {
  x ^= x >> 33;
  x *= 0xff51afd7ed558ccdull;
  x ^= x >> 33;
  x *= 0xc4ceb9fe1a85ec53ull;
  x ^= x >> 33;
  return x;
}
__host__ __device__ __forceinline__ u64 key_hash(const Key128& k) { return mix64(k.lo ^ mix64(k.hi + 0x9E3779B97F4A7C15ull)); }

// 32-bit finaliser (murmur3): the join table's bucket / Bloom hash, and the hash partitioner's
__host__ __device__ __forceinline__ uint32_t fmix32(uint32_t h) {
  h ^= h >> 16;
  h *= 0x85ebca6bu;
  h ^= h >> 13;
  h *= 0xc2b2ae35u;
  h ^= h >> 16;
  return h;
}
// Bloom word for a key: blocked filter, one 32-bit word per key, 3 bits chosen by a second multiplicative hash so
// that they are independent of the word index (the low bits of the bucket hash).  Sized at >= 16 bits per build key
// it stays L2-resident where the 8-byte-slot table (16 B per key at load 0.5) does not, and a probe that misses it
// (most rows of a selective join: Q3's lineitem side matches ~1%) never touches the table.
__device__ __forceinline__ uint32_t bloom_bits(uint32_t h) {
  uint32_t g = h * 0x9E3779B1u;
  return (1u << (g >> 27)) | (1u << ((g >> 22) & 31u)) | (1u << ((g >> 17) & 31u));
}

// Hash partitioning (RepartitionExec): 32-bit hash and multiply-shift range reduction, 7 integer instructions per row
// where `key_hash(k) % n` costs two 64-bit mixes and a 64-bit division.  An Int32 / Date32 key column and its packed
// Key128 form (lo = zero-extended key, hi = 0) hash alike, so both sides of a join co-locate whichever path packed
// them.  The destination uses the HIGH bits of the hash; the join table built behind the exchange indexes with the
// low bits, so the keys of one partition still spread over all its buckets.
__host__ __device__ __forceinline__ uint32_t part_hash(uint32_t k) { return fmix32(k); }
__host__ __device__ __forceinline__ uint32_t part_hash(const Key128& k) {
  return fmix32((uint32_t)k.lo ^ (uint32_t)(k.lo >> 32) * 0x85EBCA6Bu ^ (uint32_t)k.hi * 0xC2B2AE35u ^
                (uint32_t)(k.hi >> 32) * 0x27D4EB2Fu);
}
// Bucket of the 32-bit-key join table: the TOP bits of a second multiplicative mix of the key's hash.  Top bits, so
// that keys pre-sorted by the top bits of the same value fill the table slice by slice (join.cu builds large tables that
// way: each slice stays in L2 while it is filled); a second mix, because after an exchange all keys of a rank share the
// top bits of fmix32 itself (part_dest) and must still spread over the whole table.
__host__ __device__ __forceinline__ uint32_t table_hash(uint32_t h) { return h * 0x9E3779B1u; }
__host__ __device__ __forceinline__ int part_dest(uint32_t h, int n_out) { return (int)(((u64)h * (u64)(uint32_t)n_out) >> 32); }

// ---- streaming (read-once) loads: bypass L1 allocation ----
__device__ __forceinline__ double2 ld_stream_f64x2(const double* p) {
  double2 r;
  asm volatile("ld.global.nc.L1::no_allocate.v2.f64 {%0,%1}, [%2];" : "=d"(r.x), "=d"(r.y) : "l"(p));
  return r;
}
__device__ __forceinline__ int4 ld_stream_i32x4(const int32_t* p) {
  int4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.s32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}

// ---- atomics chosen for Blackwell: 128-bit CAS to claim a slot, result-less red.add to accumulate ----
__device__ __forceinline__ Key128 atomic_cas_key(Key128* addr, Key128 expect, Key128 desired) {
  Key128 old;
  asm volatile(
      "{\n .reg .b128 e, d, o;\n mov.b128 e, {%2, %3};\n mov.b128 d, {%4, %5};\n"
      " atom.global.relaxed.gpu.cas.b128 o, [%6], e, d;\n mov.b128 {%0, %1}, o;\n}\n"
      : "=l"(old.lo), "=l"(old.hi)
      : "l"(expect.lo), "l"(expect.hi), "l"(desired.lo), "l"(desired.hi), "l"(addr)
      : "memory");
  return old;
}
__device__ __forceinline__ void red_add_f64(double* p, double v) {
  asm volatile("red.global.add.f64 [%0], %1;" ::"l"(p), "d"(v) : "memory");
}
__device__ __forceinline__ void red_add_u64(u64* p, u64 v) {
  asm volatile("red.global.add.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ Key128 ld_key_volatile(const Key128* p) {
  Key128 k;
  asm volatile("ld.volatile.global.v2.u64 {%0,%1}, [%2];" : "=l"(k.lo), "=l"(k.hi) : "l"(p));
  return k;
}

// ---- group table in HBM: keys[cap] (EMPTY-initialised), acc[na][cap] f64 sums, cnt[nc][cap] u64 sums ----
struct TableRef {
  Key128* keys;
  double* acc;
  u64* cnt;
  u64 cap;        // power of two
  int* n_groups;  // number of claimed slots
  uint32_t* slot_list;  // slot_list[g] = slot of the g-th claimed group (claim order), or null: the emit step reads
                        // its groups from this list instead of compacting the table
};

// find or claim the slot of `key`; returns slot index, or ~0 when the table is full
__device__ __forceinline__ u64 table_upsert(const TableRef& t, const Key128& key) {
  u64 mask = t.cap - 1;
  u64 h = key_hash(key) & mask;
  const Key128 empty{KEY_EMPTY, KEY_EMPTY};
  for (u64 probe = 0; probe < t.cap; ++probe) {
    Key128 cur = ld_key_volatile(&t.keys[h]);
    if (key_eq(cur, key)) return h;
    if (key_eq(cur, empty)) {
      Key128 old = atomic_cas_key(&t.keys[h], empty, key);
      if (key_eq(old, empty)) {
        // one counter update per group of lanes that claimed together, not one same-address reduction per claim
        const unsigned together = __activemask();
        const int leader = __ffs(together) - 1, lane = threadIdx.x & 31;
        int base = 0;
        if (lane == leader) base = atomicAdd(t.n_groups, __popc(together));
        base = __shfl_sync(together, base, leader);
        if (t.slot_list) t.slot_list[base + __popc(together & ((1u << lane) - 1u))] = (uint32_t)h;
        return h;
      }
      if (key_eq(old, key)) return h;
    }
    h = (h + 1) & mask;
  }
  return ~0ull;
}
// the same without the group counter and the slot list: *claimed = this call took an empty slot.  For callers that
// count and list their claims themselves (agg_global_kernel stages a tile's claims in shared memory and reserves their
// places in the list with ONE update of the counter: at a million groups the per-claim form spends its time queueing on
// that one address in the L2).
__device__ __forceinline__ u64 table_upsert_quiet(const TableRef& t, const Key128& key, bool* claimed) {
  u64 mask = t.cap - 1;
  u64 h = key_hash(key) & mask;
  const Key128 empty{KEY_EMPTY, KEY_EMPTY};
  *claimed = false;
  for (u64 probe = 0; probe < t.cap; ++probe) {
    Key128 cur = ld_key_volatile(&t.keys[h]);
    if (key_eq(cur, key)) return h;
    if (key_eq(cur, empty)) {
      Key128 old = atomic_cas_key(&t.keys[h], empty, key);
      if (key_eq(old, empty)) {
        *claimed = true;
        return h;
      }
      if (key_eq(old, key)) return h;
    }
    h = (h + 1) & mask;
  }
  return ~0ull;
}
// read-only lookup; ~0 when absent
__device__ __forceinline__ u64 table_find(const Key128* keys, u64 cap, const Key128& key) {
  u64 mask = cap - 1;
  u64 h = key_hash(key) & mask;
  for (u64 probe = 0; probe < cap; ++probe) {
    Key128 cur = keys[h];
    if (key_eq(cur, key)) return h;
    if (cur.lo == KEY_EMPTY && cur.hi == KEY_EMPTY) return ~0ull;
    h = (h + 1) & mask;
  }
  return ~0ull;
}

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) v += __shfl_xor_sync(0xffffffffu, v, d);
  return v;
}
__device__ __forceinline__ u64 warp_sum(u64 v) {
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) v += __shfl_xor_sync(0xffffffffu, v, d);
  return v;
}

// ---- short strings held in registers: first <=15 bytes little-endian + length ----
template <class W> struct StrValT {  // W = u128 (<= 15 bytes) or u64 (<= 7 bytes: the narrow fused kernels)
  W bytes;
  int len;
};
using StrVal = StrValT<u128>;
__device__ __forceinline__ StrVal load_str(const int32_t* off, const uint8_t* chr, int64_t row) {
  int32_t b = off[row], e = off[row + 1];
  StrVal s;
  s.len = e - b;
  s.bytes = 0;
  int m = s.len < 15 ? s.len : 15;
  for (int q = 0; q < m; ++q) s.bytes |= (u128)chr[b + q] << (8 * q);
  return s;
}

}  // namespace b2
