// parquet.cu — GPU-side column decode of Parquet files: the leaf the reference reaches through DataFusion's ParquetExec
// (rust/core/src/serde/physical_plan/from_proto.rs:111-121: ParquetExec::try_from_files(filenames, projection, None,
// batch_size, num_partitions); written by the benchmark's `convert` command, rust/benchmarks/tpch/src/main.rs).
// with this the file's bytes cross PCIe once, still encoded (and still compressed), and every value is
// materialised in HBM by the GPU.
//
// The host reads only metadata (parquet_meta.cpp: the footer and one small Thrift header per page).  The column chunks of
// the requested row groups are copied to one device buffer as they lie in the file; three launches decode ALL pages of ALL
// projected columns at once, one CTA per page:
//   (1) pq_snappy_kernel   compressed pages -> scratch area of the same buffer (one warp per page; every lane parses the
//                          element tags, the copies are spread over the lanes)
//   (2) pq_dict_kernel     dictionary pages -> aligned value table / (position, length) table
//   (3) pq_data_kernel     data pages (V1 and V2): definition levels (RLE / bit-packed hybrid) -> validity bytes; values
//                          PLAIN or dictionary-coded (hybrid runs of bit-packed indices) -> output rows; pages with NULLs
//                          decode densely and are then spread over their rows in place, back to front
// Utf8 columns come out of (3) as (position, length) per row; a scan gives Arrow offsets and one kernel moves the bytes.
// Large PLAIN pages of fixed-width REQUIRED columns are cut into 64 KB slices on the host (values are independent), so a
// file of few, large pages still fills the machine.
#include <algorithm>
#include <cstring>

#include "ops.h"
#include "parquet.h"
#include "scan.cuh"

namespace b2 {
using namespace pq;

constexpr int PQ_THREADS = 256;
constexpr int64_t PQ_SLICE = 64 << 10;

struct PqPage {
  uint32_t src_off, src_len;    // the stored (possibly compressed) bytes of the value section
  uint32_t data_off, data_len;  // the uncompressed value section (== src when stored uncompressed); V1: levels first
  uint32_t lvl_off, lvl_len;    // V2 only: definition levels (never compressed)
  int32_t num_values;
  int32_t col;                  // index into the column table
  int64_t row_off;              // first output row
  int32_t dict;                 // dictionary of this page's chunk, -1 = none
  uint8_t type, encoding, compressed, slice;  // slice: values only (no level section even when the column is OPTIONAL)
};
struct PqDict {
  uint32_t off;  // byte offset into the dictionary buffer: entries of 8 bytes (value, or {uint32 position, int32 length})
  int32_t n;
};
struct PqCol {
  int32_t phys, dtype, max_def, width;
  void* out;
  uint8_t* valid;
  uint32_t* src;
  int32_t* len;
  unsigned long long* nulls;
};
enum { PQ_OK = 0, PQ_BAD_LEVELS = 1, PQ_BAD_VALUES = 2, PQ_BAD_INDEX = 3, PQ_BAD_SNAPPY = 4, PQ_BAD_DICT = 5 };

__device__ __forceinline__ void pq_fail(int* status, int code, int page) {
  if (atomicCAS(&status[0], 0, code) == 0) status[1] = page;
}
__device__ __forceinline__ uint32_t pq_le32(const uint8_t* p) {
  return (uint32_t)p[0] | (uint32_t)p[1] << 8 | (uint32_t)p[2] << 16 | (uint32_t)p[3] << 24;
}

// ---- (1) snappy: one warp per page (one CTA of 32 threads).  Format: varint uncompressed length, then elements tagged by
// their low two bits: 00 literal, 01 / 10 / 11 copy with a 1 / 2 / 4 byte offset.  The element stream is a chain (where an
// element starts depends on the one before), and the elements of dictionary pages are 2-8 bytes long, so a page decodes
// at the rate the chain is followed.  Three things keep that rate up:
//   * both ends of the chain live in shared memory: the compressed bytes are staged ahead of the parser in a 4 KB ring
//     (16-byte loads issued one refill early), the last 32 KB of output in a window ring that copies read from and that
//     goes to HBM in 16-byte stores;
//   * the chain is followed 32 input bytes at a time: every lane decodes the element that WOULD start at its byte, the
//     lanes that really start one are found by pointer doubling over those speculative "next" links (5 rounds of
//     shuffle + warp-OR), and a warp scan of their lengths gives every element its output position;
//   * the elements of one such batch are copied together: short ones that need nothing from the batch itself by their
//     own lane, the rest in stream order by the whole warp (an overlapping copy, offset < length, repeats its first
//     `offset` bytes, so lane i reads byte i mod offset).
// Literals with an explicit length (> 60 bytes) end a batch and are moved by the whole warp; above 1 KB straight from
// HBM to HBM in 16-byte pieces (incompressible pages are chains of 64 KB literals).
constexpr uint32_t SN_RING = 4096, SN_FLUSH = 2048, SN_LONG = 1024;
// The window size is a template parameter: the standard compressor works in 64 KB blocks, so copies reach back up to
// 64 KB; pages longer than SN_SMALL_PAGE get the 64 KB window (3 pages in flight per SM), the many short ones the 32 KB
// window (6 per SM), launched side by side on two streams.
constexpr uint32_t SN_WIN_SMALL = 32768, SN_WIN_LARGE = 65536, SN_SMALL_PAGE = 256 << 10;
constexpr uint32_t SN_BATCH_SPAN = 2048 + 64;  // a batch writes < 2 KB: positions within WIN - this of its start are still in the window

__device__ __forceinline__ uint4 sn_load16(const uint8_t* buf, uint32_t off, uint32_t buf_len) {
  return off + 16u <= buf_len ? *reinterpret_cast<const uint4*>(buf + off) : make_uint4(0, 0, 0, 0);
}
// window -> HBM for output positions [from, to); `out` is 16-byte aligned, so position i is aligned when i is
template <uint32_t SN_WIN>
__device__ __forceinline__ void sn_flush(const uint8_t* win, uint8_t* out, uint32_t from, uint32_t to, int lane) {
  uint32_t head = (from + 15u) & ~15u;
  if (head > to) head = to;
  for (uint32_t i = from + lane; i < head; i += 32) out[i] = win[i & (SN_WIN - 1)];
  const uint32_t mid = to & ~15u;
  for (uint32_t i = head + lane * 16; i < mid; i += 512)
    *reinterpret_cast<uint4*>(out + i) = *reinterpret_cast<const uint4*>(win + (i & (SN_WIN - 1)));
  const uint32_t tail = mid > head ? mid : head;
  for (uint32_t i = tail + lane; i < to; i += 32) out[i] = win[i & (SN_WIN - 1)];
}
// HBM -> HBM, any alignment on both sides: bytes up to a 16-byte boundary of dst, then aligned words + funnel shift
__device__ __forceinline__ void sn_copy_long(uint8_t* dst, const uint8_t* src, uint32_t len, int lane) {
  uint32_t head = (16u - (uint32_t)(reinterpret_cast<uintptr_t>(dst) & 15u)) & 15u;
  if (head > len) head = len;
  for (uint32_t i = lane; i < head; i += 32) dst[i] = src[i];
  const uint8_t* s = src + head;
  uint8_t* d = dst + head;
  const uint32_t n16 = (len - head) >> 4;
  const int shift = (int)(reinterpret_cast<uintptr_t>(s) & 3u) * 8;
  const uint32_t* sw = reinterpret_cast<const uint32_t*>(s - (shift >> 3));
#pragma unroll 4
  for (uint32_t k = lane; k < n16; k += 32) {
    const uint32_t* w = sw + 4 * k;
    const uint32_t w0 = w[0], w1 = w[1], w2 = w[2], w3 = w[3], w4 = w[4];
    uint4 v;
    v.x = __funnelshift_r(w0, w1, shift);
    v.y = __funnelshift_r(w1, w2, shift);
    v.z = __funnelshift_r(w2, w3, shift);
    v.w = __funnelshift_r(w3, w4, shift);
    *reinterpret_cast<uint4*>(d + 16 * k) = v;
  }
  for (uint32_t i = head + 16 * n16 + lane; i < len; i += 32) dst[i] = src[i];
}
// output position p as seen from a batch that starts at `pos`: recent positions are in the window, older ones (and
// everything before `win_from`, the end of the last literal that bypassed the window) have reached HBM
template <uint32_t SN_WIN>
__device__ __forceinline__ uint8_t sn_read(const uint8_t* win, const uint8_t* out, uint32_t p, uint32_t pos, uint32_t win_from) {
  return (p >= win_from && (int32_t)(pos - p) <= (int32_t)(SN_WIN - SN_BATCH_SPAN)) ? win[p & (SN_WIN - 1)] : out[p];
}

template <uint32_t SN_WIN>
__global__ void __launch_bounds__(32)
pq_snappy_kernel(uint8_t* buf, uint32_t buf_len, const PqPage* __restrict__ pages, const int* __restrict__ list, int* status) {
  extern __shared__ __align__(16) uint8_t sn_smem[];
  uint8_t* ring = sn_smem;
  uint8_t* win = sn_smem + SN_RING;
  constexpr unsigned FULL = 0xffffffffu;
  constexpr uint32_t RM = SN_RING - 1, WM = SN_WIN - 1;
  const int pi = list[blockIdx.x];
  const PqPage pg = pages[pi];
  const uint32_t lane = threadIdx.x;
  uint32_t in = pg.src_off;
  const uint32_t in_end = in + pg.src_len;
  uint8_t* out = buf + pg.data_off;
  const uint32_t out_len = pg.data_len;
  // input staging: ring[a & RM] = buf[a] for a in [in, staged)
  uint32_t staged = in & ~15u;
  uint4 pend = sn_load16(buf, staged + lane * 16, buf_len);
  auto stage = [&]() {
    if (in > staged) {  // a long literal was copied straight from HBM: restart behind it
      staged = in & ~15u;
      pend = sn_load16(buf, staged + lane * 16, buf_len);
    }
    while (staged < in_end && (int32_t)(staged - in) <= (int32_t)(SN_RING - 512)) {
      *reinterpret_cast<uint4*>(ring + ((staged + lane * 16) & RM)) = pend;
      staged += 512;
      pend = sn_load16(buf, staged + lane * 16, buf_len);
    }
    __syncwarp();
  };
  stage();
  uint32_t ulen = 0;
  for (int sh = 0;; sh += 7) {
    if (in >= in_end || sh > 28) return pq_fail(status, PQ_BAD_SNAPPY, pi);
    uint32_t b = ring[in & RM];
    ++in;
    ulen |= (b & 0x7F) << sh;
    if (!(b & 0x80)) break;
  }
  if (ulen != out_len) return pq_fail(status, PQ_BAD_SNAPPY, pi);
  uint32_t pos = 0, flushed = 0, win_from = 0;
  while (in < in_end) {
    stage();
    if (pos - flushed >= SN_FLUSH) {
      const uint32_t to = pos & ~15u;
      sn_flush<SN_WIN>(win, out, flushed, to, lane);
      flushed = to;
    }
    // every lane decodes the element that would start at its byte
    const uint32_t a = in + lane;
    const uint32_t t = ring[a & RM];
    const uint32_t b1 = ring[(a + 1) & RM], b2 = ring[(a + 2) & RM], b3 = ring[(a + 3) & RM], b4 = ring[(a + 4) & RM];
    // (no branches: the 32 lanes hold all four kinds of element)
    const uint32_t kind = t & 3, t2 = t >> 2;
    const uint32_t w4 = b1 | b2 << 8 | b3 << 16 | b4 << 24;
    const bool lit = kind == 0;
    const bool big = lit && t2 >= 60;  // explicit length, stored minus one in the next 1-4 bytes: kept that way until checked
    const uint32_t nb = t2 - 59;
    const uint32_t len = lit ? (big ? (nb < 4 ? w4 & ((1u << (8 * nb)) - 1u) : w4) : t2 + 1) : (kind == 1 ? 4 + (t2 & 7) : t2 + 1);
    const uint32_t hdr = lit ? (big ? 1 + nb : 1) : (kind == 3 ? 5 : kind + 1);
    const uint32_t off = lit ? 0 : (kind == 1 ? ((t >> 5) << 8 | b1) : (kind == 2 ? (w4 & 0xFFFFu) : w4));
    if (__shfl_sync(FULL, (int)big, 0)) {
      // a long literal at the head of the stream: the whole warp moves it
      uint32_t l = __shfl_sync(FULL, len, 0);
      const uint32_t h = __shfl_sync(FULL, hdr, 0), avail = in_end - in;
      if (h > avail || l >= avail - h || l >= out_len - pos) return pq_fail(status, PQ_BAD_SNAPPY, pi);
      l += 1;
      const uint32_t src = in + h;
      if (l > SN_LONG) {
        sn_flush<SN_WIN>(win, out, flushed, pos, lane);
        sn_copy_long(out + pos, buf + src, l, lane);
        flushed = win_from = pos + l;
      } else if (src + l <= staged) {
        for (uint32_t i = lane; i < l; i += 32) win[(pos + i) & WM] = ring[(src + i) & RM];
      } else {
        for (uint32_t i = lane; i < l; i += 32) win[(pos + i) & WM] = buf[src + i];
      }
      in = src + l;
      pos += l;
      __syncwarp();
      continue;
    }
    // which lanes start an element: reachability from lane 0 over the "next element" links, by pointer doubling.
    // An element takes at least 2 input bytes, so at most 15 links lie inside the 32 bytes: 4 rounds (1 + 2 + 4 + 8).
    const bool valid = a < in_end;
    const uint32_t adv = hdr + (lit && !big ? len : 0);
    uint32_t jump = valid && !big ? min(lane + adv, 32u) : 32u;
    uint32_t reach = 1;
#pragma unroll
    for (int r = 0; r < 4; ++r) {
      reach |= __reduce_or_sync(FULL, ((reach >> lane) & 1u) && jump < 32u ? 1u << jump : 0u);
      const uint32_t j2 = __shfl_sync(FULL, jump, jump & 31u);
      jump = jump < 32u ? j2 : 32u;
    }
    reach &= __ballot_sync(FULL, valid);
    const uint32_t bigmask = __ballot_sync(FULL, big) & reach;
    uint32_t exit_rel;
    if (bigmask) {  // the batch ends in front of the first long literal (never lane 0 here)
      exit_rel = __ffs(bigmask) - 1;
      reach &= (1u << exit_rel) - 1u;
    } else {
      exit_rel = __shfl_sync(FULL, lane + adv, 31 - __clz(reach));
    }
    const bool is_elem = (reach >> lane) & 1u;
    uint32_t incl = is_elem ? len : 0u;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      uint32_t v = __shfl_up_sync(FULL, incl, d);
      if ((int)lane >= d) incl += v;
    }
    const uint32_t tot = __shfl_sync(FULL, incl, 31);
    const uint32_t d0 = pos + incl - (is_elem ? len : 0u);  // this element's output position
    const bool err = is_elem && (hdr > in_end - a || (lit ? len > in_end - a - hdr : (off == 0 || off > d0)));
    if (__any_sync(FULL, err) || tot > out_len - pos) return pq_fail(status, PQ_BAD_SNAPPY, pi);
    const uint32_t from = d0 - off;
    // Short elements that read nothing this batch writes (and do not wrap in either ring): each by its own lane, all at
    // once, up to 12 independent byte loads then stores.  Source as an offset into sn_smem: input ring or window.
    const uint32_t dst_at = SN_RING + (d0 & WM);
    const uint32_t src_at = lit ? ((a + hdr) & RM) : SN_RING + (from & WM);
    const bool alone = is_elem && len <= 12 && dst_at + len <= SN_RING + SN_WIN &&
                       (lit ? src_at + len <= SN_RING
                            : (off >= len && from + len <= pos && from >= win_from &&
                               pos - from <= SN_WIN - SN_BATCH_SPAN && src_at + len <= SN_RING + SN_WIN));
    if (alone) {
      uint8_t v[12];
#pragma unroll
      for (int i = 0; i < 12; ++i)
        if (i < (int)len) v[i] = sn_smem[src_at + i];
#pragma unroll
      for (int i = 0; i < 12; ++i)
        if (i < (int)len) sn_smem[dst_at + i] = v[i];
    }
    __syncwarp();
    // the others in stream order, each by the whole warp
    uint32_t rest = __ballot_sync(FULL, is_elem && !alone);
    const uint32_t meta = len | (lit ? 0x100u : 0u), what = lit ? a + hdr : off;
    while (rest) {
      const int e = __ffs(rest) - 1;
      rest &= rest - 1;
      const uint32_t em = __shfl_sync(FULL, meta, e), ed = __shfl_sync(FULL, d0, e), ew = __shfl_sync(FULL, what, e);
      const uint32_t el = em & 0xFFu;
      if (em & 0x100u) {
        for (uint32_t i = lane; i < el; i += 32) win[(ed + i) & WM] = ring[(ew + i) & RM];
      } else {
        const uint32_t ef = ed - ew;
        if (ew >= el) {
          for (uint32_t i = lane; i < el; i += 32) win[(ed + i) & WM] = sn_read<SN_WIN>(win, out, ef + i, pos, win_from);
        } else {
          for (uint32_t i = lane; i < el; i += 32) win[(ed + i) & WM] = sn_read<SN_WIN>(win, out, ef + i % ew, pos, win_from);
        }
      }
      __syncwarp();
    }
    pos += tot;
    in += exit_rel;
  }
  __syncwarp();
  sn_flush<SN_WIN>(win, out, flushed, pos, lane);
  if (pos != out_len) pq_fail(status, PQ_BAD_SNAPPY, pi);
}

// ---- RLE / bit-packed hybrid (levels and dictionary indices): <varint header> then, header & 1: (header >> 1) groups of
// 8 values, `bw` bits each, LSB first; else: one value in ceil(bw / 8) bytes repeated (header >> 1) times.  Every thread
// walks the run headers (uniform, a few bytes per run); the values of a run are spread over the CTA.  No barrier inside.
template <class F>
__device__ bool pq_hybrid(const uint8_t* p, const uint8_t* end, int bw, int64_t count, F emit) {
  const int tid = threadIdx.x, nt = blockDim.x;
  const uint32_t mask = bw >= 32 ? 0xFFFFFFFFu : ((1u << bw) - 1u);
  int64_t done = 0;
  while (done < count) {
    uint32_t h = 0;
    for (int sh = 0;; sh += 7) {
      if (p >= end || sh > 28) return false;
      uint8_t b = *p++;
      h |= (uint32_t)(b & 0x7F) << sh;
      if (!(b & 0x80)) break;
    }
    const int64_t run = h >> 1;
    if (run == 0) return false;
    if (h & 1) {
      int64_t take = run * 8 < count - done ? run * 8 : count - done;
      int64_t have = end - p;
      if ((take * bw + 7) / 8 > have) return false;  // the padding of a final group may be cut short, its values not
      for (int64_t k = tid; k < take; k += nt) {
        int64_t bit = k * bw;
        const uint8_t* q = p + (bit >> 3);
        int s = (int)(bit & 7), nb = (s + bw + 7) >> 3;
        uint64_t w = 0;
        for (int b = 0; b < nb; ++b) w |= (uint64_t)q[b] << (8 * b);
        emit(done + k, (uint32_t)(w >> s) & mask);
      }
      int64_t nbytes = run * bw;
      p += nbytes < have ? nbytes : have;
      done += take;
    } else {
      int vb = (bw + 7) >> 3;
      if (vb > end - p) return false;
      uint32_t v = 0;
      for (int b = 0; b < vb; ++b) v |= (uint32_t)p[b] << (8 * b);
      p += vb;
      v &= mask;
      int64_t take = run < count - done ? run : count - done;
      for (int64_t k = tid; k < take; k += nt) emit(done + k, v);
      done += take;
    }
  }
  return true;
}

// exclusive prefix of one flag per thread over the CTA (PQ_THREADS threads); *total = number of set flags
__device__ __forceinline__ int pq_block_prefix(bool flag, int* warp_sums, int* total) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  unsigned m = __ballot_sync(0xffffffffu, flag);
  if (lane == 0) warp_sums[warp] = __popc(m);
  __syncthreads();
  int pre = __popc(m & ((1u << lane) - 1u)), t = 0;
  for (int w = 0; w < PQ_THREADS / 32; ++w) {
    int s = warp_sums[w];
    if (w < warp) pre += s;
    t += s;
  }
  *total = t;
  __syncthreads();
  return pre;
}

// a[0, non_null) holds the values of the valid rows in order: spread them over rows [0, n), NULL rows get 0.
// Tiles of PQ_THREADS rows from the back: a row's value sits at an index <= the row, and every tile reads before it writes.
template <class T>
__device__ void pq_spread(T* a, const uint8_t* valid, int64_t n, int64_t non_null, int* warp_sums) {
  int64_t running = non_null;
  for (int64_t t = (n + PQ_THREADS - 1) / PQ_THREADS - 1; t >= 0; --t) {
    int64_t row = t * PQ_THREADS + threadIdx.x;
    bool v = row < n && valid[row];
    int total;
    int pre = pq_block_prefix(v, warp_sums, &total);
    int64_t base = running - total;
    T x = v ? a[base + pre] : T(0);
    __syncthreads();
    if (row < n) a[row] = x;
    running = base;
    __syncthreads();
  }
}

// n bytes (a multiple of 4) from any address to a 4-byte aligned one: aligned word loads + funnel shift
__device__ __forceinline__ void pq_copy_words(uint32_t* dst, const uint8_t* src, int64_t nbytes) {
  const int shift = (int)(reinterpret_cast<uintptr_t>(src) & 3);
  const uint32_t* base = reinterpret_cast<const uint32_t*>(src - shift);
  const int64_t words = nbytes >> 2;
  if (shift == 0) {
    for (int64_t i = threadIdx.x; i < words; i += blockDim.x) dst[i] = base[i];
  } else {
    for (int64_t i = threadIdx.x; i < words; i += blockDim.x) dst[i] = __funnelshift_r(base[i], base[i + 1], shift * 8);
  }
}

// ---- (2) dictionary pages (PLAIN): fixed width -> aligned 8-byte entries; BYTE_ARRAY -> {position, length} entries
__global__ void pq_dict_kernel(const uint8_t* __restrict__ buf, const PqPage* __restrict__ pages, const int* __restrict__ list,
                               const PqCol* __restrict__ cols, const PqDict* __restrict__ dicts, uint8_t* dictbuf, int* status) {
  const int pi = list[blockIdx.x];
  const PqPage pg = pages[pi];
  const PqCol col = cols[pg.col];
  const PqDict d = dicts[pg.dict];
  const uint8_t* p = buf + pg.data_off;
  const uint8_t* end = p + pg.data_len;
  if (col.phys == PT_BYTE_ARRAY) {
    if (threadIdx.x) return;
    uint2* e = reinterpret_cast<uint2*>(dictbuf + d.off);
    for (int i = 0; i < d.n; ++i) {
      if (end - p < 4) return pq_fail(status, PQ_BAD_DICT, pi);
      uint32_t l = pq_le32(p);
      p += 4;
      if (l > (uint32_t)(end - p) || l > 0x7FFFFFFFu) return pq_fail(status, PQ_BAD_DICT, pi);
      e[i] = make_uint2((uint32_t)(p - buf), l);
      p += l;
    }
    return;
  }
  const int w = col.phys == PT_INT32 ? 4 : 8;
  if ((int64_t)d.n * w > end - p) {
    if (threadIdx.x == 0) pq_fail(status, PQ_BAD_DICT, pi);
    return;
  }
  unsigned long long* e = reinterpret_cast<unsigned long long*>(dictbuf + d.off);
  for (int i = threadIdx.x; i < d.n; i += blockDim.x) {
    unsigned long long v = 0;
    for (int b = 0; b < w; ++b) v |= (unsigned long long)p[(int64_t)i * w + b] << (8 * b);
    e[i] = v;
  }
}

// ---- (3) data pages.  One CTA per page (or per 64 KB slice of a PLAIN fixed-width page).
__global__ void __launch_bounds__(PQ_THREADS)
pq_data_kernel(const uint8_t* __restrict__ buf, const PqPage* __restrict__ pages, const int* __restrict__ list,
               const PqCol* __restrict__ cols, const PqDict* __restrict__ dicts, const uint8_t* __restrict__ dictbuf,
               int* status) {
  __shared__ int warp_sums[PQ_THREADS / 32];
  __shared__ long long s_count;
  __shared__ int s_bad;
  const int pi = list[blockIdx.x];
  const PqPage pg = pages[pi];
  const PqCol col = cols[pg.col];
  const int tid = threadIdx.x;
  const uint8_t* p = buf + pg.data_off;
  const uint8_t* end = p + pg.data_len;
  const int64_t n = pg.num_values;
  int64_t non_null = n;
  if (tid == 0) {
    s_count = 0;
    s_bad = 0;
  }
  __syncthreads();
  // definition levels -> validity bytes (flat schema: max level 1, one bit per value)
  if (col.max_def > 0 && !pg.slice) {
    const uint8_t *lp, *lend;
    if (pg.type == PAGE_DATA_V2) {
      lp = buf + pg.lvl_off;
      lend = lp + pg.lvl_len;
    } else {
      if (end - p < 4) {
        if (tid == 0) pq_fail(status, PQ_BAD_LEVELS, pi);
        return;
      }
      uint32_t l = pq_le32(p);
      p += 4;
      if (l > (uint32_t)(end - p)) {
        if (tid == 0) pq_fail(status, PQ_BAD_LEVELS, pi);
        return;
      }
      lp = p;
      lend = p + l;
      p = lend;
    }
    uint8_t* valid = col.valid + pg.row_off;
    if (!pq_hybrid(lp, lend, 1, n, [&](int64_t i, uint32_t v) { valid[i] = (uint8_t)v; })) {
      if (tid == 0) pq_fail(status, PQ_BAD_LEVELS, pi);
      return;
    }
    __syncthreads();
    long long c = 0;
    for (int64_t i = tid; i < n; i += PQ_THREADS) c += valid[i];
    for (int d = 16; d > 0; d >>= 1) c += __shfl_xor_sync(0xffffffffu, c, d);
    if ((tid & 31) == 0 && c) atomicAdd(reinterpret_cast<unsigned long long*>(&s_count), (unsigned long long)c);
    __syncthreads();
    non_null = s_count;
    if (tid == 0 && non_null < n) atomicAdd(col.nulls, (unsigned long long)(n - non_null));
  }
  // values, densely: the i-th stored value goes to row_off + i
  const int64_t r0 = pg.row_off;
  bool ok = true;
  const bool dict_coded = pg.encoding == ENC_PLAIN_DICT || pg.encoding == ENC_RLE_DICT;
  if (non_null == 0) {
    // nothing stored
  } else if (dict_coded) {
    if (p >= end || pg.dict < 0) {
      ok = false;
    } else {
      const int bw = *p++;
      const PqDict d = dicts[pg.dict];
      if (bw > 32) {
        ok = false;
      } else if (col.phys == PT_BYTE_ARRAY) {
        const uint2* e = reinterpret_cast<const uint2*>(dictbuf + d.off);
        ok = pq_hybrid(p, end, bw, non_null, [&](int64_t i, uint32_t v) {
          if (v >= (uint32_t)d.n) {
            s_bad = 1;
            v = 0;
          }
          uint2 x = e[v];
          col.src[r0 + i] = x.x;
          col.len[r0 + i] = (int32_t)x.y;
        });
      } else if (col.width == 4) {
        const unsigned long long* e = reinterpret_cast<const unsigned long long*>(dictbuf + d.off);
        uint32_t* out = static_cast<uint32_t*>(col.out) + r0;
        ok = pq_hybrid(p, end, bw, non_null, [&](int64_t i, uint32_t v) {
          if (v >= (uint32_t)d.n) {
            s_bad = 1;
            v = 0;
          }
          out[i] = (uint32_t)e[v];
        });
      } else {
        const unsigned long long* e = reinterpret_cast<const unsigned long long*>(dictbuf + d.off);
        unsigned long long* out = static_cast<unsigned long long*>(col.out) + r0;
        ok = pq_hybrid(p, end, bw, non_null, [&](int64_t i, uint32_t v) {
          if (v >= (uint32_t)d.n) {
            s_bad = 1;
            v = 0;
          }
          out[i] = e[v];
        });
      }
    }
  } else if (col.phys == PT_BOOLEAN) {
    uint8_t* out = static_cast<uint8_t*>(col.out) + r0;
    if (pg.encoding == ENC_RLE) {  // data page V2 booleans: <uint32 length> hybrid runs of 1-bit values
      if (end - p < 4) {
        ok = false;
      } else {
        uint32_t l = pq_le32(p);
        p += 4;
        ok = l <= (uint32_t)(end - p) && pq_hybrid(p, p + l, 1, non_null, [&](int64_t i, uint32_t v) { out[i] = (uint8_t)v; });
      }
    } else if ((non_null + 7) / 8 > end - p) {
      ok = false;
    } else {
      for (int64_t i = tid; i < non_null; i += PQ_THREADS) out[i] = (p[i >> 3] >> (i & 7)) & 1;
    }
  } else if (col.phys == PT_BYTE_ARRAY) {
    // PLAIN: <uint32 length> bytes, back to back: the positions form a chain, walked by one thread
    if (tid == 0) {
      for (int64_t i = 0; i < non_null; ++i) {
        if (end - p < 4) {
          s_bad = 2;
          break;
        }
        uint32_t l = pq_le32(p);
        p += 4;
        if (l > (uint32_t)(end - p)) {
          s_bad = 2;
          break;
        }
        col.src[r0 + i] = (uint32_t)(p - buf);
        col.len[r0 + i] = (int32_t)l;
        p += l;
      }
    }
  } else {
    const int w = col.width;
    if (non_null * w > end - p) ok = false;
    else pq_copy_words(reinterpret_cast<uint32_t*>(static_cast<uint8_t*>(col.out) + r0 * w), p, non_null * w);
  }
  __syncthreads();
  if (!ok || s_bad) {
    if (tid == 0) pq_fail(status, !ok || s_bad == 2 ? PQ_BAD_VALUES : PQ_BAD_INDEX, pi);
    return;
  }
  if (non_null == n) return;
  // NULLs: spread the dense values over their rows
  const uint8_t* valid = col.valid + r0;
  if (col.phys == PT_BYTE_ARRAY) {
    pq_spread(col.src + r0, valid, n, non_null, warp_sums);
    pq_spread(col.len + r0, valid, n, non_null, warp_sums);
  } else if (col.width == 1) {
    pq_spread(static_cast<uint8_t*>(col.out) + r0, valid, n, non_null, warp_sums);
  } else if (col.width == 4) {
    pq_spread(static_cast<uint32_t*>(col.out) + r0, valid, n, non_null, warp_sums);
  } else {
    pq_spread(static_cast<unsigned long long*>(col.out) + r0, valid, n, non_null, warp_sums);
  }
}

struct PqLenIn {
  const int32_t* v;
  __device__ __forceinline__ int operator()(int64_t i) const { return v[i]; }
};
struct PqOffsetsOut {
  int32_t* off;
  int64_t n;
  __device__ __forceinline__ void operator()(int64_t i, int v, int64_t pre) const {
    off[i] = (int32_t)pre;
    if (i == n - 1) off[n] = (int32_t)(pre + v);
  }
};
// Utf8: a half-warp per row moves the bytes (TPC-H strings are 1-79 bytes; one lane per row would serialise on the long ones)
__global__ void pq_copy_chars_kernel(const uint8_t* __restrict__ buf, const uint32_t* __restrict__ src,
                                     const int32_t* __restrict__ off, int64_t n, uint8_t* __restrict__ chr) {
  const int sub = threadIdx.x & 15;
  int64_t r = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 4;
  const int64_t stride = ((int64_t)gridDim.x * blockDim.x) >> 4;
  for (; r < n; r += stride) {
    int32_t o = off[r], len = off[r + 1] - o;
    const uint8_t* s = buf + src[r];
    for (int q = sub; q < len; q += 16) chr[o + q] = s[q];
  }
}

namespace {
inline uint32_t align16(uint64_t v) { return (uint32_t)((v + 15) & ~(uint64_t)15); }
}  // namespace

BatchPtr parquet_decode(Ctx* ctx, const uint8_t* file, size_t len, const std::vector<int>& projection, int rg_begin,
                        int rg_end) {
  FileMeta meta = parse_footer(file, len);
  Schema schema = arrow_schema(meta, projection);  // refuses nested schemas and unsupported column types
  const int n_rg = (int)meta.row_groups.size();
  if (rg_end < 0) rg_end = n_rg;
  if (rg_begin < 0 || rg_begin > rg_end || rg_end > n_rg) fail(B2_ERR_INVALID_ARG, "Parquet row group range out of bounds");
  int64_t rows = 0;
  for (int g = rg_begin; g < rg_end; ++g) rows += meta.row_groups[g].num_rows;
  if (rows >= (1ll << 31) - 1) fail(B2_ERR_NOT_IMPLEMENTED, "Parquet: more than 2^31-2 rows in one decode call (pass fewer row groups)");
  if (rows == 0 || projection.empty()) {
    auto e = empty_batch(ctx, schema);
    if (projection.empty()) e->num_rows = rows;
    return e;
  }
  const int ncols = (int)projection.size();

  // ---- page table (host): walk the page headers of every projected chunk; lay the chunks, then the scratch area, out
  std::vector<PqPage> pages;
  std::vector<PqDict> dicts;
  std::vector<int> snappy_list, dict_list, data_list;
  struct Copy {
    uint64_t file_off, bytes;
    uint32_t dev_off;
  };
  std::vector<Copy> copies;
  uint64_t dev_bytes = 0, dict_bytes = 0;
  std::vector<int> need_scratch;  // compressed pages get their scratch offsets after all chunks are placed
  int64_t row_base = 0;
  for (int g = rg_begin; g < rg_end; ++g) {
    const RowGroupMeta& rg = meta.row_groups[g];
    for (int c = 0; c < ncols; ++c) {
      const ChunkMeta& ck = rg.cols[projection[c]];
      const LeafColumn& leaf = meta.columns[projection[c]];
      if (ck.num_values != rg.num_rows) fail(B2_ERR_ARROW, "Parquet: a flat column chunk whose num_values differs from the row count");
      if (ck.codec != CODEC_NONE && ck.codec != CODEC_SNAPPY)
        fail(B2_ERR_NOT_IMPLEMENTED, std::string("Parquet: ") + codec_name(ck.codec) + " compressed column '" + leaf.name +
                                         "' (UNCOMPRESSED and SNAPPY are decoded)");
      if (ck.num_values == 0) continue;
      auto infos = walk_pages(file, len, ck);
      uint32_t chunk_dev = align16(dev_bytes);
      dev_bytes = (uint64_t)chunk_dev + (uint64_t)ck.total_compressed;
      if (dev_bytes >= (1ull << 32) - 64) fail(B2_ERR_NOT_IMPLEMENTED, "Parquet: 4 GiB or more of column chunks in one decode call (pass fewer row groups)");
      copies.push_back({(uint64_t)ck.start(), (uint64_t)ck.total_compressed, chunk_dev});
      int dict_id = -1;
      int64_t row = row_base;
      for (const PageInfo& pi : infos) {
        PqPage pg;
        memset(&pg, 0, sizeof pg);
        pg.type = (uint8_t)pi.type;
        pg.encoding = (uint8_t)pi.encoding;
        pg.col = c;
        pg.num_values = pi.num_values;
        pg.dict = -1;
        uint32_t at = chunk_dev + (uint32_t)(pi.payload - ck.start());
        uint32_t lv = pi.type == PAGE_DATA_V2 ? (uint32_t)(pi.def_len + pi.rep_len) : 0;
        if (pi.type == PAGE_DATA_V2 && pi.rep_len) fail(B2_ERR_NOT_IMPLEMENTED, "Parquet: repetition levels (nested data)");
        pg.lvl_off = at;
        pg.lvl_len = pi.type == PAGE_DATA_V2 ? (uint32_t)pi.def_len : 0;
        pg.src_off = at + lv;
        pg.src_len = (uint32_t)pi.comp_size - lv;
        pg.data_len = (uint32_t)pi.uncomp_size - lv;
        pg.compressed = ck.codec != CODEC_NONE && (pi.type != PAGE_DATA_V2 || pi.v2_compressed) && pg.data_len > 0;
        pg.data_off = pg.src_off;
        if (!pg.compressed && pg.data_len && pg.src_len != pg.data_len)
          fail(B2_ERR_ARROW, "Parquet: an uncompressed page whose two sizes differ");
        const int idx = (int)pages.size();
        if (pi.type == PAGE_DICT) {
          if (dict_id >= 0) fail(B2_ERR_ARROW, "Parquet: two dictionary pages in one column chunk");
          if (pi.encoding != ENC_PLAIN && pi.encoding != ENC_PLAIN_DICT)
            fail(B2_ERR_NOT_IMPLEMENTED, std::string("Parquet: dictionary page encoded as ") + encoding_name(pi.encoding));
          if (leaf.phys == PT_BOOLEAN) fail(B2_ERR_ARROW, "Parquet: dictionary page in a BOOLEAN column");
          dict_id = (int)dicts.size();
          dicts.push_back({(uint32_t)dict_bytes, pi.num_values});
          dict_bytes += (uint64_t)pi.num_values * 8;
          pg.dict = dict_id;
          pages.push_back(pg);
          dict_list.push_back(idx);
          if (pg.compressed) need_scratch.push_back(idx);
          continue;
        }
        // data page
        const bool dict_coded = pi.encoding == ENC_PLAIN_DICT || pi.encoding == ENC_RLE_DICT;
        if (dict_coded) {
          if (dict_id < 0) fail(B2_ERR_ARROW, "Parquet: dictionary-coded page without a dictionary page before it");
        } else if (!(pi.encoding == ENC_PLAIN || (pi.encoding == ENC_RLE && leaf.phys == PT_BOOLEAN))) {
          fail(B2_ERR_NOT_IMPLEMENTED, std::string("Parquet: ") + encoding_name(pi.encoding) + " encoded pages in column '" +
                                           leaf.name + "' (PLAIN, PLAIN_DICTIONARY and RLE_DICTIONARY are decoded)");
        }
        if (leaf.repetition == REP_OPTIONAL && pi.type == PAGE_DATA && pi.def_encoding != ENC_RLE)
          fail(B2_ERR_NOT_IMPLEMENTED, std::string("Parquet: definition levels encoded as ") + encoding_name(pi.def_encoding));
        pg.dict = dict_id;
        pg.row_off = row;
        if (row + pi.num_values > row_base + rg.num_rows) fail(B2_ERR_ARROW, "Parquet: pages hold more values than the row group has rows");
        const int w = leaf.phys == PT_INT32 ? 4 : (leaf.phys == PT_INT64 || leaf.phys == PT_DOUBLE) ? 8 : 0;
        const bool cut = w && pi.encoding == ENC_PLAIN && leaf.repetition == REP_REQUIRED && (int64_t)pg.data_len > 2 * PQ_SLICE;
        if (pg.compressed) need_scratch.push_back(idx);
        if (cut) {
          // the page itself is only decompressed (if needed); its values are decoded by independent slices
          pages.push_back(pg);
          const int64_t per = PQ_SLICE / w;
          for (int64_t v0 = 0; v0 < pi.num_values; v0 += per) {
            PqPage sl = pg;
            sl.slice = 1;
            sl.compressed = 0;
            sl.dict = -(idx + 2);  // parent page: data_off is patched from it once the scratch area is laid out
            sl.num_values = (int32_t)std::min<int64_t>(per, pi.num_values - v0);
            sl.row_off = row + v0;
            sl.src_off = (uint32_t)(v0 * w);  // relative to the parent's data for now
            sl.data_len = (uint32_t)(sl.num_values * w);
            data_list.push_back((int)pages.size());
            pages.push_back(sl);
          }
        } else {
          pages.push_back(pg);
          data_list.push_back(idx);
        }
        row += pi.num_values;
      }
      if (row != row_base + rg.num_rows) fail(B2_ERR_ARROW, "Parquet: pages hold fewer values than the row group has rows");
    }
    row_base += rg.num_rows;
  }
  // scratch area for the decompressed pages, behind the chunks
  for (int idx : need_scratch) {
    PqPage& pg = pages[idx];
    uint32_t at = align16(dev_bytes);
    dev_bytes = (uint64_t)at + pg.data_len;
    if (dev_bytes >= (1ull << 32) - 64) fail(B2_ERR_NOT_IMPLEMENTED, "Parquet: 4 GiB or more of decompressed pages in one decode call (pass fewer row groups)");
    pg.data_off = at;
    snappy_list.push_back(idx);
  }
  // largest first: a page is one chain of elements, so the longest ones decide when the launch ends
  std::stable_sort(snappy_list.begin(), snappy_list.end(), [&](int a, int b) { return pages[a].data_len > pages[b].data_len; });
  for (PqPage& pg : pages) {
    if (!pg.slice) continue;
    const PqPage& parent = pages[-pg.dict - 2];
    pg.data_off = parent.data_off + pg.src_off;
    pg.src_off = pg.data_off;
    pg.src_len = pg.data_len;
    pg.dict = -1;
  }

  // ---- device: the chunks as they lie in the file, the tables, the output columns
  auto buf = dev_alloc(ctx, (size_t)dev_bytes + 64);
  for (const Copy& cp : copies)
    B2_CUDA(cudaMemcpyAsync(buf->as<uint8_t>() + cp.dev_off, file + cp.file_off, (size_t)cp.bytes, cudaMemcpyHostToDevice, ctx->stream));
  auto out = std::make_shared<Batch>();
  out->ctx = ctx;
  out->schema = schema;
  out->num_rows = rows;
  out->cols.resize(ncols);
  std::vector<PqCol> cols(ncols);
  std::vector<BufPtr> srcs(ncols), lens(ncols);
  auto counters = dev_alloc_zero(ctx, (size_t)(ncols + 1) * 8);  // [0] = status (code, page), then NULLs per column
  for (int c = 0; c < ncols; ++c) {
    const LeafColumn& leaf = meta.columns[projection[c]];
    Column& col = out->cols[c];
    col.dtype = leaf.dtype;
    col.length = rows;
    PqCol& k = cols[c];
    memset(&k, 0, sizeof k);
    k.phys = leaf.phys;
    k.dtype = leaf.dtype;
    k.max_def = leaf.repetition == REP_OPTIONAL ? 1 : 0;
    k.width = (int)dtype_width(leaf.dtype);
    k.nulls = counters->as<unsigned long long>() + 1 + c;
    if (leaf.dtype == B2_UTF8) {
      srcs[c] = dev_alloc(ctx, (size_t)(rows + 1) * 4);
      lens[c] = dev_alloc(ctx, (size_t)(rows + 1) * 4);
      k.src = srcs[c]->as<uint32_t>();
      k.len = lens[c]->as<int32_t>();
    } else {
      col.data = dev_alloc(ctx, (size_t)rows * k.width + 16);
      k.out = col.data->ptr;
    }
    if (k.max_def) {
      col.valid = dev_alloc(ctx, (size_t)rows + 16);
      k.valid = col.valid->as<uint8_t>();
    }
  }
  auto upload = [&](const void* host, size_t bytes) {
    auto d = dev_alloc(ctx, bytes + 16);
    if (bytes) B2_CUDA(cudaMemcpyAsync(d->ptr, host, bytes, cudaMemcpyHostToDevice, ctx->stream));
    return d;
  };
  auto d_pages = upload(pages.data(), pages.size() * sizeof(PqPage));
  auto d_cols = upload(cols.data(), cols.size() * sizeof(PqCol));
  auto d_dicts = upload(dicts.data(), dicts.size() * sizeof(PqDict));
  std::vector<int> lists;
  lists.insert(lists.end(), snappy_list.begin(), snappy_list.end());
  lists.insert(lists.end(), dict_list.begin(), dict_list.end());
  lists.insert(lists.end(), data_list.begin(), data_list.end());
  auto d_lists = upload(lists.data(), lists.size() * sizeof(int));
  auto dictbuf = dev_alloc(ctx, (size_t)dict_bytes + 16);
  int* status = counters->as<int>();
  const int* l_snappy = d_lists->as<int>();
  const int* l_dict = l_snappy + snappy_list.size();
  const int* l_data = l_dict + dict_list.size();
  if (!snappy_list.empty()) {
    // snappy_list is sorted by size: [0, n_large) use the 64 KB window on the main stream, the rest the 32 KB window on
    // the second stream, both after the copies above and both before the decode below
    size_t n_large = 0;
    while (n_large < snappy_list.size() && pages[snappy_list[n_large]].data_len > SN_SMALL_PAGE) ++n_large;
    const size_t n_small = snappy_list.size() - n_large;
    const uint32_t buf_len = (uint32_t)(dev_bytes + 64) & ~15u;
    static bool attr_set = false;
    if (!attr_set) {
      B2_CUDA(cudaFuncSetAttribute(pq_snappy_kernel<SN_WIN_LARGE>, cudaFuncAttributeMaxDynamicSharedMemorySize, SN_RING + SN_WIN_LARGE));
      attr_set = true;
    }
    cudaEvent_t fork = nullptr, join = nullptr;
    if (n_small && n_large) {
      B2_CUDA(cudaEventCreateWithFlags(&fork, cudaEventDisableTiming));
      B2_CUDA(cudaEventCreateWithFlags(&join, cudaEventDisableTiming));
      B2_CUDA(cudaEventRecord(fork, ctx->stream));
      B2_CUDA(cudaStreamWaitEvent(ctx->copy_stream, fork, 0));
    }
    if (n_large) {
      pq_snappy_kernel<SN_WIN_LARGE><<<(unsigned)n_large, 32, SN_RING + SN_WIN_LARGE, ctx->stream>>>(
          buf->as<uint8_t>(), buf_len, d_pages->as<PqPage>(), l_snappy, status);
      B2_LAUNCHED(ctx);
    }
    if (n_small) {
      pq_snappy_kernel<SN_WIN_SMALL><<<(unsigned)n_small, 32, SN_RING + SN_WIN_SMALL, n_large ? ctx->copy_stream : ctx->stream>>>(
          buf->as<uint8_t>(), buf_len, d_pages->as<PqPage>(), l_snappy + n_large, status);
      B2_LAUNCHED(ctx);
    }
    if (fork) {
      B2_CUDA(cudaEventRecord(join, ctx->copy_stream));
      B2_CUDA(cudaStreamWaitEvent(ctx->stream, join, 0));
      cudaEventDestroy(fork);
      cudaEventDestroy(join);
    }
  }
  if (!dict_list.empty()) {
    pq_dict_kernel<<<(unsigned)dict_list.size(), PQ_THREADS, 0, ctx->stream>>>(
        buf->as<uint8_t>(), d_pages->as<PqPage>(), l_dict, d_cols->as<PqCol>(), d_dicts->as<PqDict>(), dictbuf->as<uint8_t>(), status);
    B2_LAUNCHED(ctx);
  }
  if (!data_list.empty()) {
    pq_data_kernel<<<(unsigned)data_list.size(), PQ_THREADS, 0, ctx->stream>>>(
        buf->as<uint8_t>(), d_pages->as<PqPage>(), l_data, d_cols->as<PqCol>(), d_dicts->as<PqDict>(), dictbuf->as<uint8_t>(), status);
    B2_LAUNCHED(ctx);
  }
  std::vector<int64_t> host_counters(ncols + 1);
  B2_READBACK(ctx, host_counters.data(), counters->ptr, (size_t)(ncols + 1) * 8);
  {
    int st[2];
    memcpy(st, &host_counters[0], 8);
    if (st[0]) {
      static const char* what[] = {"", "corrupt definition levels", "corrupt values", "dictionary index out of range",
                                   "corrupt snappy stream", "corrupt dictionary page"};
      const PqPage& pg = pages[st[1] >= 0 && st[1] < (int)pages.size() ? st[1] : 0];
      fail(B2_ERR_ARROW, std::string("Parquet: ") + what[st[0] < 6 ? st[0] : 2] + " in a page of column '" +
                             schema.fields[pg.col].name + "' (rows from " + std::to_string(pg.row_off) + ")");
    }
  }
  for (int c = 0; c < ncols; ++c) {
    Column& col = out->cols[c];
    if (col.valid && host_counters[1 + c] == 0) col.valid.reset();  // OPTIONAL in the file, no NULL in these rows
    if (col.dtype != B2_UTF8) continue;
    col.data = dev_alloc_zero(ctx, (size_t)(rows + 1) * 4);
    int64_t total = 0;
    device_scan(ctx, PqLenIn{lens[c]->as<int32_t>()}, rows, PqOffsetsOut{col.data->as<int32_t>(), rows}, &total);
    if (total >= (1ll << 31)) fail(B2_ERR_EXECUTION, "Utf8 column beyond 2 GiB in one Parquet decode call (pass fewer row groups)");
    col.chars_len = total;
    col.chars = dev_alloc(ctx, (size_t)total + 16);
    pq_copy_chars_kernel<<<grid_for(ctx, 8, rows * 16, 256), 256, 0, ctx->stream>>>(
        buf->as<uint8_t>(), srcs[c]->as<uint32_t>(), col.data->as<int32_t>(), rows, col.chars->as<uint8_t>());
    B2_LAUNCHED(ctx);
  }
  B2_SYNC(ctx);  // `file` may be unmapped by the caller as soon as this returns
  return out;
}

}  // namespace b2
