// eval.cu — column-at-a-time expression evaluation over an HBM batch: the unfused path, with the same
// execution model as DataFusion's PhysicalExpr::evaluate: every node produces a full
// column; literals are broadcast as scalars.  Arithmetic is IEEE with one rounding per operation.
//
// Vocabulary = LogicalExprNode (rust/core/proto/ballista.proto:14-45) with the operator table of
// rust/core/src/serde/logical_plan/from_proto.rs:937-957: column, literal, And Or Eq NotEq Lt LtEq Gt GtEq Plus Minus
// Multiply Divide Like, Not, IsNull / IsNotNull, Between / InList (lowered at decode), Case, Cast, Negative, and the
// numeric / length / case-mapping scalar functions of ballista.proto:80-114.
//
// NULLs: a result row is NULL when any operand row is (arrow 4.0's arithmetic / comparison / `and` / `or` kernels,
// which DataFusion@46161d2 calls, are not Kleene); IS [NOT] NULL reads the validity; CASE treats a NULL condition as
// false and yields NULL when no branch fires and there is no ELSE.
#include <limits>
#include <type_traits>

#include "expr.h"
#include "scan.cuh"
#include "table.cuh"

namespace b2 {

template <class T> struct Operand {
  const T* p;
  T s;
  __device__ __forceinline__ T get(int64_t i) const { return p ? p[i] : s; }
};

struct OpAdd { template <class T> __device__ static T run(T a, T b) { return a + b; } };
struct OpSub { template <class T> __device__ static T run(T a, T b) { return a - b; } };
struct OpMul { template <class T> __device__ static T run(T a, T b) { return a * b; } };
template <> __device__ double OpAdd::run<double>(double a, double b) { return __dadd_rn(a, b); }
template <> __device__ double OpSub::run<double>(double a, double b) { return __dsub_rn(a, b); }
template <> __device__ double OpMul::run<double>(double a, double b) { return __dmul_rn(a, b); }
struct OpLt { template <class T> __device__ static uint8_t run(T a, T b) { return a < b; } };
struct OpLe { template <class T> __device__ static uint8_t run(T a, T b) { return a <= b; } };
struct OpGt { template <class T> __device__ static uint8_t run(T a, T b) { return a > b; } };
struct OpGe { template <class T> __device__ static uint8_t run(T a, T b) { return a >= b; } };
struct OpEq { template <class T> __device__ static uint8_t run(T a, T b) { return a == b; } };
struct OpNe { template <class T> __device__ static uint8_t run(T a, T b) { return a != b; } };
struct OpAnd { template <class T> __device__ static uint8_t run(T a, T b) { return (a != 0) & (b != 0); } };
struct OpOr { template <class T> __device__ static uint8_t run(T a, T b) { return (a != 0) | (b != 0); } };

template <class T, class TO, class OP>
__global__ void binary_kernel(Operand<T> a, Operand<T> b, int64_t n, TO* out) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  for (; i < n; i += (int64_t)gridDim.x * blockDim.x) out[i] = (TO)OP::run(a.get(i), b.get(i));
}
__global__ void fdiv_kernel(Operand<double> a, Operand<double> b, int64_t n, double* out) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  for (; i < n; i += (int64_t)gridDim.x * blockDim.x) out[i] = __ddiv_rn(a.get(i), b.get(i));
}
// Integer division fails the query on a zero divisor (arrow's `divide` kernel returns ArrowError::DivideByZero) and on
// MIN / -1; rows that are NULL on either side are not looked at.  status: 1 = zero divisor, 2 = overflow.
template <class T>
__global__ void idiv_kernel(Operand<T> a, Operand<T> b, const uint8_t* valid, int64_t n, T* out, int* status) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  int bad = 0;
  for (; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    T x = a.get(i), y = b.get(i);
    if (valid && !valid[i]) {
      out[i] = 0;
      continue;
    }
    if (y == 0) {
      bad |= 1;
      out[i] = 0;
    } else if (y == (T)-1) {
      if (x == (T)((T)1 << (8 * sizeof(T) - 1))) bad |= 2;
      out[i] = (T)(0 - (typename std::make_unsigned<T>::type)x);
    } else {
      out[i] = x / y;
    }
  }
  if (bad) atomicOr(status, bad);
}
template <class TI, class TO> __global__ void cast_kernel(const TI* in, int64_t n, TO* out) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  for (; i < n; i += (int64_t)gridDim.x * blockDim.x) out[i] = (TO)in[i];
}
// Float64 -> integer casts saturate and map NaN to 0 (Rust `as`, which arrow's cast kernel uses)
template <class TO> __global__ void cast_f64_int_kernel(const double* in, int64_t n, TO* out, double lo, double hi, TO top) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  for (; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    double v = in[i];
    out[i] = v != v ? (TO)0 : (v <= lo ? (TO)lo : (v >= hi ? top : (TO)v));
  }
}
template <class T> __global__ void neg_kernel(const T* in, int64_t n, T* out) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  for (; i < n; i += (int64_t)gridDim.x * blockDim.x) out[i] = -in[i];
}
__global__ void not_kernel(const uint8_t* in, int64_t n, uint8_t* out) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  for (; i < n; i += (int64_t)gridDim.x * blockDim.x) out[i] = !in[i];
}
template <class T> __global__ void fill_kernel(T v, int64_t n, T* out) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  for (; i < n; i += (int64_t)gridDim.x * blockDim.x) out[i] = v;
}

// ---- Utf8 comparisons: = <> < <= > >= by unsigned byte order (Rust str ordering), column vs literal / column ----
__device__ __forceinline__ bool apply_cmp(int op, int c) {
  switch (op) {
    case B2_OP_EQ: return c == 0;
    case B2_OP_NOTEQ: return c != 0;
    case B2_OP_LT: return c < 0;
    case B2_OP_LTEQ: return c <= 0;
    case B2_OP_GT: return c > 0;
    default: return c >= 0;
  }
}
__device__ __forceinline__ int bytes_cmp(const uint8_t* a, int la, const uint8_t* b, int lb) {
  int m = la < lb ? la : lb;
  for (int q = 0; q < m; ++q)
    if (a[q] != b[q]) return a[q] < b[q] ? -1 : 1;
  return la == lb ? 0 : (la < lb ? -1 : 1);
}
// b_off == nullptr: the right side is the literal (b_chr, lit_len); `swap`: the literal was the LEFT operand
__global__ void str_cmp_kernel(const int32_t* a_off, const uint8_t* a_chr, const int32_t* b_off, const uint8_t* b_chr,
                               int lit_len, int op, bool swap, int64_t n, uint8_t* out) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  for (; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    int32_t ab = a_off[i], la = a_off[i + 1] - ab;
    int c;
    if (b_off) {
      int32_t bb = b_off[i];
      c = bytes_cmp(a_chr + ab, la, b_chr + bb, b_off[i + 1] - bb);
    } else {
      c = bytes_cmp(a_chr + ab, la, b_chr, lit_len);
    }
    out[i] = apply_cmp(op, swap ? -c : c);
  }
}
// SQL LIKE against a literal pattern: % = any run of characters, _ = exactly one character (one UTF-8 code point);
// no escape character (arrow 4.0's like_utf8_scalar knows none either)
__global__ void str_like_kernel(const int32_t* off, const uint8_t* chr, const uint8_t* pat, int pl, int64_t n,
                                uint8_t* out) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  for (; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const uint8_t* s = chr + off[i];
    const int sl = off[i + 1] - off[i];
    int si = 0, pi = 0, star = -1, mark = 0;
    bool ok = true;
    while (si < sl) {
      if (pi < pl && pat[pi] == '%') {
        star = pi++;
        mark = si;
      } else if (pi < pl && pat[pi] == '_') {
        ++si;
        while (si < sl && (s[si] & 0xC0) == 0x80) ++si;  // rest of a multi-byte code point
        ++pi;
      } else if (pi < pl && pat[pi] == s[si]) {
        ++si;
        ++pi;
      } else if (star >= 0) {
        pi = star + 1;
        ++mark;
        while (mark < sl && (s[mark] & 0xC0) == 0x80) ++mark;
        si = mark;
      } else {
        ok = false;
        break;
      }
    }
    while (ok && pi < pl && pat[pi] == '%') ++pi;
    out[i] = ok && pi == pl;
  }
}
__global__ void str_len_kernel(const int32_t* off, int64_t n, int32_t* out) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  for (; i < n; i += (int64_t)gridDim.x * blockDim.x) out[i] = off[i + 1] - off[i];
}
__global__ void str_case_kernel(const uint8_t* in, int64_t nbytes, bool upper, uint8_t* out, int* status) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  bool non_ascii = false;
  for (; i < nbytes; i += (int64_t)gridDim.x * blockDim.x) {
    uint8_t c = in[i];
    non_ascii |= c >= 0x80;
    if (upper) out[i] = (c >= 'a' && c <= 'z') ? c - 32 : c;
    else out[i] = (c >= 'A' && c <= 'Z') ? c + 32 : c;
  }
  if (non_ascii) atomicOr(status, 4);
}
// trim / ltrim / rtrim (ScalarFunction TRIM / LTRIM / RTRIM).  The reference's DataFusion revision sits inside the
// weeks in which these moved from Rust's str::trim* (every Unicode White_Space character) to the Postgres forms
// (U+0020 only), and its source is not on disk.  Both agree whenever the character left at a trimmed edge is no kind
// of white space, so: U+0020 is trimmed, and a row whose trimmed edge stops at U+0009..U+000D or at a byte >= 0x80
// (possibly a non-ASCII White_Space code point) raises status bit 8 - refused, not guessed.
__device__ __forceinline__ bool edge_in_doubt(uint8_t c) { return (c >= 0x09 && c <= 0x0D) || c >= 0x80; }
__global__ void str_trim_kernel(const int32_t* off, const uint8_t* chr, const uint8_t* valid, int64_t n, bool left,
                                bool right, int32_t* begin, int32_t* len, int* status) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  bool doubt = false;
  for (; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    int32_t b = off[i], e = off[i + 1];
    const bool live = !valid || valid[i];
    if (left) {
      while (b < e && chr[b] == 0x20) ++b;
      doubt |= live && b < e && edge_in_doubt(chr[b]);
    }
    if (right) {
      while (e > b && chr[e - 1] == 0x20) --e;
      doubt |= live && e > b && edge_in_doubt(chr[e - 1]);
    }
    begin[i] = b;
    len[i] = e - b;
  }
  if (doubt) atomicOr(status, 8);
}
__global__ void str_copy_ranges_kernel(const uint8_t* in_chr, const int32_t* begin, const int32_t* out_off, int64_t n,
                                       uint8_t* out_chr) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  for (; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const int32_t b = begin[i], o = out_off[i], l = out_off[i + 1] - o;
    for (int32_t q = 0; q < l; ++q) out_chr[o + q] = in_chr[b + q];
  }
}
struct TrimOffOut {
  int32_t* off;
  int64_t n;
  __device__ __forceinline__ void operator()(int64_t i, int v, int64_t pre) const {
    off[i] = (int32_t)pre;
    if (i == n - 1) off[n] = (int32_t)(pre + v);
  }
};
// md5(Utf8) -> 32 lower-case hex digits (ScalarFunction MD5; RFC 1321).  One thread per string: the message is
// consumed in 64-byte blocks straight from the character buffer, the padded tail block(s) are assembled in registers.
__device__ __forceinline__ uint32_t md5_rotl(uint32_t x, int c) { return (x << c) | (x >> (32 - c)); }
__device__ void md5_block(uint32_t st[4], const uint32_t m[16]) {
  const uint32_t K[64] = {
      0xd76aa478, 0xe8c7b756, 0x242070db, 0xc1bdceee, 0xf57c0faf, 0x4787c62a, 0xa8304613, 0xfd469501, 0x698098d8,
      0x8b44f7af, 0xffff5bb1, 0x895cd7be, 0x6b901122, 0xfd987193, 0xa679438e, 0x49b40821, 0xf61e2562, 0xc040b340,
      0x265e5a51, 0xe9b6c7aa, 0xd62f105d, 0x02441453, 0xd8a1e681, 0xe7d3fbc8, 0x21e1cde6, 0xc33707d6, 0xf4d50d87,
      0x455a14ed, 0xa9e3e905, 0xfcefa3f8, 0x676f02d9, 0x8d2a4c8a, 0xfffa3942, 0x8771f681, 0x6d9d6122, 0xfde5380c,
      0xa4beea44, 0x4bdecfa9, 0xf6bb4b60, 0xbebfbc70, 0x289b7ec6, 0xeaa127fa, 0xd4ef3085, 0x04881d05, 0xd9d4d039,
      0xe6db99e5, 0x1fa27cf8, 0xc4ac5665, 0xf4292244, 0x432aff97, 0xab9423a7, 0xfc93a039, 0x655b59c3, 0x8f0ccc92,
      0xffeff47d, 0x85845dd1, 0x6fa87e4f, 0xfe2ce6e0, 0xa3014314, 0x4e0811a1, 0xf7537e82, 0xbd3af235, 0x2ad7d2bb,
      0xeb86d391};
  const int R[64] = {7, 12, 17, 22, 7, 12, 17, 22, 7, 12, 17, 22, 7, 12, 17, 22, 5, 9,  14, 20, 5, 9,
                     14, 20, 5, 9,  14, 20, 5, 9,  14, 20, 4, 11, 16, 23, 4, 11, 16, 23, 4, 11, 16, 23,
                     4, 11, 16, 23, 6, 10, 15, 21, 6, 10, 15, 21, 6, 10, 15, 21, 6, 10, 15, 21};
  uint32_t a = st[0], b = st[1], c = st[2], d = st[3];
  for (int i = 0; i < 64; ++i) {
    uint32_t f;
    int g;
    if (i < 16) { f = (b & c) | (~b & d); g = i; }
    else if (i < 32) { f = (d & b) | (~d & c); g = (5 * i + 1) & 15; }
    else if (i < 48) { f = b ^ c ^ d; g = (3 * i + 5) & 15; }
    else { f = c ^ (b | ~d); g = (7 * i) & 15; }
    const uint32_t t = d;
    d = c;
    c = b;
    b = b + md5_rotl(a + f + K[i] + m[g], R[i]);
    a = t;
  }
  st[0] += a; st[1] += b; st[2] += c; st[3] += d;
}
__global__ void md5_kernel(const int32_t* off, const uint8_t* chr, int64_t n, int32_t* out_off, uint8_t* out_chr) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  for (; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const uint8_t* p = chr + off[i];
    const uint32_t len = (uint32_t)(off[i + 1] - off[i]);
    uint32_t st[4] = {0x67452301u, 0xefcdab89u, 0x98badcfeu, 0x10325476u};
    uint32_t m[16];
    uint32_t done = 0;
    for (; done + 64 <= len; done += 64) {
      for (int w = 0; w < 16; ++w)
        m[w] = (uint32_t)p[done + 4 * w] | ((uint32_t)p[done + 4 * w + 1] << 8) | ((uint32_t)p[done + 4 * w + 2] << 16) |
               ((uint32_t)p[done + 4 * w + 3] << 24);
      md5_block(st, m);
    }
    // tail: the remaining bytes, 0x80, zeros, the bit length in the last 8 bytes of a block (a second block when the
    // remainder is 56 bytes or more)
    const uint32_t rem = len - done;
    for (int blk = 0; blk < (rem >= 56 ? 2 : 1); ++blk) {
      for (int w = 0; w < 16; ++w) {
        uint32_t word = 0;
        for (int q = 0; q < 4; ++q) {
          const uint32_t pos = (uint32_t)blk * 64 + 4 * w + q;  // byte position within the tail
          uint32_t byte = 0;
          if (pos < rem) byte = p[done + pos];
          else if (pos == rem) byte = 0x80;
          word |= byte << (8 * q);
        }
        m[w] = word;
      }
      if (blk == (rem >= 56 ? 1 : 0)) {
        m[14] = len << 3;
        m[15] = len >> 29;
      }
      md5_block(st, m);
    }
    out_off[i] = (int32_t)(i * 32);
    if (i == n - 1) out_off[n] = (int32_t)(n * 32);
    for (int w = 0; w < 4; ++w)
      for (int q = 0; q < 4; ++q) {
        const uint32_t byte = (st[w] >> (8 * q)) & 0xFFu;
        const char* hex = "0123456789abcdef";
        out_chr[i * 32 + w * 8 + q * 2] = (uint8_t)hex[byte >> 4];
        out_chr[i * 32 + w * 8 + q * 2 + 1] = (uint8_t)hex[byte & 15];
      }
  }
}
// a Utf8 literal broadcast to a column (CASE branches, projections)
__global__ void str_fill_kernel(const uint8_t* lit, int len, int64_t n, int32_t* off, uint8_t* chr) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  for (; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    off[i] = (int32_t)(i * len);
    for (int q = 0; q < len; ++q) chr[i * len + q] = lit[q];
    if (i == n - 1) off[n] = (int32_t)(n * len);
  }
}

// ---- scalar functions on Float64 (ScalarFunction, ballista.proto:80-114) ----
enum {
  FN_SQRT = 0, FN_SIN, FN_COS, FN_TAN, FN_ASIN, FN_ACOS, FN_ATAN, FN_EXP, FN_LOG, FN_LOG2, FN_LOG10, FN_FLOOR, FN_CEIL,
  FN_ROUND, FN_TRUNC, FN_ABS, FN_SIGNUM, FN_OCTETLENGTH, FN_CONCAT, FN_LOWER, FN_UPPER, FN_TRIM, FN_LTRIM, FN_RTRIM,
  FN_TOTIMESTAMP, FN_ARRAY, FN_NULLIF, FN_DATETRUNC, FN_MD5
};
__global__ void math_kernel(int fn, const double* in, int64_t n, double* out) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  for (; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    double x = in[i], r;
    switch (fn) {
      case FN_SQRT: r = sqrt(x); break;
      case FN_SIN: r = sin(x); break;
      case FN_COS: r = cos(x); break;
      case FN_TAN: r = tan(x); break;
      case FN_ASIN: r = asin(x); break;
      case FN_ACOS: r = acos(x); break;
      case FN_ATAN: r = atan(x); break;
      case FN_EXP: r = exp(x); break;
      case FN_LOG: r = log(x); break;
      case FN_LOG2: r = log2(x); break;
      case FN_LOG10: r = log10(x); break;
      case FN_FLOOR: r = floor(x); break;
      case FN_CEIL: r = ceil(x); break;
      case FN_ROUND: r = round(x); break;  // half away from zero, as f64::round
      case FN_TRUNC: r = trunc(x); break;
      case FN_ABS: r = fabs(x); break;
      default: r = x != x ? x : (signbit(x) ? -1.0 : 1.0); break;  // f64::signum: +-0.0 -> +-1.0, NaN -> NaN
    }
    out[i] = r;
  }
}

// ---- CASE: out = cond (true and not NULL) ? a : b, validity likewise ----
template <class T>
__global__ void select_kernel(const uint8_t* cond, const uint8_t* cond_valid, Operand<T> a, const uint8_t* a_valid,
                              bool a_null, Operand<T> b, const uint8_t* b_valid, bool b_null, int64_t n, T* out,
                              uint8_t* out_valid) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  for (; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    bool c = cond[i] != 0 && (!cond_valid || cond_valid[i]);
    bool v = c ? (!a_null && (!a_valid || a_valid[i])) : (!b_null && (!b_valid || b_valid[i]));
    out[i] = v ? (c ? a.get(i) : b.get(i)) : T();
    if (out_valid) out_valid[i] = v;
  }
}

static Column new_col(Ctx* ctx, int dtype, int64_t n) {
  Column c;
  c.dtype = dtype;
  c.length = n;
  c.data = dev_alloc(ctx, (size_t)n * dtype_width(dtype) + 16);
  return c;
}

template <class T, class TO, class OP>
static void launch_binary(Ctx* ctx, Operand<T> a, Operand<T> b, int64_t n, TO* out) {
  if (!n) return;
  binary_kernel<T, TO, OP><<<grid_for(ctx, 8, n, 256), 256, 0, ctx->stream>>>(a, b, n, out);
  B2_LAUNCHED(ctx);
}

template <class T> static Operand<T> operand_of(const ExprPtr& e, const Column& c) {
  Operand<T> o;
  if (e->is_literal()) {
    o.p = nullptr;
    o.s = e->dtype == B2_FLOAT64 ? (T)e->f : (T)e->i;
  } else {
    o.p = c.values<T>();
    o.s = T();
  }
  return o;
}

static void check_eval_status(Ctx* ctx, const BufPtr& st) {
  int host = 0;
  B2_READBACK(ctx, &host, st->ptr, 4);
  if (host & 1) fail(B2_ERR_EXECUTION, "Divide by zero");
  if (host & 2) fail(B2_ERR_EXECUTION, "integer overflow in division");
  if (host & 4) fail(B2_ERR_NOT_IMPLEMENTED, "lower() / upper() on non-ASCII text");
  if (host & 16)
    fail(B2_ERR_NOT_IMPLEMENTED, "concat() with a NULL argument (the reference revision's NULL rule for concat is not known)");
  if (host & 8) fail(B2_ERR_NOT_IMPLEMENTED, "trim() of a string whose trimmed edge is a tab / line break or a non-ASCII character (the reference revision's "
                   "white-space set is not known)");
}

template <class T>
static Column eval_binary_typed(Ctx* ctx, int op, const ExprPtr& ea, const Column& ca, const ExprPtr& eb,
                                const Column& cb, int64_t n, int arith_dtype, const BufPtr& valid) {
  Operand<T> a = operand_of<T>(ea, ca), b = operand_of<T>(eb, cb);
  bool cmp = op >= B2_OP_EQ && op <= B2_OP_GTEQ;
  if (cmp) {
    Column out = new_col(ctx, B2_BOOL, n);
    uint8_t* o = out.data->as<uint8_t>();
    switch (op) {
      case B2_OP_EQ: launch_binary<T, uint8_t, OpEq>(ctx, a, b, n, o); break;
      case B2_OP_NOTEQ: launch_binary<T, uint8_t, OpNe>(ctx, a, b, n, o); break;
      case B2_OP_LT: launch_binary<T, uint8_t, OpLt>(ctx, a, b, n, o); break;
      case B2_OP_LTEQ: launch_binary<T, uint8_t, OpLe>(ctx, a, b, n, o); break;
      case B2_OP_GT: launch_binary<T, uint8_t, OpGt>(ctx, a, b, n, o); break;
      case B2_OP_GTEQ: launch_binary<T, uint8_t, OpGe>(ctx, a, b, n, o); break;
    }
    return out;
  }
  Column out = new_col(ctx, arith_dtype, n);
  T* o = out.data->as<T>();
  switch (op) {
    case B2_OP_PLUS: launch_binary<T, T, OpAdd>(ctx, a, b, n, o); break;
    case B2_OP_MINUS: launch_binary<T, T, OpSub>(ctx, a, b, n, o); break;
    case B2_OP_MULTIPLY: launch_binary<T, T, OpMul>(ctx, a, b, n, o); break;
    case B2_OP_DIVIDE:
      if (!n) break;
      if constexpr (std::is_same<T, double>::value) {
        fdiv_kernel<<<grid_for(ctx, 8, n, 256), 256, 0, ctx->stream>>>(a, b, n, o);
        B2_LAUNCHED(ctx);
      } else {
        BufPtr st = dev_alloc_zero(ctx, 16);
        idiv_kernel<T><<<grid_for(ctx, 8, n, 256), 256, 0, ctx->stream>>>(a, b, valid ? valid->as<uint8_t>() : nullptr, n,
                                                                         o, st->as<int>());
        B2_LAUNCHED(ctx);
        check_eval_status(ctx, st);
      }
      break;
    default: fail(B2_ERR_NOT_IMPLEMENTED, "binary operator");
  }
  return out;
}

static BufPtr upload_bytes(Ctx* ctx, const std::string& s) {
  auto d = dev_alloc(ctx, s.size() + 16);
  // the literal's bytes live in the expression node, which outlives the asynchronous copy
  if (!s.empty()) B2_CUDA(cudaMemcpyAsync(d->ptr, s.data(), s.size(), cudaMemcpyHostToDevice, ctx->stream));
  return d;
}

static Column eval_case(Ctx* ctx, const ExprPtr& e, const Batch& batch);
static Column eval_scalar_fn(Ctx* ctx, const ExprPtr& e, const Batch& batch);

Column eval_expr(Ctx* ctx, const ExprPtr& e, const Batch& batch) {
  int64_t n = batch.num_rows;
  int grid = grid_for(ctx, 8, n, 256);
  switch (e->kind) {
    case B2_EXPR_COLUMN:
      if (!batch.cols[e->col].data)
        fail(B2_ERR_EXECUTION, "internal: column '" + batch.schema.fields[e->col].name + "' was pruned upstream");
      return batch.cols[e->col];
    case B2_EXPR_LITERAL: {
      if (e->dtype == B2_UTF8) {
        Column out;
        out.dtype = B2_UTF8;
        out.length = n;
        int64_t total = n * (int64_t)e->s.size();
        if (total >= (1ll << 31)) fail(B2_ERR_EXECUTION, "Utf8 literal column beyond 2 GiB");
        out.data = dev_alloc_zero(ctx, (size_t)(n + 1) * 4);
        out.chars = dev_alloc(ctx, (size_t)total + 16);
        out.chars_len = total;
        if (n) {
          BufPtr lit = upload_bytes(ctx, e->s);
          str_fill_kernel<<<grid, 256, 0, ctx->stream>>>(lit->as<uint8_t>(), (int)e->s.size(), n,
                                                        out.data->as<int32_t>(), out.chars->as<uint8_t>());
          B2_LAUNCHED(ctx);
        }
        return out;
      }
      Column out = new_col(ctx, e->dtype, n);
      if (n) {
        switch (dtype_width(e->dtype)) {
          case 1: fill_kernel<uint8_t><<<grid, 256, 0, ctx->stream>>>((uint8_t)e->i, n, out.data->as<uint8_t>()); break;
          case 4: fill_kernel<int32_t><<<grid, 256, 0, ctx->stream>>>((int32_t)e->i, n, out.data->as<int32_t>()); break;
          default:
            if (e->dtype == B2_FLOAT64) fill_kernel<double><<<grid, 256, 0, ctx->stream>>>(e->f, n, out.data->as<double>());
            else fill_kernel<int64_t><<<grid, 256, 0, ctx->stream>>>(e->i, n, out.data->as<int64_t>());
        }
        B2_LAUNCHED(ctx);
      }
      return out;
    }
    case B2_EXPR_IS_NULL:
    case B2_EXPR_IS_NOT_NULL: {
      Column c = eval_expr(ctx, e->ch[0], batch);
      Column out;
      out.dtype = B2_BOOL;
      out.length = n;
      out.data = e->kind == B2_EXPR_IS_NULL ? valid_not(ctx, c.validity(), n) : valid_copy(ctx, c.validity(), n);
      return out;
    }
    case B2_EXPR_NOT: {
      Column c = eval_expr(ctx, e->ch[0], batch);
      Column out = new_col(ctx, B2_BOOL, n);
      out.valid = c.valid;
      if (n) {
        not_kernel<<<grid, 256, 0, ctx->stream>>>(c.values<uint8_t>(), n, out.data->as<uint8_t>());
        B2_LAUNCHED(ctx);
      }
      return out;
    }
    case B2_EXPR_NEGATIVE: {
      Column c = eval_expr(ctx, e->ch[0], batch);
      Column out = new_col(ctx, e->dtype, n);
      out.valid = c.valid;
      if (n) {
        if (e->dtype == B2_FLOAT64) neg_kernel<double><<<grid, 256, 0, ctx->stream>>>(c.values<double>(), n, out.data->as<double>());
        else if (e->dtype == B2_INT32) neg_kernel<int32_t><<<grid, 256, 0, ctx->stream>>>(c.values<int32_t>(), n, out.data->as<int32_t>());
        else neg_kernel<int64_t><<<grid, 256, 0, ctx->stream>>>(c.values<int64_t>(), n, out.data->as<int64_t>());
        B2_LAUNCHED(ctx);
      }
      return out;
    }
    case B2_EXPR_CAST: {
      Column c = eval_expr(ctx, e->ch[0], batch);
      int from = c.dtype, to = e->dtype;
      if (from == to) return c;
      if ((from == B2_DATE32 && to == B2_INT32) || (from == B2_INT32 && to == B2_DATE32) ||
          (from == B2_INT64 && to == B2_UINT64) || (from == B2_UINT64 && to == B2_INT64) ||
          (from == B2_BOOL && to == B2_UINT8)) {
        c.dtype = to;  // same bits
        return c;
      }
      Column out = new_col(ctx, to, n);
      out.valid = c.valid;
      if (n) {
#define B2_CAST(FROM, TFROM, TO, TTO)                                                                             \
  if (from == FROM && to == TO) {                                                                                 \
    cast_kernel<TFROM, TTO><<<grid, 256, 0, ctx->stream>>>(c.values<TFROM>(), n, out.data->as<TTO>());            \
    done = true;                                                                                                  \
  }
        bool done = false;
        B2_CAST(B2_INT32, int32_t, B2_INT64, int64_t)
        B2_CAST(B2_INT32, int32_t, B2_UINT64, u64)
        B2_CAST(B2_INT32, int32_t, B2_FLOAT64, double)
        B2_CAST(B2_DATE32, int32_t, B2_INT64, int64_t)
        B2_CAST(B2_INT64, int64_t, B2_FLOAT64, double)
        B2_CAST(B2_INT64, int64_t, B2_INT32, int32_t)
        B2_CAST(B2_INT64, int64_t, B2_DATE32, int32_t)
        B2_CAST(B2_UINT64, u64, B2_FLOAT64, double)
        B2_CAST(B2_UINT64, u64, B2_INT32, int32_t)
        B2_CAST(B2_UINT8, uint8_t, B2_INT64, int64_t)
        B2_CAST(B2_UINT8, uint8_t, B2_INT32, int32_t)
        B2_CAST(B2_UINT8, uint8_t, B2_UINT64, u64)
        B2_CAST(B2_UINT8, uint8_t, B2_FLOAT64, double)
        B2_CAST(B2_BOOL, uint8_t, B2_INT32, int32_t)
        B2_CAST(B2_BOOL, uint8_t, B2_INT64, int64_t)
        B2_CAST(B2_BOOL, uint8_t, B2_UINT64, u64)
        B2_CAST(B2_BOOL, uint8_t, B2_FLOAT64, double)
#undef B2_CAST
        if (!done && from == B2_FLOAT64 && to == B2_INT64) {
          cast_f64_int_kernel<int64_t><<<grid, 256, 0, ctx->stream>>>(c.values<double>(), n, out.data->as<int64_t>(),
                                                                     -9223372036854775808.0, 9223372036854775808.0, INT64_MAX);
          done = true;
        }
        if (!done && from == B2_FLOAT64 && to == B2_INT32) {
          cast_f64_int_kernel<int32_t><<<grid, 256, 0, ctx->stream>>>(c.values<double>(), n, out.data->as<int32_t>(),
                                                                     -2147483648.0, 2147483647.0, INT32_MAX);
          done = true;
        }
        if (!done)
          fail(B2_ERR_NOT_IMPLEMENTED, std::string("runtime CAST ") + dtype_name(from) + " -> " + dtype_name(to));
        B2_LAUNCHED(ctx);
      }
      return out;
    }
    case B2_EXPR_CASE: return eval_case(ctx, e, batch);
    case B2_EXPR_SCALAR_FN: return eval_scalar_fn(ctx, e, batch);
    case B2_EXPR_BINARY: {
      const ExprPtr &ea = e->ch[0], &eb = e->ch[1];
      int t = ea->dtype;  // operands were coerced to a common type at decode
      if (t == B2_UTF8) {
        Column out = new_col(ctx, B2_BOOL, n);
        if (e->op == B2_OP_LIKE) {
          if (!eb->is_literal()) fail(B2_ERR_NOT_IMPLEMENTED, "LIKE with a non-literal pattern");
          Column c = eval_expr(ctx, ea, batch);
          out.valid = c.valid;
          if (n) {
            BufPtr pat = upload_bytes(ctx, eb->s);
            str_like_kernel<<<grid, 256, 0, ctx->stream>>>(c.values<int32_t>(), c.chars->as<uint8_t>(),
                                                          pat->as<uint8_t>(), (int)eb->s.size(), n, out.data->as<uint8_t>());
            B2_LAUNCHED(ctx);
          }
          return out;
        }
        if (ea->is_literal() || eb->is_literal()) {
          const bool swap = ea->is_literal();
          const ExprPtr& lit = swap ? ea : eb;
          Column c = eval_expr(ctx, swap ? eb : ea, batch);
          out.valid = c.valid;
          if (n) {
            BufPtr dlit = upload_bytes(ctx, lit->s);
            str_cmp_kernel<<<grid, 256, 0, ctx->stream>>>(c.values<int32_t>(), c.chars->as<uint8_t>(), nullptr,
                                                         dlit->as<uint8_t>(), (int)lit->s.size(), e->op, swap, n,
                                                         out.data->as<uint8_t>());
            B2_LAUNCHED(ctx);
          }
        } else {
          Column a = eval_expr(ctx, ea, batch), b = eval_expr(ctx, eb, batch);
          out.valid = valid_and(ctx, a.valid, b.valid, n);
          if (n) {
            str_cmp_kernel<<<grid, 256, 0, ctx->stream>>>(a.values<int32_t>(), a.chars->as<uint8_t>(),
                                                         b.values<int32_t>(), b.chars->as<uint8_t>(), 0, e->op, false, n,
                                                         out.data->as<uint8_t>());
            B2_LAUNCHED(ctx);
          }
        }
        return out;
      }
      Column ca, cb;
      if (!ea->is_literal()) ca = eval_expr(ctx, ea, batch);
      if (!eb->is_literal()) cb = eval_expr(ctx, eb, batch);
      BufPtr valid = valid_and(ctx, ca.valid, cb.valid, n);
      Column out;
      if (e->op == B2_OP_AND || e->op == B2_OP_OR) {
        out = new_col(ctx, B2_BOOL, n);
        Operand<uint8_t> a = operand_of<uint8_t>(ea, ca), b = operand_of<uint8_t>(eb, cb);
        if (e->op == B2_OP_AND) launch_binary<uint8_t, uint8_t, OpAnd>(ctx, a, b, n, out.data->as<uint8_t>());
        else launch_binary<uint8_t, uint8_t, OpOr>(ctx, a, b, n, out.data->as<uint8_t>());
      } else {
        switch (t) {
          case B2_INT32: case B2_DATE32: out = eval_binary_typed<int32_t>(ctx, e->op, ea, ca, eb, cb, n, e->dtype, valid); break;
          case B2_INT64: out = eval_binary_typed<int64_t>(ctx, e->op, ea, ca, eb, cb, n, e->dtype, valid); break;
          case B2_FLOAT64: out = eval_binary_typed<double>(ctx, e->op, ea, ca, eb, cb, n, e->dtype, valid); break;
          default: fail(B2_ERR_NOT_IMPLEMENTED, std::string("binary operator on ") + dtype_name(t));
        }
      }
      out.valid = valid;
      return out;
    }
  }
  fail(B2_ERR_NOT_IMPLEMENTED, "expression kind " + std::to_string(e->kind));
}

// A Boolean predicate as a filter mask: NULL counts as false (FilterExec), so the result has no validity
Column eval_predicate(Ctx* ctx, const ExprPtr& pred, const Batch& batch) {
  Column m = eval_expr(ctx, pred, batch);
  if (m.valid) {
    m.data = valid_and(ctx, m.data, m.valid, batch.num_rows);
    m.valid = nullptr;
  }
  return m;
}

// CASE [base] WHEN w1 THEN t1 ... [ELSE e] END.  children: [base?] (when, then)* [else?]; op bit 0 = base present,
// bit 1 = ELSE present.  With a base the WHEN values were rewritten to `base = when` at decode.  Branches are folded
// from the last to the first: out = when_k ? then_k : out.
template <class T>
static Column case_typed(Ctx* ctx, const ExprPtr& e, const Batch& batch, size_t first, size_t n_when, bool has_else) {
  const int64_t n = batch.num_rows;
  Column acc;  // value so far (else branch, or NULL)
  bool acc_null = true, acc_lit = false;
  ExprPtr acc_e;
  if (has_else) {
    acc_e = e->ch.back();
    acc_null = false;
    if (acc_e->is_literal()) acc_lit = true;
    else acc = eval_expr(ctx, acc_e, batch);
  }
  for (size_t k = n_when; k-- > 0;) {
    const ExprPtr& w = e->ch[first + 2 * k];
    const ExprPtr& t = e->ch[first + 2 * k + 1];
    Column cond = eval_expr(ctx, w, batch);
    Column tv;
    if (!t->is_literal()) tv = eval_expr(ctx, t, batch);
    Operand<T> a = operand_of<T>(t, tv);
    Operand<T> b;
    if (acc_null) b = Operand<T>{nullptr, T()};
    else if (acc_lit) b = operand_of<T>(acc_e, acc);
    else b = Operand<T>{acc.values<T>(), T()};
    Column out = new_col(ctx, e->dtype, n);
    const bool need_valid = acc_null || acc.valid || tv.valid;
    if (need_valid) out.valid = dev_alloc(ctx, (size_t)n + 16);
    if (n) {
      select_kernel<T><<<grid_for(ctx, 8, n, 256), 256, 0, ctx->stream>>>(
          cond.values<uint8_t>(), cond.validity(), a, tv.validity(), false, b, acc_null ? nullptr : acc.validity(),
          acc_null, n, out.data->as<T>(), need_valid ? out.valid->as<uint8_t>() : nullptr);
      B2_LAUNCHED(ctx);
    }
    acc = out;
    acc_null = false;
    acc_lit = false;
  }
  return acc;
}
// CASE producing Utf8: the first WHEN whose condition holds (true and not NULL) picks its THEN string, else the ELSE
// string, else NULL.  One pass picks the branch and its length per row, one scan makes the offsets, one pass copies.
constexpr int CASE_STR_MAX = 8;
struct CaseStrArgs {
  int n_when;
  bool has_else;
  const uint8_t* cond[CASE_STR_MAX];
  const uint8_t* cval[CASE_STR_MAX];
  const int32_t* off[CASE_STR_MAX + 1];  // THEN strings, then the ELSE string
  const uint8_t* chr[CASE_STR_MAX + 1];
  const uint8_t* val[CASE_STR_MAX + 1];
};
__global__ void case_str_pick_kernel(const __grid_constant__ CaseStrArgs a, int64_t n, uint8_t* sel, int32_t* len,
                                     uint8_t* out_valid) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  for (; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    int s = -1;
    for (int k = 0; k < a.n_when && s < 0; ++k)
      if (a.cond[k][i] != 0 && (!a.cval[k] || a.cval[k][i])) s = k;
    if (s < 0 && a.has_else) s = a.n_when;
    const bool v = s >= 0 && (!a.val[s] || a.val[s][i]);
    sel[i] = v ? (uint8_t)s : (uint8_t)255;
    len[i] = v ? a.off[s][i + 1] - a.off[s][i] : 0;
    out_valid[i] = v;
  }
}
__global__ void case_str_copy_kernel(const __grid_constant__ CaseStrArgs a, int64_t n, const uint8_t* sel,
                                     const int32_t* out_off, uint8_t* out_chr) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  for (; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const int s = sel[i];
    if (s == 255) continue;
    const int32_t b = a.off[s][i], o = out_off[i], l = out_off[i + 1] - o;
    for (int32_t q = 0; q < l; ++q) out_chr[o + q] = a.chr[s][b + q];
  }
}
static Column case_utf8(Ctx* ctx, const ExprPtr& e, const Batch& batch, size_t n_when, bool has_else) {
  const int64_t n = batch.num_rows;
  if (n_when > (size_t)CASE_STR_MAX) fail(B2_ERR_NOT_IMPLEMENTED, "CASE producing Utf8 with more than 8 WHEN branches");
  CaseStrArgs a;
  memset(&a, 0, sizeof a);
  a.n_when = (int)n_when;
  a.has_else = has_else;
  std::vector<Column> keep;  // conditions and branch strings (literals broadcast) stay alive until the kernels are queued
  for (size_t k = 0; k < n_when; ++k) {
    keep.push_back(eval_expr(ctx, e->ch[2 * k], batch));
    a.cond[k] = keep.back().values<uint8_t>();
    a.cval[k] = keep.back().validity();
    keep.push_back(eval_expr(ctx, e->ch[2 * k + 1], batch));
    a.off[k] = keep.back().values<int32_t>();
    a.chr[k] = keep.back().chars->as<uint8_t>();
    a.val[k] = keep.back().validity();
  }
  if (has_else) {
    keep.push_back(eval_expr(ctx, e->ch.back(), batch));
    a.off[n_when] = keep.back().values<int32_t>();
    a.chr[n_when] = keep.back().chars->as<uint8_t>();
    a.val[n_when] = keep.back().validity();
  }
  Column out;
  out.dtype = B2_UTF8;
  out.length = n;
  out.data = dev_alloc_zero(ctx, (size_t)(n + 1) * 4);
  out.chars = dev_alloc(ctx, 16);
  out.chars_len = 0;
  if (!n) return out;
  out.valid = dev_alloc(ctx, (size_t)n + 16);
  BufPtr sel = dev_alloc(ctx, (size_t)n + 16), len = dev_alloc(ctx, (size_t)n * 4);
  const int grid = grid_for(ctx, 8, n, 256);
  case_str_pick_kernel<<<grid, 256, 0, ctx->stream>>>(a, n, sel->as<uint8_t>(), len->as<int32_t>(), out.valid->as<uint8_t>());
  B2_LAUNCHED(ctx);
  int64_t total = 0;
  device_scan(ctx, I32In{len->as<int32_t>()}, n, TrimOffOut{out.data->as<int32_t>(), n}, &total);
  if (total >= (1ll << 31)) fail(B2_ERR_EXECUTION, "Utf8 column beyond 2 GiB after CASE");
  out.chars_len = total;
  out.chars = dev_alloc(ctx, (size_t)total + 16);
  if (total) {
    case_str_copy_kernel<<<grid, 256, 0, ctx->stream>>>(a, n, sel->as<uint8_t>(), out.data->as<int32_t>(), out.chars->as<uint8_t>());
    B2_LAUNCHED(ctx);
  }
  return out;
}

// CONCAT(s1, ..., sk), k <= 8.  The reference's DataFusion revision sits between "NULL if any argument is NULL" and the
// Postgres rule "NULL arguments are skipped" and cannot be read here; both agree on rows without a NULL argument, so
// those are concatenated and a row WITH a NULL argument raises status bit 16: refused, not guessed.
struct ConcatArgs {
  int n;
  const int32_t* off[8];
  const uint8_t* chr[8];
  const uint8_t* val[8];
};
__global__ void concat_len_kernel(const __grid_constant__ ConcatArgs a, int64_t n, int32_t* len, int* status) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  bool null_arg = false;
  for (; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    int32_t l = 0;
    for (int k = 0; k < a.n; ++k) {
      null_arg |= a.val[k] && !a.val[k][i];
      l += a.off[k][i + 1] - a.off[k][i];
    }
    len[i] = l;
  }
  if (null_arg) atomicOr(status, 16);
}
__global__ void concat_copy_kernel(const __grid_constant__ ConcatArgs a, int64_t n, const int32_t* out_off, uint8_t* out_chr) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  for (; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    int32_t o = out_off[i];
    for (int k = 0; k < a.n; ++k) {
      const int32_t b = a.off[k][i], l = a.off[k][i + 1] - b;
      for (int32_t q = 0; q < l; ++q) out_chr[o + q] = a.chr[k][b + q];
      o += l;
    }
  }
}
static Column eval_concat(Ctx* ctx, const ExprPtr& e, const Batch& batch) {
  const int64_t n = batch.num_rows;
  ConcatArgs a;
  memset(&a, 0, sizeof a);
  a.n = (int)e->ch.size();
  std::vector<Column> keep;
  for (int k = 0; k < a.n; ++k) {
    keep.push_back(eval_expr(ctx, e->ch[k], batch));
    a.off[k] = keep.back().values<int32_t>();
    a.chr[k] = keep.back().chars->as<uint8_t>();
    a.val[k] = keep.back().validity();
  }
  Column out;
  out.dtype = B2_UTF8;
  out.length = n;
  out.data = dev_alloc_zero(ctx, (size_t)(n + 1) * 4);
  out.chars = dev_alloc(ctx, 16);
  out.chars_len = 0;
  if (!n) return out;
  BufPtr st = dev_alloc_zero(ctx, 16), len = dev_alloc(ctx, (size_t)n * 4);
  const int grid = grid_for(ctx, 8, n, 256);
  concat_len_kernel<<<grid, 256, 0, ctx->stream>>>(a, n, len->as<int32_t>(), st->as<int>());
  B2_LAUNCHED(ctx);
  int64_t total = 0;
  device_scan(ctx, I32In{len->as<int32_t>()}, n, TrimOffOut{out.data->as<int32_t>(), n}, &total);
  check_eval_status(ctx, st);
  if (total >= (1ll << 31)) fail(B2_ERR_EXECUTION, "Utf8 column beyond 2 GiB after concat");
  out.chars_len = total;
  out.chars = dev_alloc(ctx, (size_t)total + 16);
  if (total) {
    concat_copy_kernel<<<grid, 256, 0, ctx->stream>>>(a, n, out.data->as<int32_t>(), out.chars->as<uint8_t>());
    B2_LAUNCHED(ctx);
  }
  return out;
}

static Column eval_case(Ctx* ctx, const ExprPtr& e, const Batch& batch) {
  const bool has_base = e->op & 1, has_else = e->op & 2;
  (void)has_base;  // folded into the WHEN conditions at decode
  size_t n_when = (e->ch.size() - (has_else ? 1 : 0)) / 2;
  if (e->dtype == B2_UTF8) return case_utf8(ctx, e, batch, n_when, has_else);
  switch (e->dtype) {
    case B2_INT32: case B2_DATE32: return case_typed<int32_t>(ctx, e, batch, 0, n_when, has_else);
    case B2_INT64: case B2_UINT64: return case_typed<int64_t>(ctx, e, batch, 0, n_when, has_else);
    case B2_FLOAT64: return case_typed<double>(ctx, e, batch, 0, n_when, has_else);
    case B2_BOOL: case B2_UINT8: return case_typed<uint8_t>(ctx, e, batch, 0, n_when, has_else);
  }
  fail(B2_ERR_NOT_IMPLEMENTED, std::string("CASE producing ") + dtype_name(e->dtype));
}

// NULLIF(a, b): children = [a, a = b] (expr.cpp); out = (a = b holds) ? NULL : a
template <class T> static Column nullif_typed(Ctx* ctx, const Column& a, const Column& cond, int dtype, int64_t n) {
  Column out = new_col(ctx, dtype, n);
  out.valid = dev_alloc(ctx, (size_t)n + 16);
  if (n) {
    select_kernel<T><<<grid_for(ctx, 8, n, 256), 256, 0, ctx->stream>>>(
        cond.values<uint8_t>(), cond.validity(), Operand<T>{nullptr, T()}, nullptr, true, Operand<T>{a.values<T>(), T()},
        a.validity(), false, n, out.data->as<T>(), out.valid->as<uint8_t>());
    B2_LAUNCHED(ctx);
  }
  return out;
}

static Column eval_scalar_fn(Ctx* ctx, const ExprPtr& e, const Batch& batch) {
  const int64_t n = batch.num_rows;
  const int grid = grid_for(ctx, 8, n, 256);
  if (e->op == FN_CONCAT) return eval_concat(ctx, e, batch);
  Column c = eval_expr(ctx, e->ch[0], batch);
  if (e->op == FN_NULLIF) {
    Column cond = eval_expr(ctx, e->ch[1], batch);
    switch (e->dtype) {
      case B2_INT32: return nullif_typed<int32_t>(ctx, c, cond, e->dtype, n);
      case B2_INT64: case B2_UINT64: return nullif_typed<int64_t>(ctx, c, cond, e->dtype, n);
      case B2_FLOAT64: return nullif_typed<double>(ctx, c, cond, e->dtype, n);
      default: return nullif_typed<uint8_t>(ctx, c, cond, e->dtype, n);
    }
  }
  if (e->op <= FN_SIGNUM) {
    Column out = new_col(ctx, B2_FLOAT64, n);
    out.valid = c.valid;
    if (n) {
      math_kernel<<<grid, 256, 0, ctx->stream>>>(e->op, c.values<double>(), n, out.data->as<double>());
      B2_LAUNCHED(ctx);
    }
    return out;
  }
  if (e->op == FN_OCTETLENGTH) {
    Column out = new_col(ctx, B2_INT32, n);
    out.valid = c.valid;
    if (n) {
      str_len_kernel<<<grid, 256, 0, ctx->stream>>>(c.values<int32_t>(), n, out.data->as<int32_t>());
      B2_LAUNCHED(ctx);
    }
    return out;
  }
  if (e->op == FN_LOWER || e->op == FN_UPPER) {
    Column out = c;  // same offsets
    out.chars = dev_alloc(ctx, (size_t)c.chars_len + 16);
    if (c.chars_len) {
      BufPtr st = dev_alloc_zero(ctx, 16);
      str_case_kernel<<<grid_for(ctx, 8, c.chars_len, 256), 256, 0, ctx->stream>>>(
          c.chars->as<uint8_t>(), c.chars_len, e->op == FN_UPPER, out.chars->as<uint8_t>(), st->as<int>());
      B2_LAUNCHED(ctx);
      check_eval_status(ctx, st);
    }
    return out;
  }
  if (e->op == FN_MD5) {
    if (n * 32 >= (1ll << 31)) fail(B2_ERR_EXECUTION, "Utf8 column beyond 2 GiB after md5");
    Column out;
    out.dtype = B2_UTF8;
    out.length = n;
    out.valid = c.valid;  // md5(NULL) is NULL; the digest computed over a NULL row's bytes is never looked at
    out.data = dev_alloc_zero(ctx, (size_t)(n + 1) * 4);
    out.chars_len = n * 32;
    out.chars = dev_alloc(ctx, (size_t)n * 32 + 16);
    if (n) {
      md5_kernel<<<grid, 256, 0, ctx->stream>>>(c.values<int32_t>(), c.chars->as<uint8_t>(), n, out.data->as<int32_t>(),
                                               out.chars->as<uint8_t>());
      B2_LAUNCHED(ctx);
    }
    return out;
  }
  if (e->op == FN_TRIM || e->op == FN_LTRIM || e->op == FN_RTRIM) {
    Column out = new_col(ctx, B2_UTF8, n);
    out.valid = c.valid;
    out.data = dev_alloc_zero(ctx, (size_t)(n + 1) * 4);
    out.chars_len = 0;
    out.chars = dev_alloc(ctx, 16);
    if (n) {
      BufPtr st = dev_alloc_zero(ctx, 16);
      BufPtr begin = dev_alloc(ctx, (size_t)n * 4), len = dev_alloc(ctx, (size_t)n * 4);
      str_trim_kernel<<<grid, 256, 0, ctx->stream>>>(c.values<int32_t>(), c.chars->as<uint8_t>(), c.validity(), n, e->op != FN_RTRIM,
                                                    e->op != FN_LTRIM, begin->as<int32_t>(), len->as<int32_t>(),
                                                    st->as<int>());
      B2_LAUNCHED(ctx);
      int64_t total = 0;
      device_scan(ctx, I32In{len->as<int32_t>()}, n, TrimOffOut{out.data->as<int32_t>(), n}, &total);
      check_eval_status(ctx, st);
      out.chars_len = total;
      out.chars = dev_alloc(ctx, (size_t)total + 16);
      if (total) {
        str_copy_ranges_kernel<<<grid, 256, 0, ctx->stream>>>(c.chars->as<uint8_t>(), begin->as<int32_t>(),
                                                           out.data->as<int32_t>(), n, out.chars->as<uint8_t>());
        B2_LAUNCHED(ctx);
      }
    }
    return out;
  }
  fail(B2_ERR_NOT_IMPLEMENTED, "scalar function " + std::to_string(e->op));
}

}  // namespace b2
