// fused.cuh — compile-time expression templates for fused scan pipelines.
//
// A fused kernel is instantiated for a *shape*: predicate / key / aggregate-argument expression trees over
// typed column slots (f<k> Float64, i<k> Int32|Date32, s<k> Utf8) and literal slots (LF/LI/LS).  Which
// columns are bound to the slots and the literal values are runtime parameters, so `l_shipdate <= X` hits
// the same kernel for every X.  Every template has sig() producing exactly the text expr_signature()
// (expr.cpp) produces for a decoded runtime expression; the registry is keyed on that text.
//
// Arithmetic is IEEE Float64 with one rounding per operation (__dmul_rn/__dadd_rn: never contracted into
// FMA), as DataFusion's column-at-a-time kernels compute it.
#pragma once
#ifndef __CUDACC_RTC__
#include <string>
#endif

#include "table.cuh"

namespace b2 {
namespace fx {

constexpr int MAX_F = 6, MAX_I = 4, MAX_S = 2, MAX_LF = 8, MAX_LI = 8, MAX_LS = 2;

// values of one row's bound slots, in registers.  W is the string word: u128 holds the first 15 bytes,
// u64 the first 7 (narrow kernels; longer key strings raise the overflow flag and the host reruns wide)
template <class W_> struct RowT {
  using W = W_;
  double f[MAX_F];
  int32_t i[MAX_I];
  StrValT<W_> s[MAX_S];
};
using Row = RowT<u128>;
struct Lits {
  double f[MAX_LF];
  int64_t i[MAX_LI];
  StrVal s[MAX_LS];
};

constexpr int cmax(int a, int b) { return a > b ? a : b; }

// ---- leaves ----
template <int K> struct F {  // Float64 column slot
  static constexpr int NF = K + 1, NI = 0, NS = 0, NLS = 0;
  static constexpr bool IS_STR = false;
  template <class R> __device__ static double eval(const R& r, const Lits&) { return r.f[K]; }
  B2_SIG(return "f" + std::to_string(K);)
};
template <int K> struct I {  // Int32 / Date32 column slot
  static constexpr int NF = 0, NI = K + 1, NS = 0, NLS = 0;
  static constexpr bool IS_STR = false;
  template <class R> __device__ static int32_t eval(const R& r, const Lits&) { return r.i[K]; }
  B2_SIG(return "i" + std::to_string(K);)
};
template <int K> struct S {  // Utf8 column slot
  static constexpr int NF = 0, NI = 0, NS = K + 1, NLS = 0;
  static constexpr bool IS_STR = true;
  template <class R> __device__ static StrValT<typename R::W> eval(const R& r, const Lits&) { return r.s[K]; }
  B2_SIG(return "s" + std::to_string(K);)
};
template <int K> struct LF {
  static constexpr int NF = 0, NI = 0, NS = 0, NLS = 0;
  static constexpr bool IS_STR = false;
  template <class R> __device__ static double eval(const R&, const Lits& l) { return l.f[K]; }
  B2_SIG(return "LF" + std::to_string(K);)
};
template <int K> struct LI {
  static constexpr int NF = 0, NI = 0, NS = 0, NLS = 0;
  static constexpr bool IS_STR = false;
  template <class R> __device__ static int64_t eval(const R&, const Lits& l) { return l.i[K]; }
  B2_SIG(return "LI" + std::to_string(K);)
};
template <int K> struct LS {
  static constexpr int NF = 0, NI = 0, NS = 0, NLS = K + 1;
  static constexpr bool IS_STR = true;
  template <class R> __device__ static StrValT<typename R::W> eval(const R&, const Lits& l) {
    return StrValT<typename R::W>{(typename R::W)l.s[K].bytes, l.s[K].len};
  }
  B2_SIG(return "LS" + std::to_string(K);)
};
struct True {  // absent predicate
  static constexpr int NF = 0, NI = 0, NS = 0, NLS = 0;
  static constexpr bool IS_STR = false;
  template <class R> __device__ static bool eval(const R&, const Lits&) { return true; }
  B2_SIG(return "true";)
};

// ---- operators ----
__device__ __forceinline__ double op_add(double a, double b) { return __dadd_rn(a, b); }
__device__ __forceinline__ double op_sub(double a, double b) { return __dsub_rn(a, b); }
__device__ __forceinline__ double op_mul(double a, double b) { return __dmul_rn(a, b); }
__device__ __forceinline__ double op_div(double a, double b) { return __ddiv_rn(a, b); }
template <class W> __device__ __forceinline__ bool str_eq(const StrValT<W>& a, const StrValT<W>& b) {
  return a.len == b.len && a.bytes == b.bytes;
}

#define B2_FX_BINARY(NAME, TEXT, RET, BODY)                                                   \
  template <class A, class B> struct NAME {                                                   \
    static constexpr int NF = cmax(A::NF, B::NF), NI = cmax(A::NI, B::NI), NS = cmax(A::NS, B::NS), NLS = cmax(A::NLS, B::NLS); \
    static constexpr bool IS_STR = false;                                                     \
    template <class R> __device__ static RET eval(const R& r, const Lits& l) {                \
      auto a = A::eval(r, l);                                                                 \
      auto b = B::eval(r, l);                                                                 \
      return BODY;                                                                            \
    }                                                                                         \
    B2_SIG(return std::string(TEXT "(") + A::sig() + "," + B::sig() + ")";) \
  };
B2_FX_BINARY(Add, "add", double, op_add(a, b))
B2_FX_BINARY(Sub, "sub", double, op_sub(a, b))
B2_FX_BINARY(Mul, "mul", double, op_mul(a, b))
B2_FX_BINARY(Div, "div", double, op_div(a, b))
B2_FX_BINARY(Lt, "lt", bool, (a < b))
B2_FX_BINARY(Le, "le", bool, (a <= b))
B2_FX_BINARY(Gt, "gt", bool, (a > b))
B2_FX_BINARY(Ge, "ge", bool, (a >= b))
B2_FX_BINARY(Eq, "eq", bool, (a == b))
B2_FX_BINARY(Ne, "ne", bool, (a != b))
B2_FX_BINARY(SEq, "eq", bool, str_eq(a, b))
B2_FX_BINARY(SNe, "ne", bool, !str_eq(a, b))
B2_FX_BINARY(And, "and", bool, (a & b))
B2_FX_BINARY(Or, "or", bool, (a | b))
#undef B2_FX_BINARY

template <class A> struct Not {
  static constexpr int NF = A::NF, NI = A::NI, NS = A::NS, NLS = A::NLS;
  static constexpr bool IS_STR = false;
  template <class R> __device__ static bool eval(const R& r, const Lits& l) { return !A::eval(r, l); }
  B2_SIG(return "not(" + A::sig() + ")";)
};
template <class A> struct Neg {
  static constexpr int NF = A::NF, NI = A::NI, NS = A::NS, NLS = A::NLS;
  static constexpr bool IS_STR = false;
  template <class R> __device__ static auto eval(const R& r, const Lits& l) -> decltype(A::eval(r, l)) { return -A::eval(r, l); }
  B2_SIG(return "neg(" + A::sig() + ")";)
};
template <class A> struct ToF64 {
  static constexpr int NF = A::NF, NI = A::NI, NS = A::NS, NLS = A::NLS;
  static constexpr bool IS_STR = false;
  template <class R> __device__ static double eval(const R& r, const Lits& l) { return (double)A::eval(r, l); }
  B2_SIG(return "f64(" + A::sig() + ")";)
};

// ---- type lists ----
template <class... T> struct List {
  static constexpr int N = sizeof...(T);
};
template <class L> struct ListTraits;
template <> struct ListTraits<List<>> {
  static constexpr int NF = 0, NI = 0, NS = 0, NLS = 0;
  B2_SIG(return "";)
};
template <class H, class... T> struct ListTraits<List<H, T...>> {
  using Rest = ListTraits<List<T...>>;
  static constexpr int NF = cmax(H::NF, Rest::NF), NI = cmax(H::NI, Rest::NI), NS = cmax(H::NS, Rest::NS), NLS = cmax(H::NLS, Rest::NLS);
  B2_SIG(return H::sig() + (sizeof...(T) ? "," + Rest::sig() : std::string());)
};

// evaluate a list of Float64 expressions into out[]
template <class L> struct EvalF64;
template <> struct EvalF64<List<>> {
  template <class R> __device__ static void run(const R&, const Lits&, double*) {}
};
template <class H, class... T> struct EvalF64<List<H, T...>> {
  template <class R> __device__ static void run(const R& r, const Lits& l, double* out) {
    out[0] = H::eval(r, l);
    EvalF64<List<T...>>::run(r, l, out + 1);
  }
};

// ---- key packing: ints take 4 bytes, each Utf8 key takes `SB` bytes (1 length byte + SB-1 chars),
//      placed in list order from byte 0; byte 15 stays 0.  Host mirror: KeyLayout in agg.h ----
// KW = u128 (15 payload bytes) or u64 (8 payload bytes, narrow kernels)
template <class L, int SB> struct PackKey;
template <int SB> struct PackKey<List<>, SB> {
  template <class R, class KW> __device__ static void run(const R&, const Lits&, KW&, int, bool&) {}
};
template <class H, class... T, int SB> struct PackKey<List<H, T...>, SB> {
  template <class R, class KW> __device__ static void run(const R& r, const Lits& l, KW& key, int byte_off, bool& overflow) {
    auto v = H::eval(r, l);
    int adv = put(v, key, byte_off, overflow);
    PackKey<List<T...>, SB>::run(r, l, key, byte_off + adv, overflow);
  }
  template <class KW> __device__ static int put(int32_t v, KW& key, int off, bool&) {
    key |= (KW)(uint32_t)v << (8 * off);
    return 4;
  }
  template <class W, class KW> __device__ static int put(const StrValT<W>& s, KW& key, int off, bool& overflow) {
    if (s.len > SB - 1) overflow = true;
    const W m = (SB - 1 >= (int)sizeof(W)) ? ~(W)0 : (((W)1 << (8 * (SB - 1))) - 1);
    key |= ((KW)(uint8_t)s.len | ((KW)(s.bytes & m) << 8)) << (8 * off);
    return SB;
  }
};
// re-pack a narrow (u64, string parts of SBN bytes) key into the wide layout (u128, string parts of SBW bytes)
template <class L, int SBN, int SBW> struct WidenKey;
template <int SBN, int SBW> struct WidenKey<List<>, SBN, SBW> {
  __device__ static void run(u64, int, u128&, int) {}
};
template <class H, class... T, int SBN, int SBW> struct WidenKey<List<H, T...>, SBN, SBW> {
  __device__ static void run(u64 nk, int off_n, u128& wk, int off_w) {
    constexpr int bn = H::IS_STR ? SBN : 4, bw = H::IS_STR ? SBW : 4;
    const u64 part = bn >= 8 ? nk : ((nk >> (8 * off_n)) & ((1ull << (8 * (bn & 7))) - 1));
    wk |= (u128)part << (8 * off_w);
    WidenKey<List<T...>, SBN, SBW>::run(nk, off_n + bn, wk, off_w + bw);
  }
};
template <class L> struct CountStr;
template <> struct CountStr<List<>> { static constexpr int N = 0, INTS = 0; };
template <class H, class... T> struct CountStr<List<H, T...>> {
  static constexpr bool is_str = H::IS_STR;
  static constexpr int N = CountStr<List<T...>>::N + (is_str ? 1 : 0);
  static constexpr int INTS = CountStr<List<T...>>::INTS + (is_str ? 0 : 1);
};
constexpr int str_budget(int n_int, int n_str) { return n_str ? ((15 - 4 * n_int) / n_str > 15 ? 15 : (15 - 4 * n_int) / n_str) : 0; }
// narrow keys: 8 bytes in total, strings of at most 7 characters (+ 1 length byte)
constexpr int str_budget_narrow(int n_int, int n_str) { return n_str ? ((8 - 4 * n_int) / n_str > 8 ? 8 : (8 - 4 * n_int) / n_str) : 0; }
constexpr bool narrow_key_ok(int n_int, int n_str) { return 4 * n_int <= 8 && (n_str == 0 || str_budget_narrow(n_int, n_str) >= 2); }

}  // namespace fx
}  // namespace b2
