// eval.cu — column-at-a-time expression evaluation over an HBM batch: the unfused path, with the same
// execution model as DataFusion's PhysicalExpr::evaluate: every node produces a full
// column; literals are broadcast as scalars.  Arithmetic is IEEE with one rounding per operation.
#include "expr.h"
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
struct OpDiv { template <class T> __device__ static T run(T a, T b) { return b == 0 ? 0 : a / b; } };
template <> __device__ double OpAdd::run<double>(double a, double b) { return __dadd_rn(a, b); }
template <> __device__ double OpSub::run<double>(double a, double b) { return __dsub_rn(a, b); }
template <> __device__ double OpMul::run<double>(double a, double b) { return __dmul_rn(a, b); }
template <> __device__ double OpDiv::run<double>(double a, double b) { return __ddiv_rn(a, b); }
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
template <class TI, class TO> __global__ void cast_kernel(const TI* in, int64_t n, TO* out) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  for (; i < n; i += (int64_t)gridDim.x * blockDim.x) out[i] = (TO)in[i];
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
// Utf8 equality: column vs literal, or column vs column
__global__ void str_eq_lit_kernel(const int32_t* off, const uint8_t* chr, int64_t n, const uint8_t* lit, int lit_len,
                                  bool negate, uint8_t* out) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  for (; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    int32_t b = off[i], e = off[i + 1];
    bool eq = (e - b) == lit_len;
    for (int q = 0; eq && q < lit_len; ++q) eq = chr[b + q] == lit[q];
    out[i] = eq != negate;
  }
}
__global__ void str_eq_col_kernel(const int32_t* off_a, const uint8_t* chr_a, const int32_t* off_b,
                                  const uint8_t* chr_b, int64_t n, bool negate, uint8_t* out) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  for (; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    int32_t ba = off_a[i], la = off_a[i + 1] - ba, bb = off_b[i], lb = off_b[i + 1] - bb;
    bool eq = la == lb;
    for (int q = 0; eq && q < la; ++q) eq = chr_a[ba + q] == chr_b[bb + q];
    out[i] = eq != negate;
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

template <class T>
static Column eval_binary_typed(Ctx* ctx, int op, const ExprPtr& ea, const Column& ca, const ExprPtr& eb,
                                const Column& cb, int64_t n, int arith_dtype) {
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
    case B2_OP_DIVIDE: launch_binary<T, T, OpDiv>(ctx, a, b, n, o); break;
    default: fail(B2_ERR_NOT_IMPLEMENTED, "binary operator");
  }
  return out;
}

Column eval_expr(Ctx* ctx, const ExprPtr& e, const Batch& batch) {
  int64_t n = batch.num_rows;
  int grid = grid_for(ctx, 8, n, 256);
  switch (e->kind) {
    case B2_EXPR_COLUMN: return batch.cols[e->col];
    case B2_EXPR_LITERAL: {
      if (e->dtype == B2_UTF8) fail(B2_ERR_NOT_IMPLEMENTED, "materialising a Utf8 literal column");
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
    case B2_EXPR_NOT: {
      Column c = eval_expr(ctx, e->ch[0], batch);
      Column out = new_col(ctx, B2_BOOL, n);
      if (n) {
        not_kernel<<<grid, 256, 0, ctx->stream>>>(c.values<uint8_t>(), n, out.data->as<uint8_t>());
        B2_LAUNCHED(ctx);
      }
      return out;
    }
    case B2_EXPR_NEGATIVE: {
      Column c = eval_expr(ctx, e->ch[0], batch);
      Column out = new_col(ctx, e->dtype, n);
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
      if ((from == B2_DATE32 && to == B2_INT32) || (from == B2_INT32 && to == B2_DATE32)) {
        c.dtype = to;
        return c;
      }
      Column out = new_col(ctx, to, n);
      if (n) {
        if (from == B2_INT32 && to == B2_INT64) cast_kernel<int32_t, int64_t><<<grid, 256, 0, ctx->stream>>>(c.values<int32_t>(), n, out.data->as<int64_t>());
        else if (from == B2_INT32 && to == B2_FLOAT64) cast_kernel<int32_t, double><<<grid, 256, 0, ctx->stream>>>(c.values<int32_t>(), n, out.data->as<double>());
        else if (from == B2_INT64 && to == B2_FLOAT64) cast_kernel<int64_t, double><<<grid, 256, 0, ctx->stream>>>(c.values<int64_t>(), n, out.data->as<double>());
        else if (from == B2_UINT8 && to == B2_INT64) cast_kernel<uint8_t, int64_t><<<grid, 256, 0, ctx->stream>>>(c.values<uint8_t>(), n, out.data->as<int64_t>());
        else fail(B2_ERR_NOT_IMPLEMENTED, "runtime CAST");
        B2_LAUNCHED(ctx);
      }
      return out;
    }
    case B2_EXPR_BINARY: {
      const ExprPtr &ea = e->ch[0], &eb = e->ch[1];
      int t = ea->dtype;  // operands were coerced to a common type at decode
      if (t == B2_UTF8) {
        bool neg = e->op == B2_OP_NOTEQ;
        Column out = new_col(ctx, B2_BOOL, n);
        if (!n) return out;
        if (ea->is_literal() || eb->is_literal()) {
          const ExprPtr& lit = ea->is_literal() ? ea : eb;
          Column c = eval_expr(ctx, ea->is_literal() ? eb : ea, batch);
          auto dlit = dev_alloc(ctx, lit->s.size() + 16);
          // literal bytes must outlive the async copy: keep them in the expression node
          B2_CUDA(cudaMemcpyAsync(dlit->ptr, lit->s.data(), lit->s.size(), cudaMemcpyHostToDevice, ctx->stream));
          str_eq_lit_kernel<<<grid, 256, 0, ctx->stream>>>(c.values<int32_t>(), c.chars->as<uint8_t>(), n,
                                                          dlit->as<uint8_t>(), (int)lit->s.size(), neg,
                                                          out.data->as<uint8_t>());
        } else {
          Column a = eval_expr(ctx, ea, batch), b = eval_expr(ctx, eb, batch);
          str_eq_col_kernel<<<grid, 256, 0, ctx->stream>>>(a.values<int32_t>(), a.chars->as<uint8_t>(),
                                                          b.values<int32_t>(), b.chars->as<uint8_t>(), n, neg,
                                                          out.data->as<uint8_t>());
        }
        B2_LAUNCHED(ctx);
        return out;
      }
      Column ca, cb;
      if (!ea->is_literal()) ca = eval_expr(ctx, ea, batch);
      if (!eb->is_literal()) cb = eval_expr(ctx, eb, batch);
      if (e->op == B2_OP_AND || e->op == B2_OP_OR) {
        Column out = new_col(ctx, B2_BOOL, n);
        Operand<uint8_t> a = operand_of<uint8_t>(ea, ca), b = operand_of<uint8_t>(eb, cb);
        if (e->op == B2_OP_AND) launch_binary<uint8_t, uint8_t, OpAnd>(ctx, a, b, n, out.data->as<uint8_t>());
        else launch_binary<uint8_t, uint8_t, OpOr>(ctx, a, b, n, out.data->as<uint8_t>());
        return out;
      }
      switch (t) {
        case B2_INT32: case B2_DATE32: return eval_binary_typed<int32_t>(ctx, e->op, ea, ca, eb, cb, n, e->dtype);
        case B2_INT64: return eval_binary_typed<int64_t>(ctx, e->op, ea, ca, eb, cb, n, e->dtype);
        case B2_FLOAT64: return eval_binary_typed<double>(ctx, e->op, ea, ca, eb, cb, n, e->dtype);
        default: fail(B2_ERR_NOT_IMPLEMENTED, std::string("binary operator on ") + dtype_name(t));
      }
    }
  }
  fail(B2_ERR_NOT_IMPLEMENTED, "expression kind " + std::to_string(e->kind));
}

}  // namespace b2
