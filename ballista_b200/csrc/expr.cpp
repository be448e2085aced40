// expr.cpp — decode / type / coerce / fold / canonicalise expressions.
#include "expr.h"

#include <cstdio>
#include <cstring>

namespace b2 {

static bool is_int(int dt) { return dt == B2_INT32 || dt == B2_INT64 || dt == B2_UINT64 || dt == B2_UINT8; }
static bool is_num(int dt) { return is_int(dt) || dt == B2_FLOAT64; }
static int num_rank(int dt) {
  switch (dt) {
    case B2_UINT8: return 1;
    case B2_INT32: return 2;
    case B2_INT64: return 3;
    case B2_UINT64: return 3;
    case B2_FLOAT64: return 4;
  }
  return 0;
}

int32_t parse_date32(const std::string& s) {
  int y, m, d;
  if (sscanf(s.c_str(), "%d-%d-%d", &y, &m, &d) != 3)
    fail(B2_ERR_EXECUTION, "cannot cast '" + s + "' to Date32");
  // days from civil (proleptic Gregorian)
  y -= m <= 2;
  int era = (y >= 0 ? y : y - 399) / 400;
  unsigned yoe = (unsigned)(y - era * 400);
  unsigned doy = (153 * (m + (m > 2 ? -3 : 9)) + 2) / 5 + d - 1;
  unsigned doe = yoe * 365 + yoe / 4 - yoe / 100 + doy;
  return era * 146097 + (int)doe - 719468;
}

static ExprPtr mk_lit_i(int dt, int64_t v) {
  auto e = std::make_shared<Expr>();
  e->kind = B2_EXPR_LITERAL;
  e->dtype = dt;
  e->i = v;
  e->f = (double)v;
  return e;
}
static ExprPtr mk_lit_f(double v) {
  auto e = std::make_shared<Expr>();
  e->kind = B2_EXPR_LITERAL;
  e->dtype = B2_FLOAT64;
  e->f = v;
  return e;
}

static ExprPtr make_cast(ExprPtr c, int target) {
  if (c->dtype == target) return c;
  if (c->is_literal()) {
    if (c->dtype == B2_UTF8 && target == B2_DATE32) return mk_lit_i(B2_DATE32, parse_date32(c->s));
    if (is_int(c->dtype) && target == B2_FLOAT64) return mk_lit_f((double)c->i);
    if (is_int(c->dtype) && (is_int(target) || target == B2_DATE32)) return mk_lit_i(target, c->i);
    if (c->dtype == B2_DATE32 && is_int(target)) return mk_lit_i(target, c->i);
    if (c->dtype == B2_FLOAT64 && is_int(target)) return mk_lit_i(target, (int64_t)c->f);
    fail(B2_ERR_NOT_IMPLEMENTED,
         std::string("CAST of ") + dtype_name(c->dtype) + " literal to " + dtype_name(target));
  }
  auto in = [&](std::initializer_list<int> l) {
    for (int t : l)
      if (t == target) return true;
    return false;
  };
  bool ok = false;  // the pairs eval.cu has a kernel (or a reinterpretation) for
  switch (c->dtype) {
    case B2_INT32: ok = in({B2_INT64, B2_UINT64, B2_FLOAT64, B2_DATE32}); break;
    case B2_DATE32: ok = in({B2_INT32, B2_INT64}); break;
    case B2_INT64: ok = in({B2_FLOAT64, B2_INT32, B2_DATE32, B2_UINT64}); break;
    case B2_UINT64: ok = in({B2_FLOAT64, B2_INT32, B2_INT64}); break;
    case B2_UINT8: ok = in({B2_INT64, B2_INT32, B2_UINT64, B2_FLOAT64}); break;
    case B2_BOOL: ok = in({B2_INT32, B2_INT64, B2_UINT64, B2_FLOAT64, B2_UINT8}); break;
    case B2_FLOAT64: ok = in({B2_INT64, B2_INT32}); break;
  }
  if (!ok)
    fail(B2_ERR_NOT_IMPLEMENTED, std::string("CAST ") + dtype_name(c->dtype) + " -> " + dtype_name(target));
  auto e = std::make_shared<Expr>();
  e->kind = B2_EXPR_CAST;
  e->dtype = target;
  e->ch = {c};
  return e;
}

static ExprPtr make_binary(int op, ExprPtr a, ExprPtr b);

static ExprPtr fold_binary(int op, const ExprPtr& a, const ExprPtr& b, int t) {
  // both literals of common type t
  bool cmp = op >= B2_OP_EQ && op <= B2_OP_GTEQ;
  if (t == B2_FLOAT64) {
    double x = a->f, y = b->f;
    switch (op) {
      case B2_OP_PLUS: return mk_lit_f(x + y);
      case B2_OP_MINUS: return mk_lit_f(x - y);
      case B2_OP_MULTIPLY: return mk_lit_f(x * y);
      case B2_OP_DIVIDE: return mk_lit_f(x / y);
      case B2_OP_EQ: return mk_lit_i(B2_BOOL, x == y);
      case B2_OP_NOTEQ: return mk_lit_i(B2_BOOL, x != y);
      case B2_OP_LT: return mk_lit_i(B2_BOOL, x < y);
      case B2_OP_LTEQ: return mk_lit_i(B2_BOOL, x <= y);
      case B2_OP_GT: return mk_lit_i(B2_BOOL, x > y);
      case B2_OP_GTEQ: return mk_lit_i(B2_BOOL, x >= y);
    }
  } else if (t == B2_UTF8) {
    int c = a->s.compare(b->s);
    switch (op) {
      case B2_OP_EQ: return mk_lit_i(B2_BOOL, c == 0);
      case B2_OP_NOTEQ: return mk_lit_i(B2_BOOL, c != 0);
      case B2_OP_LT: return mk_lit_i(B2_BOOL, c < 0);
      case B2_OP_LTEQ: return mk_lit_i(B2_BOOL, c <= 0);
      case B2_OP_GT: return mk_lit_i(B2_BOOL, c > 0);
      case B2_OP_GTEQ: return mk_lit_i(B2_BOOL, c >= 0);
    }
  } else if (t == B2_BOOL) {
    switch (op) {
      case B2_OP_AND: return mk_lit_i(B2_BOOL, a->i && b->i);
      case B2_OP_OR: return mk_lit_i(B2_BOOL, a->i || b->i);
    }
  } else {
    int64_t x = a->i, y = b->i;
    switch (op) {
      case B2_OP_PLUS: return mk_lit_i(t, (int64_t)((uint64_t)x + (uint64_t)y));
      case B2_OP_MINUS: return mk_lit_i(t, (int64_t)((uint64_t)x - (uint64_t)y));
      case B2_OP_MULTIPLY: return mk_lit_i(t, (int64_t)((uint64_t)x * (uint64_t)y));
      case B2_OP_DIVIDE:
        if (y == 0) fail(B2_ERR_EXECUTION, "Divide by zero");
        return mk_lit_i(t, x / y);
      case B2_OP_EQ: return mk_lit_i(B2_BOOL, x == y);
      case B2_OP_NOTEQ: return mk_lit_i(B2_BOOL, x != y);
      case B2_OP_LT: return mk_lit_i(B2_BOOL, x < y);
      case B2_OP_LTEQ: return mk_lit_i(B2_BOOL, x <= y);
      case B2_OP_GT: return mk_lit_i(B2_BOOL, x > y);
      case B2_OP_GTEQ: return mk_lit_i(B2_BOOL, x >= y);
    }
  }
  (void)cmp;
  return nullptr;
}

static ExprPtr make_binary(int op, ExprPtr a, ExprPtr b) {
  auto e = std::make_shared<Expr>();
  e->kind = B2_EXPR_BINARY;
  e->op = op;
  bool logical = op == B2_OP_AND || op == B2_OP_OR;
  bool cmp = op >= B2_OP_EQ && op <= B2_OP_GTEQ;
  bool arith = op >= B2_OP_PLUS && op <= B2_OP_DIVIDE;
  int t = 0;
  if (op == B2_OP_LIKE) {
    if (a->dtype != B2_UTF8 || b->dtype != B2_UTF8) fail(B2_ERR_EXECUTION, "LIKE needs Utf8 operands");
    if (!b->is_literal()) fail(B2_ERR_NOT_IMPLEMENTED, "LIKE with a non-literal pattern");
    if (a->is_literal()) fail(B2_ERR_NOT_IMPLEMENTED, "LIKE over a literal value");
    e->dtype = B2_BOOL;
    e->ch = {a, b};
    return e;
  }
  if (logical) {
    if (a->dtype != B2_BOOL || b->dtype != B2_BOOL)
      fail(B2_ERR_EXECUTION, "AND/OR need Boolean operands");
    t = B2_BOOL;
    e->dtype = B2_BOOL;
  } else if (cmp || arith) {
    if (is_num(a->dtype) && is_num(b->dtype)) {
      t = num_rank(a->dtype) >= num_rank(b->dtype) ? a->dtype : b->dtype;
      if (t == B2_UINT8) t = B2_INT64;
      // an integer literal against an Int32 column stays Int32 when it fits (the reference coerces the
      // literal's side with a CAST; the comparison result is identical)
      if (t == B2_INT64) {
        const ExprPtr& lit = a->is_literal() ? a : b;
        const ExprPtr& other = a->is_literal() ? b : a;
        if (lit->is_literal() && !other->is_literal() && other->dtype == B2_INT32 && cmp &&
            lit->i >= INT32_MIN && lit->i <= INT32_MAX)
          t = B2_INT32;
      }
    } else if (a->dtype == B2_DATE32 || b->dtype == B2_DATE32) {
      // the reference has none either: DataFusion@46161d2's binary coercion accepts numeric operands only
      if (arith) fail(B2_ERR_NOT_IMPLEMENTED, "arithmetic on Date32");
      t = B2_DATE32;
    } else if (a->dtype == B2_UTF8 && b->dtype == B2_UTF8) {
      if (!cmp) fail(B2_ERR_EXECUTION, "arithmetic on Utf8");
      t = B2_UTF8;  // byte-wise ordering (Rust str comparison)
    } else if (a->dtype == B2_BOOL && b->dtype == B2_BOOL && (op == B2_OP_EQ || op == B2_OP_NOTEQ)) {
      fail(B2_ERR_NOT_IMPLEMENTED, "Boolean equality");
    } else {
      fail(B2_ERR_EXECUTION, std::string("cannot coerce ") + dtype_name(a->dtype) + " and " +
                                 dtype_name(b->dtype) + " for a binary operator");
    }
    a = make_cast(a, t);
    b = make_cast(b, t);
    e->dtype = cmp ? B2_BOOL : t;
  } else {
    fail(B2_ERR_NOT_IMPLEMENTED, "binary operator " + std::to_string(op) + " (LIKE etc.)");
  }
  if (a->is_literal() && b->is_literal()) {
    if (auto f = fold_binary(op, a, b, t)) return f;
  }
  e->ch = {a, b};
  return e;
}

static ExprPtr make_not(ExprPtr c) {
  if (c->dtype != B2_BOOL) fail(B2_ERR_EXECUTION, "NOT needs a Boolean operand");
  if (c->is_literal()) return mk_lit_i(B2_BOOL, !c->i);
  auto e = std::make_shared<Expr>();
  e->kind = B2_EXPR_NOT;
  e->dtype = B2_BOOL;
  e->ch = {c};
  return e;
}

ExprPtr decode_expr(const b2_expr* ex, const Schema& input) {
  if (!ex || !ex->nodes || ex->n_nodes <= 0) fail(B2_ERR_INVALID_ARG, "empty expression");
  std::vector<ExprPtr> st;
  auto pop = [&]() {
    if (st.empty()) fail(B2_ERR_INVALID_ARG, "malformed expression (stack underflow)");
    ExprPtr e = st.back();
    st.pop_back();
    return e;
  };
  for (int k = 0; k < ex->n_nodes; ++k) {
    const b2_expr_node& n = ex->nodes[k];
    switch (n.kind) {
      case B2_EXPR_COLUMN: {
        if (n.col < 0 || n.col >= (int)input.fields.size())
          fail(B2_ERR_INVALID_ARG, "column index " + std::to_string(n.col) + " out of range");
        auto e = std::make_shared<Expr>();
        e->kind = B2_EXPR_COLUMN;
        e->col = n.col;
        e->dtype = input.fields[n.col].dtype;
        st.push_back(e);
        break;
      }
      case B2_EXPR_LITERAL: {
        auto e = std::make_shared<Expr>();
        e->kind = B2_EXPR_LITERAL;
        e->dtype = n.dtype;
        switch (n.dtype) {
          case B2_FLOAT64: e->f = n.f; break;
          case B2_UTF8: e->s = n.s ? n.s : ""; break;
          case B2_INT32: case B2_INT64: case B2_UINT64: case B2_DATE32: case B2_BOOL: case B2_UINT8:
            e->i = n.i;
            e->f = (double)n.i;
            break;
          default: fail(B2_ERR_NOT_IMPLEMENTED, "literal type " + std::to_string(n.dtype));
        }
        st.push_back(e);
        break;
      }
      case B2_EXPR_BINARY: {
        ExprPtr b = pop(), a = pop();
        st.push_back(make_binary(n.op, a, b));
        break;
      }
      case B2_EXPR_NOT: st.push_back(make_not(pop())); break;
      case B2_EXPR_IS_NULL:
      case B2_EXPR_IS_NOT_NULL: {
        ExprPtr c = pop();
        if (c->is_literal()) {  // literals on this path are never NULL
          st.push_back(mk_lit_i(B2_BOOL, n.kind == B2_EXPR_IS_NOT_NULL));
          break;
        }
        auto e = std::make_shared<Expr>();
        e->kind = n.kind;
        e->dtype = B2_BOOL;
        e->ch = {c};
        st.push_back(e);
        break;
      }
      case B2_EXPR_CAST: st.push_back(make_cast(pop(), n.dtype)); break;
      case B2_EXPR_NEGATIVE: {
        ExprPtr c = pop();
        if (!is_num(c->dtype)) fail(B2_ERR_EXECUTION, "negation of a non-numeric value");
        if (c->is_literal()) {
          st.push_back(c->dtype == B2_FLOAT64 ? mk_lit_f(-c->f) : mk_lit_i(c->dtype, -c->i));
        } else {
          auto e = std::make_shared<Expr>();
          e->kind = B2_EXPR_NEGATIVE;
          e->dtype = c->dtype;
          e->ch = {c};
          st.push_back(e);
        }
        break;
      }
      case B2_EXPR_BETWEEN: {  // x BETWEEN lo AND hi  ==  x >= lo AND x <= hi   
        ExprPtr hi = pop(), lo = pop(), x = pop();
        ExprPtr r = make_binary(B2_OP_AND, make_binary(B2_OP_GTEQ, x, lo), make_binary(B2_OP_LTEQ, x, hi));
        st.push_back(n.op ? make_not(r) : r);
        break;
      }
      case B2_EXPR_IN_LIST: {
        int nl = n.nchildren - 1;
        if (nl < 1 || (int)st.size() < n.nchildren) fail(B2_ERR_INVALID_ARG, "malformed IN list");
        std::vector<ExprPtr> items(nl);
        for (int q = nl - 1; q >= 0; --q) items[q] = pop();
        ExprPtr x = pop();
        ExprPtr r = make_binary(B2_OP_EQ, x, items[0]);
        for (int q = 1; q < nl; ++q) r = make_binary(B2_OP_OR, r, make_binary(B2_OP_EQ, x, items[q]));
        st.push_back(n.op ? make_not(r) : r);
        break;
      }
      case B2_EXPR_CASE: {
        // children: [base] (when, then)* [else]   (CaseNode, ballista.proto:141-150)
        const bool has_base = n.op & 1, has_else = n.op & 2;
        int nc = n.nchildren;
        int pairs = nc - (has_base ? 1 : 0) - (has_else ? 1 : 0);
        if (pairs < 2 || (pairs & 1) || (int)st.size() < nc) fail(B2_ERR_INVALID_ARG, "malformed CASE expression");
        std::vector<ExprPtr> items(nc);
        for (int q = nc - 1; q >= 0; --q) items[q] = pop();
        size_t first = has_base ? 1 : 0;
        auto e = std::make_shared<Expr>();
        e->kind = B2_EXPR_CASE;
        e->op = has_else ? 2 : 0;
        // result type: the common type of the THEN / ELSE values
        int t = 0;
        auto widen = [&](const ExprPtr& v) {
          if (!t) t = v->dtype;
          else if (t != v->dtype) {
            if (!is_num(t) || !is_num(v->dtype))
              fail(B2_ERR_EXECUTION, std::string("CASE branches of types ") + dtype_name(t) + " and " + dtype_name(v->dtype));
            t = num_rank(t) >= num_rank(v->dtype) ? t : v->dtype;
          }
        };
        for (int q = 0; q < pairs / 2; ++q) widen(items[first + 2 * q + 1]);
        if (has_else) widen(items.back());
        if (t == B2_UINT8) t = B2_INT64;
        for (int q = 0; q < pairs / 2; ++q) {
          ExprPtr w = items[first + 2 * q];
          if (has_base) w = make_binary(B2_OP_EQ, items[0], w);  // CASE x WHEN v  ==  CASE WHEN x = v
          if (w->dtype != B2_BOOL) fail(B2_ERR_EXECUTION, "CASE WHEN needs a Boolean condition");
          e->ch.push_back(w);
          e->ch.push_back(make_cast(items[first + 2 * q + 1], t));
        }
        if (has_else) e->ch.push_back(make_cast(items.back(), t));
        if (t == B2_UTF8 && pairs / 2 > 8) fail(B2_ERR_NOT_IMPLEMENTED, "CASE producing Utf8 with more than 8 WHEN branches");
        e->dtype = t;
        st.push_back(e);
        break;
      }
      case B2_EXPR_SCALAR_FN: {
        if (n.op == 26 && n.nchildren == 2 && st.size() >= 2) {
          // NULLIF(a, b) (DataFusion nullif_func: eq(a, b), then arrow's nullif): NULL where a = b holds, else a.  A NULL
          // on either side makes the comparison NULL, which counts as "does not hold".  Same-typed arguments only (the
          // reference's Uniform(2, ...) signature leaves such arguments as they are; mixed types would be coerced by
          // rules that are not restated here).
          ExprPtr b = pop(), a = pop();
          const int t = a->dtype;
          const bool ok_type = t == B2_INT32 || t == B2_INT64 || t == B2_UINT64 || t == B2_FLOAT64 || t == B2_BOOL || t == B2_UINT8;
          if (!ok_type || b->dtype != t)
            fail(B2_ERR_NOT_IMPLEMENTED, std::string("nullif over ") + dtype_name(a->dtype) + " and " + dtype_name(b->dtype) +
                                             " (implemented: two arguments of one numeric or Boolean type)");
          if (a->is_literal()) fail(B2_ERR_NOT_IMPLEMENTED, "nullif over a literal first argument");
          auto e = std::make_shared<Expr>();
          e->kind = B2_EXPR_SCALAR_FN;
          e->op = n.op;
          e->dtype = t;
          e->ch = {a, make_binary(B2_OP_EQ, a, b)};
          st.push_back(e);
          break;
        }
        if (n.op == 18 && n.nchildren >= 1 && n.nchildren <= 8 && (int)st.size() >= n.nchildren) {
          // CONCAT(s1, s2, ...) over Utf8 arguments (the reference's Variadic(Utf8) signature would cast others to Utf8:
          // not restated).  See eval.cu for the NULL rule.
          std::vector<ExprPtr> args(n.nchildren);
          for (int q = n.nchildren - 1; q >= 0; --q) args[q] = pop();
          bool any_col = false;
          for (auto& a : args) {
            if (a->dtype != B2_UTF8) fail(B2_ERR_NOT_IMPLEMENTED, std::string("concat over a ") + dtype_name(a->dtype) + " argument");
            any_col = any_col || !a->is_literal();
          }
          if (!any_col) fail(B2_ERR_NOT_IMPLEMENTED, "concat over literals only");
          auto e = std::make_shared<Expr>();
          e->kind = B2_EXPR_SCALAR_FN;
          e->op = n.op;
          e->dtype = B2_UTF8;
          e->ch = args;
          st.push_back(e);
          break;
        }
        if (n.nchildren != 1 || st.empty())
          fail(B2_ERR_NOT_IMPLEMENTED, "scalar function " + std::to_string(n.op) + " with " + std::to_string(n.nchildren) +
                                           " arguments");
        ExprPtr c = pop();
        auto e = std::make_shared<Expr>();
        e->kind = B2_EXPR_SCALAR_FN;
        e->op = n.op;
        if (n.op >= 0 && n.op <= 16) {  // SQRT .. SIGNUM: Float64 in, Float64 out (math_unary_function)
          if (!is_num(c->dtype)) fail(B2_ERR_EXECUTION, "math function over a non-numeric value");
          c = make_cast(c, B2_FLOAT64);
          e->dtype = B2_FLOAT64;
        } else if (n.op == 17) {  // OCTETLENGTH
          if (c->dtype != B2_UTF8) fail(B2_ERR_EXECUTION, "octet_length needs a Utf8 argument");
          e->dtype = B2_INT32;
        } else if (n.op == 19 || n.op == 20) {  // LOWER / UPPER
          if (c->dtype != B2_UTF8) fail(B2_ERR_EXECUTION, "lower / upper need a Utf8 argument");
          e->dtype = B2_UTF8;
        } else if (n.op == 28) {  // MD5: Utf8 -> Utf8 (32 hex digits)
          if (c->dtype != B2_UTF8) fail(B2_ERR_EXECUTION, "md5 needs a Utf8 argument");
          e->dtype = B2_UTF8;
        } else if (n.op >= 21 && n.op <= 23) {  // TRIM / LTRIM / RTRIM
          if (c->dtype != B2_UTF8) fail(B2_ERR_EXECUTION, "trim / ltrim / rtrim need a Utf8 argument");
          e->dtype = B2_UTF8;
        } else {
          fail(B2_ERR_NOT_IMPLEMENTED, "scalar function " + std::to_string(n.op) +
                                           " (implemented: sqrt sin cos tan asin acos atan exp ln log2 log10 floor ceil "
                                           "round trunc abs signum octet_length lower upper trim ltrim rtrim nullif md5 concat)");
        }
        if (c->is_literal() && c->dtype != B2_FLOAT64) fail(B2_ERR_NOT_IMPLEMENTED, "scalar function over a literal");
        e->ch = {c};
        st.push_back(e);
        break;
      }
      default: fail(B2_ERR_INVALID_ARG, "unknown expression node kind " + std::to_string(n.kind));
    }
  }
  if (st.size() != 1) fail(B2_ERR_INVALID_ARG, "malformed expression (unbalanced)");
  ExprPtr top = st[0];
  if (ex->name) top->name = ex->name;
  return top;
}

static const char* op_text(int op) {
  switch (op) {
    case B2_OP_AND: return "and";
    case B2_OP_OR: return "or";
    case B2_OP_EQ: return "eq";
    case B2_OP_NOTEQ: return "ne";
    case B2_OP_LTEQ: return "le";
    case B2_OP_LT: return "lt";
    case B2_OP_GT: return "gt";
    case B2_OP_GTEQ: return "ge";
    case B2_OP_PLUS: return "add";
    case B2_OP_MINUS: return "sub";
    case B2_OP_MULTIPLY: return "mul";
    case B2_OP_DIVIDE: return "div";
    case B2_OP_LIKE: return "like";
  }
  return "?";
}

std::string expr_to_string(const ExprPtr& e, const Schema& input) {
  char buf[64];
  switch (e->kind) {
    case B2_EXPR_COLUMN: return input.fields[e->col].name;
    case B2_EXPR_LITERAL:
      if (e->dtype == B2_FLOAT64) { snprintf(buf, sizeof buf, "%.17g", e->f); return buf; }
      if (e->dtype == B2_UTF8) return "'" + e->s + "'";
      return std::to_string(e->i);
    case B2_EXPR_BINARY:
      return std::string(op_text(e->op)) + "(" + expr_to_string(e->ch[0], input) + "," +
             expr_to_string(e->ch[1], input) + ")";
    case B2_EXPR_NOT: return "not(" + expr_to_string(e->ch[0], input) + ")";
    case B2_EXPR_CAST: return std::string("cast_") + dtype_name(e->dtype) + "(" + expr_to_string(e->ch[0], input) + ")";
    case B2_EXPR_NEGATIVE: return "neg(" + expr_to_string(e->ch[0], input) + ")";
    case B2_EXPR_IS_NULL: return "isnull(" + expr_to_string(e->ch[0], input) + ")";
    case B2_EXPR_IS_NOT_NULL: return "isnotnull(" + expr_to_string(e->ch[0], input) + ")";
    case B2_EXPR_CASE:
    case B2_EXPR_SCALAR_FN: {
      std::string t = e->kind == B2_EXPR_CASE ? "case" + std::to_string(e->op) + "(" : "fn" + std::to_string(e->op) + "(";
      for (size_t k = 0; k < e->ch.size(); ++k) t += (k ? "," : "") + expr_to_string(e->ch[k], input);
      return t + ")";
    }
  }
  return "?";
}

// Signature grammar (must match the sig() of the device expression templates in fused.cuh):
//   f<k> i<k> s<k>        column slots (Float64 / Int32|Date32 / Utf8)
//   LF<k> LI<k> LS<k>     literal slots
//   op(a,b)               binary; and/or chains are kept as nested binaries in source order
//   not(a) neg(a) f64(a)  unary (f64 = cast of an int slot to Float64)
std::string expr_signature(const ExprPtr& e, SlotBinder& b) {
  switch (e->kind) {
    case B2_EXPR_COLUMN:
      if (e->dtype == B2_FLOAT64) return "f" + std::to_string(b.slot_for(b.f64_cols, e->col));
      if (e->dtype == B2_INT32 || e->dtype == B2_DATE32) return "i" + std::to_string(b.slot_for(b.i32_cols, e->col));
      if (e->dtype == B2_UTF8) return "s" + std::to_string(b.slot_for(b.str_cols, e->col));
      return "x" + std::to_string(e->dtype);  // no fused support
    case B2_EXPR_LITERAL:
      // equal literal values share a slot, so `p*(1-d)` and `p*(1-d)*(1+t)` expose their common subexpression
      if (e->dtype == B2_FLOAT64) {
        for (size_t k = 0; k < b.f64_lits.size(); ++k)
          if (memcmp(&b.f64_lits[k], &e->f, sizeof(double)) == 0) return "LF" + std::to_string(k);
        b.f64_lits.push_back(e->f);
        return "LF" + std::to_string(b.f64_lits.size() - 1);
      }
      if (e->dtype == B2_UTF8) {
        for (size_t k = 0; k < b.str_lits.size(); ++k)
          if (b.str_lits[k] == e->s) return "LS" + std::to_string(k);
        b.str_lits.push_back(e->s);
        return "LS" + std::to_string(b.str_lits.size() - 1);
      }
      for (size_t k = 0; k < b.i64_lits.size(); ++k)
        if (b.i64_lits[k] == e->i) return "LI" + std::to_string(k);
      b.i64_lits.push_back(e->i);
      return "LI" + std::to_string(b.i64_lits.size() - 1);
    case B2_EXPR_BINARY:
      if (e->op == B2_OP_AND) {
        // conjunctions are flattened and re-nested left-associatively so that every association of the
        // same conjunct sequence maps to one kernel shape
        std::vector<ExprPtr> leaves;
        std::vector<ExprPtr> work{e};
        while (!work.empty()) {
          ExprPtr x = work.back();
          work.pop_back();
          if (x->kind == B2_EXPR_BINARY && x->op == B2_OP_AND) {
            work.push_back(x->ch[1]);
            work.push_back(x->ch[0]);
          } else {
            leaves.push_back(x);
          }
        }
        std::string acc = expr_signature(leaves[0], b);
        for (size_t q = 1; q < leaves.size(); ++q) acc = "and(" + acc + "," + expr_signature(leaves[q], b) + ")";
        return acc;
      }
      {
        std::string l = expr_signature(e->ch[0], b);
        std::string r = expr_signature(e->ch[1], b);
        return std::string(op_text(e->op)) + "(" + l + "," + r + ")";
      }
    case B2_EXPR_NOT: return "not(" + expr_signature(e->ch[0], b) + ")";
    case B2_EXPR_NEGATIVE: return "neg(" + expr_signature(e->ch[0], b) + ")";
    case B2_EXPR_CAST:
      if (e->dtype == B2_FLOAT64) return "f64(" + expr_signature(e->ch[0], b) + ")";
      if (e->dtype == B2_INT32 || e->dtype == B2_DATE32) return expr_signature(e->ch[0], b);  // reinterpret
      return "cast" + std::to_string(e->dtype) + "(" + expr_signature(e->ch[0], b) + ")";
    default: {  // IS [NOT] NULL, CASE, scalar functions: no fused form
      std::string t = "x" + std::to_string(e->kind) + "_" + std::to_string(e->op) + "(";
      for (size_t k = 0; k < e->ch.size(); ++k) t += (k ? "," : "") + expr_signature(e->ch[k], b);
      return t + ")";
    }
  }
  return "?";
}

void expr_columns(const ExprPtr& e, std::vector<char>& used) {
  if (!e) return;
  if (e->kind == B2_EXPR_COLUMN && e->col >= 0) {
    if ((int)used.size() <= e->col) used.resize(e->col + 1, 0);
    used[e->col] = 1;
  }
  for (auto& c : e->ch) expr_columns(c, used);
}

}  // namespace b2
