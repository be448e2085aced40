// expr.h — physical expressions: decoded from the flat C-ABI IR (mirror of LogicalExprNode,
// rust/core/proto/ballista.proto:14-45), typed and coerced the way the reference recompiles them on the
// executor (compile_expr, rust/core/src/serde/physical_plan/from_proto.rs:348-364), constant-folded with
// IEEE f64 arithmetic, and canonicalised into a signature used to select a fused kernel.
#pragma once
#include <map>
#include <memory>
#include <string>
#include <vector>

#include "core.h"

namespace b2 {

struct Expr;
using ExprPtr = std::shared_ptr<Expr>;

struct Expr {
  int kind = 0;   // b2_expr_kind (BETWEEN / IN_LIST are lowered to AND/OR of comparisons at decode)
  int op = 0;     // b2_binary_op
  int dtype = 0;  // result type
  int col = -1;   // COLUMN
  int64_t i = 0;  // integer / date literal
  double f = 0;   // Float64 literal
  std::string s;  // Utf8 literal
  std::vector<ExprPtr> ch;
  std::string name;  // output field name (top level only)
  bool is_literal() const { return kind == B2_EXPR_LITERAL; }
};

// decode + type + coerce + fold.  Throws NotImplemented for vocabulary outside the TPC-H path.
ExprPtr decode_expr(const b2_expr* e, const Schema& input);
// human-readable form (Debug name)
std::string expr_to_string(const ExprPtr& e, const Schema& input);
// "yyyy-mm-dd" -> days since 1970-01-01 (CAST(Utf8 AS Date32))
int32_t parse_date32(const std::string& s);

// Slot binding shared by all expressions of one fused pipeline: input columns are numbered per type in
// first-encounter order; literals likewise.  The signature text is what the kernel registry is keyed on;
// literals and column bindings stay runtime parameters.
struct SlotBinder {
  std::vector<int> f64_cols, i32_cols, str_cols;  // input-schema index per slot
  std::vector<double> f64_lits;
  std::vector<int64_t> i64_lits;
  std::vector<std::string> str_lits;
  int slot_for(std::vector<int>& v, int col) {
    for (size_t k = 0; k < v.size(); ++k)
      if (v[k] == col) return (int)k;
    v.push_back(col);
    return (int)v.size() - 1;
  }
};
std::string expr_signature(const ExprPtr& e, SlotBinder& b);

// marks used[c] = 1 for every input column the expression reads (used is resized as needed)
void expr_columns(const ExprPtr& e, std::vector<char>& used);

// column-at-a-time evaluation over an HBM batch (the unfused path; eval.cu)
Column eval_expr(Ctx* ctx, const ExprPtr& e, const Batch& batch);
// a Boolean predicate as a filter mask (one byte per row, NULL -> 0)
Column eval_predicate(Ctx* ctx, const ExprPtr& pred, const Batch& batch);

}  // namespace b2
