// proto.cpp — wire compatibility for the physical-plan subtree: decodes a serialized `PhysicalPlanNode`
// (rust/core/proto/ballista.proto:294-422) with its `LogicalExprNode` / `ScalarValue` / `ArrowType` / `Schema`
// payloads (ballista.proto:14-161, 611-795) straight into the operator tree, doing what the executor does in
// rust/core/src/serde/physical_plan/from_proto.rs:58-346 — so a stage plan as the scheduler stores and ships it
// runs unmodified.  A hand-written proto3 reader (varint / length-delimited / fixed32 / fixed64): `protoc` and a
// C++ protobuf runtime are not part of the build.
//
// Leaves (CsvScanExecNode, ParquetScanExecNode, ShuffleReaderExecNode, UnresolvedShuffleExecNode, EmptyExecNode) are
// not decoded into readers: the i-th leaf in pre-order (left to right) is bound to the i-th operator the caller
// passes (a MemoryExec fed by the CPU reader / the previous stage's output), see INTEGRATION.md.
#include <cstring>
#include <deque>
#include <sstream>

#include "ops.h"

namespace b2 {
namespace {

struct Slice {
  const uint8_t* p = nullptr;
  const uint8_t* e = nullptr;
  bool empty() const { return p >= e; }
  std::string str() const { return std::string(reinterpret_cast<const char*>(p), (size_t)(e - p)); }
};
[[noreturn]] void bad(const std::string& what) { fail(B2_ERR_INVALID_ARG, "malformed plan protobuf: " + what); }

uint64_t varint(Slice& s) {
  uint64_t v = 0;
  for (int shift = 0; shift < 70; shift += 7) {
    if (s.p >= s.e) bad("truncated varint");
    uint8_t b = *s.p++;
    v |= (uint64_t)(b & 0x7F) << shift;
    if (!(b & 0x80)) return v;
  }
  bad("varint longer than 10 bytes");
}
struct Field {
  uint32_t number = 0;
  int wire = 0;     // 0 varint, 1 fixed64, 2 length-delimited, 5 fixed32
  uint64_t v = 0;   // varint / fixed value
  Slice sub;        // length-delimited payload
};
bool next_field(Slice& s, Field* f) {
  if (s.empty()) return false;
  uint64_t tag = varint(s);
  f->number = (uint32_t)(tag >> 3);
  f->wire = (int)(tag & 7);
  f->v = 0;
  f->sub = Slice();
  switch (f->wire) {
    case 0: f->v = varint(s); break;
    case 1:
      if (s.e - s.p < 8) bad("truncated fixed64");
      memcpy(&f->v, s.p, 8);
      s.p += 8;
      break;
    case 2: {
      uint64_t n = varint(s);
      if ((uint64_t)(s.e - s.p) < n) bad("truncated length-delimited field");
      f->sub.p = s.p;
      f->sub.e = s.p + n;
      s.p += n;
      break;
    }
    case 5: {
      if (s.e - s.p < 4) bad("truncated fixed32");
      uint32_t x;
      memcpy(&x, s.p, 4);
      f->v = x;
      s.p += 4;
      break;
    }
    default: bad("unsupported wire type " + std::to_string(f->wire));
  }
  return true;
}
// the single populated member of a oneof-only message (PhysicalPlanNode, LogicalExprNode, ScalarValue, ArrowType)
Field only_field(Slice s, const char* what) {
  Field f, last;
  bool any = false;
  while (next_field(s, &f)) {
    last = f;
    any = true;
  }
  if (!any) fail(B2_ERR_INVALID_ARG, std::string(what) + " is empty (no oneof member set)");
  return last;
}

// ---- ArrowType (ballista.proto:739-775) -> b2_dtype ----
int dtype_of_arrow_type(Slice s) {
  Field f = only_field(s, "ArrowType");
  switch (f.number) {
    case 2: return B2_BOOL;
    case 3: return B2_UINT8;
    case 8: return B2_INT32;
    case 9: return B2_UINT64;
    case 10: return B2_INT64;
    case 13: return B2_FLOAT64;
    case 14: return B2_UTF8;
    case 17: return B2_DATE32;
    default: fail(B2_ERR_NOT_IMPLEMENTED, "ArrowType variant " + std::to_string(f.number) + " (outside the TPC-H path types)");
  }
}
Schema decode_schema(Slice s) {  // Schema{columns=1}: Field{name=1, arrow_type=2, nullable=3}
  Schema out;
  Field f;
  while (next_field(s, &f)) {
    if (f.number != 1 || f.wire != 2) continue;
    b2::Field fld;
    Slice fs = f.sub;
    Field g;
    while (next_field(fs, &g)) {
      if (g.number == 1 && g.wire == 2) fld.name = g.sub.str();
      else if (g.number == 2 && g.wire == 2) fld.dtype = dtype_of_arrow_type(g.sub);
      else if (g.number == 3) fld.nullable = g.v != 0;
    }
    out.fields.push_back(fld);
  }
  return out;
}

// ---- LogicalExprNode -> the flat post-order IR of include/ballista_b200.h ----
struct ExprBuilder {
  const Schema& schema;
  std::vector<b2_expr_node> nodes;
  std::deque<std::string> strings;  // stable storage for Utf8 literals
  std::string alias;
  explicit ExprBuilder(const Schema& s) : schema(s) {}

  b2_expr_node blank(int kind) {
    b2_expr_node n;
    memset(&n, 0, sizeof n);
    n.kind = kind;
    n.col = -1;
    return n;
  }
  Slice member(Slice msg, uint32_t number, const char* what, bool required = true) {
    Field f;
    Slice found;
    bool ok = false;
    while (next_field(msg, &f))
      if (f.number == number && f.wire == 2) {
        found = f.sub;
        ok = true;
      }
    if (!ok && required) fail(B2_ERR_INVALID_ARG, std::string(what) + " is missing");
    return found;
  }
  void literal(Slice sv) {  // ScalarValue (ballista.proto:686-709)
    Field f = only_field(sv, "ScalarValue");
    b2_expr_node n = blank(B2_EXPR_LITERAL);
    switch (f.number) {
      case 1: n.dtype = B2_BOOL; n.i = f.v != 0; break;
      case 2: case 3:
        strings.push_back(f.sub.str());
        n.dtype = B2_UTF8;
        n.s = strings.back().c_str();
        break;
      case 4: case 5: case 6: n.dtype = B2_INT32; n.i = (int64_t)(int32_t)f.v; break;
      case 7: n.dtype = B2_INT64; n.i = (int64_t)f.v; break;
      case 8: n.dtype = B2_UINT8; n.i = (int64_t)f.v; break;
      case 9: case 10: n.dtype = B2_INT64; n.i = (int64_t)f.v; break;
      case 11: n.dtype = B2_UINT64; n.i = (int64_t)f.v; break;
      case 12: {
        float x;
        uint32_t bits = (uint32_t)f.v;
        memcpy(&x, &bits, 4);
        n.dtype = B2_FLOAT64;
        n.f = (double)x;
        break;
      }
      case 13: n.dtype = B2_FLOAT64; memcpy(&n.f, &f.v, 8); break;
      case 14: n.dtype = B2_DATE32; n.i = (int64_t)(int32_t)f.v; break;
      default: fail(B2_ERR_NOT_IMPLEMENTED, "ScalarValue variant " + std::to_string(f.number));
    }
    nodes.push_back(n);
  }
  static int binary_op(const std::string& op) {  // logical_plan/from_proto.rs:937-957
    static const char* names[] = {"And", "Or", "Eq", "NotEq", "LtEq", "Lt", "Gt", "GtEq", "Plus", "Minus", "Multiply",
                                  "Divide", "Like"};
    for (int k = 0; k < 13; ++k)
      if (op == names[k]) return B2_OP_AND + k;
    fail(B2_ERR_INVALID_ARG, "Unsupported binary operator '" + op + "'");
  }
  void unary(Slice msg, int kind, const char* what) {
    emit(member(msg, 1, what));
    nodes.push_back(blank(kind));
  }
  void emit(Slice expr) {
    Field f = only_field(expr, "LogicalExprNode");
    switch (f.number) {
      case 1: {  // column_name
        std::string name = f.sub.str();
        int idx = schema.index_of(name);
        if (idx < 0) {
          std::string valid;
          for (auto& fl : schema.fields) valid += (valid.empty() ? "" : ", ") + fl.name;
          fail(B2_ERR_INVALID_ARG, "No field named '" + name + "'. Valid fields are " + valid);
        }
        b2_expr_node n = blank(B2_EXPR_COLUMN);
        n.col = idx;
        nodes.push_back(n);
        break;
      }
      case 2: {  // AliasNode{expr=1, alias=2}
        Slice m = f.sub;
        Field g;
        Slice inner;
        while (next_field(m, &g)) {
          if (g.number == 1 && g.wire == 2) inner = g.sub;
          if (g.number == 2 && g.wire == 2) alias = g.sub.str();
        }
        emit(inner);
        break;
      }
      case 3: literal(f.sub); break;
      case 4: {  // BinaryExprNode{l=1, r=2, op=3}
        Slice m = f.sub;
        Field g;
        Slice l, r;
        std::string op;
        while (next_field(m, &g)) {
          if (g.number == 1 && g.wire == 2) l = g.sub;
          if (g.number == 2 && g.wire == 2) r = g.sub;
          if (g.number == 3 && g.wire == 2) op = g.sub.str();
        }
        emit(l);
        emit(r);
        b2_expr_node n = blank(B2_EXPR_BINARY);
        n.op = binary_op(op);
        nodes.push_back(n);
        break;
      }
      case 5: fail(B2_ERR_INVALID_ARG, "aggregate expression outside HashAggregateExecNode.aggr_expr");
      case 6: unary(f.sub, B2_EXPR_IS_NULL, "IsNull.expr"); break;
      case 7: unary(f.sub, B2_EXPR_IS_NOT_NULL, "IsNotNull.expr"); break;
      case 8: unary(f.sub, B2_EXPR_NOT, "Not.expr"); break;
      case 9: {  // BetweenNode{expr=1, negated=2, low=3, high=4}
        Slice m = f.sub;
        Field g;
        Slice x, lo, hi;
        bool neg = false;
        while (next_field(m, &g)) {
          if (g.number == 1 && g.wire == 2) x = g.sub;
          if (g.number == 2) neg = g.v != 0;
          if (g.number == 3 && g.wire == 2) lo = g.sub;
          if (g.number == 4 && g.wire == 2) hi = g.sub;
        }
        emit(x);
        emit(lo);
        emit(hi);
        b2_expr_node n = blank(B2_EXPR_BETWEEN);
        n.op = neg;
        nodes.push_back(n);
        break;
      }
      case 10: {  // CaseNode{expr=1, when_then_expr=2 repeated WhenThen{when_expr=1, then_expr=2}, else_expr=3}
        Slice m = f.sub;
        Field g;
        Slice base, els;
        bool has_base = false, has_else = false;
        std::vector<Slice> wt;
        while (next_field(m, &g)) {
          if (g.number == 1 && g.wire == 2) { base = g.sub; has_base = true; }
          if (g.number == 2 && g.wire == 2) wt.push_back(g.sub);
          if (g.number == 3 && g.wire == 2) { els = g.sub; has_else = true; }
        }
        if (wt.empty()) fail(B2_ERR_INVALID_ARG, "CaseNode without WHEN / THEN");
        if (has_base) emit(base);
        for (auto& w : wt) {
          emit(member(w, 1, "WhenThen.when_expr"));
          emit(member(w, 2, "WhenThen.then_expr"));
        }
        if (has_else) emit(els);
        b2_expr_node n = blank(B2_EXPR_CASE);
        n.op = (has_base ? 1 : 0) | (has_else ? 2 : 0);
        n.nchildren = (has_base ? 1 : 0) + 2 * (int)wt.size() + (has_else ? 1 : 0);
        nodes.push_back(n);
        break;
      }
      case 11: {  // CastNode{expr=1, arrow_type=2}
        emit(member(f.sub, 1, "CastNode.expr"));
        b2_expr_node n = blank(B2_EXPR_CAST);
        n.dtype = dtype_of_arrow_type(member(f.sub, 2, "CastNode.arrow_type"));
        nodes.push_back(n);
        break;
      }
      case 12: fail(B2_ERR_INVALID_ARG, "sort expression outside SortExecNode.expr");
      case 13: unary(f.sub, B2_EXPR_NEGATIVE, "NegativeNode.expr"); break;
      case 14: {  // InListNode{expr=1, list=2 repeated, negated=3}
        Slice m = f.sub;
        Field g;
        Slice x;
        std::vector<Slice> list;
        bool neg = false;
        while (next_field(m, &g)) {
          if (g.number == 1 && g.wire == 2) x = g.sub;
          if (g.number == 2 && g.wire == 2) list.push_back(g.sub);
          if (g.number == 3) neg = g.v != 0;
        }
        emit(x);
        for (auto& it : list) emit(it);
        b2_expr_node n = blank(B2_EXPR_IN_LIST);
        n.nchildren = 1 + (int)list.size();
        n.op = neg;
        nodes.push_back(n);
        break;
      }
      case 16: {  // ScalarFunctionNode{fun=1 (enum, default SQRT = 0), expr=2 repeated}
        Slice m = f.sub;
        Field g;
        int fun = 0;
        std::vector<Slice> args;
        while (next_field(m, &g)) {
          if (g.number == 1 && g.wire == 0) fun = (int)g.v;
          if (g.number == 2 && g.wire == 2) args.push_back(g.sub);
        }
        for (auto& a : args) emit(a);
        b2_expr_node n = blank(B2_EXPR_SCALAR_FN);
        n.op = fun;
        n.nchildren = (int)args.size();
        nodes.push_back(n);
        break;
      }
      default: fail(B2_ERR_NOT_IMPLEMENTED, "LogicalExprNode variant " + std::to_string(f.number));
    }
  }
};

// compile_expr (from_proto.rs:348-364): LogicalExprNode + input schema -> typed / coerced / folded expression
ExprPtr compile_expr(Slice expr, const Schema& schema, const std::string& name) {
  ExprBuilder b(schema);
  b.emit(expr);
  std::string nm = !name.empty() ? name : b.alias;
  b2_expr e;
  e.nodes = b.nodes.data();
  e.n_nodes = (int32_t)b.nodes.size();
  e.name = nm.empty() ? nullptr : nm.c_str();
  return decode_expr(&e, schema);
}

// ---- PhysicalPlanNode ----
struct PlanDecoder {
  Ctx* ctx;                     // null: describe only
  const std::vector<OpPtr>* leaves;
  size_t next_leaf = 0;
  std::ostringstream text;      // pre-order description (one node per line, two spaces per level)
  int leaf_count = 0;

  void line(int depth, const std::string& s) { text << std::string((size_t)depth * 2, ' ') << s << "\n"; }

  OpPtr leaf(int depth, const char* kind, const Schema* declared) {
    line(depth, std::string(kind) + " [leaf " + std::to_string(leaf_count) + "]");
    ++leaf_count;
    if (!ctx) return nullptr;
    if (next_leaf >= leaves->size())
      fail(B2_ERR_INVALID_ARG, "plan has more leaf nodes than operators were supplied (" + std::to_string(leaves->size()) + ")");
    OpPtr op = (*leaves)[next_leaf++];
    if (declared && !declared->fields.empty()) {  // the bound source must produce the columns the plan was built on
      Schema got = op->schema();
      if (got.fields.size() != declared->fields.size())
        fail(B2_ERR_INVALID_ARG, std::string(kind) + ": bound source has " + std::to_string(got.fields.size()) +
                                     " columns, the plan declares " + std::to_string(declared->fields.size()));
      for (size_t i = 0; i < got.fields.size(); ++i)
        if (got.fields[i].name != declared->fields[i].name)
          fail(B2_ERR_INVALID_ARG, std::string(kind) + ": column " + std::to_string(i) + " is '" + got.fields[i].name +
                                       "', the plan declares '" + declared->fields[i].name + "'");
    }
    return op;
  }
  // members of a node message, gathered first: inputs are decoded before the expressions that need their schema
  struct Members {
    std::vector<std::pair<uint32_t, Field>> all;
    Slice one(uint32_t number) const {
      Slice s;
      for (auto& kv : all)
        if (kv.first == number && kv.second.wire == 2) s = kv.second.sub;
      return s;
    }
    bool has(uint32_t number) const {
      for (auto& kv : all)
        if (kv.first == number) return true;
      return false;
    }
    uint64_t num(uint32_t number, uint64_t dflt = 0) const {
      uint64_t v = dflt;
      for (auto& kv : all)
        if (kv.first == number && kv.second.wire != 2) v = kv.second.v;
      return v;
    }
    std::vector<Slice> many(uint32_t number) const {
      std::vector<Slice> v;
      for (auto& kv : all)
        if (kv.first == number && kv.second.wire == 2) v.push_back(kv.second.sub);
      return v;
    }
    std::vector<uint64_t> packed(uint32_t number) const {  // repeated scalar: packed or one varint per element
      std::vector<uint64_t> v;
      for (auto& kv : all) {
        if (kv.first != number) continue;
        if (kv.second.wire == 2) {
          Slice s = kv.second.sub;
          while (!s.empty()) v.push_back(varint(s));
        } else {
          v.push_back(kv.second.v);
        }
      }
      return v;
    }
  };
  static Members members(Slice s) {
    Members m;
    Field f;
    while (next_field(s, &f)) m.all.emplace_back(f.number, f);
    return m;
  }
  OpPtr input_of(const Members& m, uint32_t number, int depth, const char* what) {
    if (!m.has(number)) fail(B2_ERR_INVALID_ARG, std::string(what) + " is missing its input");
    return node(m.one(number), depth);
  }
  std::vector<std::string> names(const Members& m, uint32_t number) {
    std::vector<std::string> v;
    for (auto& s : m.many(number)) v.push_back(s.str());
    return v;
  }

  OpPtr node(Slice plan, int depth) {
    Field f = only_field(plan, "PhysicalPlanNode");
    if (f.wire != 2) bad("PhysicalPlanNode member is not a message");
    Members m = members(f.sub);
    switch (f.number) {
      case 1: return leaf(depth, "ParquetScan", nullptr);
      case 2: {  // CsvScanExecNode{projection=2, schema=3}: the projected file schema is what the leaf produces
        Schema file = decode_schema(m.one(3)), proj;
        for (uint64_t i : m.packed(2)) {
          if (i >= file.fields.size()) fail(B2_ERR_INVALID_ARG, "CsvScan projection index out of range");
          proj.fields.push_back(file.fields[(size_t)i]);
        }
        return leaf(depth, "CsvScan", m.has(2) ? &proj : &file);
      }
      case 3: return leaf(depth, "Empty", nullptr);
      case 10: {
        Schema s = decode_schema(m.one(2));
        return leaf(depth, "ShuffleReader", &s);
      }
      case 15: {
        Schema s = decode_schema(m.one(2));
        return leaf(depth, "UnresolvedShuffle", &s);
      }
      case 4: {  // ProjectionExecNode{input=1, expr=2, expr_name=3}   from_proto.rs:69-80
        line(depth, "ProjectionExec");
        OpPtr in = input_of(m, 1, depth + 1, "ProjectionExecNode");
        if (!ctx) return nullptr;
        Schema s = in->schema();
        auto exprs = m.many(2);
        auto nm = names(m, 3);
        std::vector<ExprPtr> ev;
        for (size_t i = 0; i < exprs.size(); ++i) ev.push_back(compile_expr(exprs[i], s, i < nm.size() ? nm[i] : ""));
        return make_projection(ctx, in, std::move(ev));
      }
      case 6: case 7: {  // Global / LocalLimitExecNode{input=1, limit=2}   from_proto.rs:165-171
        line(depth, f.number == 6 ? "GlobalLimitExec" : "LocalLimitExec");
        OpPtr in = input_of(m, 1, depth + 1, "LimitExecNode");
        if (!ctx) return nullptr;
        return make_limit(ctx, in, (int64_t)m.num(2), f.number == 6);
      }
      case 8: return hash_aggregate(m, depth);
      case 9: {  // HashJoinExecNode{left=1, right=2, on=3{left=1,right=2}, join_type=4}   from_proto.rs:253-276
        line(depth, "HashJoinExec");
        OpPtr l = input_of(m, 1, depth + 1, "HashJoinExecNode.left");
        OpPtr r = input_of(m, 2, depth + 1, "HashJoinExecNode.right");
        std::vector<std::string> on_l, on_r;
        for (auto& on : m.many(3)) {
          Members om = members(on);
          on_l.push_back(om.one(1).str());
          on_r.push_back(om.one(2).str());
        }
        if (!ctx) return nullptr;
        return make_hash_join(ctx, l, r, on_l, on_r, (int)m.num(4));
      }
      case 11: {  // SortExecNode{input=1, expr=2: LogicalExprNode.sort{expr=1, asc=2, nulls_first=3}}   :291-330
        line(depth, "SortExec");
        OpPtr in = input_of(m, 1, depth + 1, "SortExecNode");
        if (!ctx) return nullptr;
        Schema s = in->schema();
        std::vector<ExprPtr> ev;
        std::vector<char> desc;
        for (auto& e : m.many(2)) {
          Field sf = only_field(e, "LogicalExprNode");
          if (sf.number != 12) fail(B2_ERR_INVALID_ARG, "physical_plan::from_proto() SortExecNode expects sort expressions");
          Members sm = members(sf.sub);
          ev.push_back(compile_expr(sm.one(1), s, ""));
          desc.push_back(sm.num(2) == 0);  // asc = false (absent) -> descending
        }
        return make_sort(ctx, in, std::move(ev), std::move(desc));
      }
      case 12: {  // CoalesceBatchesExecNode{input=1, target_batch_size=2}   :122-128
        line(depth, "CoalesceBatchesExec");
        OpPtr in = input_of(m, 1, depth + 1, "CoalesceBatchesExecNode");
        if (!ctx) return nullptr;
        return make_coalesce(ctx, in, (int64_t)m.num(2));
      }
      case 13: {  // FilterExecNode{input=1, expr=2}   :81-92
        line(depth, "FilterExec");
        OpPtr in = input_of(m, 1, depth + 1, "FilterExecNode");
        if (!m.has(2)) fail(B2_ERR_INVALID_ARG, "filter (FilterExecNode) in PhysicalPlanNode is missing.");
        if (!ctx) return nullptr;
        return make_filter(ctx, in, compile_expr(m.one(2), in->schema(), ""));
      }
      case 14: {  // MergeExecNode{input=1}   :129-132
        line(depth, "MergeExec");
        OpPtr in = input_of(m, 1, depth + 1, "MergeExecNode");
        if (!ctx) return nullptr;
        return make_merge(ctx, in);
      }
      case 16: {  // RepartitionExecNode{input=1, round_robin=2 | hash=3{hash_expr=1, partition_count=2} | unknown=4}   :133-164
        line(depth, "RepartitionExec");
        OpPtr in = input_of(m, 1, depth + 1, "RepartitionExecNode");
        if (!ctx) return nullptr;
        if (m.has(3)) {
          Members hm = members(m.one(3));
          Schema s = in->schema();
          std::vector<ExprPtr> keys;
          for (auto& e : hm.many(1)) keys.push_back(compile_expr(e, s, ""));
          return make_repartition(ctx, in, B2_PART_HASH, std::move(keys), (int)hm.num(2));
        }
        if (m.has(2)) return make_repartition(ctx, in, B2_PART_ROUND_ROBIN, {}, (int)m.num(2));
        fail(B2_ERR_NOT_IMPLEMENTED, "RepartitionExec with unknown partitioning");
      }
      default:
        fail(B2_ERR_NOT_IMPLEMENTED, "physical_plan::from_proto() Unsupported physical plan variant " + std::to_string(f.number));
    }
  }

  // HashAggregateExecNode{group_expr=1, aggr_expr=2, mode=3, input=4, group_expr_name=5, aggr_expr_name=6,
  // input_schema=7}   from_proto.rs:173-252
  OpPtr hash_aggregate(const Members& m, int depth) {
    line(depth, "HashAggregateExec");
    OpPtr in = input_of(m, 4, depth + 1, "HashAggregateExecNode");
    if (!ctx) return nullptr;
    int mode = (int)m.num(3);
    if (mode != B2_AGG_PARTIAL && mode != B2_AGG_FINAL)
      fail(B2_ERR_INVALID_ARG, "Received a HashAggregateNode message with unknown AggregateMode " + std::to_string(mode));
    if (!m.has(7)) fail(B2_ERR_INVALID_ARG, "input_schema in HashAggregateNode is missing.");
    Schema input_schema = decode_schema(m.one(7));
    Schema s = in->schema();
    auto gexprs = m.many(1), aexprs = m.many(2);
    auto gnames = names(m, 5), anames = names(m, 6);
    std::vector<ExprPtr> groups;
    for (size_t i = 0; i < gexprs.size(); ++i) groups.push_back(compile_expr(gexprs[i], s, i < gnames.size() ? gnames[i] : ""));
    std::vector<AggSpec> aggs;
    for (size_t i = 0; i < aexprs.size(); ++i) {
      Field af = only_field(aexprs[i], "LogicalExprNode");
      if (af.number != 5) fail(B2_ERR_INVALID_ARG, "Invalid expression for HashAggregateExec");
      Members am = members(af.sub);  // AggregateExprNode{aggr_function=1, expr=2}
      AggSpec sp;
      switch (am.num(1)) {  // enum AggregateFunction { MIN, MAX, SUM, AVG, COUNT }
        case 2: sp.fn = B2_AGG_SUM; break;
        case 3: sp.fn = B2_AGG_AVG; break;
        case 4: sp.fn = B2_AGG_COUNT; break;
        default: fail(B2_ERR_NOT_IMPLEMENTED, "aggregate function MIN / MAX (only Sum, Avg, Count cross the Ballista wire, to_proto.rs:352-363)");
      }
      sp.name = i < anames.size() ? anames[i] : "";
      sp.arg = compile_expr(am.one(2), input_schema, sp.name);
      aggs.push_back(sp);
    }
    return make_hash_aggregate(ctx, in, mode, std::move(groups), std::move(aggs), &input_schema);
  }
};

}  // namespace

OpPtr plan_from_proto(Ctx* ctx, const uint8_t* buf, size_t len, const std::vector<OpPtr>& leaves) {
  PlanDecoder d{ctx, &leaves};
  OpPtr root = d.node(Slice{buf, buf + len}, 0);
  if (d.next_leaf != leaves.size())
    fail(B2_ERR_INVALID_ARG, "plan has " + std::to_string(d.next_leaf) + " leaf nodes but " + std::to_string(leaves.size()) +
                                 " operators were supplied");
  return root;
}
std::string plan_proto_describe(const uint8_t* buf, size_t len, int* n_leaves) {
  PlanDecoder d{nullptr, nullptr};
  d.node(Slice{buf, buf + len}, 0);
  if (n_leaves) *n_leaves = d.leaf_count;
  return d.text.str();
}

}  // namespace b2
