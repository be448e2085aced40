// ops_agg.cu — HashAggregateExec (Partial / Final; grouped and ungrouped).  SURVEY §8a rows B, C, D and
// Appendix A.2.  Constructed where from_proto.rs:173-252 calls HashAggregateExec::try_new.
//
// Partial: when the input is FilterExec (optionally under CoalesceBatchesExec) the predicate is fused into
// the aggregate's scan, and when the (predicate, keys, arguments) shape is in the kernel registry the whole
// scan -> filter -> expression -> partial-aggregate pipeline is ONE kernel with nothing materialised.
// Otherwise the unfused path evaluates mask / key / argument columns first (eval.cu) and feeds the same
// aggregate kernels through the materialised source.
#include <cstdlib>

#include "ops.h"

namespace b2 {

bool op_is_filter(Op* op, ExprPtr* pred);
bool op_is_coalesce(Op* op);

struct HashAggregateExec : Op {
  int mode = B2_AGG_PARTIAL;
  std::vector<ExprPtr> groups;
  std::vector<AggSpec> aggs;
  Schema out;
  KeyLayout layout;
  // Partial: distinct argument expressions (CSE across aggregates) and each aggregate's accumulator index
  std::vector<ExprPtr> args;
  std::vector<int> agg_acc;  // per aggregate: index into args (SUM/AVG), -1 for COUNT
  bool disable_fusion = false;

  const char* name() const override { return "HashAggregateExec"; }
  Schema schema() const override { return out; }

  struct S : Stream {
    HashAggregateExec* op;
    int partition;
    bool done = false;
    BatchPtr next() override {
      if (done) return nullptr;
      done = true;
      return op->mode == B2_AGG_PARTIAL ? op->run_partial(partition) : op->run_final(partition);
    }
  };
  StreamPtr execute(int partition) override {
    auto s = std::make_unique<S>();
    s->op = this;
    s->partition = partition;
    return s;
  }

  BatchPtr emit(AggRunner& r, bool partial, const std::vector<int>& sum_idx, const std::vector<int>& cnt_idx);
  BatchPtr run_partial(int partition);
  BatchPtr run_final(int partition);
};

// ---- output assembly shared by both modes.  sum_idx / cnt_idx: per aggregate accumulator indices ----
BatchPtr HashAggregateExec::emit(AggRunner& r, bool partial, const std::vector<int>& sum_idx,
                                 const std::vector<int>& cnt_idx) {
  BufPtr slots;
  int64_t ng = r.finish(&slots);
  auto b = std::make_shared<Batch>();
  b->ctx = ctx;
  b->schema = out;
  if (groups.empty() && ng == 0) {
    // ungrouped aggregate over empty input still yields one row; sums are 0, counts 0
    // (the reference yields NULL sums; this path has no validity bitmaps)
    b->num_rows = 1;
    for (auto& f : out.fields) {
      Column c;
      c.dtype = f.dtype;
      c.length = 1;
      c.data = dev_alloc_zero(ctx, 16);
      b->cols.push_back(c);
    }
    return b;
  }
  b->num_rows = ng;
  EmitPlan ep(ctx, *r.main, slots, ng);
  for (size_t g = 0; g < groups.size(); ++g) ep.key(layout.parts[g]);
  for (size_t a = 0; a < aggs.size(); ++a) {
    switch (aggs[a].fn) {
      case B2_AGG_SUM: ep.sum(sum_idx[a]); break;
      case B2_AGG_COUNT: ep.count(cnt_idx[a]); break;
      case B2_AGG_AVG:
        if (partial) {
          ep.count(cnt_idx[a]);
          ep.sum(sum_idx[a]);
        } else {
          ep.avg(sum_idx[a], cnt_idx[a]);
        }
        break;
    }
  }
  b->cols = ep.run();
  if (partial && !groups.empty()) b->unique_key_cols = (int)groups.size();
  return b;
}

BatchPtr HashAggregateExec::run_partial(int partition) {
  {  // column pruning: tell the input which columns the group / argument expressions read
    std::vector<char> used(children[0]->schema().fields.size(), 0);
    for (auto& g : groups) expr_columns(g, used);
    for (auto& a : args) expr_columns(a, used);
    children[0]->require(used);
  }
  // look through CoalesceBatchesExec for a FilterExec whose predicate can ride in the aggregate's scan
  Op* src = children[0].get();
  ExprPtr pred;
  {
    Op* probe = src;
    if (op_is_coalesce(probe)) probe = probe->children[0].get();
    ExprPtr p;
    if (op_is_filter(probe, &p)) {
      pred = p;
      src = probe->children[0].get();
    }
  }
  if (pred && pred->is_literal()) {
    if (!pred->i) fail(B2_ERR_NOT_IMPLEMENTED, "constant-false filter under an aggregate");
    pred = nullptr;
  }
  StreamPtr in = src->execute(partition);
  std::unique_ptr<AggRunner> runner;
  std::string path;
  std::vector<BatchPtr> keep_alive;  // kernels are asynchronous: inputs must outlive them
  while (BatchPtr b = in->next()) {
    if (b->num_rows == 0) continue;
    AggLaunch l;
    std::string sig;
    std::vector<Column> temps;
    BufPtr keybuf;
    bool fused = !disable_fusion && !getenv("B2_DISABLE_FUSION") && bind_fused_agg(ctx, pred, groups, args, *b, &l, &sig);
    if (fused) {
      if (path.empty()) path = "fused:" + sig + (l.tma ? (l.narrow ? "[tma,narrow]" : "[tma]") : "[ld]");
    } else {
      if (path.empty()) path = "unfused";
      const uint8_t* mask = nullptr;
      if (pred) {
        temps.push_back(eval_expr(ctx, pred, *b));
        mask = temps.back().values<uint8_t>();
      }
      std::vector<const double*> vals;
      for (auto& a : args) {
        temps.push_back(eval_expr(ctx, a, *b));
        vals.push_back(temps.back().values<double>());
      }
      const Key128* keys = nullptr;
      BufPtr st = dev_alloc_zero(ctx, 16);
      if (!groups.empty()) {
        std::vector<Column> kc;
        for (auto& g : groups) kc.push_back(eval_expr(ctx, g, *b));
        keybuf = pack_keys(ctx, layout, kc, b->num_rows, st->as<int>());
        int host = 0;
        if (layout_has_utf8(layout)) B2_READBACK(ctx, &host, st->ptr, 4);
        if (host) fail(B2_ERR_NOT_IMPLEMENTED, "a Utf8 group key value is longer than the packed-key budget");
        keys = keybuf->as<Key128>();
      }
      l = bind_column_agg(ctx, keys, mask, vals, {}, /*count_rows=*/true);
    }
    if (!runner) runner = std::make_unique<AggRunner>(ctx, l.na, l.nc, l.keyed);
    if (runner->na != l.na || runner->nc != l.nc)
      fail(B2_ERR_EXECUTION, "internal: aggregate source shape changed between batches");
    runner->consume(l, b->num_rows);  // synchronises on the status word, so temporaries may go
  }
  if (!runner) runner = std::make_unique<AggRunner>(ctx, (int)(args.empty() ? 1 : args.size()), 1, !groups.empty());
  last_path = (path.empty() ? std::string("empty") : path) + (runner->path.empty() ? "" : "/" + runner->path);
  std::vector<int> sum_idx(aggs.size(), 0), cnt_idx(aggs.size(), 0);
  for (size_t a = 0; a < aggs.size(); ++a) sum_idx[a] = agg_acc[a] < 0 ? 0 : agg_acc[a];
  return emit(*runner, true, sum_idx, cnt_idx);
}

BatchPtr HashAggregateExec::run_final(int partition) {
  // input = partial-state batches: group columns first, then state columns positionally
  StreamPtr in = children[0]->execute(partition);
  size_t ng = groups.size();
  std::vector<int> sum_col(aggs.size(), -1), cnt_col(aggs.size(), -1);
  std::vector<int> sum_idx(aggs.size(), 0), cnt_idx(aggs.size(), 0);
  int col = (int)ng, n_sum = 0, n_cnt = 0;
  for (size_t a = 0; a < aggs.size(); ++a) {
    switch (aggs[a].fn) {
      case B2_AGG_SUM: sum_col[a] = col++; sum_idx[a] = n_sum++; break;
      case B2_AGG_COUNT: cnt_col[a] = col++; cnt_idx[a] = n_cnt++; break;
      case B2_AGG_AVG: cnt_col[a] = col++; sum_col[a] = col++; cnt_idx[a] = n_cnt++; sum_idx[a] = n_sum++; break;
    }
  }
  Schema in_schema = children[0]->schema();
  if ((int)in_schema.fields.size() != col)
    fail(B2_ERR_EXECUTION, "Final aggregate: input has " + std::to_string(in_schema.fields.size()) +
                               " columns, expected " + std::to_string(col) + " (groups + accumulator state)");
  // The whole input is ONE Partial output over these group keys (a stage with a single partition, or states that an
  // exchange already co-located and a Partial re-aggregated): every key occurs once, so the merge is the identity and
  // SUM / COUNT states are the results.  AVG (a division) and computed group expressions take the general path.
  auto next_nonempty = [&]() {
    BatchPtr b;
    while ((b = in->next()) && b->num_rows == 0) {}
    return b;
  };
  BatchPtr first = next_nonempty(), second = first ? next_nonempty() : nullptr;
  bool identity = first && !second && ng > 0 && first->unique_key_cols == (int)ng && !getenv("B2_DISABLE_FUSION");
  for (size_t g = 0; g < ng && identity; ++g) identity = groups[g]->kind == B2_EXPR_COLUMN && groups[g]->col == (int)g;
  for (size_t a = 0; a < aggs.size() && identity; ++a) identity = aggs[a].fn != B2_AGG_AVG;
  if (identity) {
    auto o = std::make_shared<Batch>();
    o->ctx = ctx;
    o->schema = out;
    o->num_rows = first->num_rows;
    for (size_t g = 0; g < ng; ++g) o->cols.push_back(first->cols[g]);
    for (size_t a = 0; a < aggs.size(); ++a) o->cols.push_back(first->cols[sum_col[a] >= 0 ? sum_col[a] : cnt_col[a]]);
    last_path = "identity(unique partial states)";
    return o;
  }
  std::unique_ptr<AggRunner> runner;
  std::vector<BatchPtr> head;
  if (first) head.push_back(first);
  if (second) head.push_back(second);
  size_t head_i = 0;
  auto next_batch = [&]() -> BatchPtr {
    if (head_i < head.size()) return head[head_i++];
    return in->next();
  };
  while (BatchPtr b = next_batch()) {
    if (b->num_rows == 0) continue;
    std::vector<const double*> vals;
    std::vector<const u64*> cnts;
    for (size_t a = 0; a < aggs.size(); ++a) {
      if (sum_col[a] >= 0) {
        if (b->cols[sum_col[a]].dtype != B2_FLOAT64) fail(B2_ERR_NOT_IMPLEMENTED, "non-Float64 sum state");
        vals.push_back(b->cols[sum_col[a]].values<double>());
      }
    }
    for (size_t a = 0; a < aggs.size(); ++a)
      if (cnt_col[a] >= 0) {
        if (b->cols[cnt_col[a]].dtype != B2_UINT64) fail(B2_ERR_EXECUTION, "count state must be UInt64");
        cnts.push_back(b->cols[cnt_col[a]].values<u64>());
      }
    // note: vals/cnts were pushed in aggregate order per kind, matching sum_idx / cnt_idx
    const Key128* keys = nullptr;
    BufPtr keybuf;
    if (ng) {
      std::vector<Column> kc;
      for (auto& g : groups) kc.push_back(eval_expr(ctx, g, *b));
      BufPtr st = dev_alloc_zero(ctx, 16);
      keybuf = pack_keys(ctx, layout, kc, b->num_rows, st->as<int>());
      keys = keybuf->as<Key128>();
    }
    AggLaunch l = bind_column_agg(ctx, keys, nullptr, vals, cnts, /*count_rows=*/false);
    if (!runner) runner = std::make_unique<AggRunner>(ctx, l.na, l.nc, l.keyed);
    // partial states carry one row per group of their partition: a large state batch means more groups than the
    // per-CTA tables of the low-cardinality kernel hold, so go straight to the HBM table
    if (ng && b->num_rows > 16384) runner->force_global = true;
    runner->consume(l, b->num_rows);
  }
  if (!runner) runner = std::make_unique<AggRunner>(ctx, n_sum ? n_sum : 1, n_cnt ? n_cnt : 1, ng > 0);
  last_path = "unfused/" + runner->path;
  return emit(*runner, false, sum_idx, cnt_idx);
}

OpPtr make_hash_aggregate(Ctx* ctx, OpPtr input, int mode, std::vector<ExprPtr> groups, std::vector<AggSpec> aggs,
                          const Schema* input_schema) {
  auto op = std::make_shared<HashAggregateExec>();
  op->ctx = ctx;
  op->children = {input};
  op->mode = mode;
  Schema in = input->schema();
  std::vector<int> kd;
  for (auto& g : groups) {
    kd.push_back(g->dtype);
    op->out.fields.push_back(Field{g->name.empty() ? expr_to_string(g, in) : g->name, g->dtype, false});
  }
  op->layout = make_key_layout(kd);
  const Schema& arg_schema = input_schema ? *input_schema : in;
  for (auto& a : aggs) {
    int acc = -1;
    if (a.fn == B2_AGG_SUM || a.fn == B2_AGG_AVG) {
      ExprPtr arg = a.arg;
      if (arg->dtype != B2_FLOAT64) {
        if (a.fn == B2_AGG_SUM)
          fail(B2_ERR_NOT_IMPLEMENTED, std::string("SUM over ") + dtype_name(arg->dtype) +
                                           " (only Float64 sums occur on the TPC-H path)");
        // AVG casts its input to Float64 (avg_return_type is always Float64)
        auto c = std::make_shared<Expr>();
        c->kind = B2_EXPR_CAST;
        c->dtype = B2_FLOAT64;
        c->ch = {arg};
        if (!(arg->dtype == B2_INT32 || arg->dtype == B2_INT64))
          fail(B2_ERR_NOT_IMPLEMENTED, std::string("AVG over ") + dtype_name(arg->dtype));
        arg = c;
      }
      if (mode == B2_AGG_PARTIAL) {
        std::string text = expr_to_string(arg, arg_schema);
        for (size_t k = 0; k < op->args.size(); ++k)
          if (expr_to_string(op->args[k], arg_schema) == text) acc = (int)k;
        if (acc < 0) {
          op->args.push_back(arg);
          acc = (int)op->args.size() - 1;
        }
      }
    } else if (a.fn != B2_AGG_COUNT) {
      fail(B2_ERR_NOT_IMPLEMENTED, "aggregate function " + std::to_string(a.fn) +
                                       " (only Sum, Avg, Count cross the Ballista wire, to_proto.rs:352-363)");
    }
    op->agg_acc.push_back(acc);
    if (mode == B2_AGG_PARTIAL) {
      switch (a.fn) {
        case B2_AGG_SUM: op->out.fields.push_back(Field{a.name + "[sum]", B2_FLOAT64, true}); break;
        case B2_AGG_COUNT: op->out.fields.push_back(Field{a.name + "[count]", B2_UINT64, true}); break;
        case B2_AGG_AVG:
          op->out.fields.push_back(Field{a.name + "[count]", B2_UINT64, true});
          op->out.fields.push_back(Field{a.name + "[sum]", B2_FLOAT64, true});
          break;
      }
    } else {
      op->out.fields.push_back(Field{a.name, a.fn == B2_AGG_COUNT ? B2_UINT64 : B2_FLOAT64, true});
    }
  }
  op->groups = std::move(groups);
  op->aggs = std::move(aggs);
  return op;
}

void hash_aggregate_disable_fusion(Op* op, bool off) {
  if (auto* h = dynamic_cast<HashAggregateExec*>(op)) h->disable_fusion = off;
}

}  // namespace b2
