// ops_join.cu — HashJoinExec (collect-left build + probe) and RepartitionExec (hash / round-robin; with a
// communicator attached the hash repartition IS the in-box shuffle).  SURVEY §8a rows F, G, H.
// Constructed where from_proto.rs:253-276 and :133-164 construct the DataFusion operators.
#include <cstdlib>

#include "comm.h"
#include "ops.h"

namespace b2 {

// ------------------------------------------------------------------------------------------------------
// HashJoinExec
// ------------------------------------------------------------------------------------------------------
bool op_is_filter(Op* op, ExprPtr* pred);
bool op_is_coalesce(Op* op);
bool op_is_exchange(Op* op);  // RepartitionExec whose output partitions live one per rank (communicator attached)

// FilterExec (optionally under CoalesceBatchesExec) directly below a join side: returns the filter's input
// and its predicate, so the join evaluates the mask itself and never compacts the side
static Op* peel_filter(Op* op, ExprPtr* pred) {
  Op* probe = op;
  if (op_is_coalesce(probe)) probe = probe->children[0].get();
  ExprPtr p;
  if (op_is_filter(probe, &p) && !p->is_literal()) {
    *pred = p;
    return probe->children[0].get();
  }
  return op;
}

struct HashJoinExec : Op {
  std::vector<int> on_left, on_right;  // column indices
  int join_type = B2_JOIN_INNER;
  Schema out;
  std::vector<int> right_keep;  // right columns that survive the duplicate-name rule
  KeyLayout layout;
  bool k32 = false;             // one Int32 / Date32 key per side: 8-byte-slot table, fused filters
  std::vector<char> need;       // output columns a parent will read (empty = all)
  // build side, collected once (the reference rebuilds it per task; result-neutral)
  BatchPtr left_batch;
  std::shared_ptr<JoinTable> table;
  BufPtr left_keys;
  bool semi_offered = false;

  const char* name() const override { return "HashJoinExec"; }
  Schema schema() const override { return out; }
  void partitioning(int* scheme, int* n) const override { children[1]->partitioning(scheme, n); }
  bool per_rank_partitions() const override { return children[1]->per_rank_partitions(); }

  bool fuse() const { return k32 && !getenv("B2_DISABLE_FUSION"); }
  bool needed(size_t out_col) const { return need.empty() || (out_col < need.size() && need[out_col]); }

  void require(const std::vector<char>& n) override {
    need = n;
    size_t nl = children[0]->schema().fields.size(), nr = children[1]->schema().fields.size();
    std::vector<char> l(nl, 0), r(nr, 0);
    for (size_t c = 0; c < nl; ++c) l[c] = needed(c);
    for (size_t k = 0; k < right_keep.size(); ++k) r[right_keep[k]] = needed(nl + k);
    for (int c : on_left) l[c] = 1;
    for (int c : on_right) r[c] = 1;
    children[0]->require(l);
    children[1]->require(r);
  }

  BufPtr pack(const Batch& b, const std::vector<int>& on) {
    std::vector<Column> kc;
    for (int c : on) kc.push_back(b.cols[c]);
    BufPtr st = dev_alloc_zero(ctx, 16);
    BufPtr keys = pack_keys(ctx, layout, kc, b.num_rows, st->as<int>());
    int host = 0;
    if (layout_has_utf8(layout)) B2_READBACK(ctx, &host, st->ptr, 4);
    if (host) fail(B2_ERR_NOT_IMPLEMENTED, "a Utf8 join key value is longer than the packed-key budget");
    return keys;
  }
  void build(int partition) {
    if (table) return;
    Op* l = children[0].get();
    ExprPtr pred;
    if (fuse()) l = peel_filter(l, &pred);
    std::vector<BatchPtr> all;
    int np = l->n_partitions();
    // Partitioned join: when the build side arrives through an exchange (a hash repartition whose output partition
    // r lives on rank r), only this rank's partition is built - the probe side was shuffled on the same key.
    // Otherwise collect-left, as the reference: MergeExec(left) over all partitions.
    bool partitioned = l->per_rank_partitions();
    for (int p = 0; p < np; ++p) {
      if (partitioned && p != partition) continue;
      StreamPtr s = l->execute(p);
      while (BatchPtr b = s->next())
        if (b->num_rows) all.push_back(b);
    }
    if (all.empty()) {
      left_batch = empty_batch(ctx, l->schema());
    } else {
      left_batch = concat_batches(ctx, l->schema(), all);
    }
    if (k32) {
      Column mask;
      RowPred rp;
      if (pred && left_batch->num_rows && !compile_row_pred(pred, *left_batch, &rp))
        mask = eval_expr(ctx, pred, *left_batch);
      table = join_build32(ctx, left_batch->cols[on_left[0]].values<int32_t>(), rp,
                           mask.data ? mask.values<uint8_t>() : nullptr, left_batch->num_rows);
      // the mask buffer is freed stream-ordered after the build kernel
    } else {
      left_keys = pack(*left_batch, on_left);
      table = join_build(ctx, left_keys->as<Key128>(), left_batch->num_rows);
    }
  }
  struct S : Stream {
    HashJoinExec* op;
    StreamPtr in;
    ExprPtr pred;  // fused right-side FilterExec predicate (k32 path)
    BatchPtr next() override {
      BatchPtr r = in->next();
      if (!r) return nullptr;
      Ctx* ctx = op->ctx;
      BufPtr li, ri;
      int64_t m;
      if (op->k32) {
        Column mask;
        RowPred rp;
        if (pred && r->num_rows && !compile_row_pred(pred, *r, &rp)) mask = eval_expr(ctx, pred, *r);
        m = join_probe32(ctx, *op->table, r->cols[op->on_right[0]].values<int32_t>(), rp,
                         mask.data ? mask.values<uint8_t>() : nullptr, r->num_rows, &li, &ri);
      } else {
        BufPtr pk = op->pack(*r, op->on_right);
        m = join_probe(ctx, *op->table, pk->as<Key128>(), r->num_rows, &li, &ri);
      }
      auto o = std::make_shared<Batch>();
      o->ctx = ctx;
      o->schema = op->out;
      o->num_rows = m;
      size_t oc = 0;
      auto elided = [&](const Column& c) {
        Column e;
        e.dtype = c.dtype;
        e.length = m;
        return e;
      };
      for (auto& c : op->left_batch->cols) {
        o->cols.push_back(op->needed(oc) ? take_column(ctx, c, li->as<uint32_t>(), m) : elided(c));
        ++oc;
      }
      for (int c : op->right_keep) {
        o->cols.push_back(op->needed(oc) ? take_column(ctx, r->cols[c], ri->as<uint32_t>(), m) : elided(r->cols[c]));
        ++oc;
      }
      return o;
    }
  };
  // Partitioned join over an exchange: publish the Bloom filter of this rank's build partition to every rank, so the
  // probe side's split drops, before the exchange, the rows whose destination cannot hold their key.  Collective:
  // every rank runs the same plan and takes the same branch.
  void offer_semi_filter() {
    if (semi_offered || !k32 || !table || !ctx->comm || ctx->nranks < 2 || getenv("B2_DISABLE_SEMI_FILTER")) return;
    semi_offered = true;
    Op* l = children[0].get();
    ExprPtr unused;
    if (fuse()) l = peel_filter(l, &unused);
    if (!l->per_rank_partitions() || !children[1]->accepts_semi_filter(on_right[0])) return;
    const int nr = ctx->nranks;
    std::vector<int64_t> words = allgather_counts(ctx, std::vector<int64_t>{(int64_t)table->bloom_words});
    int64_t stride = 0;
    for (int64_t w : words) stride = std::max(stride, w);
    auto f = std::make_shared<SemiFilter>();
    f->n = nr;
    f->stride = (uint32_t)stride;
    for (int r = 0; r < nr; ++r) f->wmask[r] = (uint32_t)(words[r] - 1);
    BufPtr send = table->bloom;
    if ((int64_t)table->bloom_words < stride) {
      send = dev_alloc_zero(ctx, (size_t)stride * 4);
      B2_CUDA(cudaMemcpyAsync(send->ptr, table->bloom->ptr, (size_t)table->bloom_words * 4, cudaMemcpyDeviceToDevice,
                              ctx->stream));
    }
    f->words = dev_alloc(ctx, (size_t)stride * 4 * nr);
    B2_NCCL(ncclAllGather(send->ptr, f->words->ptr, (size_t)stride, ncclUint32, ctx->comm->nccl, ctx->stream));
    ctx->bytes_sent += stride * 4 * (nr - 1);
    children[1]->set_semi_filter(f);
  }
  StreamPtr execute(int partition) override {
    build(partition);
    offer_semi_filter();
    auto s = std::make_unique<S>();
    s->op = this;
    Op* r = children[1].get();
    if (fuse()) r = peel_filter(r, &s->pred);
    s->in = r->execute(partition);
    last_path = k32 ? (fuse() ? "k32:fused-filter" : "k32") : "key128";
    return s;
  }
};

OpPtr make_hash_join(Ctx* ctx, OpPtr left, OpPtr right, std::vector<std::string> on_l, std::vector<std::string> on_r,
                     int join_type) {
  if (join_type != B2_JOIN_INNER)
    fail(B2_ERR_NOT_IMPLEMENTED,
         "Left/Right outer joins need null-extended rows; only Inner is implemented (all TPC-H Q3/Q5 joins)");
  if (on_l.empty() || on_l.size() != on_r.size()) fail(B2_ERR_INVALID_ARG, "join needs matching on-column lists");
  auto op = std::make_shared<HashJoinExec>();
  op->ctx = ctx;
  op->children = {left, right};
  op->join_type = join_type;
  Schema ls = left->schema(), rs = right->schema();
  std::vector<int> kd;
  for (size_t i = 0; i < on_l.size(); ++i) {
    int a = ls.index_of(on_l[i]), b = rs.index_of(on_r[i]);
    if (a < 0) fail(B2_ERR_INVALID_ARG, "join column '" + on_l[i] + "' not in the left schema");
    if (b < 0) fail(B2_ERR_INVALID_ARG, "join column '" + on_r[i] + "' not in the right schema");
    int ta = ls.fields[a].dtype, tb = rs.fields[b].dtype;
    bool i32 = (ta == B2_INT32 || ta == B2_DATE32) && (tb == B2_INT32 || tb == B2_DATE32);
    if (ta != tb && !i32) fail(B2_ERR_EXECUTION, "join key types differ for '" + on_l[i] + "' / '" + on_r[i] + "'");
    op->on_left.push_back(a);
    op->on_right.push_back(b);
    kd.push_back(ta);
  }
  op->layout = make_key_layout(kd);
  op->k32 = kd.size() == 1 && (kd[0] == B2_INT32 || kd[0] == B2_DATE32);
  // output = left fields ++ right fields minus right on-columns named like their left partner (A.3)
  op->out = ls;
  for (size_t c = 0; c < rs.fields.size(); ++c) {
    bool dup = false;
    for (size_t i = 0; i < on_l.size(); ++i)
      if ((int)c == op->on_right[i] && on_l[i] == on_r[i]) dup = true;
    if (dup) continue;
    op->right_keep.push_back((int)c);
    op->out.fields.push_back(rs.fields[c]);
  }
  return op;
}

// ------------------------------------------------------------------------------------------------------
// RepartitionExec
// ------------------------------------------------------------------------------------------------------
struct RepartitionExec : Op {
  int scheme = B2_PART_HASH;
  std::vector<ExprPtr> keys;
  int n_out = 1;
  KeyLayout layout;
  bool computed = false;
  std::vector<std::vector<BatchPtr>> outputs;  // per output partition
  std::vector<char> need;                      // output columns a parent will read (empty = all)

  const char* name() const override { return "RepartitionExec"; }
  Schema schema() const override { return children[0]->schema(); }
  void partitioning(int* s, int* n) const override {
    *s = scheme;
    *n = n_out;
  }
  std::shared_ptr<SemiFilter> semi;            // Bloom filters of the join this exchange feeds (probe side), or null
  bool exchange_mode() const { return scheme == B2_PART_HASH && ctx->comm && n_out == ctx->nranks; }
  bool accepts_semi_filter(int key_col) const override {
    if (computed || !exchange_mode() || n_out > 32 || keys.size() != 1) return false;
    const Expr& k = *keys[0];
    return k.kind == B2_EXPR_COLUMN && k.col == key_col && (k.dtype == B2_INT32 || k.dtype == B2_DATE32);
  }
  void set_semi_filter(std::shared_ptr<SemiFilter> f) override { semi = std::move(f); }
  bool per_rank_partitions() const override { return exchange_mode(); }
  // a FilterExec below a hash repartition rides in the split kernels (rows it rejects get no destination)
  bool fuse() const { return scheme == B2_PART_HASH && n_out <= 32 && !getenv("B2_DISABLE_FUSION"); }

  void require(const std::vector<char>& n) override {
    need = n;
    if (n.empty()) return children[0]->require(n);
    std::vector<char> c = n;
    c.resize(schema().fields.size(), 0);
    for (auto& k : keys) expr_columns(k, c);
    children[0]->require(c);
  }
  // columns nobody reads downstream are not moved: they stay elided in the output
  BatchPtr prune(const BatchPtr& b) const {
    if (need.empty()) return b;
    auto o = std::make_shared<Batch>(*b);
    for (size_t c = 0; c < o->cols.size(); ++c)
      if (!(c < need.size() && need[c])) {
        Column e;
        e.dtype = o->cols[c].dtype;
        e.length = o->cols[c].length;
        o->cols[c] = e;
      }
    return o;
  }

  // phase 1 of the hash split of `b` under the optional fused predicate (destinations + histogram + scan)
  SplitState split_begin(const Batch& b, const ExprPtr& pred) {
    std::vector<Column> kc;
    for (auto& k : keys) kc.push_back(eval_expr(ctx, k, b));
    Column mask;
    RowPred rp;
    if (pred && b.num_rows && !compile_row_pred(pred, b, &rp)) mask = eval_expr(ctx, pred, b);
    const uint8_t* m = mask.data ? mask.values<uint8_t>() : nullptr;
    if (kc.size() == 1 && (kc[0].dtype == B2_INT32 || kc[0].dtype == B2_DATE32)) {
      // one Int32 key: hashed straight from the column (same value as its packed form), nothing materialised
      return hash_split_begin(ctx, nullptr, kc[0].values<int32_t>(), b.num_rows, n_out, &rp, m, semi.get());
    }
    BufPtr stt = dev_alloc_zero(ctx, 16);
    BufPtr pk = pack_keys(ctx, layout, kc, b.num_rows, stt->as<int>());
    SplitState st = hash_split_begin(ctx, pk->as<Key128>(), nullptr, b.num_rows, n_out, &rp, m);
    int host = 0;
    if (layout_has_utf8(layout)) B2_READBACK(ctx, &host, stt->ptr, 4);
    if (host) fail(B2_ERR_NOT_IMPLEMENTED, "a Utf8 partitioning key value is longer than the packed-key budget");
    return st;
  }
  // stable destination-major permutation of the surviving rows of `b` + rows per destination
  BufPtr split_perm(const Batch& b, const ExprPtr& pred, std::vector<int64_t>* counts) {
    if (n_out <= 32) {
      SplitState st = split_begin(b, pred);
      BufPtr perm = hash_split_perm(ctx, st);
      hash_split_counts(ctx, st, counts);
      return perm;
    }
    std::vector<Column> kc;
    for (auto& k : keys) kc.push_back(eval_expr(ctx, k, b));
    BufPtr st = dev_alloc_zero(ctx, 16);
    BufPtr pk = pack_keys(ctx, layout, kc, b.num_rows, st->as<int>());
    BufPtr perm = hash_partition_indices(ctx, pk->as<Key128>(), nullptr, b.num_rows, n_out, counts);
    int host = 0;
    B2_READBACK(ctx, &host, st->ptr, 4);
    if (host) fail(B2_ERR_NOT_IMPLEMENTED, "a Utf8 partitioning key value is longer than the packed-key budget");
    return perm;
  }
  void compute() {
    if (computed) return;
    outputs.assign(n_out, {});
    Op* in = children[0].get();
    ExprPtr pred;
    if (fuse()) in = peel_filter(in, &pred);
    int np = in->n_partitions();
    const bool own_only = in->per_rank_partitions();  // input sits above another exchange: only partition `rank` is here
    if (exchange_mode()) {
      // one collective per execution: all local input is concatenated, split, and exchanged once
      std::vector<BatchPtr> all;
      for (int p = 0; p < np; ++p) {
        if (own_only && p != ctx->rank) continue;
        StreamPtr s = in->execute(p);
        while (BatchPtr b = s->next())
          if (b->num_rows) all.push_back(b);
      }
      BatchPtr local;
      if (all.empty()) {
        local = empty_batch(ctx, schema());
      } else {
        local = concat_batches(ctx, schema(), all);
      }
      BatchPtr moved = prune(local);  // predicate / key columns are still read from `local`
      BatchPtr got;
      if (n_out <= 32) {
        SplitState st = split_begin(*local, pred);
        got = comm_shuffle_p2p(ctx, *moved, st);
        if (got) {
          last_path = pred ? "nvlink-peer-pull+fused-filter" : "nvlink-peer-pull";
        } else {
          std::vector<int64_t> counts;
          BufPtr perm = hash_split_perm(ctx, st);
          hash_split_counts(ctx, st, &counts);
          int64_t m = 0;
          for (int64_t c : counts) m += c;
          BatchPtr permuted = take_batch(ctx, *moved, perm->as<uint32_t>(), m);
          got = comm_all_to_all(ctx, *permuted, counts);
          last_path = "nccl-all-to-all";
        }
      } else {
        std::vector<int64_t> counts;
        BufPtr perm = split_perm(*local, pred, &counts);
        BatchPtr permuted = take_batch(ctx, *moved, perm->as<uint32_t>(), local->num_rows);
        got = comm_all_to_all(ctx, *permuted, counts);
        last_path = "nccl-all-to-all";
      }
      outputs[ctx->rank].push_back(got);
    } else {
      int64_t k = 0;
      for (int p = 0; p < np; ++p) {
        if (own_only && p != ctx->rank) continue;
        StreamPtr s = in->execute(p);
        while (BatchPtr b = s->next()) {
          if (scheme == B2_PART_ROUND_ROBIN) {
            outputs[k++ % n_out].push_back(b);
            continue;
          }
          if (!b->num_rows) continue;
          std::vector<int64_t> counts;
          BufPtr perm = split_perm(*b, pred, &counts);
          int64_t m = 0;
          for (int64_t c : counts) m += c;
          BatchPtr permuted = take_batch(ctx, *prune(b), perm->as<uint32_t>(), m);
          int64_t off = 0;
          for (int d = 0; d < n_out; ++d) {
            if (counts[d]) outputs[d].push_back(slice_batch(ctx, *permuted, off, off + counts[d]));
            off += counts[d];
          }
        }
      }
      last_path = pred ? "local+fused-filter" : "local";
    }
    computed = true;
  }
  struct S : Stream {
    RepartitionExec* op;
    int part;
    size_t pos = 0;
    BatchPtr next() override {
      auto& v = op->outputs[part];
      return pos < v.size() ? v[pos++] : nullptr;
    }
  };
  StreamPtr execute(int partition) override {
    if (partition < 0 || partition >= n_out) fail(B2_ERR_INVALID_ARG, "RepartitionExec: partition out of range");
    if (exchange_mode() && partition != ctx->rank)
      fail(B2_ERR_INVALID_ARG, "with a communicator attached, rank r executes output partition r");
    compute();
    auto s = std::make_unique<S>();
    s->op = this;
    s->part = partition;
    return s;
  }
};

OpPtr make_repartition(Ctx* ctx, OpPtr input, int scheme, std::vector<ExprPtr> keys, int n_out) {
  if (n_out < 1) fail(B2_ERR_INVALID_ARG, "RepartitionExec needs n >= 1");
  if (scheme != B2_PART_HASH && scheme != B2_PART_ROUND_ROBIN)
    fail(B2_ERR_NOT_IMPLEMENTED, "RepartitionExec supports Hash and RoundRobinBatch partitioning");
  auto op = std::make_shared<RepartitionExec>();
  op->ctx = ctx;
  op->children = {input};
  op->scheme = scheme;
  op->n_out = n_out;
  if (scheme == B2_PART_HASH) {
    if (keys.empty()) fail(B2_ERR_INVALID_ARG, "hash repartition needs key expressions");
    std::vector<int> kd;
    for (auto& k : keys) kd.push_back(k->dtype);
    op->layout = make_key_layout(kd);
  }
  op->keys = std::move(keys);
  return op;
}

bool op_is_exchange(Op* op) {
  auto* r = dynamic_cast<RepartitionExec*>(op);
  return r && r->exchange_mode();
}

void repartition_reset(Op* op) {
  if (auto* r = dynamic_cast<RepartitionExec*>(op)) {
    r->computed = false;
    r->outputs.clear();
  }
}

}  // namespace b2
