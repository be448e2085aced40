// ops_join.cu — HashJoinExec (collect-left build + probe) and RepartitionExec (hash / round-robin; with a
// communicator attached the hash repartition IS the in-box shuffle).  SURVEY §8a rows F, G, H.
// Constructed where from_proto.rs:253-276 and :133-164 construct the DataFusion operators.
#include <cstdlib>
#include <cstring>

#include "comm.h"
#include "ops.h"
#include "scan.cuh"

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

__global__ void mark_rows_kernel(const uint32_t* idx, int64_t m, uint8_t* flags) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  for (; i < m; i += (int64_t)gridDim.x * blockDim.x) flags[idx[i]] = 1;
}

struct HashJoinExec : Op {
  std::vector<int> on_left, on_right;  // column indices
  int join_type = B2_JOIN_INNER;
  Schema out;
  // output = kept left columns ++ kept right columns.  Inner / Left drop a right `on` column named like its left
  // partner, Right drops the left one instead (DataFusion's build_join_schema)
  std::vector<int> left_keep, right_keep;
  KeyLayout layout;
  bool k32 = false;             // Int32 / Date32 keys only: 8-byte-slot table on the first pair, fused filters;
                                // further pairs are verified on the candidate pairs (JoinVerify)
  std::vector<char> need;       // output columns a parent will read (empty = all)
  // build side, collected once per version of the left subtree (the reference rebuilds it per task; result-neutral,
  // SURVEY App. C.10)
  BatchPtr left_batch;
  std::shared_ptr<JoinTable> table;
  BufPtr left_keys;
  bool semi_offered = false;
  bool partitioned = false;     // both inputs arrive through exchanges on the join keys: partition-wise join
  uint64_t built_version = 0;
  ExprPtr left_pred;            // FilterExec predicate peeled off the left side (folded into the build / probe kernels)

  const char* name() const override { return "HashJoinExec"; }
  Schema schema() const override { return out; }
  void partitioning(int* scheme, int* n) const override { children[1]->partitioning(scheme, n); }
  bool per_rank_partitions() const override { return children[1]->per_rank_partitions(); }
  // rows stay in the partition of their probe-side row: the output is exchanged on the right key columns
  bool exchange_on(std::vector<int>* key_cols, int* n_out) const override {
    std::vector<int> rk;
    if (!children[1]->exchange_on(&rk, n_out)) return false;
    key_cols->clear();
    for (int c : rk) {
      int o = -1;
      for (size_t k = 0; k < right_keep.size(); ++k)
        if (right_keep[k] == c) o = (int)(left_keep.size() + k);
      if (o < 0 && join_type != B2_JOIN_RIGHT)  // dropped as a duplicate of its (equal-valued) left partner
        for (size_t i = 0; i < on_right.size(); ++i)
          if (on_right[i] == c)
            for (size_t k = 0; k < left_keep.size(); ++k)
              if (left_keep[k] == on_left[i]) o = (int)k;
      if (o < 0) return false;
      key_cols->push_back(o);
    }
    return true;
  }

  // A FilterExec below a side is folded into the build / probe kernels, except on the side whose unmatched rows an
  // outer join must enumerate (there the filter is materialised, so that "row of the input" means "row that passed")
  bool fuse_left() const { return k32 && join_type != B2_JOIN_LEFT && !getenv("B2_DISABLE_FUSION"); }
  bool fuse_right() const { return k32 && join_type != B2_JOIN_RIGHT && !getenv("B2_DISABLE_FUSION"); }
  bool needed(size_t out_col) const { return need.empty() || (out_col < need.size() && need[out_col]); }

  void require(const std::vector<char>& n) override {
    need = n;
    size_t nl = children[0]->schema().fields.size(), nr = children[1]->schema().fields.size();
    std::vector<char> l(nl, 0), r(nr, 0);
    for (size_t k = 0; k < left_keep.size(); ++k) l[left_keep[k]] = needed(k);
    for (size_t k = 0; k < right_keep.size(); ++k) r[right_keep[k]] = needed(left_keep.size() + k);
    for (int c : on_left) l[c] = 1;
    for (int c : on_right) r[c] = 1;
    children[0]->require(l);
    children[1]->require(r);
  }

  // packed keys of one side; rows with a NULL key part get a per-side tag in byte 15, so they equal no key of the
  // other side (null keys never match)
  BufPtr pack(const Batch& b, const std::vector<int>& on, int null_tag) {
    std::vector<Column> kc;
    for (int c : on) kc.push_back(b.cols[c]);
    BufPtr st = dev_alloc_zero(ctx, 16);
    BufPtr keys = pack_keys(ctx, layout, kc, b.num_rows, st->as<int>(), null_tag);
    int host = 0;
    if (layout_has_utf8(layout)) B2_READBACK(ctx, &host, st->ptr, 4);
    if (host) fail(B2_ERR_NOT_IMPLEMENTED, "a Utf8 join key value is longer than the packed-key budget");
    return keys;
  }
  // fused-filter mask of one side of the 32-bit-key path, with the key column's NULLs folded in
  Column side_mask(const ExprPtr& pred, const Batch& b, int key_col, RowPred* rp) {
    Column mask;
    if (pred && b.num_rows && !compile_row_pred(pred, b, rp)) mask = eval_predicate(ctx, pred, b);
    const Column& kc = b.cols[key_col];
    if (kc.valid) {
      if (mask.data) mask.data = valid_and(ctx, mask.data, kc.valid, b.num_rows);
      else mask.data = kc.valid;
    }
    return mask;
  }
  void build(int partition) {
    const uint64_t v = children[0]->version();
    if (left_batch && built_version == v) return;  // (the table itself is built on first use: see ensure_table)
    table.reset();
    left_batch.reset();
    left_keys.reset();
    semi_offered = false;
    built_version = v;
    Op* l = children[0].get();
    left_pred = nullptr;
    if (fuse_left()) l = peel_filter(l, &left_pred);
    std::vector<BatchPtr> all;
    int np = l->n_partitions();
    // Partitioned join: when BOTH inputs arrive through exchanges (hash repartitions whose output partition r lives on
    // rank r) on this join's keys with one partition count, only this rank's build partition is needed - every probe
    // row was sent to the rank that holds its matches.  A per-rank build side under any other probe side is
    // all-gathered: collect-left, as the reference's MergeExec(left) over all partitions.
    const bool per_rank = l->per_rank_partitions();
    partitioned = false;
    if (per_rank) {
      std::vector<int> lk, rk;
      int ln = 0, rn = 0;
      partitioned = children[0]->exchange_on(&lk, &ln) && children[1]->exchange_on(&rk, &rn) && ln == rn &&
                    lk == on_left && rk == on_right;
    }
    for (int p = 0; p < np; ++p) {
      if (per_rank && p != ctx->rank) continue;  // partition r of an exchange lives on rank r
      StreamPtr s = l->execute(p);
      while (BatchPtr b = s->next())
        if (b->num_rows) all.push_back(b);
    }
    if (all.empty()) {
      left_batch = empty_batch(ctx, l->schema());
    } else {
      left_batch = concat_batches(ctx, l->schema(), all);
    }
    if (per_rank && !partitioned) {
      if (!ctx->comm) fail(B2_ERR_EXECUTION, "HashJoinExec: per-rank build partitions without a communicator");
      left_batch = comm_allgather_batch(ctx, left_batch);
    }
  }
  // hash table over one side's batch (the collected left side, or - swapped - a small right batch)
  std::shared_ptr<JoinTable> build_table(const Batch& b, const ExprPtr& pred, const std::vector<int>& on, int null_tag,
                                         BufPtr* keys_keep) {
    if (k32) {
      RowPred rp;
      Column mask = side_mask(pred, b, on[0], &rp);
      // the mask buffer is freed stream-ordered after the build kernel
      return join_build32(ctx, b.cols[on[0]].values<int32_t>(), rp, mask.data ? mask.values<uint8_t>() : nullptr,
                          b.num_rows);
    }
    *keys_keep = pack(b, on, null_tag);
    return join_build(ctx, (*keys_keep)->as<Key128>(), b.num_rows);
  }
  void ensure_table() {
    if (!table) table = build_table(*left_batch, left_pred, on_left, 0x81, &left_keys);
  }
  // the key pairs after the first (k32 with several key columns): checked on the candidate pairs inside join_probe32
  static JoinVerify verify_of(const Batch& build_b, const std::vector<int>& build_on, const Batch& b,
                              const std::vector<int>& on) {
    JoinVerify v;
    memset(&v, 0, sizeof v);
    v.n = (int)on.size() - 1;
    for (int k = 0; k < v.n; ++k) {
      const Column& bc = build_b.cols[build_on[k + 1]];
      const Column& pc = b.cols[on[k + 1]];
      v.bcol[k] = bc.values<int32_t>();
      v.pcol[k] = pc.values<int32_t>();
      v.bval[k] = bc.validity();
      v.pval[k] = pc.validity();
    }
    return v;
  }
  // pairs (row of `b`, build row) of one probe batch against `t` (built over `build_b` on `build_on`)
  int64_t probe_table(JoinTable& t, const Batch& build_b, const std::vector<int>& build_on, const Batch& b,
                      const ExprPtr& pred, const std::vector<int>& on, int null_tag, BufPtr* build_idx,
                      BufPtr* probe_idx) {
    if (k32) {
      RowPred rp;
      Column mask = side_mask(pred, b, on[0], &rp);
      JoinVerify v = verify_of(build_b, build_on, b, on);
      return join_probe32(ctx, t, b.cols[on[0]].values<int32_t>(), rp, mask.data ? mask.values<uint8_t>() : nullptr,
                          b.num_rows, build_idx, probe_idx, v.n ? &v : nullptr);
    }
    BufPtr pk = pack(b, on, null_tag);
    return join_probe(ctx, t, pk->as<Key128>(), b.num_rows, build_idx, probe_idx);
  }
  struct S : Stream {
    HashJoinExec* op;
    StreamPtr in;
    ExprPtr pred;        // fused right-side FilterExec predicate (k32 path)
    BufPtr left_matched; // Left join: one byte per build row, set by every probe batch of this stream
    bool tail_done = false;
    bool decided = false, swapped_done = false;
    BatchPtr first, second;  // one batch of look-ahead on the probe side

    BatchPtr assemble(const Batch* r, const BufPtr& li, const BufPtr& ri, int64_t m, bool null_left, bool null_right) {
      Ctx* ctx = op->ctx;
      auto o = std::make_shared<Batch>();
      o->ctx = ctx;
      o->schema = op->out;
      o->num_rows = m;
      size_t oc = 0;
      auto elided = [&](int dtype) {
        Column e;
        e.dtype = dtype;
        e.length = m;
        return e;
      };
      // the columns of one side that a parent reads are gathered together (one launch per side for fixed-width data)
      std::vector<const Column*> lsrc, rsrc;
      std::vector<size_t> lpos, rpos;
      for (int c : op->left_keep) {
        const Column& col = op->left_batch->cols[c];
        if (!op->needed(oc)) o->cols.push_back(elided(col.dtype));
        else if (!li) o->cols.push_back(null_column(ctx, col.dtype, m));
        else {
          lsrc.push_back(&col);
          lpos.push_back(o->cols.size());
          o->cols.push_back(Column());
        }
        ++oc;
      }
      Schema rs = op->children[1]->schema();
      for (int c : op->right_keep) {
        if (!op->needed(oc)) o->cols.push_back(elided(rs.fields[c].dtype));
        else if (!ri) o->cols.push_back(null_column(ctx, rs.fields[c].dtype, m));
        else {
          rsrc.push_back(&r->cols[c]);
          rpos.push_back(o->cols.size());
          o->cols.push_back(Column());
        }
        ++oc;
      }
      if (!lsrc.empty()) {
        std::vector<Column> g = take_columns(ctx, lsrc, li->as<uint32_t>(), m, null_left);
        for (size_t k = 0; k < g.size(); ++k) o->cols[lpos[k]] = std::move(g[k]);
      }
      if (!rsrc.empty()) {
        std::vector<Column> g = take_columns(ctx, rsrc, ri->as<uint32_t>(), m, null_right);
        for (size_t k = 0; k < g.size(); ++k) o->cols[rpos[k]] = std::move(g[k]);
      }
      return o;
    }
    BatchPtr pull() {
      if (first) {
        BatchPtr b = first;
        first = second;
        second = nullptr;
        return b;
      }
      return in->next();
    }
    BatchPtr next() override {
      Ctx* ctx = op->ctx;
      if (!decided) {
        decided = true;
        while ((first = in->next()) && first->num_rows == 0) {}
        if (first)
          while ((second = in->next()) && second->num_rows == 0) {}
        // Side selection at run time (inner joins): the reference always builds on the left (collect-left), which in a
        // left-deep plan is the accumulated join result - Q5 joins millions of rows against the 25 nations and then
        // against one region, every probe row walking a chain of 10^5 build rows.  When the whole right side is ONE
        // batch far smaller than the left, the table is built over it and the collected left side probes: the same
        // row set (order is unobservable), a table that fits L2 and no chains.
        if (op->join_type == B2_JOIN_INNER && !op->table && first && !second && !getenv("B2_DISABLE_JOIN_SWAP") &&
            first->num_rows * 4 <= op->left_batch->num_rows) {
          swapped_done = true;
          BatchPtr r = first;
          first = nullptr;
          BufPtr keep, bi, pi;
          auto t = op->build_table(*r, pred, op->on_right, 0x82, &keep);
          int64_t m = op->probe_table(*t, *r, op->on_right, *op->left_batch, op->left_pred, op->on_left, 0x81, &bi, &pi);
          op->last_path += "/build-right";
          return assemble(r.get(), pi, bi, m, false, false);
        }
        op->ensure_table();
      }
      if (swapped_done) return nullptr;
      BatchPtr r = pull();
      if (!r) {
        // Left join: build rows no probe batch of this stream matched, null-extended on the right (emitted once, at
        // the end of the stream; DataFusion@46161d2 emits them per probe batch, which repeats them - SURVEY A.3
        // states the intended result and that is what is produced here)
        if (op->join_type == B2_JOIN_LEFT && !tail_done) {
          tail_done = true;
          int64_t nb = op->left_batch->num_rows, m = 0;
          if (!nb) return nullptr;
          BufPtr un = valid_not(ctx, left_matched ? left_matched->as<uint8_t>() : nullptr, nb);
          if (!left_matched) B2_CUDA(cudaMemsetAsync(un->ptr, 1, (size_t)nb, ctx->stream));
          BufPtr idx = mask_to_indices(ctx, un->as<uint8_t>(), nb, &m);
          if (!m) return nullptr;
          return assemble(nullptr, idx, nullptr, m, false, false);
        }
        return nullptr;
      }
      BufPtr li, ri;
      int64_t m = op->probe_table(*op->table, *op->left_batch, op->on_left, *r, pred, op->on_right, 0x82, &li, &ri);
      if (op->join_type == B2_JOIN_LEFT && op->left_batch->num_rows) {
        if (!left_matched) left_matched = dev_alloc_zero(ctx, (size_t)op->left_batch->num_rows + 16);
        if (m) {
          mark_rows_kernel<<<grid_for(ctx, 8, m, 256), 256, 0, ctx->stream>>>(li->as<uint32_t>(), m,
                                                                             left_matched->as<uint8_t>());
          B2_LAUNCHED(ctx);
        }
      }
      bool null_left = false;
      if (op->join_type == B2_JOIN_RIGHT && r->num_rows) {
        // probe rows without a match follow the matched pairs of their batch, null-extended on the left
        BufPtr hit = dev_alloc_zero(ctx, (size_t)r->num_rows + 16);
        if (m) {
          mark_rows_kernel<<<grid_for(ctx, 8, m, 256), 256, 0, ctx->stream>>>(ri->as<uint32_t>(), m, hit->as<uint8_t>());
          B2_LAUNCHED(ctx);
        }
        BufPtr un = valid_not(ctx, hit->as<uint8_t>(), r->num_rows);
        int64_t u = 0;
        BufPtr uidx = mask_to_indices(ctx, un->as<uint8_t>(), r->num_rows, &u);
        if (u) {
          BufPtr li2 = dev_alloc(ctx, (size_t)(m + u + 1) * 4), ri2 = dev_alloc(ctx, (size_t)(m + u + 1) * 4);
          if (m) {
            B2_CUDA(cudaMemcpyAsync(li2->ptr, li->ptr, (size_t)m * 4, cudaMemcpyDeviceToDevice, ctx->stream));
            B2_CUDA(cudaMemcpyAsync(ri2->ptr, ri->ptr, (size_t)m * 4, cudaMemcpyDeviceToDevice, ctx->stream));
          }
          B2_CUDA(cudaMemsetAsync(li2->as<uint32_t>() + m, 0xFF, (size_t)u * 4, ctx->stream));
          B2_CUDA(cudaMemcpyAsync(ri2->as<uint32_t>() + m, uidx->ptr, (size_t)u * 4, cudaMemcpyDeviceToDevice, ctx->stream));
          li = li2;
          ri = ri2;
          m += u;
          null_left = true;
        }
      }
      return assemble(r.get(), li, ri, m, null_left, false);
    }
  };
  // Partitioned join over an exchange: publish the Bloom filter of this rank's build partition to every rank, so the
  // probe side's split drops, before the exchange, the rows whose destination cannot hold their key.  Collective:
  // every rank runs the same plan and takes the same branch.  Inner joins only (an outer join keeps unmatched rows).
  void offer_semi_filter() {
    if (semi_offered || !k32 || !ctx->comm || ctx->nranks < 2 || getenv("B2_DISABLE_SEMI_FILTER")) return;
    semi_offered = true;
    if (!partitioned || join_type != B2_JOIN_INNER || !children[1]->accepts_semi_filter(on_right[0])) return;
    ensure_table();  // the filter is the Bloom filter of the left side's table: no side swapping for this join
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
    if (fuse_right()) r = peel_filter(r, &s->pred);
    s->in = r->execute(partition);
    last_path = std::string(k32 ? (fuse_left() || fuse_right() ? "k32:fused-filter" : "k32") : "key128") +
                (k32 && on_left.size() > 1 ? "+verify" : "") + (partitioned ? "/partitioned" : "");
    return s;
  }
};

OpPtr make_hash_join(Ctx* ctx, OpPtr left, OpPtr right, std::vector<std::string> on_l, std::vector<std::string> on_r,
                     int join_type) {
  if (join_type != B2_JOIN_INNER && join_type != B2_JOIN_LEFT && join_type != B2_JOIN_RIGHT)
    fail(B2_ERR_INVALID_ARG, "unknown JoinType " + std::to_string(join_type));  // ballista.proto:266-270
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
  op->k32 = !kd.empty() && (int)kd.size() <= JOIN_VERIFY_MAX + 1 && !getenv("B2_DISABLE_K32_VERIFY");
  for (int t : kd) op->k32 = op->k32 && (t == B2_INT32 || t == B2_DATE32);
  if (kd.size() == 1) op->k32 = kd[0] == B2_INT32 || kd[0] == B2_DATE32;
  // output = left fields ++ right fields; an `on` column named like its partner appears once: Inner / Left keep the
  // left one, Right keeps the right one (A.3).  An outer join makes the null-extended side's fields nullable.
  auto is_dup = [&](const std::vector<int>& on_side, int c) {
    for (size_t i = 0; i < on_l.size(); ++i)
      if (c == on_side[i] && on_l[i] == on_r[i]) return true;
    return false;
  };
  for (size_t c = 0; c < ls.fields.size(); ++c) {
    if (join_type == B2_JOIN_RIGHT && is_dup(op->on_left, (int)c)) continue;
    op->left_keep.push_back((int)c);
    Field f = ls.fields[c];
    if (join_type == B2_JOIN_RIGHT) f.nullable = true;
    op->out.fields.push_back(f);
  }
  for (size_t c = 0; c < rs.fields.size(); ++c) {
    if (join_type != B2_JOIN_RIGHT && is_dup(op->on_right, (int)c)) continue;
    op->right_keep.push_back((int)c);
    Field f = rs.fields[c];
    if (join_type == B2_JOIN_LEFT) f.nullable = true;
    op->out.fields.push_back(f);
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
  uint64_t computed_version = 0;               // version of the input subtree the cached outputs were computed from
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
    if (!exchange_mode() || n_out > 32 || keys.size() != 1) return false;
    const Expr& k = *keys[0];
    return k.kind == B2_EXPR_COLUMN && k.col == key_col && (k.dtype == B2_INT32 || k.dtype == B2_DATE32);
  }
  void set_semi_filter(std::shared_ptr<SemiFilter> f) override { semi = std::move(f); }
  bool per_rank_partitions() const override { return exchange_mode(); }
  bool exchange_on(std::vector<int>* key_cols, int* n) const override {
    if (!exchange_mode()) return false;
    key_cols->clear();
    for (auto& k : keys) {
      if (k->kind != B2_EXPR_COLUMN) return false;
      key_cols->push_back(k->col);
    }
    *n = n_out;
    return true;
  }
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
    if (pred && b.num_rows && !compile_row_pred(pred, b, &rp)) mask = eval_predicate(ctx, pred, b);
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
    // the outputs of one input version are computed once and served partition by partition; re-feeding a leaf
    // (b2_memory_exec_clear / add) moves the version and the next execute() recomputes, as the reference does per task
    const uint64_t v = children[0]->version();
    if (computed && computed_version == v) return;
    computed_version = v;
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
        bool pushed = false;
        got = comm_shuffle_p2p(ctx, *moved, st, &pushed);
        if (got) {
          last_path = std::string(pushed ? "nvlink-peer-push" : "nvlink-peer-pull") + (pred ? "+fused-filter" : "");
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

}  // namespace b2
