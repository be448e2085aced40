// ops.h — physical operators (the ExecutionPlan trait objects behind b2_op) and the helper entry points
// of the kernel files.
#pragma once
#include <memory>
#include <string>
#include <vector>

#include "agg.h"
#include "core.h"
#include "expr.h"

namespace b2 {

// ---- kernel-file helpers ----
BufPtr mask_to_indices(Ctx* ctx, const uint8_t* mask, int64_t n, int64_t* count);        // take.cu
// idx 0xFFFFFFFF = null-extended row (outer joins); pass null_extended when the index vector may hold such entries
Column take_column(Ctx* ctx, const Column& c, const uint32_t* idx, int64_t m, bool null_extended = false);  // take.cu
// the same for several columns through one index vector: fixed-width columns without NULLs share one launch
std::vector<Column> take_columns(Ctx* ctx, const std::vector<const Column*>& cols, const uint32_t* idx, int64_t m,
                                 bool null_extended = false);
Column null_column(Ctx* ctx, int dtype, int64_t n);                                       // take.cu
BatchPtr take_batch(Ctx* ctx, const Batch& b, const uint32_t* idx, int64_t m);            // take.cu
BatchPtr concat_batches(Ctx* ctx, const Schema& schema, const std::vector<BatchPtr>& b);  // take.cu
BatchPtr empty_batch(Ctx* ctx, const Schema& schema);                                      // take.cu
BatchPtr slice_batch(Ctx* ctx, const Batch& b, int64_t begin, int64_t end);               // repartition.cu
BatchPtr tpch_generate(Ctx* ctx, const std::string& table, double sf, int64_t row_begin, int64_t row_end,
                       const char* const* columns);                                       // tpch_gen.cu
int64_t tpch_num_rows(const std::string& table, double sf);

// join.cu
// Bloom filters of a partitioned join's build keys, one per destination rank (filter d covers the build keys that
// live on rank d; word = fmix32(key) & wmask[d], bits = bloom_bits).  The probe side's hash split tests every row
// against the filter of its destination and gives the row no destination on a miss: rows that cannot match never
// cross NVLink (Q3: ~98 % of the lineitem rows that pass the date filter).
struct SemiFilter {
  BufPtr words;          // n filters, `stride` 32-bit words apart
  uint32_t stride = 0;
  uint32_t wmask[32] = {0};
  int n = 0;
};

struct JoinTable {
  Ctx* ctx;
  u64 cap = 0;
  int64_t n_build = 0;
  BufPtr keys, head, count, next;
  // 32-bit-key variant (join_build32): slots[cap] = (key << 32) | head_row; flags = {dup flag, -, count(8 B)}
  bool k32 = false;
  BufPtr slots, flags, bloom;  // bloom: blocked Bloom filter over the build keys (32-bit words, 3 bits per key)
  u64 bloom_words = 0;
  int bshift = 28;  // bucket = table_hash(fmix32(key)) >> bshift
  bool dups_known = false, has_dups = false;
};
// in-place exclusive scan of v[0..n) on the context's stream; v[n] = total (single CTA for short inputs, chunked above)
void scan_i64(Ctx* ctx, int64_t* v, int64_t n);
std::shared_ptr<JoinTable> join_build(Ctx* ctx, const Key128* keys, int64_t n);
// probe: returns number of output pairs; left_idx / right_idx are uint32 row indices in probe order
int64_t join_probe(Ctx* ctx, const JoinTable& t, const Key128* probe_keys, int64_t n_probe, BufPtr* left_idx,
                   BufPtr* right_idx);

// A FilterExec predicate evaluated inside the join kernels: a conjunction of up to 4 comparisons of an
// Int32 / Date32 / Float64 column with a literal (every TPC-H date / key / quantity filter).  Anything else is
// evaluated column-at-a-time into a byte mask first (eval.cu) and passed as `mask`.
struct RowPred {  // every term is a closed range test lo <= column[i] <= hi (strict bounds are tightened on the host)
  int n = 0;
  const void* col[4];
  int is_f64[4];
  int ilo[4];            // Int32 / Date32 terms: (uint32)(x - ilo) <= iwidth
  unsigned iwidth[4];
  double flo[4], fhi[4];  // Float64 terms
};
bool compile_row_pred(const ExprPtr& pred, const Batch& batch, RowPred* out);
// single Int32 / Date32 key: fused filter (RowPred and / or mask, both optional) + build / probe over the
// uncompacted key column; row ids are those of the unfiltered batch
std::shared_ptr<JoinTable> join_build32(Ctx* ctx, const int32_t* key, const RowPred& pred, const uint8_t* mask,
                                        int64_t n);
// Further Int32 / Date32 key pairs of a join whose 32-bit table is keyed on the FIRST pair: a candidate pair (build row,
// probe row) stays when every further pair agrees and neither side is NULL there (validity bytes, optional)
constexpr int JOIN_VERIFY_MAX = 4;
struct JoinVerify {
  int n = 0;
  const int32_t* bcol[JOIN_VERIFY_MAX];
  const int32_t* pcol[JOIN_VERIFY_MAX];
  const uint8_t* bval[JOIN_VERIFY_MAX];
  const uint8_t* pval[JOIN_VERIFY_MAX];
};
int64_t join_probe32(Ctx* ctx, JoinTable& t, const int32_t* key, const RowPred& pred, const uint8_t* mask,
                     int64_t n_probe, BufPtr* left_idx, BufPtr* right_idx, const JoinVerify* verify = nullptr);

// repartition.cu: rows -> destination partition by hash of packed key; returns the stable permutation
// (destination-major) and fills counts[n_out]
// `key32` (optional): the single Int32 / Date32 key column, hashed as its packed form without materialising keys
// `pred` / `mask` (optional, n_out <= 32 only): a fused FilterExec; rejected rows appear in no destination
BufPtr hash_partition_indices(Ctx* ctx, const Key128* keys, const int32_t* key32, int64_t n, int n_out,
                              std::vector<int64_t>* counts, const RowPred* pred = nullptr,
                              const uint8_t* mask = nullptr);

// two-phase form used by the shuffle: phase 1 = destination bytes + per-tile histogram + scan (no host sync)
struct SplitState {
  BufPtr dest, hist;
  int64_t n = 0, n_tiles = 0;
  int n_out = 0;
};
// table_order: split by the top bits of table_hash (the join table's bucket order) instead of the exchange's hash
SplitState hash_split_begin(Ctx* ctx, const Key128* keys, const int32_t* key32, int64_t n, int n_out,
                            const RowPred* pred = nullptr, const uint8_t* mask = nullptr,
                            const SemiFilter* semi = nullptr, bool table_order = false);
void hash_split_starts_device(Ctx* ctx, const SplitState& st, int64_t* out_dev);  // n_out + 1 slice starts
void hash_split_counts(Ctx* ctx, const SplitState& st, std::vector<int64_t>* counts);
BufPtr hash_split_perm(Ctx* ctx, const SplitState& st);
// moves fixed-width columns to dst[c] + base[c * n_out + d] + k * width[c] (k = rank inside destination d's slice)
void hash_split_move(Ctx* ctx, const SplitState& st, int n_cols, const void* const* src, void* const* dst,
                     const int* width, const int64_t* base_dev, uint32_t* perm_or_null, int64_t n_live = -1);
// push form (the exchange's remote stores): addr_dev[c * n_out + d] = absolute address of this source's first element of
// column c at destination d; a tile's values are staged destination-major in shared memory and written in whole runs
void hash_split_push(Ctx* ctx, const SplitState& st, int n_cols, const void* const* src, const int* width,
                     const int64_t* addr_dev, int64_t n_live = -1);
// allocation-free split of one Int32 / Date32 key column: scratch in a file-scope device array (repartition.cu), at most
// hash_split_i32_static_max_rows() rows per call; perm = destination-major stable row ids, starts = n_out + 1 slice starts
void hash_split_i32_static(int device, cudaStream_t stream, const int32_t* keys, int64_t n, int n_out, uint32_t* perm,
                           int64_t* starts);
int64_t hash_split_i32_static_max_rows();

// SortExec support (repartition.cu): order-preserving 64-bit key words per sort column, stable LSD radix sort
void sort_key_words(Ctx* ctx, const Column& c, bool descending, std::vector<BufPtr>* words, std::vector<int>* bits,
                    int* status_dev);
BufPtr radix_sort_order(Ctx* ctx, const std::vector<const u64*>& words, const std::vector<int>& bits, int64_t n);

// comm.cu
void comm_unique_id(void* id128);
void comm_init(Ctx* ctx, int rank, int nranks, const void* id128);
void comm_destroy(Ctx* ctx);
// allow_lazy: a lazy `local` (device-side row count) may yield a lazy result (nothing read back on the hop)
BatchPtr comm_allgather_batch(Ctx* ctx, const BatchPtr& local, bool allow_lazy = false);
// all-to-all of a destination-major permuted batch: slice d (counts[d] rows) goes to rank d
BatchPtr comm_all_to_all(Ctx* ctx, const Batch& permuted, const std::vector<int64_t>& counts);
// shuffle.cu: the same exchange over CUDA-IPC peer memory (columns moved into a peer-mapped arena by the split's
// second phase + one NVLink pull kernel).  Takes the unpermuted batch and phase 1 of its split.
// nullptr = peer mapping unavailable.
BatchPtr comm_shuffle_p2p(Ctx* ctx, const Batch& local, const SplitState& st, bool* pushed = nullptr);

// ---- RecordBatchStream ----
struct Stream {
  virtual ~Stream() {}
  virtual BatchPtr next() = 0;  // nullptr at end of stream; never a lazy batch
  // may return a batch whose row count is still on the device (Batch::lazy); only consumers that work from the
  // device-side count ask for it
  virtual BatchPtr next_lazy() { return next(); }
};
using StreamPtr = std::unique_ptr<Stream>;

// ---- ExecutionPlan ----
struct Op : std::enable_shared_from_this<Op> {
  Ctx* ctx = nullptr;
  std::vector<std::shared_ptr<Op>> children;
  std::string last_path = "unexecuted";
  virtual ~Op() {}
  virtual const char* name() const = 0;
  virtual Schema schema() const = 0;
  virtual void partitioning(int* scheme, int* n) const {
    children[0]->partitioning(scheme, n);
  }
  virtual StreamPtr execute(int partition) = 0;
  // Column pruning at execution time: a parent announces which of this operator's output columns it will
  // read (empty = all).  Operators that assemble their output by gathers (HashJoinExec) leave the others
  // elided (Column::data == nullptr); the schema is unchanged.  Default: children produce everything.
  virtual void require(const std::vector<char>& need) { (void)need; }
  // Semi-join reduction across an exchange (see SemiFilter): a partitioned hash join offers the Bloom filters of
  // its build partitions to its probe-side child; only a hash RepartitionExec on exactly that key column accepts
  virtual bool accepts_semi_filter(int key_col) const { (void)key_col; return false; }
  virtual void set_semi_filter(std::shared_ptr<SemiFilter> f) { (void)f; }
  // true when output partition r exists only on rank r (the plan sits above an exchange): consumers then
  // execute their own rank's partition instead of iterating over all of them
  virtual bool per_rank_partitions() const { return !children.empty() && children[0]->per_rank_partitions(); }
  // When this operator's output partitions are those of a hash exchange over plain columns (a communicator-attached
  // RepartitionExec, seen through operators that neither move rows nor renumber columns): the key columns and the
  // partition count.  A HashJoinExec joins partition-wise only when BOTH inputs report exchanges on its own keys.
  virtual bool exchange_on(std::vector<int>* key_cols, int* n_out) const {
    (void)key_cols;
    (void)n_out;
    return false;
  }
  // Changes whenever the data a leaf holds changes (MemoryExec: add / clear); operators that cache work across
  // execute() calls (join build table, repartition outputs) recompute when their subtree's version moved.
  virtual uint64_t version() const {
    uint64_t v = 1469598103934665603ull;
    for (auto& c : children) v = (v ^ c->version()) * 1099511628211ull;
    return v;
  }
  int n_partitions() const {
    int s, n;
    partitioning(&s, &n);
    return n;
  }
};
using OpPtr = std::shared_ptr<Op>;

struct AggSpec {
  int fn;
  ExprPtr arg;
  std::string name;
};

OpPtr make_memory_exec(Ctx* ctx, const Schema& schema, int n_partitions);
void memory_exec_add_host(Op* op, int partition, ArrowArray* batch);
void memory_exec_add_device(Op* op, int partition, BatchPtr batch);
void memory_exec_clear(Op* op);
OpPtr make_filter(Ctx* ctx, OpPtr input, ExprPtr pred);
OpPtr make_coalesce(Ctx* ctx, OpPtr input, int64_t target);
OpPtr make_merge(Ctx* ctx, OpPtr input);
OpPtr make_projection(Ctx* ctx, OpPtr input, std::vector<ExprPtr> exprs);
OpPtr make_sort(Ctx* ctx, OpPtr input, std::vector<ExprPtr> exprs, std::vector<char> descending);
OpPtr make_limit(Ctx* ctx, OpPtr input, int64_t limit, bool global);
OpPtr make_hash_aggregate(Ctx* ctx, OpPtr input, int mode, std::vector<ExprPtr> groups, std::vector<AggSpec> aggs,
                          const Schema* input_schema);
OpPtr make_hash_join(Ctx* ctx, OpPtr left, OpPtr right, std::vector<std::string> on_left,
                     std::vector<std::string> on_right, int join_type);
OpPtr make_repartition(Ctx* ctx, OpPtr input, int scheme, std::vector<ExprPtr> keys, int n_out);

// csv.cu: delimiter-separated text (dbgen .tbl) -> projected columns in HBM; `text` holds whole rows
BatchPtr csv_decode(Ctx* ctx, const Schema& file_schema, const std::vector<int>& projection, char delimiter,
                    bool has_header, const char* host_text, size_t len);
// proto.cpp: serialized PhysicalPlanNode (ballista.proto:294-422) -> operator tree; leaves bound in pre-order
OpPtr plan_from_proto(Ctx* ctx, const uint8_t* buf, size_t len, const std::vector<OpPtr>& leaves);
std::string plan_proto_describe(const uint8_t* buf, size_t len, int* n_leaves);

}  // namespace b2

struct b2_op { b2::OpPtr op; };
struct b2_stream { b2::StreamPtr s; };
