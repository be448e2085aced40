// agg.h — host side of the hash aggregate: key layout, HBM group table, fused-kernel registry, runner.
#pragma once
#include <functional>
#include <map>
#include <string>
#include <vector>

#include "core.h"
#include "expr.h"
#include "table.cuh"

namespace b2 {

// byte layout of a packed key (mirror of fx::PackKey): parts in group-expression order
struct KeyPart {
  int dtype;
  int byte_off;
  int bytes;  // Int32/Date32: 4, Int64: 8, Utf8: 1 length byte + (bytes-1) chars
};
struct KeyLayout {
  std::vector<KeyPart> parts;
};
KeyLayout make_key_layout(const std::vector<int>& dtypes);  // throws NotImplemented when > 15 bytes
// only a Utf8 part can overflow its budget: integer-only keys need no status readback after pack_keys
inline bool layout_has_utf8(const KeyLayout& l) {
  for (auto& p : l.parts)
    if (p.dtype == B2_UTF8) return true;
  return false;
}

struct GroupTable {
  Ctx* ctx;
  u64 cap;
  int na, nc;
  BufPtr keys, acc, cnt, n_groups, slot_list;
  TableRef ref() const {
    return TableRef{keys->as<Key128>(), acc->as<double>(), cnt->as<u64>(), cap, n_groups->as<int>(),
                    slot_list->as<uint32_t>()};
  }
};
std::shared_ptr<GroupTable> make_group_table(Ctx* ctx, u64 cap, int na, int nc, bool init = true);
void reset_group_table(GroupTable& t);

// type-erased launchers of the two aggregate kernels for one bound source
struct AggLaunch {
  std::function<void(int64_t n, TableRef tbl, int* status)> small;
  std::function<void(int64_t row_begin, int64_t row_end, int64_t n, TableRef tbl, int* status)> global;
  int na = 0, nc = 0;
  bool keyed = false;
  bool narrow = false;
  bool tma = false;  // `small` is the TMA-fed bandwidth kernel (agg_tma.cuh)
  // when `small` is a narrow-key kernel (8-byte keys, Utf8 values of <= 7 bytes) and reports
  // AGG_KEY_TOO_WIDE, the batch is redone with this wide (15-byte) variant
  std::function<void(int64_t n, TableRef tbl, int* status)> small_wide;
};

// ---- fused-kernel registry: signature text -> instantiated kernels ----
struct FusedParams;
struct FusedEntry {
  std::string sig;
  int nf, ni, ns, na;
  bool keyed;
  std::function<AggLaunch(Ctx*, const FusedParams&)> bind;
};
const std::map<std::string, FusedEntry>& fused_registry();

// Try to bind a fused kernel for (predicate?, group exprs, distinct aggregate args) over `batch`.
// Returns false when the shape is not in the registry (the caller then runs the unfused path).
bool bind_fused_agg(Ctx* ctx, const ExprPtr& pred, const std::vector<ExprPtr>& groups,
                    const std::vector<ExprPtr>& args, const Batch& batch, AggLaunch* out, std::string* sig);
// Materialised source: keys (null = ungrouped), mask (null = all rows), value / count columns.
// `cnts`: one u64 column per count accumulator; a null entry counts 1 per row (COUNT(*), Partial mode).
AggLaunch bind_column_agg(Ctx* ctx, const Key128* keys, const uint8_t* mask, const std::vector<const double*>& vals,
                          const std::vector<const u64*>& cnts);
// accumulator counts of the instantiated materialised-source kernels (what bind_column_agg pads a request to)
inline int agg_padded_na(int n) { return n <= 1 ? 1 : n <= 2 ? 2 : n <= 4 ? 4 : 8; }
inline int agg_padded_nc(int n) { return n <= 1 ? 1 : n <= 4 ? 4 : 12; }
// pack evaluated group columns into a Key128 column
// null_tag: byte 15 of the key of a row with a NULL key part (joins give each side its own tag, so that null keys
// match nothing); 0 = NULL key parts are refused (group / partition keys)
BufPtr pack_keys(Ctx* ctx, const KeyLayout& layout, const std::vector<Column>& cols, int64_t n, int* status_dev,
                 int null_tag = 0);

// Strategy state of one HashAggregateExec execution (one partition)
struct AggRunner {
  Ctx* ctx;
  int na, nc;
  bool keyed;
  std::shared_ptr<GroupTable> main, delta;
  BufPtr status;  // int[4]: status, spare
  bool force_global = false;
  bool narrow_overflowed = false;  // a narrow-key kernel overflowed once: later batches go straight to wide keys
  int64_t group_bound = 0;  // upper bound on groups in `main`
  bool groups_exact = false;  // group_bound is main's exact group count and nothing ran on the table since (finish() skips its readback)
  bool speculative = false;   // consume_unchecked() ran: status / group count are still on the device
  std::string path;         // diagnostics
  AggRunner(Ctx* c, int na_, int nc_, bool keyed_);
  void consume(const AggLaunch& l, int64_t n);
  // The whole input is ONE batch: the low-cardinality kernel writes straight into `main` and nothing is read back;
  // the emit step returns status and group count with its own (single) readback, and the caller redoes the batch
  // through consume() when the status says so.  False when this runner cannot speculate (global path forced).
  bool consume_unchecked(const AggLaunch& l, int64_t n);
  void reset();  // forget everything accumulated so far (after a failed speculation)
  // number of groups in `main` (synchronises unless already known)
  int64_t finish();
  void check_status(int st);
 private:
  void ensure_main(int64_t extra_groups);
};

// Output assembly (agg.cu): every output column of one aggregate is gathered from the table by ONE launch
// (blockIdx.y = column) through the table's slot list; status, group count and the character totals of all Utf8 key
// columns come back in ONE readback.
struct EmitPlan {
  struct EmitDesc {
    int kind, byte_off;
    const void *src, *src2;
    const u64* vsrc;  // validity = (vsrc[slot] > 0), or null: always valid
  };
  Ctx* ctx;
  const GroupTable& table;
  int64_t bound;  // rows the output columns are allocated for (>= the group count, or the emit reports failure)
  std::vector<EmitDesc> descs;
  std::vector<Column> cols;
  // lazy results: header + all column buffers are carved out of one arena, so the export copies once
  BufPtr arena, hdr;
  size_t arena_used = 0;
  EmitPlan(Ctx* ctx, const GroupTable& t, int64_t bound, bool use_arena = false);
  void key(const KeyPart& part);
  void sum(int acc_idx, int vcnt_idx = -1);
  void isum(int cnt_idx, int dtype, int vcnt_idx = -1);  // Int64 / UInt64 sums live in the count accumulators
  void count(int cnt_idx);
  void avg(int acc_idx, int cnt_idx, bool track_null);
  // Runs the emit.  exact_ng >= 0: the host already knows the group count (no status to check).  Otherwise the count
  // and `status_dev` are read back with the Utf8 totals; returns false (and *status_out) when the speculation
  // failed: bad status or more groups than `bound`.  On success *ng_out rows are valid in every column.
  bool run(int64_t exact_ng, const int* status_dev, int64_t* ng_out, int* status_out);
  // The same launch with nothing read back: columns keep `bound` rows, the returned header (int64 in HBM: status,
  // group count, character totals by column) is what Batch::lazy carries.
  BufPtr run_lazy(const int* status_dev);
 private:
  void push(int kind, int byte_off, const void* src, const void* src2, const u64* vsrc, Column col);
  BufPtr alloc(size_t bytes);
  Column new_fixed(int dtype);
};

// ---- small Final aggregate: partial states of a few rows (one batch per partition or per rank) merged by ONE
// single-CTA kernel that packs keys, claims groups and accumulates straight from the state columns, reading each
// batch's row count from its device-side header when the batch is lazy.  (agg.cu) ----
constexpr int TINY_BATCHES = 4;        // per launch
constexpr int64_t TINY_MAX_ROWS = 8192;  // all batches together
constexpr u64 TINY_CAP = 2048;         // slots of the group table (the emit bound is 1024 groups)
constexpr int64_t SPEC_BOUND = 1024;   // groups a speculative (lazy) aggregate result may hold
struct TinyCol {
  const void* data;
  const uint8_t* valid;
  const uint8_t* chars;
};
struct TinyBatch {
  const int64_t* hdr;  // lazy input: [0] status, [1] rows; null = `rows` is exact
  int64_t rows, bound;
  int redo_if_empty;
  TinyCol key[8], val[8], cnt[12];
};
struct TinyArgs {
  int nb, nkeys, na, nc, init;
  int single;  // this launch holds every input batch: the hashed part of the table is sized from the rows present
  int part_dtype[8], part_off[8], part_bytes[8];
  int vslot[8];    // count slot that receives the number of non-NULL rows of value a, or -1
  int cvslot[12];  // likewise for an integer-sum column
  TinyBatch b[TINY_BATCHES];
};
void final_tiny_launch(Ctx* ctx, const TinyArgs& a, const GroupTable& t, int* status);

}  // namespace b2
