// oracle.cpp — CPU restatement of the reference's executor-side operator algorithms.  TEST INFRASTRUCTURE:
// only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load this.
// The product (ballista_b200/) never links, imports or calls anything in oracle/.
//
// PARITY UNPINNED: the algorithms live in the un-vendored dependency datafusion/arrow 4.0.0-SNAPSHOT @
// apache/arrow 46161d26fd439c17cd7c7d2a52d5f8e0b1c14726 (rust/core/Cargo.toml:24-26, rust/Cargo.lock:78-80,
// 498-500), which is not on disk, and the reference's own tests assert no query result.
// What is restated here is that revision's published algorithm, anchored on the
// reference's call sites: FilterExec from_proto.rs:81-92, HashAggregateExec from_proto.rs:173-252,
// HashJoinExec from_proto.rs:253-276, RepartitionExec from_proto.rs:133-164, batch size 32768
// from_proto.rs:100-102.  tests/ cross-checks it against pyarrow Acero and exactly rounded sums.
//
// Restated semantics:
//   A.1 FilterExec            : per batch, Boolean mask -> every column compacted, order kept.
//   A.2 HashAggregateExec     : per batch, per row a byte key (native-endian values; Utf8 = 8-byte length
//                               prefix + bytes) looked up in a hash map; per group the batch's rows are
//                               taken and summed left-to-right (arrow compute::sum without simd), the
//                               result added to the group's running accumulator; COUNT adds the row count.
//   A.3 HashJoinExec          : collect-left build of key -> row indices (insertion order); probe rows in
//                               order; emits (left,right) index pairs.
//   A.4 RepartitionExec(Hash) : per batch, hash(key) % n, per destination `take` in arrival order.
//   A.5 expressions           : column-at-a-time temporaries, IEEE f64, one rounding per operation.
#include <algorithm>
#include <cstdint>
#include <cstring>
#include <string>
#include <thread>
#include <unordered_map>
#include <vector>

#include "../ballista_b200/csrc/tpch_spec.h"  // the workload definition (data only, no operator code)

namespace {

enum { T_INT32 = 2, T_INT64 = 3, T_UINT64 = 4, T_FLOAT64 = 5, T_DATE32 = 6, T_UTF8 = 7 };  // == b2_dtype

struct KeyCol {
  int32_t dtype;
  int32_t pad;
  const void* data;      // values, or int32 offsets for Utf8
  const uint8_t* chars;  // Utf8 bytes
};

inline void append_key(std::string& k, const KeyCol& c, int64_t row) {
  switch (c.dtype) {
    case T_INT32: case T_DATE32: k.append((const char*)((const int32_t*)c.data + row), 4); break;
    case T_INT64: case T_UINT64: case T_FLOAT64: k.append((const char*)((const int64_t*)c.data + row), 8); break;
    case T_UTF8: {
      const int32_t* off = (const int32_t*)c.data;
      uint64_t len = (uint64_t)(off[row + 1] - off[row]);
      k.append((const char*)&len, 8);
      k.append((const char*)c.chars + off[row], len);
      break;
    }
  }
}

// ---- A.2 grouped accumulation ----
struct AggState {
  int n_vals = 0, n_cnts = 0;
  std::unordered_map<std::string, int64_t> map;  // key bytes -> group index
  std::vector<int64_t> rep_row;                  // first input row of each group (global row index)
  std::vector<double> sums;                      // [group][n_vals]
  std::vector<uint64_t> cnts;                    // [group][max(n_cnts,1)]
};

// one batch [b0,b1) of the input: rows -> groups, then per group take + sequential sum
void agg_batch(AggState& st, int64_t b0, int64_t b1, int64_t row_base, int n_keys, const KeyCol* keys, int n_vals,
               const double* const* vals, int n_cnts, const uint64_t* const* cnts) {
  int nc = std::max(n_cnts, 1);
  std::vector<int64_t> touched;                  // groups seen in this batch, in first-seen order
  std::vector<std::vector<int64_t>> rows;        // per touched group: row indices of this batch
  std::unordered_map<int64_t, size_t> slot;      // group -> index into touched/rows
  std::string key;
  for (int64_t r = b0; r < b1; ++r) {
    key.clear();
    for (int k = 0; k < n_keys; ++k) append_key(key, keys[k], r);
    auto it = st.map.find(key);
    int64_t g;
    if (it == st.map.end()) {
      g = (int64_t)st.rep_row.size();
      st.map.emplace(key, g);
      st.rep_row.push_back(row_base + r);
      st.sums.resize(st.sums.size() + (size_t)st.n_vals, 0.0);
      st.cnts.resize(st.cnts.size() + (size_t)nc, 0);
    } else {
      g = it->second;
    }
    auto s = slot.find(g);
    if (s == slot.end()) {
      slot.emplace(g, touched.size());
      touched.push_back(g);
      rows.emplace_back();
      rows.back().push_back(r);
    } else {
      rows[s->second].push_back(r);
    }
  }
  std::vector<double> taken;
  for (size_t t = 0; t < touched.size(); ++t) {
    int64_t g = touched[t];
    const std::vector<int64_t>& idx = rows[t];
    for (int v = 0; v < n_vals; ++v) {
      taken.resize(idx.size());
      for (size_t q = 0; q < idx.size(); ++q) taken[q] = vals[v][idx[q]];  // take
      double s = 0.0;
      for (double x : taken) s += x;                                        // compute::sum, left to right
      st.sums[(size_t)g * st.n_vals + v] += s;                              // accumulator += batch sum
    }
    if (n_cnts == 0) {
      st.cnts[(size_t)g * nc] += idx.size();
    } else {
      for (int c = 0; c < n_cnts; ++c) {
        uint64_t s = 0;
        for (int64_t q : idx) s += cnts[c][q];
        st.cnts[(size_t)g * nc + c] += s;
      }
    }
  }
}

}  // namespace

extern "C" {

// ---------------------------------------------------------------------------------------------------
// generic primitives (driven from Python with numpy-evaluated, already filtered columns)
// ---------------------------------------------------------------------------------------------------
void* oracle_agg_create(int n_vals, int n_cnts) {
  auto* st = new AggState;
  st->n_vals = n_vals;
  st->n_cnts = n_cnts;
  return st;
}
// feed n rows as batches of batch_size (HashAggregateExec consumes its input batch by batch)
void oracle_agg_update(void* h, int64_t n, int64_t batch_size, int64_t row_base, int n_keys, const KeyCol* keys,
                       const double* const* vals, const uint64_t* const* cnts) {
  auto& st = *(AggState*)h;
  if (n_keys == 0 && n > 0 && st.rep_row.empty()) {  // ungrouped: a single accumulator set
    st.map.emplace(std::string(), 0);
    st.rep_row.push_back(row_base);
    st.sums.resize((size_t)st.n_vals, 0.0);
    st.cnts.resize((size_t)std::max(st.n_cnts, 1), 0);
  }
  for (int64_t b0 = 0; b0 < n; b0 += batch_size)
    agg_batch(st, b0, std::min(n, b0 + batch_size), row_base, n_keys, keys, st.n_vals, vals, st.n_cnts, cnts);
}
int64_t oracle_agg_num_groups(void* h) { return (int64_t)((AggState*)h)->rep_row.size(); }
void oracle_agg_result(void* h, int64_t* rep_rows, double* sums, uint64_t* cnts) {
  auto& st = *(AggState*)h;
  std::copy(st.rep_row.begin(), st.rep_row.end(), rep_rows);
  std::copy(st.sums.begin(), st.sums.end(), sums);
  std::copy(st.cnts.begin(), st.cnts.end(), cnts);
}
void oracle_agg_destroy(void* h) { delete (AggState*)h; }

// A.3: returns the number of pairs; call again with buffers of that size to fill them (two-call protocol)
int64_t oracle_hash_join(int64_t n_left, int n_keys, const KeyCol* lkeys, int64_t n_right, const KeyCol* rkeys,
                         int64_t* left_idx, int64_t* right_idx, int64_t capacity) {
  std::unordered_map<std::string, std::vector<int64_t>> map;
  std::string key;
  for (int64_t r = 0; r < n_left; ++r) {
    key.clear();
    for (int k = 0; k < n_keys; ++k) append_key(key, lkeys[k], r);
    map[key].push_back(r);
  }
  int64_t m = 0;
  for (int64_t r = 0; r < n_right; ++r) {
    key.clear();
    for (int k = 0; k < n_keys; ++k) append_key(key, rkeys[k], r);
    auto it = map.find(key);
    if (it == map.end()) continue;
    for (int64_t l : it->second) {
      if (m < capacity) {
        left_idx[m] = l;
        right_idx[m] = r;
      }
      ++m;
    }
  }
  return m;
}

// The same join, one pass: the pairs are kept inside a handle and fetched once their number is known (the two-call
// protocol above joins twice).  Keys of fixed-width columns totalling <= 16 bytes - every TPC-H join - go through an
// open-addressing table of (key -> first / last build row) with a `next` chain per build row, which is the same
// structure (hash of the key -> build row indices in insertion order, probe rows in order) without a heap allocation
// per key; anything else falls back to the byte-string map.
struct JoinResult {
  std::vector<int64_t> l, r;
};
static inline uint64_t mix64_(uint64_t x) {
  x ^= x >> 33;
  x *= 0xff51afd7ed558ccdull;
  x ^= x >> 33;
  x *= 0xc4ceb9fe1a85ec53ull;
  x ^= x >> 33;
  return x;
}
static int fixed_key_bytes(int n_keys, const KeyCol* keys) {
  int total = 0;
  for (int k = 0; k < n_keys; ++k) {
    switch (keys[k].dtype) {
      case T_INT32: case T_DATE32: total += 4; break;
      case T_INT64: case T_UINT64: case T_FLOAT64: total += 8; break;
      default: return -1;
    }
  }
  return total;
}
static inline void fixed_key(int n_keys, const KeyCol* keys, int64_t row, uint64_t out[2]) {
  unsigned char buf[16] = {0};
  int off = 0;
  for (int k = 0; k < n_keys; ++k) {
    if (keys[k].dtype == T_INT32 || keys[k].dtype == T_DATE32) {
      memcpy(buf + off, (const int32_t*)keys[k].data + row, 4);
      off += 4;
    } else {
      memcpy(buf + off, (const int64_t*)keys[k].data + row, 8);
      off += 8;
    }
  }
  memcpy(out, buf, 16);
}
void* oracle_hash_join_run(int64_t n_left, int n_keys, const KeyCol* lkeys, int64_t n_right, const KeyCol* rkeys) {
  auto* res = new JoinResult;
  int kb = fixed_key_bytes(n_keys, lkeys);
  if (kb < 0 || kb > 16 || fixed_key_bytes(n_keys, rkeys) != kb) {
    std::unordered_map<std::string, std::vector<int64_t>> map;
    std::string key;
    for (int64_t r = 0; r < n_left; ++r) {
      key.clear();
      for (int k = 0; k < n_keys; ++k) append_key(key, lkeys[k], r);
      map[key].push_back(r);
    }
    for (int64_t r = 0; r < n_right; ++r) {
      key.clear();
      for (int k = 0; k < n_keys; ++k) append_key(key, rkeys[k], r);
      auto it = map.find(key);
      if (it == map.end()) continue;
      for (int64_t l : it->second) {
        res->l.push_back(l);
        res->r.push_back(r);
      }
    }
    return res;
  }
  struct Slot {
    uint64_t k0, k1;
    int64_t first, last;
  };
  uint64_t cap = 16;
  while (cap < 2 * (uint64_t)n_left + 16) cap <<= 1;
  std::vector<Slot> slots(cap, Slot{0, 0, -1, -1});
  std::vector<int64_t> next((size_t)n_left, -1);
  uint64_t key[2];
  for (int64_t r = 0; r < n_left; ++r) {
    fixed_key(n_keys, lkeys, r, key);
    uint64_t h = mix64_(key[0] ^ mix64_(key[1] + 0x9E3779B97F4A7C15ull)) & (cap - 1);
    while (true) {
      Slot& sl = slots[h];
      if (sl.first < 0) {
        sl = Slot{key[0], key[1], r, r};
        break;
      }
      if (sl.k0 == key[0] && sl.k1 == key[1]) {
        next[(size_t)sl.last] = r;  // insertion order
        sl.last = r;
        break;
      }
      h = (h + 1) & (cap - 1);
    }
  }
  for (int64_t r = 0; r < n_right; ++r) {
    fixed_key(n_keys, rkeys, r, key);
    uint64_t h = mix64_(key[0] ^ mix64_(key[1] + 0x9E3779B97F4A7C15ull)) & (cap - 1);
    while (true) {
      const Slot& sl = slots[h];
      if (sl.first < 0) break;
      if (sl.k0 == key[0] && sl.k1 == key[1]) {
        for (int64_t l = sl.first; l >= 0; l = next[(size_t)l]) {
          res->l.push_back(l);
          res->r.push_back(r);
        }
        break;
      }
      h = (h + 1) & (cap - 1);
    }
  }
  return res;
}
int64_t oracle_join_size(void* h) { return (int64_t)((JoinResult*)h)->l.size(); }
void oracle_join_fetch(void* h, int64_t* left_idx, int64_t* right_idx) {
  auto* res = (JoinResult*)h;
  std::copy(res->l.begin(), res->l.end(), left_idx);
  std::copy(res->r.begin(), res->r.end(), right_idx);
  delete res;
}

// A.4: destination of every row.  The reference hashes with ahash (seeds fixed); no consumer observes which
// partition a key lands in, only that equal keys co-locate, so the oracle uses FNV-1a.
void oracle_hash_partition(int64_t n, int n_keys, const KeyCol* keys, int n_out, int32_t* dest) {
  std::string key;
  for (int64_t r = 0; r < n; ++r) {
    key.clear();
    for (int k = 0; k < n_keys; ++k) append_key(key, keys[k], r);
    uint64_t h = 1469598103934665603ull;
    for (unsigned char c : key) h = (h ^ c) * 1099511628211ull;
    dest[r] = (int32_t)(h % (uint64_t)n_out);
  }
}

// ---------------------------------------------------------------------------------------------------
// workload generation on the host (same closed-form spec the CUDA generator compiles)
// ---------------------------------------------------------------------------------------------------
void oracle_gen_lineitem(double sf, int64_t r0, int64_t n, int32_t* orderkey, int32_t* suppkey, double* quantity,
                         double* extendedprice, double* discount, double* tax, int32_t* shipdate,
                         int32_t* commitdate, int32_t* receiptdate, int32_t* rf_off, uint8_t* rf_chr,
                         int32_t* ls_off, uint8_t* ls_chr, int n_threads) {
  auto work = [&](int64_t a, int64_t b) {
    for (int64_t i = a; i < b; ++i) {
      b2spec::LineItem li = b2spec::lineitem_row(r0 + i, sf);
      if (orderkey) orderkey[i] = li.l_orderkey;
      if (suppkey) suppkey[i] = li.l_suppkey;
      if (quantity) quantity[i] = li.l_quantity;
      if (extendedprice) extendedprice[i] = li.l_extendedprice;
      if (discount) discount[i] = li.l_discount;
      if (tax) tax[i] = li.l_tax;
      if (shipdate) shipdate[i] = li.l_shipdate;
      if (commitdate) commitdate[i] = li.l_commitdate;
      if (receiptdate) receiptdate[i] = li.l_receiptdate;
      if (rf_off) { rf_off[i] = (int32_t)i; rf_chr[i] = (uint8_t)li.l_returnflag; }
      if (ls_off) { ls_off[i] = (int32_t)i; ls_chr[i] = (uint8_t)li.l_linestatus; }
    }
  };
  if (n_threads < 1) n_threads = 1;
  std::vector<std::thread> th;
  for (int t = 0; t < n_threads; ++t) th.emplace_back(work, n * t / n_threads, n * (t + 1) / n_threads);
  for (auto& t : th) t.join();
  if (rf_off) rf_off[n] = (int32_t)n;
  if (ls_off) ls_off[n] = (int32_t)n;
}
void oracle_gen_orders(double sf, int64_t r0, int64_t n, int32_t* orderkey, int32_t* custkey, int32_t* orderdate,
                       int32_t* shippriority) {
  int64_t ncust = b2spec::n_customer(sf);
  for (int64_t i = 0; i < n; ++i) {
    if (orderkey) orderkey[i] = b2spec::o_orderkey(r0 + i);
    if (custkey) custkey[i] = b2spec::o_custkey(r0 + i, ncust);
    if (orderdate) orderdate[i] = b2spec::o_orderdate(r0 + i);
    if (shippriority) shippriority[i] = b2spec::o_shippriority(r0 + i);
  }
}
// returns total chars; seg_chr must hold 10*n bytes
int64_t oracle_gen_customer(double, int64_t r0, int64_t n, int32_t* custkey, int32_t* nationkey, int32_t* seg_off,
                            uint8_t* seg_chr) {
  int64_t pos = 0;
  for (int64_t i = 0; i < n; ++i) {
    if (custkey) custkey[i] = b2spec::c_custkey(r0 + i);
    if (nationkey) nationkey[i] = b2spec::c_nationkey(r0 + i);
    if (seg_off) {
      int id = b2spec::c_mktsegment_id(r0 + i);
      int len = b2spec::mktsegment_len(id);
      seg_off[i] = (int32_t)pos;
      memcpy(seg_chr + pos, b2spec::mktsegment_name(id), (size_t)len);
      pos += len;
    }
  }
  if (seg_off) seg_off[n] = (int32_t)pos;
  return pos;
}
void oracle_gen_supplier(double, int64_t r0, int64_t n, int32_t* suppkey, int32_t* nationkey) {
  for (int64_t i = 0; i < n; ++i) {
    if (suppkey) suppkey[i] = b2spec::s_suppkey(r0 + i);
    if (nationkey) nationkey[i] = b2spec::s_nationkey(r0 + i);
  }
}
const char* oracle_nation_name(int id) { return b2spec::nation_name(id); }
int oracle_nation_regionkey(int id) { return b2spec::nation_regionkey(id); }
const char* oracle_region_name(int id) { return b2spec::region_name(id); }
int64_t oracle_tpch_num_rows(const char* table, double sf) {
  if (!strcmp(table, "lineitem")) return b2spec::n_lineitem(sf);
  if (!strcmp(table, "orders")) return b2spec::n_orders(sf);
  if (!strcmp(table, "customer")) return b2spec::n_customer(sf);
  if (!strcmp(table, "supplier")) return b2spec::n_supplier(sf);
  if (!strcmp(table, "nation")) return b2spec::N_NATION;
  if (!strcmp(table, "region")) return b2spec::N_REGION;
  return -1;
}

}  // extern "C"
