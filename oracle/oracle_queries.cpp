// oracle_queries.cpp — the reference's physical plans for TPC-H Q1 / Q6 executed the
// way DataFusion@46161d2 executes them on the CPU: one task per partition (rust/scheduler/src/lib.rs:381-399),
// 32,768-row batches (from_proto.rs:100-102), FilterExec materialising a compacted copy of every column,
// CoalesceBatchesExec, column-at-a-time expression temporaries without common-subexpression elimination,
// HashAggregateExec(Partial) per partition, then merge + HashAggregateExec(Final).
// TEST INFRASTRUCTURE / CPU BASELINE ONLY (see oracle.cpp header).  Parity unpinned (see oracle.cpp).
#include <algorithm>
#include <cstdint>
#include <cstring>
#include <string>
#include <thread>
#include <unordered_map>
#include <vector>

namespace {

constexpr int64_t BATCH = 32768;

struct Q1Group {
  char rf, ls;
  double sum_qty = 0, sum_base = 0, sum_disc_price = 0, sum_charge = 0;
  double avg_qty_sum = 0, avg_price_sum = 0, avg_disc_sum = 0;
  uint64_t avg_qty_cnt = 0, avg_price_cnt = 0, avg_disc_cnt = 0, count = 0;
};

struct LineitemCols {
  const double *qty, *price, *disc, *tax;
  const int32_t* shipdate;
  const int32_t *rf_off, *ls_off;
  const uint8_t *rf_chr, *ls_chr;
};

// a materialised (filtered / coalesced) batch of the 7 Q1 columns
struct Q1Batch {
  std::vector<double> qty, price, disc, tax;
  std::vector<int32_t> shipdate;
  std::vector<std::string> rf, ls;  // Utf8 values
  size_t rows() const { return qty.size(); }
  void clear() { qty.clear(); price.clear(); disc.clear(); tax.clear(); shipdate.clear(); rf.clear(); ls.clear(); }
};

inline double seq_sum(const std::vector<double>& v) {
  double s = 0.0;
  for (double x : v) s += x;
  return s;
}

// HashAggregateExec(Partial) over one batch: expression temporaries, byte keys, per-group take + sum
void q1_partial_batch(const Q1Batch& b, std::unordered_map<std::string, Q1Group>& groups) {
  size_t n = b.rows();
  // aggregate input expressions, column at a time, no CSE (q1.sql:6 and :7 both evaluate price*(1-disc))
  std::vector<double> t1(n), disc_price(n), t3(n), t4(n), t5(n), charge(n);
  for (size_t i = 0; i < n; ++i) t1[i] = 1.0 - b.disc[i];
  for (size_t i = 0; i < n; ++i) disc_price[i] = b.price[i] * t1[i];
  for (size_t i = 0; i < n; ++i) t3[i] = 1.0 - b.disc[i];
  for (size_t i = 0; i < n; ++i) t4[i] = b.price[i] * t3[i];
  for (size_t i = 0; i < n; ++i) t5[i] = 1.0 + b.tax[i];
  for (size_t i = 0; i < n; ++i) charge[i] = t4[i] * t5[i];
  // rows -> groups
  std::unordered_map<std::string, std::vector<uint32_t>> idx;
  std::vector<std::string> order;
  std::string key;
  for (size_t i = 0; i < n; ++i) {
    key.clear();
    uint64_t l1 = b.rf[i].size(), l2 = b.ls[i].size();
    key.append((const char*)&l1, 8);
    key.append(b.rf[i]);
    key.append((const char*)&l2, 8);
    key.append(b.ls[i]);
    auto it = idx.find(key);
    if (it == idx.end()) {
      order.push_back(key);
      it = idx.emplace(key, std::vector<uint32_t>()).first;
    }
    it->second.push_back((uint32_t)i);
  }
  std::vector<double> taken;
  auto take_sum = [&](const std::vector<double>& col, const std::vector<uint32_t>& rows) {
    taken.resize(rows.size());
    for (size_t q = 0; q < rows.size(); ++q) taken[q] = col[rows[q]];
    return seq_sum(taken);
  };
  for (auto& k : order) {
    const auto& rows = idx[k];
    Q1Group& g = groups[k];
    g.rf = b.rf[rows[0]][0];
    g.ls = b.ls[rows[0]][0];
    g.sum_qty += take_sum(b.qty, rows);
    g.sum_base += take_sum(b.price, rows);
    g.sum_disc_price += take_sum(disc_price, rows);
    g.sum_charge += take_sum(charge, rows);
    g.avg_qty_sum += take_sum(b.qty, rows);
    g.avg_qty_cnt += rows.size();
    g.avg_price_sum += take_sum(b.price, rows);
    g.avg_price_cnt += rows.size();
    g.avg_disc_sum += take_sum(b.disc, rows);
    g.avg_disc_cnt += rows.size();
    g.count += rows.size();
  }
}

// one partition task: scan -> FilterExec -> CoalesceBatchesExec -> HashAggregateExec(Partial)
void q1_partition(const LineitemCols& c, int64_t r0, int64_t r1, int32_t date_lit,
                  std::unordered_map<std::string, Q1Group>& groups) {
  Q1Batch buf;
  for (int64_t b0 = r0; b0 < r1; b0 += BATCH) {
    int64_t b1 = std::min(r1, b0 + BATCH);
    // FilterExec: predicate -> BooleanArray, then filter_record_batch copies every column
    std::vector<uint8_t> mask((size_t)(b1 - b0));
    for (int64_t i = b0; i < b1; ++i) mask[(size_t)(i - b0)] = c.shipdate[i] <= date_lit;
    for (int64_t i = b0; i < b1; ++i) {
      if (!mask[(size_t)(i - b0)]) continue;
      buf.qty.push_back(c.qty[i]);
      buf.price.push_back(c.price[i]);
      buf.disc.push_back(c.disc[i]);
      buf.tax.push_back(c.tax[i]);
      buf.shipdate.push_back(c.shipdate[i]);
      buf.rf.emplace_back((const char*)c.rf_chr + c.rf_off[i], (size_t)(c.rf_off[i + 1] - c.rf_off[i]));
      buf.ls.emplace_back((const char*)c.ls_chr + c.ls_off[i], (size_t)(c.ls_off[i + 1] - c.ls_off[i]));
    }
    // CoalesceBatchesExec: emit once the buffered rows reach the target batch size
    if ((int64_t)buf.rows() >= BATCH) {
      q1_partial_batch(buf, groups);
      buf.clear();
    }
  }
  if (buf.rows()) q1_partial_batch(buf, groups);
}

}  // namespace

extern "C" {

// out: up to max_groups rows of [rf, ls] chars and 10 doubles/uint64 per the Q1 output schema
// (sum_qty, sum_base_price, sum_disc_price, sum_charge, avg_qty, avg_price, avg_disc as doubles; count)
int oracle_q1(int64_t n, const double* qty, const double* price, const double* disc, const double* tax,
              const int32_t* shipdate, const int32_t* rf_off, const uint8_t* rf_chr, const int32_t* ls_off,
              const uint8_t* ls_chr, int32_t date_lit, int n_partitions, int max_groups, char* out_rf,
              char* out_ls, double* out_f64 /*[g][7]*/, uint64_t* out_count) {
  LineitemCols c{qty, price, disc, tax, shipdate, rf_off, ls_off, rf_chr, ls_chr};
  if (n_partitions < 1) n_partitions = 1;
  std::vector<std::unordered_map<std::string, Q1Group>> partial((size_t)n_partitions);
  std::vector<std::thread> th;
  for (int p = 0; p < n_partitions; ++p) {
    int64_t r0 = n * p / n_partitions, r1 = n * (p + 1) / n_partitions;
    th.emplace_back([&, p, r0, r1] { q1_partition(c, r0, r1, date_lit, partial[(size_t)p]); });
  }
  for (auto& t : th) t.join();
  // MergeExec (partition order) + HashAggregateExec(Final): merge_batch adds states, evaluate() divides
  std::vector<std::string> order;
  std::unordered_map<std::string, Q1Group> fin;
  for (auto& pm : partial)
    for (auto& kv : pm) {
      auto it = fin.find(kv.first);
      if (it == fin.end()) {
        order.push_back(kv.first);
        fin.emplace(kv.first, kv.second);
      } else {
        Q1Group& g = it->second;
        const Q1Group& s = kv.second;
        g.sum_qty += s.sum_qty; g.sum_base += s.sum_base; g.sum_disc_price += s.sum_disc_price;
        g.sum_charge += s.sum_charge; g.avg_qty_sum += s.avg_qty_sum; g.avg_price_sum += s.avg_price_sum;
        g.avg_disc_sum += s.avg_disc_sum; g.avg_qty_cnt += s.avg_qty_cnt; g.avg_price_cnt += s.avg_price_cnt;
        g.avg_disc_cnt += s.avg_disc_cnt; g.count += s.count;
      }
    }
  int ng = 0;
  for (auto& k : order) {
    if (ng >= max_groups) break;
    const Q1Group& g = fin[k];
    out_rf[ng] = g.rf;
    out_ls[ng] = g.ls;
    double* o = out_f64 + (size_t)ng * 7;
    o[0] = g.sum_qty; o[1] = g.sum_base; o[2] = g.sum_disc_price; o[3] = g.sum_charge;
    o[4] = g.avg_qty_sum / (double)g.avg_qty_cnt;
    o[5] = g.avg_price_sum / (double)g.avg_price_cnt;
    o[6] = g.avg_disc_sum / (double)g.avg_disc_cnt;
    out_count[ng] = g.count;
    ++ng;
  }
  return ng;
}

// Q6: filter (5 conjuncts; BETWEEN bounds folded in f64) -> ungrouped SUM(price*discount); returns revenue,
// *passing = rows that passed the filter
double oracle_q6(int64_t n, const int32_t* shipdate, const double* disc, const double* qty, const double* price,
                 int32_t date_lo, int32_t date_hi, double disc_lo, double disc_hi, double qty_lt, int n_partitions,
                 uint64_t* passing) {
  if (n_partitions < 1) n_partitions = 1;
  std::vector<double> psum((size_t)n_partitions, 0.0);
  std::vector<uint64_t> pcnt((size_t)n_partitions, 0);
  auto task = [&](int p, int64_t r0, int64_t r1) {
    std::vector<double> bprice, bdisc, tmp;
    double acc = 0.0;
    uint64_t cnt = 0;
    auto flush = [&] {
      tmp.resize(bprice.size());
      for (size_t i = 0; i < bprice.size(); ++i) tmp[i] = bprice[i] * bdisc[i];  // expression temporary
      double s = 0.0;
      for (double x : tmp) s += x;  // compute::sum
      acc += s;                     // accumulator update
      cnt += bprice.size();
      bprice.clear();
      bdisc.clear();
    };
    for (int64_t b0 = r0; b0 < r1; b0 += BATCH) {
      int64_t b1 = std::min(r1, b0 + BATCH);
      size_t m = (size_t)(b1 - b0);
      // binary expression tree evaluated column at a time into Boolean temporaries
      std::vector<uint8_t> m1(m), m2(m), m3(m), m4(m), m5(m);
      for (size_t i = 0; i < m; ++i) m1[i] = shipdate[b0 + i] >= date_lo;
      for (size_t i = 0; i < m; ++i) m2[i] = shipdate[b0 + i] < date_hi;
      for (size_t i = 0; i < m; ++i) m3[i] = disc[b0 + i] >= disc_lo;
      for (size_t i = 0; i < m; ++i) m4[i] = disc[b0 + i] <= disc_hi;
      for (size_t i = 0; i < m; ++i) m5[i] = qty[b0 + i] < qty_lt;
      for (size_t i = 0; i < m; ++i)
        if (m1[i] & m2[i] & m3[i] & m4[i] & m5[i]) {  // filter_record_batch: copy surviving rows
          bprice.push_back(price[b0 + i]);
          bdisc.push_back(disc[b0 + i]);
        }
      if ((int64_t)bprice.size() >= BATCH) flush();  // CoalesceBatchesExec
    }
    if (!bprice.empty()) flush();
    psum[(size_t)p] = acc;
    pcnt[(size_t)p] = cnt;
  };
  std::vector<std::thread> th;
  for (int p = 0; p < n_partitions; ++p) th.emplace_back(task, p, n * p / n_partitions, n * (p + 1) / n_partitions);
  for (auto& t : th) t.join();
  double total = 0.0;
  uint64_t c = 0;
  for (int p = 0; p < n_partitions; ++p) {  // Final: merge partial sums in partition order
    total += psum[(size_t)p];
    c += pcnt[(size_t)p];
  }
  if (passing) *passing = c;
  return total;
}

}  // extern "C"
