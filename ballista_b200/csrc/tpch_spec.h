// tpch_spec.h — the synthetic TPC-H-shaped workload, as closed-form functions of (seed, row index).
//
// Every value of every table is a pure function of its row index (counter-based RNG), so any row range can
// be generated on any device and is bit-identical everywhere.  Shapes follow the benchmark schema in the
// reference (rust/benchmarks/tpch/src/main.rs:267-360) and the sparse order keys of its fixtures
// (rust/scheduler/testdata/orders/orders.tbl).  This header is the single definition of the workload; it is
// compiled into the CUDA generator (tpch_gen.cu) and into the host generator used by the oracle.
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define B2_HD __host__ __device__ __forceinline__
#define B2_UNROLL _Pragma("unroll")
#else
#define B2_HD inline
#define B2_UNROLL
#endif

namespace b2spec {

constexpr uint64_t SEED = 42;

// streams (one per generated attribute)
enum Stream : uint64_t {
  S_O_CUSTKEY = 1, S_O_ORDERDATE, S_L_COUNTROT, S_L_PARTKEY, S_L_SUPPKEY, S_L_QUANTITY, S_L_DISCOUNT,
  S_L_TAX, S_L_SHIPDELTA, S_L_COMMITDELTA, S_L_RECEIPTDELTA, S_L_RFLAG, S_C_MKTSEGMENT, S_C_NATIONKEY, S_S_NATIONKEY
};

B2_HD uint64_t splitmix64(uint64_t x) {
  uint64_t z = x + 0x9E3779B97F4A7C15ull;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}
// 64 random bits for (stream, idx)
B2_HD uint64_t rnd(uint64_t stream, uint64_t idx) {
  return splitmix64(splitmix64(SEED ^ (stream << 32)) ^ idx);
}
// uniform integer in [lo, hi] (multiply-high on the top 32 bits)
B2_HD int64_t uniform(uint64_t stream, uint64_t idx, int64_t lo, int64_t hi) {
  uint64_t r = rnd(stream, idx) >> 32;
  return lo + (int64_t)((r * (uint64_t)(hi - lo + 1)) >> 32);
}

// ---- cardinalities ----
B2_HD int64_t n_orders(double sf) { return (int64_t)(1500000.0 * sf + 0.5) & ~7ll; }  // multiple of 8
B2_HD int64_t n_lineitem(double sf) { return 4 * n_orders(sf); }
B2_HD int64_t n_customer(double sf) { return (int64_t)(150000.0 * sf + 0.5); }
B2_HD int64_t n_supplier(double sf) {
  int64_t n = (int64_t)(10000.0 * sf + 0.5);  // the range l_suppkey is drawn from (lineitem_row)
  return n < 1 ? 1 : n;
}
constexpr int64_t N_NATION = 25, N_REGION = 5;

// ---- orders ----
B2_HD int32_t o_orderkey(int64_t i) { return (int32_t)(((i >> 3) << 5) + (i & 7) + 1); }
B2_HD int32_t o_custkey(int64_t i, int64_t ncust) {
  // uniform over customer keys that are not multiples of 3 (dbgen leaves a third of customers order-less)
  int64_t k = uniform(S_O_CUSTKEY, (uint64_t)i, 0, (ncust / 3) * 2 - 1);
  return (int32_t)(3 * (k >> 1) + (k & 1) + 1);
}
B2_HD int32_t o_orderdate(int64_t i) { return (int32_t)uniform(S_O_ORDERDATE, (uint64_t)i, 8035, 10440); }
B2_HD int32_t o_shippriority(int64_t) { return 0; }

// ---- lineitem: 32 lines per 8 consecutive orders; per-order line counts are a rotation of this ----
// row r -> (order index, line number).  counts sum to 32.
B2_HD void l_locate(int64_t r, int64_t* order_idx, int32_t* linenumber) {
  const int base[8] = {1, 2, 3, 4, 5, 6, 7, 4};
  int64_t g = r >> 5;
  int j = (int)(r & 31);
  int rot = (int)(rnd(S_L_COUNTROT, (uint64_t)g) & 7);
  int s = 0, cum = 0;
B2_UNROLL
  for (int t = 0; t < 8; ++t) {
    int c = base[(t + rot) & 7];
    if (j >= cum + c && t < 7) { cum += c; s = t + 1; }
    else break;
  }
  *order_idx = g * 8 + s;
  *linenumber = j - cum + 1;
}

struct LineItem {
  int32_t l_orderkey, l_partkey, l_suppkey, l_linenumber;
  double l_quantity, l_extendedprice, l_discount, l_tax;
  int32_t l_shipdate, l_commitdate, l_receiptdate;
  char l_returnflag, l_linestatus;
};

B2_HD LineItem lineitem_row(int64_t r, double sf) {
  LineItem li;
  int64_t oi;
  l_locate(r, &oi, &li.l_linenumber);
  li.l_orderkey = o_orderkey(oi);
  int64_t nparts = (int64_t)(200000.0 * sf + 0.5);
  int64_t nsupp = (int64_t)(10000.0 * sf + 0.5);
  if (nparts < 1) nparts = 1;
  if (nsupp < 1) nsupp = 1;
  li.l_partkey = (int32_t)uniform(S_L_PARTKEY, (uint64_t)r, 1, nparts);
  li.l_suppkey = (int32_t)uniform(S_L_SUPPKEY, (uint64_t)r, 1, nsupp);
  int64_t qty = uniform(S_L_QUANTITY, (uint64_t)r, 1, 50);
  int64_t pk = li.l_partkey;
  int64_t cents = 90000 + ((pk / 10) % 20001) + 100 * (pk % 1000);
  li.l_quantity = (double)qty;
  li.l_extendedprice = (double)(qty * cents) / 100.0;  // correctly rounded 2-decimal value, as CSV parsing yields
  li.l_discount = (double)uniform(S_L_DISCOUNT, (uint64_t)r, 0, 10) / 100.0;
  li.l_tax = (double)uniform(S_L_TAX, (uint64_t)r, 0, 8) / 100.0;
  int32_t od = o_orderdate(oi);
  li.l_shipdate = od + (int32_t)uniform(S_L_SHIPDELTA, (uint64_t)r, 1, 121);
  li.l_commitdate = od + (int32_t)uniform(S_L_COMMITDELTA, (uint64_t)r, 30, 90);
  li.l_receiptdate = li.l_shipdate + (int32_t)uniform(S_L_RECEIPTDELTA, (uint64_t)r, 1, 30);
  li.l_returnflag = (li.l_receiptdate <= 9298) ? ((rnd(S_L_RFLAG, (uint64_t)r) >> 63) ? 'R' : 'A') : 'N';
  li.l_linestatus = (li.l_shipdate > 9298) ? 'O' : 'F';
  return li;
}

// ---- customer ----
B2_HD int32_t c_custkey(int64_t i) { return (int32_t)(i + 1); }
B2_HD int c_mktsegment_id(int64_t i) { return (int)uniform(S_C_MKTSEGMENT, (uint64_t)i, 0, 4); }
B2_HD int32_t c_nationkey(int64_t i) { return (int32_t)uniform(S_C_NATIONKEY, (uint64_t)i, 0, 24); }
// segment names and their lengths; offsets are a prefix sum of these lengths
B2_HD const char* mktsegment_name(int id) {
  switch (id) {
    case 0: return "AUTOMOBILE";
    case 1: return "BUILDING";
    case 2: return "FURNITURE";
    case 3: return "HOUSEHOLD";
    default: return "MACHINERY";
  }
}
B2_HD int mktsegment_len(int id) { return id == 0 ? 10 : (id == 1 ? 8 : 9); }

// ---- supplier / nation / region (TPC-H Q5: rust/benchmarks/tpch/queries/q5.sql; schemas main.rs:267-360) ----
B2_HD int32_t s_suppkey(int64_t i) { return (int32_t)(i + 1); }
B2_HD int32_t s_nationkey(int64_t i) { return (int32_t)uniform(S_S_NATIONKEY, (uint64_t)i, 0, 24); }
// the 25 nations and 5 regions of the TPC-H specification (dbgen's fixed tables)
B2_HD const char* nation_name(int id) {
  const char* names[25] = {"ALGERIA", "ARGENTINA", "BRAZIL", "CANADA", "EGYPT", "ETHIOPIA", "FRANCE", "GERMANY", "INDIA",
                           "INDONESIA", "IRAN", "IRAQ", "JAPAN", "JORDAN", "KENYA", "MOROCCO", "MOZAMBIQUE", "PERU",
                           "CHINA", "ROMANIA", "SAUDI ARABIA", "VIETNAM", "RUSSIA", "UNITED KINGDOM", "UNITED STATES"};
  return names[id];
}
B2_HD int nation_regionkey(int id) {
  const int region[25] = {0, 1, 1, 1, 4, 0, 3, 3, 2, 2, 4, 4, 2, 4, 0, 0, 0, 1, 2, 3, 4, 2, 3, 3, 1};
  return region[id];
}
B2_HD const char* region_name(int id) {
  const char* names[5] = {"AFRICA", "AMERICA", "ASIA", "EUROPE", "MIDDLE EAST"};
  return names[id];
}

}  // namespace b2spec
