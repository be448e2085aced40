// tpch_gen.cu — generate the synthetic TPC-H-shaped tables of tpch_spec.h directly in HBM.
// Workload setup, not the measured path.  Schemas: rust/benchmarks/tpch/src/main.rs:267-360.
#include <cstring>
#include <set>

#include "core.h"
#include "tpch_spec.h"

namespace b2 {

struct LineitemPtrs {
  int32_t *orderkey, *partkey, *suppkey, *linenumber, *shipdate, *commitdate, *receiptdate;
  double *quantity, *extendedprice, *discount, *tax;
  int32_t *rf_off, *ls_off;
  uint8_t *rf_chr, *ls_chr;
};

__global__ void gen_lineitem_kernel(LineitemPtrs p, int64_t row0, int64_t n, double sf) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  for (; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    b2spec::LineItem li = b2spec::lineitem_row(row0 + i, sf);
    if (p.orderkey) p.orderkey[i] = li.l_orderkey;
    if (p.partkey) p.partkey[i] = li.l_partkey;
    if (p.suppkey) p.suppkey[i] = li.l_suppkey;
    if (p.linenumber) p.linenumber[i] = li.l_linenumber;
    if (p.quantity) p.quantity[i] = li.l_quantity;
    if (p.extendedprice) p.extendedprice[i] = li.l_extendedprice;
    if (p.discount) p.discount[i] = li.l_discount;
    if (p.tax) p.tax[i] = li.l_tax;
    if (p.shipdate) p.shipdate[i] = li.l_shipdate;
    if (p.commitdate) p.commitdate[i] = li.l_commitdate;
    if (p.receiptdate) p.receiptdate[i] = li.l_receiptdate;
    if (p.rf_off) { p.rf_off[i] = (int32_t)i; p.rf_chr[i] = (uint8_t)li.l_returnflag; if (i == n - 1) p.rf_off[n] = (int32_t)n; }
    if (p.ls_off) { p.ls_off[i] = (int32_t)i; p.ls_chr[i] = (uint8_t)li.l_linestatus; if (i == n - 1) p.ls_off[n] = (int32_t)n; }
  }
}

__global__ void gen_orders_kernel(int32_t* okey, int32_t* ckey, int32_t* odate, int32_t* prio, int64_t row0,
                                  int64_t n, int64_t ncust) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  for (; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    int64_t r = row0 + i;
    if (okey) okey[i] = b2spec::o_orderkey(r);
    if (ckey) ckey[i] = b2spec::o_custkey(r, ncust);
    if (odate) odate[i] = b2spec::o_orderdate(r);
    if (prio) prio[i] = b2spec::o_shippriority(r);
  }
}

// customer: c_mktsegment is variable-length, so offsets need a prefix sum.  One thread block scans
// chunks of rows sequentially (the table is tiny next to lineitem); chunk carry is kept in a register.
__global__ void gen_customer_kernel(int32_t* ckey, int32_t* nkey, int32_t* seg_off, uint8_t* seg_chr,
                                    int64_t row0, int64_t n) {
  __shared__ int32_t warp_sums[32];
  __shared__ int32_t carry_s;
  if (threadIdx.x == 0) carry_s = 0;
  __syncthreads();
  int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
  for (int64_t base = 0; base < n; base += blockDim.x) {
    int64_t i = base + threadIdx.x;
    int id = 0, len = 0;
    if (i < n) {
      int64_t r = row0 + i;
      if (ckey) ckey[i] = b2spec::c_custkey(r);
      if (nkey) nkey[i] = b2spec::c_nationkey(r);
      id = b2spec::c_mktsegment_id(r);
      len = b2spec::mktsegment_len(id);
    }
    if (seg_off) {
      int v = len;
      for (int d = 1; d < 32; d <<= 1) {
        int t = __shfl_up_sync(0xffffffffu, v, d);
        if (lane >= d) v += t;
      }
      if (lane == 31) warp_sums[warp] = v;
      __syncthreads();
      int woff = 0;
      for (int w = 0; w < warp; ++w) woff += warp_sums[w];
      int total = 0;
      for (int w = 0; w < nwarps; ++w) total += warp_sums[w];
      int carry = carry_s;
      int excl = carry + woff + v - len;
      if (i < n) {
        seg_off[i] = excl;
        const char* nm = b2spec::mktsegment_name(id);
        for (int q = 0; q < len; ++q) seg_chr[excl + q] = (uint8_t)nm[q];
        if (i == n - 1) seg_off[n] = excl + len;
      }
      __syncthreads();
      if (threadIdx.x == 0) carry_s = carry + total;
      __syncthreads();
    }
  }
}

static bool want(const std::set<std::string>& sel, const char* name) { return sel.empty() || sel.count(name); }

static Column fixed_col(Ctx* ctx, int dtype, int64_t n) {
  Column c;
  c.dtype = dtype;
  c.length = n;
  c.data = dev_alloc(ctx, (size_t)n * dtype_width(dtype) + 16);
  return c;
}
static Column utf8_col(Ctx* ctx, int64_t n, int64_t chars) {
  Column c;
  c.dtype = B2_UTF8;
  c.length = n;
  c.data = dev_alloc_zero(ctx, (size_t)(n + 1) * 4);
  c.chars = dev_alloc(ctx, (size_t)chars + 16);
  c.chars_len = chars;
  return c;
}

__global__ void gen_supplier_kernel(int32_t* suppkey, int32_t* nationkey, int64_t r0, int64_t n) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  for (; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    if (suppkey) suppkey[i] = b2spec::s_suppkey(r0 + i);
    if (nationkey) nationkey[i] = b2spec::s_nationkey(r0 + i);
  }
}

BatchPtr tpch_generate(Ctx* ctx, const std::string& table, double sf, int64_t row_begin, int64_t row_end,
                       const char* const* columns) {
  std::set<std::string> sel;
  if (columns)
    for (const char* const* c = columns; *c; ++c) sel.insert(*c);
  int64_t n = row_end - row_begin;
  if (n < 0) fail(B2_ERR_INVALID_ARG, "row_end < row_begin");
  auto b = std::make_shared<Batch>();
  b->ctx = ctx;
  b->num_rows = n;
  auto add = [&](const char* name, Column c) {
    b->schema.fields.push_back(Field{name, c.dtype, false});
    b->cols.push_back(std::move(c));
    return &b->cols.back();
  };
  int grid = grid_for(ctx, 8, n, 256);
  if (table == "lineitem") {
    if (row_end > b2spec::n_lineitem(sf)) fail(B2_ERR_INVALID_ARG, "lineitem row range beyond table");
    if (n >= (1ll << 31) - 1) fail(B2_ERR_INVALID_ARG, "a batch holds at most 2^31-2 rows");
    LineitemPtrs p;
    memset(&p, 0, sizeof p);
    b->cols.reserve(16);
    // schema order as in main.rs:303-323 (subset)
    if (want(sel, "l_orderkey")) p.orderkey = (int32_t*)add("l_orderkey", fixed_col(ctx, B2_INT32, n))->data->ptr;
    if (want(sel, "l_partkey")) p.partkey = (int32_t*)add("l_partkey", fixed_col(ctx, B2_INT32, n))->data->ptr;
    if (want(sel, "l_suppkey")) p.suppkey = (int32_t*)add("l_suppkey", fixed_col(ctx, B2_INT32, n))->data->ptr;
    if (want(sel, "l_linenumber")) p.linenumber = (int32_t*)add("l_linenumber", fixed_col(ctx, B2_INT32, n))->data->ptr;
    if (want(sel, "l_quantity")) p.quantity = (double*)add("l_quantity", fixed_col(ctx, B2_FLOAT64, n))->data->ptr;
    if (want(sel, "l_extendedprice")) p.extendedprice = (double*)add("l_extendedprice", fixed_col(ctx, B2_FLOAT64, n))->data->ptr;
    if (want(sel, "l_discount")) p.discount = (double*)add("l_discount", fixed_col(ctx, B2_FLOAT64, n))->data->ptr;
    if (want(sel, "l_tax")) p.tax = (double*)add("l_tax", fixed_col(ctx, B2_FLOAT64, n))->data->ptr;
    if (want(sel, "l_returnflag")) {
      Column* c = add("l_returnflag", utf8_col(ctx, n, n));
      p.rf_off = (int32_t*)c->data->ptr;
      p.rf_chr = (uint8_t*)c->chars->ptr;
    }
    if (want(sel, "l_linestatus")) {
      Column* c = add("l_linestatus", utf8_col(ctx, n, n));
      p.ls_off = (int32_t*)c->data->ptr;
      p.ls_chr = (uint8_t*)c->chars->ptr;
    }
    if (want(sel, "l_shipdate")) p.shipdate = (int32_t*)add("l_shipdate", fixed_col(ctx, B2_DATE32, n))->data->ptr;
    if (want(sel, "l_commitdate")) p.commitdate = (int32_t*)add("l_commitdate", fixed_col(ctx, B2_DATE32, n))->data->ptr;
    if (want(sel, "l_receiptdate")) p.receiptdate = (int32_t*)add("l_receiptdate", fixed_col(ctx, B2_DATE32, n))->data->ptr;
    if (n) {
      gen_lineitem_kernel<<<grid, 256, 0, ctx->stream>>>(p, row_begin, n, sf);
      B2_LAUNCHED(ctx);
    }
  } else if (table == "orders") {
    if (row_end > b2spec::n_orders(sf)) fail(B2_ERR_INVALID_ARG, "orders row range beyond table");
    b->cols.reserve(8);
    int32_t *ok = nullptr, *ck = nullptr, *od = nullptr, *sp = nullptr;
    if (want(sel, "o_orderkey")) ok = (int32_t*)add("o_orderkey", fixed_col(ctx, B2_INT32, n))->data->ptr;
    if (want(sel, "o_custkey")) ck = (int32_t*)add("o_custkey", fixed_col(ctx, B2_INT32, n))->data->ptr;
    if (want(sel, "o_orderdate")) od = (int32_t*)add("o_orderdate", fixed_col(ctx, B2_DATE32, n))->data->ptr;
    if (want(sel, "o_shippriority")) sp = (int32_t*)add("o_shippriority", fixed_col(ctx, B2_INT32, n))->data->ptr;
    if (n) {
      gen_orders_kernel<<<grid, 256, 0, ctx->stream>>>(ok, ck, od, sp, row_begin, n, b2spec::n_customer(sf));
      B2_LAUNCHED(ctx);
    }
  } else if (table == "customer") {
    if (row_end > b2spec::n_customer(sf)) fail(B2_ERR_INVALID_ARG, "customer row range beyond table");
    b->cols.reserve(8);
    int32_t *ck = nullptr, *nk = nullptr, *so = nullptr;
    uint8_t* sc = nullptr;
    if (want(sel, "c_custkey")) ck = (int32_t*)add("c_custkey", fixed_col(ctx, B2_INT32, n))->data->ptr;
    if (want(sel, "c_nationkey")) nk = (int32_t*)add("c_nationkey", fixed_col(ctx, B2_INT32, n))->data->ptr;
    Column* seg = nullptr;
    if (want(sel, "c_mktsegment")) {
      seg = add("c_mktsegment", utf8_col(ctx, n, n * 10));
      so = (int32_t*)seg->data->ptr;
      sc = (uint8_t*)seg->chars->ptr;
    }
    if (n) {
      gen_customer_kernel<<<1, 1024, 0, ctx->stream>>>(ck, nk, so, sc, row_begin, n);
      B2_LAUNCHED(ctx);
      if (seg) {
        int32_t total = 0;
        B2_READBACK(ctx, &total, so + n, 4);
        seg->chars_len = total;
      }
    }
  } else if (table == "supplier") {
    if (row_end > b2spec::n_supplier(sf)) fail(B2_ERR_INVALID_ARG, "supplier row range beyond table");
    b->cols.reserve(4);
    int32_t *sk = nullptr, *nk = nullptr;
    if (want(sel, "s_suppkey")) sk = (int32_t*)add("s_suppkey", fixed_col(ctx, B2_INT32, n))->data->ptr;
    if (want(sel, "s_nationkey")) nk = (int32_t*)add("s_nationkey", fixed_col(ctx, B2_INT32, n))->data->ptr;
    if (n) {
      gen_supplier_kernel<<<grid, 256, 0, ctx->stream>>>(sk, nk, row_begin, n);
      B2_LAUNCHED(ctx);
    }
  } else if (table == "nation" || table == "region") {
    // dbgen's fixed tables (25 / 5 rows): assembled on the host, one upload per column
    const bool nat = table == "nation";
    if (row_end > (nat ? b2spec::N_NATION : b2spec::N_REGION)) fail(B2_ERR_INVALID_ARG, table + " row range beyond table");
    std::vector<int32_t> key((size_t)n), region((size_t)n), off((size_t)n + 1, 0);
    std::string chars;
    for (int64_t i = 0; i < n; ++i) {
      int id = (int)(row_begin + i);
      key[(size_t)i] = id;
      region[(size_t)i] = nat ? b2spec::nation_regionkey(id) : 0;
      chars += nat ? b2spec::nation_name(id) : b2spec::region_name(id);
      off[(size_t)i + 1] = (int32_t)chars.size();
    }
    auto put = [&](const char* name, int dtype, const void* src, size_t bytes) {
      Column c = fixed_col(ctx, dtype, n);
      if (bytes) B2_CUDA(cudaMemcpyAsync(c.data->ptr, src, bytes, cudaMemcpyHostToDevice, ctx->stream));
      add(name, std::move(c));
    };
    b->cols.reserve(4);
    if (want(sel, nat ? "n_nationkey" : "r_regionkey")) put(nat ? "n_nationkey" : "r_regionkey", B2_INT32, key.data(), (size_t)n * 4);
    if (want(sel, nat ? "n_name" : "r_name")) {
      Column c = utf8_col(ctx, n, (int64_t)chars.size());
      B2_CUDA(cudaMemcpyAsync(c.data->ptr, off.data(), (size_t)(n + 1) * 4, cudaMemcpyHostToDevice, ctx->stream));
      if (!chars.empty())
        B2_CUDA(cudaMemcpyAsync(c.chars->ptr, chars.data(), chars.size(), cudaMemcpyHostToDevice, ctx->stream));
      add(nat ? "n_name" : "r_name", std::move(c));
    }
    if (nat && want(sel, "n_regionkey")) put("n_regionkey", B2_INT32, region.data(), (size_t)n * 4);
    B2_CUDA(cudaStreamSynchronize(ctx->stream));  // the host vectors go out of scope
  } else {
    fail(B2_ERR_INVALID_ARG, "unknown TPC-H table '" + table + "' (lineitem | orders | customer | supplier | nation | region)");
  }
  return b;
}

int64_t tpch_num_rows(const std::string& table, double sf) {
  if (table == "lineitem") return b2spec::n_lineitem(sf);
  if (table == "orders") return b2spec::n_orders(sf);
  if (table == "customer") return b2spec::n_customer(sf);
  if (table == "supplier") return b2spec::n_supplier(sf);
  if (table == "nation") return b2spec::N_NATION;
  if (table == "region") return b2spec::N_REGION;
  return -1;
}

}  // namespace b2
