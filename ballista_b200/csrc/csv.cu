// csv.cu — GPU-side decode of delimiter-separated text (dbgen `.tbl`: '|' separated, no header, no quoting, a
// trailing delimiter per line) into Arrow columns in HBM: the column decode of the reference's CsvExec leaf
// (rust/core/src/serde/physical_plan/from_proto.rs:93-110: CsvExec::try_new(path, CsvReadOptions{has_header,
// delimiter, schema, file_extension}, projection, batch_size); table schemas rust/benchmarks/tpch/src/main.rs:267-360).
// with this the raw file bytes cross PCIe once and parsing stops being host work.
//
// Three steps: (1) newline positions (per-tile count -> scan -> write), (2) one thread per row walks its bytes and
// parses the projected fixed-width fields in place (Int32 / Int64 / Float64 / Date32) while recording (start, length)
// of projected Utf8 fields, (3) Utf8: lengths -> offsets (scan) -> character copy.
// Float64 text is converted exactly like a correctly-rounded parser for the inputs this path sees: up to 15
// significant digits and a decimal exponent within +-22 give mantissa / 10^k with both operands exact in binary64, so
// one IEEE division is the correctly rounded result (Clinger's fast path); anything longer is refused, not guessed.
#include <cstring>

#include "ops.h"
#include "scan.cuh"

namespace b2 {

constexpr int CSV_TILE = 4096;  // bytes per newline-count tile
constexpr int CSV_MAX_FIELDS = 64;

__global__ void csv_count_kernel(const char* __restrict__ text, int64_t len, int64_t* __restrict__ tile_cnt) {
  __shared__ int cnt;
  if (threadIdx.x == 0) cnt = 0;
  __syncthreads();
  int64_t base = (int64_t)blockIdx.x * CSV_TILE;
  int c = 0;
  for (int i = threadIdx.x; i < CSV_TILE; i += blockDim.x)
    if (base + i < len && text[base + i] == '\n') ++c;
  for (int d = 16; d > 0; d >>= 1) c += __shfl_xor_sync(0xffffffffu, c, d);
  if ((threadIdx.x & 31) == 0 && c) atomicAdd(&cnt, c);
  __syncthreads();
  if (threadIdx.x == 0) tile_cnt[blockIdx.x] = cnt;
}
// row_end[r] = byte position of the r-th newline (one warp per tile keeps the positions ordered)
__global__ void csv_positions_kernel(const char* __restrict__ text, int64_t len, const int64_t* __restrict__ tile_off,
                                     int64_t n_tiles, int64_t* __restrict__ row_end) {
  int lane = threadIdx.x & 31;
  int64_t tile = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
  if (tile >= n_tiles) return;
  int64_t base = tile * CSV_TILE, out = tile_off[tile];
  for (int i = lane; i < CSV_TILE; i += 32) {
    bool nl = base + i < len && text[base + i] == '\n';
    unsigned m = __ballot_sync(0xffffffffu, nl);
    if (nl) row_end[out + __popc(m & ((1u << lane) - 1))] = base + i;
    out += __popc(m);
  }
}

struct CsvArgs {
  int n_fields;                    // columns of the file schema
  int target[CSV_MAX_FIELDS];      // file column -> projected column index, -1 = skip
  int dtype[CSV_MAX_FIELDS];       // of the file column
  void* out[CSV_MAX_FIELDS];       // projected column: values (Utf8: uint32 start offsets into the text)
  int32_t* str_len[CSV_MAX_FIELDS];
  char delimiter;
};
enum { CSV_OK = 0, CSV_BAD_NUMBER = 1, CSV_BAD_DATE = 2, CSV_TOO_FEW_FIELDS = 3, CSV_FLOAT_TOO_LONG = 4, CSV_EMPTY_FIELD = 5 };

__device__ __forceinline__ int64_t days_from_civil(int y, int m, int d) {  // proleptic Gregorian -> days since 1970-01-01
  y -= m <= 2;
  int era = (y >= 0 ? y : y - 399) / 400;
  unsigned yoe = (unsigned)(y - era * 400);
  unsigned doy = (153u * (unsigned)(m + (m > 2 ? -3 : 9)) + 2u) / 5u + (unsigned)d - 1u;
  unsigned doe = yoe * 365u + yoe / 4u - yoe / 100u + doy;
  return (int64_t)era * 146097 + (int64_t)doe - 719468;
}
__device__ const double CSV_POW10[23] = {1e0,  1e1,  1e2,  1e3,  1e4,  1e5,  1e6,  1e7,  1e8,  1e9,  1e10, 1e11,
                                         1e12, 1e13, 1e14, 1e15, 1e16, 1e17, 1e18, 1e19, 1e20, 1e21, 1e22};

// parses text[b, e) as the given type; returns an error code
__device__ int csv_parse_field(const char* __restrict__ text, int64_t b, int64_t e, int dtype, void* out, int64_t row) {
  if (b >= e) return CSV_EMPTY_FIELD;
  if (dtype == B2_DATE32) {
    if (e - b != 10 || text[b + 4] != '-' || text[b + 7] != '-') return CSV_BAD_DATE;
    int v[8], k = 0;
    for (int64_t i = b; i < e; ++i) {
      if (i == b + 4 || i == b + 7) continue;
      int c = text[i] - '0';
      if (c < 0 || c > 9) return CSV_BAD_DATE;
      v[k++] = c;
    }
    int y = v[0] * 1000 + v[1] * 100 + v[2] * 10 + v[3], m = v[4] * 10 + v[5], d = v[6] * 10 + v[7];
    if (m < 1 || m > 12 || d < 1 || d > 31) return CSV_BAD_DATE;
    static_cast<int32_t*>(out)[row] = (int32_t)days_from_civil(y, m, d);
    return CSV_OK;
  }
  bool neg = false;
  int64_t i = b;
  if (text[i] == '-' || text[i] == '+') {
    neg = text[i] == '-';
    ++i;
  }
  if (i >= e) return CSV_BAD_NUMBER;
  if (dtype == B2_FLOAT64) {
    unsigned long long mant = 0;
    int digits = 0, scale = 0;
    bool seen_dot = false, any = false;
    for (; i < e; ++i) {
      char c = text[i];
      if (c == '.') {
        if (seen_dot) return CSV_BAD_NUMBER;
        seen_dot = true;
        continue;
      }
      int dgt = c - '0';
      if (dgt < 0 || dgt > 9) return CSV_BAD_NUMBER;  // exponents / inf / nan do not occur in .tbl files
      any = true;
      if (mant != 0 || dgt != 0) ++digits;
      if (digits > 15) return CSV_FLOAT_TOO_LONG;
      mant = mant * 10 + (unsigned)dgt;
      if (seen_dot) ++scale;
    }
    if (!any) return CSV_BAD_NUMBER;
    if (scale > 22) return CSV_FLOAT_TOO_LONG;
    double v = __ddiv_rn((double)mant, CSV_POW10[scale]);
    static_cast<double*>(out)[row] = neg ? -v : v;
    return CSV_OK;
  }
  unsigned long long acc = 0;
  for (; i < e; ++i) {
    int dgt = text[i] - '0';
    if (dgt < 0 || dgt > 9) return CSV_BAD_NUMBER;
    if (acc > ((1ull << 62) - 9) / 10) return CSV_BAD_NUMBER;  // checked before the multiply: 20-digit fields must not wrap
    acc = acc * 10 + (unsigned)dgt;
  }
  long long v = neg ? -(long long)acc : (long long)acc;
  if (dtype == B2_INT32) {
    if (v > 2147483647ll || v < -2147483648ll) return CSV_BAD_NUMBER;
    static_cast<int32_t*>(out)[row] = (int32_t)v;
  } else {
    static_cast<int64_t*>(out)[row] = v;
  }
  return CSV_OK;
}

// one thread per row
__global__ void csv_parse_rows_kernel(const char* __restrict__ text, const int64_t* __restrict__ row_end, int64_t first_row,
                                      int64_t n_rows, CsvArgs a, int* status) {
  int64_t r = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  for (; r < n_rows; r += (int64_t)gridDim.x * blockDim.x) {
    int64_t fr = first_row + r;
    int64_t b = fr == 0 ? 0 : row_end[fr - 1] + 1, e = row_end[fr];
    if (e > b && text[e - 1] == '\r') --e;
    int field = 0;
    int64_t fb = b;
    int err = CSV_OK;
    for (int64_t i = b; i <= e && field < a.n_fields; ++i) {
      if (i == e || text[i] == a.delimiter) {
        int t = a.target[field];
        if (t >= 0) {
          if (a.dtype[field] == B2_UTF8) {
            static_cast<uint32_t*>(a.out[field])[r] = (uint32_t)fb;
            a.str_len[field][r] = (int32_t)(i - fb);
          } else {
            int rc = csv_parse_field(text, fb, i, a.dtype[field], a.out[field], r);
            if (rc && !err) err = rc;
          }
        }
        ++field;
        fb = i + 1;
      }
    }
    if (field < a.n_fields && !err) err = CSV_TOO_FEW_FIELDS;
    if (err) {
      atomicCAS(&status[0], 0, err);
      atomicMin(reinterpret_cast<unsigned long long*>(status + 2), (unsigned long long)fr);
    }
  }
}
struct LenIn {
  const int32_t* v;
  __device__ __forceinline__ int operator()(int64_t i) const { return v[i]; }
};
struct OffsetsOut {
  int32_t* off;
  int64_t n;
  __device__ __forceinline__ void operator()(int64_t i, int v, int64_t pre) const {
    off[i] = (int32_t)pre;
    if (i == n - 1) off[n] = (int32_t)(pre + v);
  }
};
__global__ void csv_copy_chars_kernel(const char* __restrict__ text, const uint32_t* __restrict__ start,
                                      const int32_t* __restrict__ off, int64_t n, uint8_t* __restrict__ chr) {
  int64_t r = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  for (; r < n; r += (int64_t)gridDim.x * blockDim.x) {
    int32_t o = off[r], len = off[r + 1] - o;
    const char* src = text + start[r];
    for (int q = 0; q < len; ++q) chr[o + q] = (uint8_t)src[q];
  }
}

BatchPtr csv_decode(Ctx* ctx, const Schema& file_schema, const std::vector<int>& projection, char delimiter,
                    bool has_header, const char* host_text, size_t len) {
  if (file_schema.fields.size() > CSV_MAX_FIELDS) fail(B2_ERR_NOT_IMPLEMENTED, "more than 64 columns in a CSV schema");
  if (len >= (1ull << 32) - 16) fail(B2_ERR_NOT_IMPLEMENTED, "CSV chunks of 4 GiB or more (split the file at a row boundary)");
  Schema out_schema;
  for (int p : projection) {
    if (p < 0 || p >= (int)file_schema.fields.size()) fail(B2_ERR_INVALID_ARG, "CSV projection index out of range");
    out_schema.fields.push_back(file_schema.fields[p]);
  }
  auto out = std::make_shared<Batch>();
  out->ctx = ctx;
  out->schema = out_schema;
  // the text, newline-terminated, in HBM
  bool add_nl = len > 0 && host_text[len - 1] != '\n';
  int64_t n = (int64_t)len + (add_nl ? 1 : 0);
  auto text = dev_alloc(ctx, (size_t)n + 16);
  if (len) B2_CUDA(cudaMemcpyAsync(text->ptr, host_text, len, cudaMemcpyHostToDevice, ctx->stream));
  if (add_nl) B2_CUDA(cudaMemsetAsync(text->as<char>() + len, '\n', 1, ctx->stream));
  int64_t n_lines = 0;
  BufPtr row_end;
  if (n) {
    int64_t n_tiles = (n + CSV_TILE - 1) / CSV_TILE;
    auto tiles = dev_alloc(ctx, (size_t)(n_tiles + 2) * 8);
    csv_count_kernel<<<(unsigned)n_tiles, 256, 0, ctx->stream>>>(text->as<char>(), n, tiles->as<int64_t>());
    B2_LAUNCHED(ctx);
    scan_i64(ctx, tiles->as<int64_t>(), n_tiles);
    B2_READBACK(ctx, &n_lines, tiles->as<int64_t>() + n_tiles, 8);
    row_end = dev_alloc(ctx, (size_t)(n_lines + 1) * 8);
    if (n_lines) {
      csv_positions_kernel<<<(unsigned)((n_tiles * 32 + 255) / 256), 256, 0, ctx->stream>>>(
          text->as<char>(), n, tiles->as<int64_t>(), n_tiles, row_end->as<int64_t>());
      B2_LAUNCHED(ctx);
    }
  }
  int64_t first = has_header && n_lines > 0 ? 1 : 0;
  int64_t rows = n_lines - first;
  if (rows >= (1ll << 31) - 1) fail(B2_ERR_NOT_IMPLEMENTED, "CSV chunk beyond 2^31-2 rows");
  out->num_rows = rows;
  CsvArgs a;
  memset(&a, 0, sizeof a);
  a.n_fields = (int)file_schema.fields.size();
  a.delimiter = delimiter;
  for (int f = 0; f < a.n_fields; ++f) {
    a.target[f] = -1;
    a.dtype[f] = file_schema.fields[f].dtype;
  }
  std::vector<BufPtr> starts(projection.size()), lens(projection.size());
  out->cols.resize(projection.size());
  for (size_t p = 0; p < projection.size(); ++p) {
    int f = projection[p];
    if (a.target[f] >= 0) fail(B2_ERR_NOT_IMPLEMENTED, "a CSV column projected twice");
    a.target[f] = (int)p;
    Column& c = out->cols[p];
    c.dtype = file_schema.fields[f].dtype;
    c.length = rows;
    switch (c.dtype) {
      case B2_INT32: case B2_DATE32: case B2_INT64: case B2_FLOAT64:
        c.data = dev_alloc(ctx, (size_t)rows * dtype_width(c.dtype) + 16);
        a.out[f] = c.data->ptr;
        break;
      case B2_UTF8:
        starts[p] = dev_alloc(ctx, (size_t)(rows + 1) * 4);
        lens[p] = dev_alloc(ctx, (size_t)(rows + 1) * 4);
        a.out[f] = starts[p]->ptr;
        a.str_len[f] = lens[p]->as<int32_t>();
        break;
      default: fail(B2_ERR_NOT_IMPLEMENTED, std::string("CSV column of type ") + dtype_name(c.dtype));
    }
  }
  auto status = dev_alloc_zero(ctx, 16);
  B2_CUDA(cudaMemsetAsync(status->as<char>() + 8, 0xFF, 8, ctx->stream));
  if (rows) {
    csv_parse_rows_kernel<<<grid_for(ctx, 8, rows, 256), 256, 0, ctx->stream>>>(
        text->as<char>(), row_end->as<int64_t>(), first, rows, a, status->as<int>());
    B2_LAUNCHED(ctx);
  }
  struct {
    int code, pad;
    unsigned long long row;
  } st = {0, 0, 0};
  B2_READBACK(ctx, &st, status->ptr, 16);
  if (st.code) {
    static const char* what[] = {"", "a field is not a valid number", "a field is not a valid YYYY-MM-DD date",
                                 "fewer fields than the schema has columns", "a Float64 field has more than 15 significant digits",
                                 "an empty field in a non-nullable column"};
    fail(B2_ERR_ARROW, "Error while parsing CSV line " + std::to_string(st.row + 1) + ": " + what[st.code]);
  }
  for (size_t p = 0; p < projection.size(); ++p) {
    Column& c = out->cols[p];
    if (c.dtype != B2_UTF8) continue;
    c.data = dev_alloc_zero(ctx, (size_t)(rows + 1) * 4);
    int64_t total = 0;
    device_scan(ctx, LenIn{lens[p]->as<int32_t>()}, rows, OffsetsOut{c.data->as<int32_t>(), rows}, &total);
    if (total >= (1ll << 31)) fail(B2_ERR_EXECUTION, "Utf8 column beyond 2 GiB in one CSV chunk");
    c.chars_len = total;
    c.chars = dev_alloc(ctx, (size_t)total + 16);
    if (rows) {
      csv_copy_chars_kernel<<<grid_for(ctx, 8, rows, 256), 256, 0, ctx->stream>>>(
          text->as<char>(), starts[p]->as<uint32_t>(), c.data->as<int32_t>(), rows, c.chars->as<uint8_t>());
      B2_LAUNCHED(ctx);
    }
  }
  B2_SYNC(ctx);  // `host_text` may be released by the caller as soon as this returns
  return out;
}

}  // namespace b2
