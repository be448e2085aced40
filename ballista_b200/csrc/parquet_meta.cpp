// parquet_meta.cpp — the host half of the Parquet reader: footer (FileMetaData) and page headers, both Thrift
// compact-protocol structs, read with a small hand-written decoder (no thrift / parquet library in this image).
// Field numbers follow parquet.thrift of the Apache Parquet format specification.  The reference reaches this
// format through DataFusion's ParquetExec (rust/core/src/serde/physical_plan/from_proto.rs:111-121).
#include <cstring>

#include "parquet.h"

namespace b2 {
namespace pq {

namespace {

enum TType { T_STOP = 0, T_TRUE = 1, T_FALSE = 2, T_BYTE = 3, T_I16 = 4, T_I32 = 5, T_I64 = 6, T_DOUBLE = 7, T_BINARY = 8,
             T_LIST = 9, T_SET = 10, T_MAP = 11, T_STRUCT = 12 };

struct Thrift {
  const uint8_t* p;
  const uint8_t* end;
  const char* what;
  int depth = 0;
  [[noreturn]] void bad(const char* why) const { fail(B2_ERR_ARROW, std::string("Parquet: ") + what + ": " + why); }
  uint8_t byte() {
    if (p >= end) bad("truncated");
    return *p++;
  }
  uint64_t varint() {
    uint64_t v = 0;
    for (int shift = 0; shift < 64; shift += 7) {
      uint8_t b = byte();
      v |= (uint64_t)(b & 0x7F) << shift;
      if (!(b & 0x80)) return v;
    }
    bad("varint too long");
  }
  int64_t zigzag() {
    uint64_t v = varint();
    return (int64_t)(v >> 1) ^ -(int64_t)(v & 1);
  }
  std::string binary() {
    uint64_t n = varint();
    if (n > (uint64_t)(end - p)) bad("string beyond the end");
    std::string s(reinterpret_cast<const char*>(p), (size_t)n);
    p += n;
    return s;
  }
  void skip_bytes(uint64_t n) {
    if (n > (uint64_t)(end - p)) bad("field beyond the end");
    p += n;
  }
  // field header: returns false at STOP; `last` carries the previous field id of this struct
  bool field(int& last, int& id, int& type) {
    uint8_t h = byte();
    if (h == 0) return false;
    type = h & 0x0F;
    int delta = h >> 4;
    id = delta ? last + delta : (int)zigzag();
    last = id;
    return true;
  }
  void list_header(int& elem_type, uint64_t& n) {
    uint8_t h = byte();
    elem_type = h & 0x0F;
    n = h >> 4;
    if (n == 15) n = varint();
    if (n > (uint64_t)(end - p) + 1) bad("list longer than the message");
  }
  void skip(int type) {
    if (++depth > 64) bad("nesting too deep");
    switch (type) {
      case T_TRUE: case T_FALSE: break;
      case T_BYTE: byte(); break;
      case T_I16: case T_I32: case T_I64: varint(); break;
      case T_DOUBLE: skip_bytes(8); break;
      case T_BINARY: skip_bytes(varint()); break;
      case T_LIST: case T_SET: {
        int et;
        uint64_t n;
        list_header(et, n);
        for (uint64_t i = 0; i < n; ++i) {
          if (et == T_TRUE || et == T_FALSE) byte();  // booleans inside a list take one byte each
          else skip(et);
        }
        break;
      }
      case T_MAP: {
        uint64_t n = varint();
        if (n) {
          uint8_t kv = byte();
          for (uint64_t i = 0; i < n; ++i) {
            skip(kv >> 4);
            skip(kv & 0x0F);
          }
        }
        break;
      }
      case T_STRUCT: {
        int last = 0, id, t;
        while (field(last, id, t)) skip(t);
        break;
      }
      default: bad("unknown Thrift type");
    }
    --depth;
  }
  int64_t integer(int type) {
    if (type == T_BYTE) return (int8_t)byte();
    if (type != T_I16 && type != T_I32 && type != T_I64) bad("expected an integer field");
    return zigzag();
  }
};

struct SchemaElement {
  std::string name;
  int type = -1, type_length = 0, repetition = REP_REQUIRED, num_children = 0, converted = -1, logical = -1;
  int int_bits = 0;
  bool int_signed = true;
};

void read_logical_type(Thrift& t, SchemaElement& e) {
  int last = 0, id, ty;
  while (t.field(last, id, ty)) {
    e.logical = id;
    if (id == 10 && ty == T_STRUCT) {  // IntType { 1: bitWidth (byte), 2: isSigned }
      int l2 = 0, id2, ty2;
      while (t.field(l2, id2, ty2)) {
        if (id2 == 1) e.int_bits = (int)t.integer(ty2);
        else if (id2 == 2) e.int_signed = ty2 == T_TRUE;
        else t.skip(ty2);
      }
    } else {
      t.skip(ty);
    }
  }
}

SchemaElement read_schema_element(Thrift& t) {
  SchemaElement e;
  int last = 0, id, ty;
  while (t.field(last, id, ty)) {
    switch (id) {
      case 1: e.type = (int)t.integer(ty); break;
      case 2: e.type_length = (int)t.integer(ty); break;
      case 3: e.repetition = (int)t.integer(ty); break;
      case 4: e.name = t.binary(); break;
      case 5: e.num_children = (int)t.integer(ty); break;
      case 6: e.converted = (int)t.integer(ty); break;
      case 10:
        if (ty == T_STRUCT) read_logical_type(t, e);
        else t.skip(ty);
        break;
      default: t.skip(ty);
    }
  }
  return e;
}

ChunkMeta read_column_meta(Thrift& t) {
  ChunkMeta c;
  int last = 0, id, ty;
  while (t.field(last, id, ty)) {
    switch (id) {
      case 1: c.phys = (int)t.integer(ty); break;
      case 2: {
        int et;
        uint64_t n;
        t.list_header(et, n);
        for (uint64_t i = 0; i < n; ++i) c.encodings.push_back((int)t.integer(et));
        break;
      }
      case 4: c.codec = (int)t.integer(ty); break;
      case 5: c.num_values = t.integer(ty); break;
      case 6: c.total_uncompressed = t.integer(ty); break;
      case 7: c.total_compressed = t.integer(ty); break;
      case 9: c.data_page_offset = t.integer(ty); break;
      case 11: c.dict_page_offset = t.integer(ty); break;
      default: t.skip(ty);
    }
  }
  return c;
}

ChunkMeta read_column_chunk(Thrift& t) {
  ChunkMeta c;
  bool have = false, external = false;
  int last = 0, id, ty;
  while (t.field(last, id, ty)) {
    if (id == 1 && ty == T_BINARY) {
      external = !t.binary().empty();
    } else if (id == 3 && ty == T_STRUCT) {
      c = read_column_meta(t);
      have = true;
    } else {
      t.skip(ty);
    }
  }
  if (external) fail(B2_ERR_NOT_IMPLEMENTED, "Parquet: column chunk stored in another file");
  if (!have) t.bad("column chunk without meta data");
  return c;
}

RowGroupMeta read_row_group(Thrift& t) {
  RowGroupMeta g;
  int last = 0, id, ty;
  while (t.field(last, id, ty)) {
    if (id == 1 && ty == T_LIST) {
      int et;
      uint64_t n;
      t.list_header(et, n);
      if (et != T_STRUCT) t.bad("row group columns are not structs");
      for (uint64_t i = 0; i < n; ++i) g.cols.push_back(read_column_chunk(t));
    } else if (id == 3) {
      g.num_rows = t.integer(ty);
    } else {
      t.skip(ty);
    }
  }
  return g;
}

// the Arrow type a leaf decodes to on this path (the TPC-H schema's types, rust/benchmarks/tpch/src/main.rs:267-360)
void assign_dtype(LeafColumn& c) {
  enum { CT_UTF8 = 0, CT_ENUM = 4, CT_DECIMAL = 5, CT_DATE = 6, CT_UINT_64 = 14, CT_INT_32 = 17, CT_INT_64 = 18, CT_JSON = 19 };
  enum { LT_STRING = 1, LT_ENUM = 4, LT_DECIMAL = 5, LT_DATE = 6, LT_INTEGER = 10, LT_UNKNOWN = 11, LT_JSON = 12 };
  auto no = [&](const std::string& w) {
    c.dtype = 0;
    c.why = w;
  };
  switch (c.phys) {
    case PT_BOOLEAN: c.dtype = B2_BOOL; break;
    case PT_INT32:
      if (c.logical == LT_DATE || (c.logical < 0 && c.converted == CT_DATE)) c.dtype = B2_DATE32;
      else if (c.logical == LT_INTEGER) c.dtype = c.int_bits == 32 && c.int_signed ? B2_INT32 : 0;
      else if (c.logical < 0 && (c.converted < 0 || c.converted == CT_INT_32)) c.dtype = B2_INT32;
      if (!c.dtype) no("INT32 annotated as something other than Int32 / Date32");
      break;
    case PT_INT64:
      if (c.logical == LT_INTEGER) c.dtype = c.int_bits == 64 ? (c.int_signed ? B2_INT64 : B2_UINT64) : 0;
      else if (c.logical < 0 && (c.converted < 0 || c.converted == CT_INT_64)) c.dtype = B2_INT64;
      else if (c.logical < 0 && c.converted == CT_UINT_64) c.dtype = B2_UINT64;
      if (!c.dtype) no("INT64 annotated as something other than Int64 / UInt64 (timestamps, decimals)");
      break;
    case PT_DOUBLE: c.dtype = B2_FLOAT64; break;
    case PT_BYTE_ARRAY:
      if (c.logical == LT_DECIMAL || (c.logical < 0 && c.converted == CT_DECIMAL)) no("BYTE_ARRAY decimals");
      else c.dtype = B2_UTF8;  // STRING / ENUM / JSON / unannotated binary: the bytes, as arrow's reader returns them
      break;
    case PT_FLOAT: no("FLOAT (Float32) columns"); break;
    case PT_INT96: no("INT96 timestamps"); break;
    case PT_FIXED: no("FIXED_LEN_BYTE_ARRAY columns"); break;
    default: no("unknown physical type");
  }
}

uint32_t le32(const uint8_t* p) { return (uint32_t)p[0] | (uint32_t)p[1] << 8 | (uint32_t)p[2] << 16 | (uint32_t)p[3] << 24; }

}  // namespace

const char* codec_name(int c) {
  static const char* n[] = {"UNCOMPRESSED", "SNAPPY", "GZIP", "LZO", "BROTLI", "LZ4", "ZSTD", "LZ4_RAW"};
  return c >= 0 && c < 8 ? n[c] : "?";
}
const char* encoding_name(int e) {
  static const char* n[] = {"PLAIN", "?", "PLAIN_DICTIONARY", "RLE", "BIT_PACKED", "DELTA_BINARY_PACKED",
                            "DELTA_LENGTH_BYTE_ARRAY", "DELTA_BYTE_ARRAY", "RLE_DICTIONARY", "BYTE_STREAM_SPLIT"};
  return e >= 0 && e < 10 ? n[e] : "?";
}

FileMeta parse_footer(const uint8_t* file, size_t len) {
  if (len < 12 || !file) fail(B2_ERR_ARROW, "Parquet: file shorter than its magic and footer length");
  if (memcmp(file + len - 4, "PARE", 4) == 0) fail(B2_ERR_NOT_IMPLEMENTED, "Parquet: encrypted footer");
  if (memcmp(file, "PAR1", 4) != 0 || memcmp(file + len - 4, "PAR1", 4) != 0)
    fail(B2_ERR_ARROW, "Parquet: missing PAR1 magic");
  uint64_t flen = le32(file + len - 8);
  if (flen + 12 > len) fail(B2_ERR_ARROW, "Parquet: footer length beyond the file");
  Thrift t{file + len - 8 - flen, file + len - 8, "footer"};
  FileMeta m;
  std::vector<SchemaElement> schema;
  int last = 0, id, ty;
  while (t.field(last, id, ty)) {
    if (id == 1) {
      m.version = (int)t.integer(ty);
    } else if (id == 2 && ty == T_LIST) {
      int et;
      uint64_t n;
      t.list_header(et, n);
      if (et != T_STRUCT) t.bad("schema elements are not structs");
      for (uint64_t i = 0; i < n; ++i) schema.push_back(read_schema_element(t));
    } else if (id == 3) {
      m.num_rows = t.integer(ty);
    } else if (id == 4 && ty == T_LIST) {
      int et;
      uint64_t n;
      t.list_header(et, n);
      if (et != T_STRUCT) t.bad("row groups are not structs");
      for (uint64_t i = 0; i < n; ++i) m.row_groups.push_back(read_row_group(t));
    } else if (id == 6 && ty == T_BINARY) {
      m.created_by = t.binary();
    } else {
      t.skip(ty);
    }
  }
  if (schema.empty()) t.bad("no schema");
  // leaves in depth-first order = the order of the column chunks in every row group
  if ((size_t)schema[0].num_children != schema.size() - 1) m.flat = false;
  for (size_t i = 1; i < schema.size(); ++i) {
    const SchemaElement& e = schema[i];
    if (e.num_children > 0) {
      m.flat = false;
      continue;
    }
    LeafColumn c;
    c.name = e.name;
    c.phys = e.type;
    c.type_length = e.type_length;
    c.repetition = e.repetition;
    c.converted = e.converted;
    c.logical = e.logical;
    c.int_bits = e.int_bits;
    c.int_signed = e.int_signed;
    if (e.repetition == REP_REPEATED) m.flat = false;
    assign_dtype(c);
    m.columns.push_back(c);
  }
  int64_t rows = 0;
  for (auto& g : m.row_groups) {
    if (g.cols.size() != m.columns.size()) t.bad("a row group does not hold one chunk per leaf column");
    if (g.num_rows < 0) t.bad("negative row count");
    rows += g.num_rows;
    for (size_t c = 0; c < g.cols.size(); ++c) {
      const ChunkMeta& k = g.cols[c];
      if (k.phys != m.columns[c].phys) t.bad("chunk type differs from the schema");
      if (k.num_values < 0) t.bad("negative value count");
      if (k.num_values == 0) continue;  // a chunk without values has no pages (its offsets may be anything)
      if (k.data_page_offset < 4 || k.total_compressed < 0 || k.start() < 4 ||
          (uint64_t)k.start() + (uint64_t)k.total_compressed > len - 8)
        t.bad("column chunk outside the file");
    }
  }
  if (rows != m.num_rows) t.bad("row groups do not add up to num_rows");
  return m;
}

std::vector<PageInfo> walk_pages(const uint8_t* file, size_t len, const ChunkMeta& chunk) {
  std::vector<PageInfo> pages;
  int64_t pos = chunk.start(), stop = pos + chunk.total_compressed, seen = 0;
  while (seen < chunk.num_values && pos < stop) {
    Thrift t{file + pos, file + (stop < (int64_t)len ? stop : (int64_t)len), "page header"};
    PageInfo pg;
    pg.type = -1;
    bool have_sizes[2] = {false, false};
    int last = 0, id, ty;
    while (t.field(last, id, ty)) {
      if (id == 1) {
        pg.type = (int)t.integer(ty);
      } else if (id == 2) {
        pg.uncomp_size = (int32_t)t.integer(ty);
        have_sizes[0] = true;
      } else if (id == 3) {
        pg.comp_size = (int32_t)t.integer(ty);
        have_sizes[1] = true;
      } else if ((id == 5 || id == 7) && ty == T_STRUCT) {  // DataPageHeader / DictionaryPageHeader
        int l2 = 0, id2, ty2;
        while (t.field(l2, id2, ty2)) {
          if (id2 == 1) pg.num_values = (int32_t)t.integer(ty2);
          else if (id2 == 2) pg.encoding = (int)t.integer(ty2);
          else if (id2 == 3 && id == 5) pg.def_encoding = (int)t.integer(ty2);
          else t.skip(ty2);
        }
      } else if (id == 8 && ty == T_STRUCT) {  // DataPageHeaderV2
        int l2 = 0, id2, ty2;
        while (t.field(l2, id2, ty2)) {
          switch (id2) {
            case 1: pg.num_values = (int32_t)t.integer(ty2); break;
            case 2: pg.num_nulls = (int32_t)t.integer(ty2); break;
            case 4: pg.encoding = (int)t.integer(ty2); break;
            case 5: pg.def_len = (int32_t)t.integer(ty2); break;
            case 6: pg.rep_len = (int32_t)t.integer(ty2); break;
            case 7: pg.v2_compressed = ty2 == T_TRUE; break;
            default: t.skip(ty2);
          }
        }
      } else {
        t.skip(ty);
      }
    }
    if (pg.type < 0 || !have_sizes[0] || !have_sizes[1]) t.bad("type or sizes missing");
    pg.payload = t.p - file;
    if (pg.comp_size < 0 || pg.uncomp_size < 0 || pg.num_values < 0 || pg.payload + pg.comp_size > stop)
      t.bad("page beyond its column chunk");
    if (pg.type == PAGE_DATA_V2 && (pg.def_len < 0 || pg.rep_len < 0 || (int64_t)pg.def_len + pg.rep_len > pg.comp_size ||
                                    (int64_t)pg.def_len + pg.rep_len > pg.uncomp_size))
      t.bad("level sections longer than the page");
    pos = pg.payload + pg.comp_size;
    if (pg.type == PAGE_DATA || pg.type == PAGE_DATA_V2) seen += pg.num_values;
    else if (pg.type != PAGE_DICT) continue;  // index pages carry nothing this reader needs
    pages.push_back(pg);
  }
  if (seen != chunk.num_values) fail(B2_ERR_ARROW, "Parquet: the pages of a column chunk do not add up to its num_values");
  return pages;
}

Schema arrow_schema(const FileMeta& m, const std::vector<int>& projection) {
  if (!m.flat) fail(B2_ERR_NOT_IMPLEMENTED, "Parquet: nested or repeated columns");
  Schema s;
  for (int p : projection) {
    if (p < 0 || p >= (int)m.columns.size()) fail(B2_ERR_INVALID_ARG, "Parquet projection index out of range");
    const LeafColumn& c = m.columns[p];
    if (!c.dtype) fail(B2_ERR_NOT_IMPLEMENTED, "Parquet column '" + c.name + "': " + c.why);
    Field f;
    f.name = c.name;
    f.dtype = c.dtype;
    f.nullable = c.repetition == REP_OPTIONAL;
    s.fields.push_back(f);
  }
  return s;
}

std::string describe(const FileMeta& m, const uint8_t* file, size_t len) {
  std::string out = "parquet version=" + std::to_string(m.version) + " rows=" + std::to_string(m.num_rows) +
                    " row_groups=" + std::to_string(m.row_groups.size()) + " columns=" + std::to_string(m.columns.size()) +
                    (m.flat ? "" : " nested") + "\n";
  for (size_t c = 0; c < m.columns.size(); ++c) {
    const LeafColumn& k = m.columns[c];
    out += "column " + std::to_string(c) + " " + k.name + " phys=" + std::to_string(k.phys) +
           (k.repetition == REP_OPTIONAL ? " optional" : k.repetition == REP_REPEATED ? " repeated" : " required") + " -> " +
           (k.dtype ? dtype_name(k.dtype) : ("unsupported: " + k.why).c_str()) + "\n";
  }
  for (size_t g = 0; g < m.row_groups.size(); ++g) {
    out += "row_group " + std::to_string(g) + " rows=" + std::to_string(m.row_groups[g].num_rows) + "\n";
    for (size_t c = 0; c < m.row_groups[g].cols.size(); ++c) {
      const ChunkMeta& k = m.row_groups[g].cols[c];
      auto pages = walk_pages(file, len, k);
      int n_dict = 0, n_data = 0;
      std::string encs;
      int seen_enc = 0;
      for (auto& p : pages) {
        if (p.type == PAGE_DICT) {
          ++n_dict;
          continue;
        }
        ++n_data;
        if (!(seen_enc >> p.encoding & 1)) {
          seen_enc |= 1 << p.encoding;
          encs += std::string(encs.empty() ? "" : ",") + encoding_name(p.encoding);
        }
      }
      out += "  chunk " + std::to_string(c) + " codec=" + codec_name(k.codec) + " values=" + std::to_string(k.num_values) +
             " bytes=" + std::to_string(k.total_compressed) + " dict_pages=" + std::to_string(n_dict) +
             " data_pages=" + std::to_string(n_data) + " encodings=" + encs + "\n";
    }
  }
  return out;
}

}  // namespace pq
}  // namespace b2
