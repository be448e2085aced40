// parquet.h — Parquet file metadata as the GPU column decoder needs it (parquet_meta.cpp: host only, no CUDA).
// The reference reads Parquet through DataFusion's ParquetExec (rust/core/src/serde/physical_plan/from_proto.rs:111-121:
// ParquetExec::try_from_files(filenames, projection, None, batch_size, num_partitions)); the format itself is the
// Apache Parquet specification (parquet.thrift): footer and page headers are Thrift compact-protocol structs.
#pragma once
#include <stdint.h>
#include <string>
#include <vector>

#include "core.h"

namespace b2 {
namespace pq {

// parquet.thrift enums
enum PhysType { PT_BOOLEAN = 0, PT_INT32 = 1, PT_INT64 = 2, PT_INT96 = 3, PT_FLOAT = 4, PT_DOUBLE = 5, PT_BYTE_ARRAY = 6, PT_FIXED = 7 };
enum Encoding { ENC_PLAIN = 0, ENC_PLAIN_DICT = 2, ENC_RLE = 3, ENC_BIT_PACKED = 4, ENC_DELTA_BINARY = 5, ENC_DELTA_LENGTH = 6,
                ENC_DELTA_BYTE_ARRAY = 7, ENC_RLE_DICT = 8, ENC_BYTE_STREAM_SPLIT = 9 };
enum Codec { CODEC_NONE = 0, CODEC_SNAPPY = 1, CODEC_GZIP = 2, CODEC_LZO = 3, CODEC_BROTLI = 4, CODEC_LZ4 = 5, CODEC_ZSTD = 6, CODEC_LZ4_RAW = 7 };
enum PageType { PAGE_DATA = 0, PAGE_INDEX = 1, PAGE_DICT = 2, PAGE_DATA_V2 = 3 };
enum Repetition { REP_REQUIRED = 0, REP_OPTIONAL = 1, REP_REPEATED = 2 };

struct LeafColumn {
  std::string name;
  int phys = -1;
  int type_length = 0;
  int repetition = REP_REQUIRED;
  int converted = -1;      // ConvertedType, -1 = absent
  int logical = -1;        // field id of the LogicalType union member that is set, -1 = absent
  int int_bits = 0;        // LogicalType INTEGER
  bool int_signed = true;
  int dtype = 0;           // b2_dtype this column decodes to, 0 = not supported (see `why`)
  std::string why;
};
struct ChunkMeta {
  int phys = -1;
  int codec = 0;
  int64_t num_values = 0;
  int64_t data_page_offset = -1, dict_page_offset = -1;
  int64_t total_compressed = 0, total_uncompressed = 0;
  std::vector<int> encodings;
  int64_t start() const {  // first byte of the chunk: the dictionary page when there is one
    return dict_page_offset > 0 && dict_page_offset < data_page_offset ? dict_page_offset : data_page_offset;
  }
};
struct RowGroupMeta {
  int64_t num_rows = 0;
  std::vector<ChunkMeta> cols;
};
struct FileMeta {
  int version = 0;
  int64_t num_rows = 0;
  std::string created_by;
  bool flat = true;  // every leaf is a direct child of the root and none is repeated
  std::vector<LeafColumn> columns;
  std::vector<RowGroupMeta> row_groups;
};
struct PageInfo {
  int type = 0;             // PageType
  int64_t payload = 0;      // file offset of the bytes after the page header
  int32_t comp_size = 0, uncomp_size = 0;
  int32_t num_values = 0;
  int encoding = 0, def_encoding = ENC_RLE;
  int32_t def_len = 0, rep_len = 0;  // V2: byte lengths of the (never compressed) level sections
  int32_t num_nulls = -1;            // V2
  bool v2_compressed = true;
};

// fails with B2_ERR_ARROW on anything that is not a Parquet file this reader understands
FileMeta parse_footer(const uint8_t* file, size_t len);
// the page headers of one column chunk, in file order (dictionary page first when present)
std::vector<PageInfo> walk_pages(const uint8_t* file, size_t len, const ChunkMeta& chunk);
std::string describe(const FileMeta& m, const uint8_t* file, size_t len);
Schema arrow_schema(const FileMeta& m, const std::vector<int>& projection);
const char* codec_name(int c);
const char* encoding_name(int e);

}  // namespace pq

// parquet.cu: row groups [rg_begin, rg_end) of the projected columns, decoded on the GPU into one batch
BatchPtr parquet_decode(Ctx* ctx, const uint8_t* file, size_t len, const std::vector<int>& projection, int rg_begin,
                        int rg_end);

}  // namespace b2
