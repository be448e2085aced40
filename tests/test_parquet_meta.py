"""CPU: the host half of the Parquet reader (csrc/parquet_meta.cpp: Thrift compact footer + page headers) against pyarrow's
own metadata of files pyarrow wrote, the ParquetExec file -> partition split, and malformed input.  No device needed:
b2_parquet_describe / b2_parquet_schema run without one.  The column decode itself is tests/test_gpu_parquet.py."""
import datetime
import io
import re

import numpy as np
import pyarrow as pa
import pyarrow.parquet as pq
import pytest

import ballista_b200 as b2
from ballista_b200.parquet import split_files


def table(n=5000, seed=3, nullable=True):
    rng = np.random.default_rng(seed)
    d0 = datetime.date(1992, 1, 1)
    cols = {
        "k": pa.array(rng.integers(0, 1 << 30, n), pa.int32()),
        "q": pa.array(rng.integers(1, 51, n).astype(np.float64)),
        "flag": pa.array([("A", "N", "R")[i] for i in rng.integers(0, 3, n)]),
        "d": pa.array([d0 + datetime.timedelta(days=int(x)) for x in rng.integers(0, 2500, n)]),
        "big": pa.array(rng.integers(-(1 << 60), 1 << 60, n), pa.int64()),
        "ok": pa.array(rng.integers(0, 2, n).astype(bool)),
    }
    t = pa.table(cols)
    if not nullable:
        t = t.cast(pa.schema([pa.field(f.name, f.type, nullable=False) for f in t.schema]))
    return t


def written(t, **kw):
    bio = io.BytesIO()
    pq.write_table(t, bio, **kw)
    return bio.getvalue()


@pytest.mark.parametrize("kw", [
    {}, {"compression": "none"}, {"compression": "snappy", "data_page_version": "2.0"},
    {"use_dictionary": False, "compression": "none", "data_page_size": 4096},
    {"row_group_size": 1200, "data_page_size": 2048, "compression": "zstd"},
])
@pytest.mark.parametrize("nullable", [True, False])
def test_describe_and_schema_match_pyarrow_metadata(kw, nullable):
    t = table(nullable=nullable)
    data = written(t, **kw)
    f = b2.ParquetFile(data)
    md = pq.ParquetFile(io.BytesIO(data)).metadata
    assert (f.num_rows, f.num_row_groups, f.num_columns) == (md.num_rows, md.num_row_groups, md.num_columns)
    assert f.schema() == t.schema.remove_metadata()
    assert f.schema([2, 0]) == pa.schema([t.schema.field(2), t.schema.field(0)]).remove_metadata()
    text = f.describe()
    groups = re.findall(r"row_group (\d+) rows=(\d+)", text)
    assert [int(r) for _, r in groups] == [md.row_group(g).num_rows for g in range(md.num_row_groups)]
    chunks = re.findall(r"chunk (\d+) codec=(\w+) values=(\d+) bytes=(\d+) dict_pages=(\d+) data_pages=(\d+) encodings=(\S*)", text)
    assert len(chunks) == md.num_row_groups * md.num_columns
    i = 0
    for g in range(md.num_row_groups):
        for c in range(md.num_columns):
            col = md.row_group(g).column(c)
            _, codec, values, nbytes, n_dict, n_data, encs = chunks[i]
            i += 1
            assert codec == col.compression
            assert int(values) == col.num_values
            assert int(nbytes) == col.total_compressed_size
            assert int(n_dict) == (1 if col.has_dictionary_page else 0)
            assert int(n_data) >= 1
            for e in encs.split(","):  # every page encoding we saw is one the chunk's metadata lists
                assert e in col.encodings or (e == "RLE_DICTIONARY" and "PLAIN_DICTIONARY" in col.encodings), (e, col.encodings)


def test_many_pages_are_walked():
    t = table(20000)
    f = b2.ParquetFile(written(t, use_dictionary=False, compression="none", data_page_size=1024, row_group_size=20000))
    pages = [int(x) for x in re.findall(r"data_pages=(\d+)", f.describe())]
    assert len(pages) == 6 and min(pages[:5]) >= 20  # pyarrow closes a page every 1024 rows at this page size


def test_unsupported_column_types_are_named():
    t = pa.table({"a": pa.array([1.0, 2.0], pa.float32()), "ts": pa.array([1, 2], pa.timestamp("us")),
                  "ok": pa.array([1, 2], pa.int32()), "dec": pa.array([1, 2], pa.decimal128(10, 2)),
                  "i8": pa.array([1, 2], pa.int8()), "u32": pa.array([1, 2], pa.uint32())})
    f = b2.ParquetFile(written(t))
    assert f.schema([2]) == pa.schema([("ok", pa.int32())])
    for idx, name in ((0, "a"), (1, "ts"), (3, "dec"), (4, "i8"), (5, "u32")):
        with pytest.raises(b2.NotImplementedOnGpu, match=f"'{name}'"):
            f.schema([idx])
    with pytest.raises(b2.NotImplementedOnGpu):
        f.schema()
    assert "unsupported" in f.describe()


def test_nested_schema_is_refused():
    t = pa.table({"l": pa.array([[1, 2], [3]], pa.list_(pa.int32())), "x": pa.array([1, 2], pa.int32())})
    f = b2.ParquetFile(written(t))
    with pytest.raises(b2.NotImplementedOnGpu, match="nested"):
        f.schema()


def test_uint64_and_explicit_int_annotations():
    t = pa.table({"u": pa.array([1, 2 ** 63 + 5], pa.uint64()), "i": pa.array([-1, 5], pa.int64())})
    f = b2.ParquetFile(written(t))
    assert f.schema() == t.schema


def test_malformed_files_fail_cleanly():
    data = written(table(300))
    for bad in (b"", b"PAR1", data[:-1], b"XXXX" + data[4:], data[:-4] + b"PARX", data[:100],
                data[:-8] + (1 << 30).to_bytes(4, "little") + b"PAR1"):
        with pytest.raises(b2.BallistaError):
            b2.ParquetFile(bad)
    with pytest.raises(b2.NotImplementedOnGpu, match="encrypted"):
        b2.ParquetFile(data[:-4] + b"PARE")


def test_footer_fuzz_never_crashes():
    data = bytearray(written(table(300), row_group_size=100))
    flen = int.from_bytes(data[-8:-4], "little")
    lo = len(data) - 8 - flen
    rng = np.random.default_rng(11)
    outcomes = {"ok": 0, "error": 0}
    for _ in range(400):
        d = bytearray(data)
        for _ in range(int(rng.integers(1, 4))):
            d[int(rng.integers(lo, len(data) - 8))] = int(rng.integers(0, 256))
        try:
            f = b2.ParquetFile(bytes(d))
            f.describe()
            outcomes["ok"] += 1
        except b2.BallistaError:
            outcomes["error"] += 1
    assert outcomes["error"] > 50  # most corruptions are caught as errors, none as a crash


def test_split_files_is_datafusions():
    names = [f"f{i}" for i in range(7)]
    assert split_files(names, 1) == [names]
    assert split_files(names, 2) == [names[:4], names[4:]]
    assert split_files(names, 3) == [names[:3], names[3:6], names[6:]]
    assert split_files(names, 7) == [[n] for n in names]
    assert split_files(names, 16) == [[n] for n in names]  # never more partitions than files
    assert split_files(names[:4], 3) == [names[:2], names[2:4]]  # ceil(4/3) = 2 files per chunk -> 2 partitions
