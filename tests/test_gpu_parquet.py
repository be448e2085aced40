"""GPU: Parquet column decode (csrc/parquet.cu; the ParquetExec leaf, from_proto.rs:111-121) against pyarrow's reader on
the same bytes — bit-exact for every type — over the writer options that change the byte layout: compression, dictionary
coding (and its fallback to PLAIN), page version, page and row-group size, REQUIRED / OPTIONAL columns with and without
NULLs; then the ParquetExec leaf under TPC-H Q1."""
import datetime
import io

import numpy as np
import pyarrow as pa
import pyarrow.parquet as pq
import pytest

import ballista_b200 as b2
from ballista_b200 import tpch

pytestmark = pytest.mark.gpu


def sample(n, seed, nulls, required=False):
    """every supported type; low- and high-cardinality columns; strings from empty to long and repetitive"""
    rng = np.random.default_rng(seed)
    d0 = datetime.date(1992, 1, 1)
    words = ["", "a", "REG AIR", "DELIVER IN PERSON", "x" * 300, "ab" * 700, "furiously final requests"]
    cols = {
        "key": pa.array(rng.integers(-(1 << 31), 1 << 31, n), pa.int32()),
        "small": pa.array(rng.integers(0, 7, n), pa.int32()),
        "price": pa.array(np.round(rng.random(n) * 1e5, 2)),
        "qty": pa.array(rng.integers(1, 51, n).astype(np.float64)),
        "flag": pa.array([("A", "N", "R")[i] for i in rng.integers(0, 3, n)], pa.string()),
        "text": pa.array([words[i] + str(int(j)) * int(k) for i, j, k in
                          zip(rng.integers(0, len(words), n), rng.integers(0, 1000, n), rng.integers(0, 3, n))], pa.string()),
        "day": pa.array([d0 + datetime.timedelta(days=int(x)) for x in rng.integers(0, 2500, n)], pa.date32()),
        "big": pa.array(rng.integers(-(1 << 62), 1 << 62, n), pa.int64()),
        "ubig": pa.array(rng.integers(0, 1 << 63, n).astype(np.uint64) * np.uint64(2) + np.uint64(1)),
        "ok": pa.array(rng.integers(0, 2, n).astype(bool)),
        "const": pa.array(np.full(n, 42), pa.int32()),
    }
    if nulls:
        for i, (name, arr) in enumerate(list(cols.items())):
            frac = (0.0, 0.02, 0.5, 0.98, 1.0)[i % 5]
            mask = rng.random(n) < frac
            cols[name] = pa.array(arr.to_pylist(), arr.type, mask=mask) if n else arr
    t = pa.table(cols)
    if required:
        t = t.cast(pa.schema([pa.field(f.name, f.type, nullable=False) for f in t.schema]))
    return t


def written(t, **kw):
    bio = io.BytesIO()
    pq.write_table(t, bio, **kw)
    return bio.getvalue()


def check(ctx, data, projection=None, row_groups=None):
    f = b2.ParquetFile(data)
    got = f.read(ctx, projection, row_groups).to_arrow()
    ref = pq.ParquetFile(io.BytesIO(data))
    groups = list(range(ref.num_row_groups)) if row_groups is None else list(range(*row_groups))
    names = None if projection is None else [ref.schema_arrow.names[p] if isinstance(p, int) else p for p in projection]
    want = ref.read_row_groups(groups, columns=names).combine_chunks()
    if names is not None:
        want = want.select(names)
    assert got.schema.names == want.schema.names
    assert got.num_rows == want.num_rows
    for i, name in enumerate(want.schema.names):
        w = want.column(i).combine_chunks() if want.num_rows else pa.array([], want.schema.field(i).type)
        g = got.column(i)
        assert g.type == w.type, name
        assert g.null_count == w.null_count, name
        assert g.equals(w), f"column {name} differs"
    return got


WRITER_OPTIONS = [
    dict(compression="none", use_dictionary=False),
    dict(compression="none", use_dictionary=True),
    dict(compression="snappy", use_dictionary=False),
    dict(compression="snappy", use_dictionary=True),
    dict(compression="snappy", use_dictionary=True, data_page_version="2.0"),
    dict(compression="none", use_dictionary=False, data_page_version="2.0"),
    dict(compression="snappy", use_dictionary=True, data_page_size=2048, write_batch_size=64, row_group_size=7000),
    dict(compression="none", use_dictionary=True, data_page_size=512, write_batch_size=32, row_group_size=20000,
         dictionary_pagesize_limit=4096),  # dictionaries overflow: dictionary pages first, PLAIN pages after
    dict(compression="snappy", use_dictionary=["flag", "small", "text"], data_page_version="2.0", data_page_size=4096,
         write_batch_size=100),
]


@pytest.mark.parametrize("opts", range(len(WRITER_OPTIONS)))
@pytest.mark.parametrize("shape", ["required", "optional_no_nulls", "nulls"])
def test_decode_matches_pyarrow(ctx, opts, shape):
    t = sample(23456, seed=opts + 1, nulls=shape == "nulls", required=shape == "required")
    check(ctx, written(t, **WRITER_OPTIONS[opts]))


def test_projection_and_row_group_ranges(ctx):
    t = sample(10000, seed=9, nulls=True)
    data = written(t, row_group_size=1500, compression="snappy")
    check(ctx, data, projection=[5, 0], row_groups=(2, 5))
    check(ctx, data, projection=["day", "flag", "ok"], row_groups=(6, 7))
    check(ctx, data, projection=[3], row_groups=(0, 1))
    f = b2.ParquetFile(data)
    assert f.read(ctx, [1], (3, 3)).num_rows == 0
    with pytest.raises(b2.BallistaError):
        f.read(ctx, [1], (5, 99))
    with pytest.raises(b2.BallistaError):
        f.read(ctx, [99])


def test_empty_and_tiny_files(ctx):
    for n in (0, 1, 7, 8, 9, 255, 256, 257):
        for kw in (dict(compression="none"), dict(compression="snappy", data_page_version="2.0")):
            check(ctx, written(sample(n, seed=n, nulls=n > 1), **kw))
    t = sample(5, 1, False)
    assert check(ctx, written(t.slice(0, 0))).num_rows == 0


def test_large_plain_pages_are_sliced(ctx):
    """1 MB PLAIN pages of REQUIRED fixed-width columns are decoded by 64 KB slices, compressed or not"""
    n = 700_000
    rng = np.random.default_rng(5)
    t = pa.table({"a": pa.array(rng.random(n)), "b": pa.array(rng.integers(0, 1 << 31, n), pa.int32()),
                  "c": pa.array(np.arange(n) // 3, pa.int64())})
    t = t.cast(pa.schema([pa.field(f.name, f.type, nullable=False) for f in t.schema]))
    for comp in ("none", "snappy"):
        data = written(t, use_dictionary=False, compression=comp, data_page_size=1 << 20, write_batch_size=1 << 17)
        assert "encodings=PLAIN" in b2.ParquetFile(data).describe()
        check(ctx, data)


def test_snappy_overlapping_copies_and_long_literals(ctx):
    rng = np.random.default_rng(2)
    texts = ["ab" * int(k) + "xyz" * int(j) for k, j in zip(rng.integers(0, 3000, 400), rng.integers(0, 50, 400))]
    noise = [bytes(rng.integers(32, 127, int(k), dtype=np.uint8)).decode() for k in rng.integers(0, 5000, 400)]
    t = pa.table({"rep": texts, "noise": noise, "zeros": pa.array(np.zeros(400), pa.float64())})
    check(ctx, written(t, compression="snappy", use_dictionary=False))
    check(ctx, written(t, compression="snappy", use_dictionary=True, data_page_version="2.0"))


def test_unsupported_features_are_refused_by_name(ctx):
    t = sample(1000, 3, False)
    for codec in ("zstd", "gzip", "lz4", "brotli"):
        with pytest.raises(b2.NotImplementedOnGpu, match=codec.upper()[:3]):
            b2.ParquetFile(written(t, compression=codec)).read(ctx)
    for enc in ("DELTA_BINARY_PACKED", "BYTE_STREAM_SPLIT"):
        col = "key" if enc == "DELTA_BINARY_PACKED" else "price"
        data = written(t, use_dictionary=False, column_encoding={col: enc})
        f = b2.ParquetFile(data)
        with pytest.raises(b2.NotImplementedOnGpu, match=enc):
            f.read(ctx, [col])
        check(ctx, data, projection=["flag", "day"])  # the other columns of the same file still decode
    with pytest.raises(b2.NotImplementedOnGpu, match="DELTA"):
        b2.ParquetFile(written(t, use_dictionary=False, column_encoding={"text": "DELTA_LENGTH_BYTE_ARRAY"})).read(ctx, ["text"])


def test_corrupt_pages_fail_or_decode_but_never_crash(ctx):
    t = sample(4000, 4, True)
    data = bytearray(written(t, compression="snappy", data_page_size=2048, write_batch_size=64))
    flen = int.from_bytes(data[-8:-4], "little")
    rng = np.random.default_rng(8)
    errors = 0
    for _ in range(25):
        d = bytearray(data)
        for _ in range(20):
            d[int(rng.integers(4, len(data) - 8 - flen))] = int(rng.integers(0, 256))
        try:
            b2.ParquetFile(bytes(d)).read(ctx)
        except b2.BallistaError:
            errors += 1
    assert errors > 0
    check(ctx, bytes(data))  # the context is still usable


def q1_over(ctx, scan):
    """Partial per input partition, one Final over all the partial states (the reference's two stages)"""
    partial = tpch.q1_partial(scan)
    states = [b for p in range(scan.output_partitioning()[1]) for b in partial.collect_device(p)]
    final = tpch.q1_final(b2.MemoryExec(ctx, partial.schema(), [states]), scan.schema())
    return final.collect().to_pylist()


def lineitem_table(ctx, sf, required=True):
    li = ctx.tpch("lineitem", sf).to_arrow()
    t = pa.Table.from_batches([li])
    if not required:
        t = t.cast(pa.schema([pa.field(f.name, f.type, nullable=True) for f in t.schema]))
    return t


def test_tpch_lineitem_file_roundtrip_and_q1(ctx, tmp_path):
    """The benchmark's `--format parquet` path: lineitem files -> ParquetExec(num_partitions) -> Q1 == Q1 over the
    generated table; files written REQUIRED (the benchmark's schema, main.rs:303-343) and OPTIONAL (pyarrow's default)"""
    t = lineitem_table(ctx, 0.02)
    n = t.num_rows
    want = None
    for required in (True, False):
        tt = t if required else lineitem_table(ctx, 0.02, required=False)
        names = []
        cuts = [0, n // 3, n // 2, n]
        for i in range(3):
            path = str(tmp_path / f"part-{i}-{int(required)}.parquet")
            pq.write_table(tt.slice(cuts[i], cuts[i + 1] - cuts[i]), path, row_group_size=30000,
                           compression="snappy" if i else "none", data_page_version="2.0" if i == 2 else "1.0")
            names.append(path)
        proj = [tt.schema.names.index(c) for c in tpch.Q1_COLUMNS]
        scan = b2.ParquetExec(ctx, names, projection=proj, num_partitions=2)
        assert scan.output_partitioning()[1] == 2
        assert scan.schema().names == list(tpch.Q1_COLUMNS)
        rows = 0
        for p in range(2):
            for b in scan.execute(p):
                rows += b.num_rows
        assert rows == n
        got = q1_over(ctx, scan)
        if want is None:
            want = q1_over(ctx, tpch.scan(ctx, [[ctx.upload(t.select(list(tpch.Q1_COLUMNS)))]]))
        key = lambda r: (r["l_returnflag"], r["l_linestatus"])
        got, ref = sorted(got, key=key), sorted(want, key=key)
        assert len(got) == len(ref) == 4
        for g, w in zip(got, ref):
            assert key(g) == key(w) and g["count_order"] == w["count_order"]
            for k in w:
                if isinstance(w[k], float):
                    assert abs(g[k] - w[k]) <= 1e-9 * abs(w[k]), k
    # whole file, every column, bit-exact
    check(ctx, written(t, compression="snappy", row_group_size=50000))
