"""GPU: decode of dbgen-style .tbl text (the CsvExec leaf's column decode, from_proto.rs:93-110) against pyarrow's CSV
reader on the same bytes and against the values the text was printed from; bit-exact, Float64 included."""
import datetime
import io
import json
import os

import numpy as np
import pyarrow as pa
import pyarrow.csv
import pytest

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))
EPOCH = datetime.date(1970, 1, 1)

LINEITEM = pa.schema([("l_orderkey", pa.int32()), ("l_partkey", pa.int32()), ("l_suppkey", pa.int32()),
                      ("l_linenumber", pa.int32()), ("l_quantity", pa.float64()), ("l_extendedprice", pa.float64()),
                      ("l_discount", pa.float64()), ("l_tax", pa.float64()), ("l_returnflag", pa.utf8()),
                      ("l_linestatus", pa.utf8()), ("l_shipdate", pa.date32()), ("l_commitdate", pa.date32()),
                      ("l_receiptdate", pa.date32()), ("l_shipinstruct", pa.utf8()), ("l_shipmode", pa.utf8()),
                      ("l_comment", pa.utf8())])


def tbl_text(n, seed=1, crlf=False, trailing_newline=True):
    rng = np.random.default_rng(seed)
    d0 = datetime.date(1992, 1, 1)
    rows = []
    for i in range(n):
        qty = int(rng.integers(1, 51))
        price = int(rng.integers(90_000, 10_500_000))  # cents
        ship = d0 + datetime.timedelta(days=int(rng.integers(0, 2500)))
        comment = "".join(chr(int(c)) for c in rng.integers(97, 123, int(rng.integers(0, 40)))).replace("|", " ")
        rows.append("|".join([str(i // 4 * 32 + i % 4 + 1), str(int(rng.integers(1, 200_000))), str(int(rng.integers(1, 10_000))),
                              str(i % 7 + 1), str(qty), f"{price // 100}.{price % 100:02d}", f"0.{int(rng.integers(0, 11)):02d}",
                              f"0.0{int(rng.integers(0, 9))}", "RAN"[int(rng.integers(0, 3))], "OF"[int(rng.integers(0, 2))],
                              ship.isoformat(), (ship + datetime.timedelta(days=30)).isoformat(),
                              (ship + datetime.timedelta(days=7)).isoformat(), "DELIVER IN PERSON", "TRUCK", comment]) + "|")
    nl = "\r\n" if crlf else "\n"
    text = nl.join(rows) + (nl if trailing_newline else "")
    return text.encode()


def arrow_read(text, schema, names):
    # pyarrow sees the trailing '|' as one extra, empty column
    cols = schema.names + ["_trailing"]
    types = {f.name: f.type for f in schema}
    types["_trailing"] = pa.utf8()
    t = pa.csv.read_csv(io.BytesIO(text), read_options=pa.csv.ReadOptions(column_names=cols),
                        parse_options=pa.csv.ParseOptions(delimiter="|", quote_char=False),
                        convert_options=pa.csv.ConvertOptions(column_types=types, strings_can_be_null=False))
    return t.select(names)


@pytest.mark.parametrize("n,crlf,trailing", [(1, False, True), (2000, False, True), (30_011, True, False)])
def test_tbl_decode_matches_pyarrow(ctx, n, crlf, trailing):
    text = tbl_text(n, seed=n, crlf=crlf, trailing_newline=trailing)
    names = ["l_quantity", "l_extendedprice", "l_discount", "l_tax", "l_returnflag", "l_linestatus", "l_shipdate",
             "l_orderkey", "l_comment"]
    got = pa.Table.from_batches([ctx.read_csv(text, LINEITEM, names).to_arrow()])
    want = arrow_read(text, LINEITEM, names)
    assert got.schema.names == names and got.num_rows == n
    for c in names:
        a, b = got[c].combine_chunks(), want[c].combine_chunks()
        if pa.types.is_floating(a.type):  # bitwise: both are correctly rounded conversions of the same text
            assert np.array_equal(a.to_numpy().view(np.uint64), b.to_numpy().view(np.uint64)), c
        else:
            assert a.equals(b), c


def test_decoded_text_feeds_q1(ctx, O):
    """text -> GPU decode -> the Q1 plan: same result as the plan over the values parsed by pyarrow."""
    from ballista_b200 import MemoryExec, tpch
    text = tbl_text(20_000, seed=5)
    batch = ctx.read_csv(text, LINEITEM, tpch.Q1_COLUMNS)
    scan = tpch.scan(ctx, [[batch]])
    partial = tpch.q1_partial(scan)
    got = tpch.run_two_stage(ctx, partial, lambda s: tpch.q1_final(s, scan.schema())).to_pylist()
    t = arrow_read(text, LINEITEM, tpch.Q1_COLUMNS)
    scan2 = tpch.scan(ctx, [t.to_batches()])
    partial2 = tpch.q1_partial(scan2)
    want = tpch.run_two_stage(ctx, partial2, lambda s: tpch.q1_final(s, scan2.schema())).to_pylist()
    key = lambda r: (r["l_returnflag"], r["l_linestatus"])
    got.sort(key=key)
    want.sort(key=key)
    assert [key(r) for r in got] == [key(r) for r in want] and len(got) >= 4
    for g, w in zip(got, want):
        assert g["count_order"] == w["count_order"]
        assert abs(g["sum_charge"] - w["sum_charge"]) <= 1e-12 * abs(w["sum_charge"])


def test_reference_fixture_values_roundtrip(ctx):
    """The reference's own 10-row lineitem fixtures (values in tests/golden/ref_testdata.json), printed back to .tbl
    syntax: the decoder returns exactly those values."""
    with open(os.path.join(HERE, "golden", "ref_testdata.json")) as f:
        ref = json.load(f)
    p = ref["lineitem"]["partitions"][0]
    n = len(p["l_orderkey"])
    def cell(c, i):
        v = p.get(c)
        if v is None:
            return "x"
        v = v[i]
        if c in ("l_shipdate", "l_commitdate", "l_receiptdate"):
            return (EPOCH + datetime.timedelta(days=v)).isoformat()
        return str(v)  # decimals are kept as the fixture's own text
    text = "".join("|".join(cell(c, i) for c in LINEITEM.names) + "|\n" for i in range(n)).encode()
    have = [c for c in LINEITEM.names if c in p]
    got = ctx.read_csv(text, LINEITEM, have).to_arrow()
    for c in have:
        vals = got[c].to_pylist()
        if c in ("l_shipdate", "l_commitdate", "l_receiptdate"):
            vals = [(d - EPOCH).days for d in vals]
        want = [float(v) for v in p[c]] if pa.types.is_floating(LINEITEM.field(c).type) else p[c]
        assert vals == want, c


def test_header_empty_and_malformed(ctx):
    from ballista_b200 import BallistaError
    sch = pa.schema([("a", pa.int32()), ("b", pa.float64()), ("d", pa.date32()), ("s", pa.utf8())])
    t = ctx.read_csv(b"a,b,d,s\n-5,1.25,1999-12-31,hello\n7,0.5,2000-02-29,\n", sch, delimiter=",", has_header=True).to_arrow()
    assert t["a"].to_pylist() == [-5, 7] and t["b"].to_pylist() == [1.25, 0.5] and t["s"].to_pylist() == ["hello", ""]
    assert t["d"].to_pylist() == [datetime.date(1999, 12, 31), datetime.date(2000, 2, 29)]
    assert ctx.read_csv(b"", sch, delimiter=",").num_rows == 0
    for bad, line in ((b"1,2.0,2000-01-01,x\nzz,2.0,2000-01-01,x\n", "line 2"), (b"1,2.0,2000-13-01,x\n", "line 1"),
                      (b"1,2.0\n", "line 1"), (b"1,0.1234567890123456,2000-01-01,x\n", "15 significant")):
        with pytest.raises(BallistaError) as e:
            ctx.read_csv(bad, sch, delimiter=",")
        assert line in str(e.value)
