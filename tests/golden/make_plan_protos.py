"""Generates tests/golden/plan_protos.json: serialized `PhysicalPlanNode` messages of the benchmark stage plans,
encoded by Google's protobuf runtime from the REFERENCE's own schema (/root/reference/rust/core/proto/ballista.proto).

`protoc` is not available, so the .proto text is parsed here (proto3 subset: messages, enums, oneof, repeated) into a
FileDescriptorProto and handed to the runtime.  Run in the authoring container only (the GPU box has no
/root/reference); the JSON fixture travels.  The plans are what `rust/core/src/serde/physical_plan/to_proto.rs`
emits for the DataFusion plans of q1.sql / q3.sql / q6.sql (expression shapes: SQL literals are Int64 / Float64 /
Utf8 scalars, `date '...'` is CAST(Utf8 AS Date32), COUNT(*) is COUNT(UInt8(1))).

    python tests/golden/make_plan_protos.py
"""
import base64
import json
import os
import re

from google.protobuf import descriptor_pb2, descriptor_pool, message_factory

PROTO = "/root/reference/rust/core/proto/ballista.proto"
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "plan_protos.json")

SCALARS = {"double": 1, "float": 2, "int64": 3, "uint64": 4, "int32": 5, "bool": 8, "string": 9, "bytes": 12,
           "uint32": 13, "sint32": 17, "sint64": 18}


def parse_proto(text):
    text = re.sub(r"//[^\n]*", "", text)
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    toks = re.findall(r"\"[^\"]*\"|[A-Za-z_][A-Za-z0-9_.]*|\d+|[{}=;<>,\[\]()]", text)
    fd = descriptor_pb2.FileDescriptorProto(name="ballista.proto", package="ballista.protobuf", syntax="proto3")
    pos = 0
    enums, messages = set(), set()
    # first pass: type names
    for i, t in enumerate(toks):
        if t == "message":
            messages.add(toks[i + 1])
        elif t == "enum":
            enums.add(toks[i + 1])

    def field(msg, label_repeated, ftype, name, number, oneof_index=None):
        f = msg.field.add(name=name, number=int(number))
        f.label = 3 if label_repeated else 1
        if ftype in SCALARS:
            f.type = SCALARS[ftype]
        elif ftype in enums:
            f.type, f.type_name = 14, ".ballista.protobuf." + ftype
        elif ftype in messages:
            f.type, f.type_name = 11, ".ballista.protobuf." + ftype
        else:
            raise ValueError(f"unknown type {ftype}")
        if oneof_index is not None:
            f.oneof_index = oneof_index

    while pos < len(toks):
        t = toks[pos]
        if t in ("syntax", "package", "option"):
            while toks[pos] != ";":
                pos += 1
            pos += 1
        elif t == "enum":
            e = fd.enum_type.add(name=toks[pos + 1])
            pos += 3
            while toks[pos] != "}":
                e.value.add(name=toks[pos], number=int(toks[pos + 2]))
                pos += 4
            pos += 1
        elif t == "message":
            m = fd.message_type.add(name=toks[pos + 1])
            pos += 3
            while toks[pos] != "}":
                if toks[pos] == "oneof":
                    m.oneof_decl.add(name=toks[pos + 1])
                    idx = len(m.oneof_decl) - 1
                    pos += 3
                    while toks[pos] != "}":
                        field(m, False, toks[pos], toks[pos + 1], toks[pos + 3], idx)
                        pos += 5
                    pos += 1
                else:
                    rep = toks[pos] == "repeated"
                    if rep:
                        pos += 1
                    field(m, rep, toks[pos], toks[pos + 1], toks[pos + 3])
                    pos += 5
            pos += 1
        elif t == "service":  # the gRPC services are not needed: skip the balanced block
            while toks[pos] != "{":
                pos += 1
            depth = 0
            while True:
                depth += toks[pos] == "{"
                depth -= toks[pos] == "}"
                pos += 1
                if depth == 0:
                    break
        else:
            raise ValueError(f"unexpected token {t!r} at {pos}")
    return fd


def load():
    fd = parse_proto(open(PROTO).read())
    pool = descriptor_pool.DescriptorPool()
    pool.Add(fd)

    def cls(name):
        return message_factory.GetMessageClass(pool.FindMessageTypeByName("ballista.protobuf." + name))
    return cls


cls = load()
Plan, Expr, Scalar = cls("PhysicalPlanNode"), cls("LogicalExprNode"), cls("ScalarValue")
ArrowType, Schema = cls("ArrowType"), cls("Schema")


# ---- expression helpers (shapes of logical_plan/to_proto.rs) ----
def col(name):
    return Expr(column_name=name)


def lit(v):
    if isinstance(v, bool):
        return Expr(literal=Scalar(bool_value=v))
    if isinstance(v, int):
        return Expr(literal=Scalar(int64_value=v))
    if isinstance(v, float):
        return Expr(literal=Scalar(float64_value=v))
    return Expr(literal=Scalar(utf8_value=v))


def binary(op, l, r):
    e = Expr()
    e.binary_expr.l.CopyFrom(l)
    e.binary_expr.r.CopyFrom(r)
    e.binary_expr.op = op
    return e


def arrow_type(name):
    t = ArrowType()
    getattr(t, name).SetInParent()
    return t


def date(s):
    e = Expr()
    e.cast.expr.CopyFrom(lit(s))
    e.cast.arrow_type.CopyFrom(arrow_type("DATE32"))
    return e


def between(x, lo, hi):
    e = Expr()
    e.between.expr.CopyFrom(x)
    e.between.low.CopyFrom(lo)
    e.between.high.CopyFrom(hi)
    return e


def scalar_fn(fun, *args):
    e = Expr()
    e.scalar_function.fun = {"OCTETLENGTH": 17, "CONCAT": 18, "LOWER": 19, "UPPER": 20, "TRIM": 21, "LTRIM": 22, "RTRIM": 23,
                             "NULLIF": 26, "MD5": 28}[fun]
    for a in args:
        e.scalar_function.expr.add().CopyFrom(a)
    return e


def case(when_then, otherwise=None):
    e = Expr()
    for w, t in when_then:
        wt = e.case_.when_then_expr.add()
        wt.when_expr.CopyFrom(w)
        wt.then_expr.CopyFrom(t)
    if otherwise is not None:
        e.case_.else_expr.CopyFrom(otherwise)
    return e


def agg(fn, arg):
    e = Expr()
    e.aggregate_expr.aggr_function = {"MIN": 0, "MAX": 1, "SUM": 2, "AVG": 3, "COUNT": 4}[fn]
    e.aggregate_expr.expr.CopyFrom(arg)
    return e


def sort_expr(x, asc=True):
    e = Expr()
    e.sort.expr.CopyFrom(x)
    e.sort.asc = asc
    e.sort.nulls_first = not asc
    return e


def schema(fields):
    s = Schema()
    for name, t in fields:
        f = s.columns.add(name=name, nullable=False)
        f.arrow_type.CopyFrom(arrow_type(t))
    return s


LINEITEM = [("l_orderkey", "INT32"), ("l_partkey", "INT32"), ("l_suppkey", "INT32"), ("l_linenumber", "INT32"),
            ("l_quantity", "FLOAT64"), ("l_extendedprice", "FLOAT64"), ("l_discount", "FLOAT64"), ("l_tax", "FLOAT64"),
            ("l_returnflag", "UTF8"), ("l_linestatus", "UTF8"), ("l_shipdate", "DATE32"), ("l_commitdate", "DATE32"),
            ("l_receiptdate", "DATE32"), ("l_shipinstruct", "UTF8"), ("l_shipmode", "UTF8"), ("l_comment", "UTF8")]
ORDERS = [("o_orderkey", "INT32"), ("o_custkey", "INT32"), ("o_orderstatus", "UTF8"), ("o_totalprice", "FLOAT64"),
          ("o_orderdate", "DATE32"), ("o_orderpriority", "UTF8"), ("o_clerk", "UTF8"), ("o_shippriority", "INT32"),
          ("o_comment", "UTF8")]
CUSTOMER = [("c_custkey", "INT32"), ("c_name", "UTF8"), ("c_address", "UTF8"), ("c_nationkey", "INT32"),
            ("c_phone", "UTF8"), ("c_acctbal", "FLOAT64"), ("c_mktsegment", "UTF8"), ("c_comment", "UTF8")]


def csv_scan(table, fields, projection):
    p = Plan()
    n = p.csv_scan
    n.path = f"/data/{table}"
    n.projection.extend([[f for f, _ in fields].index(c) for c in projection])
    n.schema.CopyFrom(schema(fields))
    n.file_extension, n.has_header, n.batch_size, n.delimiter = ".tbl", False, 32768, "|"
    n.filename.append(f"/data/{table}/partition0.tbl")
    return p


def unary(kind, input, **kw):
    p = Plan()
    n = getattr(p, kind)
    n.input.CopyFrom(input)
    for k, v in kw.items():
        if isinstance(v, list):
            getattr(n, k).extend(v)
        elif hasattr(v, "SerializeToString"):
            getattr(n, k).CopyFrom(v)
        else:
            setattr(n, k, v)
    return p


def q1_aggs():
    one_minus = binary("Minus", lit(1), col("l_discount"))
    disc = binary("Multiply", col("l_extendedprice"), one_minus)
    charge = binary("Multiply", disc, binary("Plus", lit(1), col("l_tax")))
    return [("sum_qty", agg("SUM", col("l_quantity"))), ("sum_base_price", agg("SUM", col("l_extendedprice"))),
            ("sum_disc_price", agg("SUM", disc)), ("sum_charge", agg("SUM", charge)),
            ("avg_qty", agg("AVG", col("l_quantity"))), ("avg_price", agg("AVG", col("l_extendedprice"))),
            ("avg_disc", agg("AVG", col("l_discount"))),
            ("count_order", agg("COUNT", Expr(literal=Scalar(uint8_value=1))))]


def hash_aggregate(mode, groups, aggs, input, input_fields):
    p = Plan()
    n = p.hash_aggregate
    n.mode = {"PARTIAL": 0, "FINAL": 1}[mode]
    n.input.CopyFrom(input)
    for name, e in groups:
        n.group_expr.add().CopyFrom(e)
        n.group_expr_name.append(name)
    for name, e in aggs:
        n.aggr_expr.add().CopyFrom(e)
        n.aggr_expr_name.append(name)
    n.input_schema.CopyFrom(schema(input_fields))
    return p


def build():
    plans = {}
    q1_cols = ["l_quantity", "l_extendedprice", "l_discount", "l_tax", "l_returnflag", "l_linestatus", "l_shipdate"]
    q1_fields = [f for f in LINEITEM if f[0] in q1_cols]
    q1_fields.sort(key=lambda f: q1_cols.index(f[0]))
    scan = csv_scan("lineitem", LINEITEM, q1_cols)
    filt = unary("filter", scan, expr=binary("LtEq", col("l_shipdate"), date("1998-09-02")))
    coal = unary("coalesce_batches", filt, target_batch_size=4096)
    groups = [("l_returnflag", col("l_returnflag")), ("l_linestatus", col("l_linestatus"))]
    plans["q1_partial_stage"] = {"plan": hash_aggregate("PARTIAL", groups, q1_aggs(), coal, q1_fields),
                                 "leaves": ["lineitem:" + ",".join(q1_cols)],
                                 "nodes": ["HashAggregateExec", "CoalesceBatchesExec", "FilterExec", "CsvScan"]}
    # Final stage: HashAggregate(Final) <- Merge <- ShuffleReader(partial states), then Sort (q1.sql ORDER BY)
    reader = Plan()
    reader.shuffle_reader.schema.CopyFrom(Schema())
    merge = unary("merge", reader)
    final = hash_aggregate("FINAL", groups, q1_aggs(), merge, q1_fields)
    plans["q1_final_stage"] = {"plan": unary("sort", final, expr=[sort_expr(col("l_returnflag")), sort_expr(col("l_linestatus"))]),
                               "leaves": ["partial_states"],
                               "nodes": ["SortExec", "HashAggregateExec", "MergeExec", "ShuffleReader"]}
    # Q6: filter + global sum
    q6_cols = ["l_quantity", "l_extendedprice", "l_discount", "l_shipdate"]
    q6_fields = sorted([f for f in LINEITEM if f[0] in q6_cols], key=lambda f: q6_cols.index(f[0]))
    pred = binary("And", binary("And", binary("And",
                  binary("GtEq", col("l_shipdate"), date("1994-01-01")),
                  binary("Lt", col("l_shipdate"), date("1995-01-01"))),
                  between(col("l_discount"), binary("Minus", lit(0.06), lit(0.01)), binary("Plus", lit(0.06), lit(0.01)))),
                  binary("Lt", col("l_quantity"), lit(24)))
    filt6 = unary("coalesce_batches", unary("filter", csv_scan("lineitem", LINEITEM, q6_cols), expr=pred), target_batch_size=4096)
    rev = [("revenue", agg("SUM", binary("Multiply", col("l_extendedprice"), col("l_discount"))))]
    plans["q6_partial_stage"] = {"plan": hash_aggregate("PARTIAL", [], rev, filt6, q6_fields),
                                 "leaves": ["lineitem:" + ",".join(q6_cols)],
                                 "nodes": ["HashAggregateExec", "CoalesceBatchesExec", "FilterExec", "CsvScan"]}
    # Q3: (customer |x| orders) |x| lineitem -> partial aggregate; joins wrapped in CoalesceBatches like the planner does
    c_cols, o_cols = ["c_custkey", "c_mktsegment"], ["o_orderkey", "o_custkey", "o_orderdate", "o_shippriority"]
    l_cols = ["l_orderkey", "l_extendedprice", "l_discount", "l_shipdate"]
    c = unary("coalesce_batches", unary("filter", csv_scan("customer", CUSTOMER, c_cols),
                                        expr=binary("Eq", col("c_mktsegment"), lit("BUILDING"))), target_batch_size=4096)
    o = unary("coalesce_batches", unary("filter", csv_scan("orders", ORDERS, o_cols),
                                        expr=binary("Lt", col("o_orderdate"), date("1995-03-15"))), target_batch_size=4096)
    l = unary("coalesce_batches", unary("filter", csv_scan("lineitem", LINEITEM, l_cols),
                                        expr=binary("Gt", col("l_shipdate"), date("1995-03-15"))), target_batch_size=4096)

    def join(left, right, a, b):
        p = Plan()
        p.hash_join.left.CopyFrom(left)
        p.hash_join.right.CopyFrom(right)
        p.hash_join.on.add(left=a, right=b)
        p.hash_join.join_type = 0
        return unary("coalesce_batches", p, target_batch_size=4096)
    j2 = join(join(c, o, "c_custkey", "o_custkey"), l, "o_orderkey", "l_orderkey")
    q3_groups = [("l_orderkey", col("l_orderkey")), ("o_orderdate", col("o_orderdate")), ("o_shippriority", col("o_shippriority"))]
    q3_rev = [("revenue", agg("SUM", binary("Multiply", col("l_extendedprice"), binary("Minus", lit(1), col("l_discount")))))]
    joined_fields = ([f for f in CUSTOMER if f[0] in c_cols] + [f for f in ORDERS if f[0] in o_cols]
                     + [f for f in LINEITEM if f[0] in l_cols])
    plans["q3_partial_stage"] = {"plan": hash_aggregate("PARTIAL", q3_groups, q3_rev, j2, joined_fields),
                                 "leaves": ["customer:" + ",".join(c_cols), "orders:" + ",".join(o_cols),
                                            "lineitem:" + ",".join(l_cols)],
                                 "nodes": ["HashAggregateExec", "CoalesceBatchesExec", "HashJoinExec", "CoalesceBatchesExec",
                                           "HashJoinExec", "CoalesceBatchesExec", "FilterExec", "CsvScan",
                                           "CoalesceBatchesExec", "FilterExec", "CsvScan", "CoalesceBatchesExec",
                                           "FilterExec", "CsvScan"]}
    # Q3 final stage: Final aggregate, ORDER BY revenue DESC, o_orderdate, LIMIT 10
    reader3 = Plan()
    reader3.shuffle_reader.schema.CopyFrom(Schema())
    final3 = hash_aggregate("FINAL", q3_groups, q3_rev, unary("merge", reader3), joined_fields)
    top = unary("global_limit", unary("sort", final3, expr=[sort_expr(col("revenue"), False), sort_expr(col("o_orderdate"))]), limit=10)
    plans["q3_final_stage"] = {"plan": top, "leaves": ["partial_states"],
                               "nodes": ["GlobalLimitExec", "SortExec", "HashAggregateExec", "MergeExec", "ShuffleReader"]}
    # hash repartition of the filtered lineitem side (what a shuffle-writing stage would run)
    rp = Plan()
    rp.repartition.input.CopyFrom(l)
    rp.repartition.hash.hash_expr.add().CopyFrom(col("l_orderkey"))
    rp.repartition.hash.partition_count = 4
    plans["lineitem_repartition_stage"] = {"plan": rp, "leaves": ["lineitem:" + ",".join(l_cols)],
                                           "nodes": ["RepartitionExec", "CoalesceBatchesExec", "FilterExec", "CsvScan"]}
    # projection over customer with the string / NULL-producing vocabulary (ScalarFunctionNode, CaseNode)
    pc_cols = ["c_custkey", "c_nationkey", "c_mktsegment"]
    proj = Plan()
    proj.projection.input.CopyFrom(csv_scan("customer", CUSTOMER, pc_cols))
    for name, e in [("c_custkey", col("c_custkey")),
                    ("seg", scalar_fn("TRIM", col("c_mktsegment"))),
                    ("h", scalar_fn("MD5", col("c_mktsegment"))),
                    ("tag", scalar_fn("CONCAT", col("c_mktsegment"), lit("#"), scalar_fn("LOWER", col("c_mktsegment")))),
                    ("nk", scalar_fn("NULLIF", col("c_nationkey"), Expr(literal=Scalar(int32_value=3)))),
                    ("cls", case([(binary("Eq", col("c_mktsegment"), lit("BUILDING")), lit("b")),
                                  (binary("Lt", col("c_nationkey"), Expr(literal=Scalar(int32_value=5))), col("c_mktsegment"))],
                                 lit("other")))]:
        proj.projection.expr.add().CopyFrom(e)
        proj.projection.expr_name.append(name)
    plans["customer_projection_stage"] = {"plan": proj, "leaves": ["customer:" + ",".join(pc_cols)],
                                          "nodes": ["ProjectionExec", "CsvScan"]}
    return plans


if __name__ == "__main__":
    out = {"_generated_by": "tests/golden/make_plan_protos.py from rust/core/proto/ballista.proto (protobuf runtime "
                            "encoding of the reference's own message definitions)"}
    for name, p in build().items():
        out[name] = {"bytes_b64": base64.b64encode(p["plan"].SerializeToString()).decode(), "leaves": p["leaves"],
                     "nodes_preorder": p["nodes"]}
    with open(OUT, "w") as f:
        json.dump(out, f, indent=1)
    print("wrote", OUT, {k: len(v["bytes_b64"]) for k, v in out.items() if k != "_generated_by"})
