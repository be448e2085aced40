"""TPC-H query plans as the reference's planner shapes them (SQL in
rust/benchmarks/tpch/queries/q{1,3,6}.sql), assembled from the operator mirror in plan.py.

Each builder returns (partial_stage_plan, make_final) so that the Partial stage runs per partition / per GPU
and the Final stage runs over the merged partial states — the two Ballista query stages split at
rust/scheduler/src/planner.rs:149-169.
"""
import functools

import pyarrow as pa

from . import _capi as K
from .expr import avg, col, count, date, lit, sum as sum_
from .plan import (CoalesceBatchesExec, FilterExec, GlobalLimitExec, HashAggregateExec, HashJoinExec, MemoryExec,
                   MergeExec, ProjectionExec, RepartitionExec, SortExec)

Q1_COLUMNS = ["l_quantity", "l_extendedprice", "l_discount", "l_tax", "l_returnflag", "l_linestatus", "l_shipdate"]
Q6_COLUMNS = ["l_quantity", "l_extendedprice", "l_discount", "l_shipdate"]
Q3_LINEITEM_COLUMNS = ["l_orderkey", "l_extendedprice", "l_discount", "l_shipdate"]
Q3_ORDERS_COLUMNS = ["o_orderkey", "o_custkey", "o_orderdate", "o_shippriority"]
Q3_CUSTOMER_COLUMNS = ["c_custkey", "c_mktsegment"]

Q5_CUSTOMER_COLUMNS = ["c_custkey", "c_nationkey"]
Q5_ORDERS_COLUMNS = ["o_orderkey", "o_custkey", "o_orderdate"]
Q5_LINEITEM_COLUMNS = ["l_orderkey", "l_suppkey", "l_extendedprice", "l_discount"]
Q5_SUPPLIER_COLUMNS = ["s_suppkey", "s_nationkey"]
Q5_NATION_COLUMNS = ["n_nationkey", "n_name", "n_regionkey"]
Q5_REGION_COLUMNS = ["r_regionkey", "r_name"]

# algorithmic bytes per input row
Q1_BYTES_PER_ROW = 46
Q6_BYTES_PER_ROW = 28


@functools.lru_cache(maxsize=None)
def q1_aggregates():
    disc_price = col("l_extendedprice") * (lit(1) - col("l_discount"))
    charge = col("l_extendedprice") * (lit(1) - col("l_discount")) * (lit(1) + col("l_tax"))
    return [
        sum_(col("l_quantity")).alias("sum_qty"),
        sum_(col("l_extendedprice")).alias("sum_base_price"),
        sum_(disc_price).alias("sum_disc_price"),
        sum_(charge).alias("sum_charge"),
        avg(col("l_quantity")).alias("avg_qty"),
        avg(col("l_extendedprice")).alias("avg_price"),
        avg(col("l_discount")).alias("avg_disc"),
        count().alias("count_order"),
    ]


@functools.lru_cache(maxsize=None)
def _q1_groups():
    return (col("l_returnflag"), col("l_linestatus"))


@functools.lru_cache(maxsize=64)
def _q1_predicate(ship_date):
    return col("l_shipdate") <= date(ship_date)


def q1_partial(scan, ship_date="1998-09-02"):
    """HashAggregateExec(Partial) <- CoalesceBatchesExec <- FilterExec(l_shipdate <= d) <- scan."""
    filt = CoalesceBatchesExec(FilterExec(_q1_predicate(ship_date), scan))
    return HashAggregateExec("Partial", _q1_groups(), q1_aggregates(), filt)


def q1_final(partial_states, input_schema):
    """HashAggregateExec(Final) <- MergeExec <- partial states (any plan producing the Partial schema)."""
    return HashAggregateExec("Final", _q1_groups(), q1_aggregates(), MergeExec(partial_states), input_schema=input_schema)


@functools.lru_cache(maxsize=None)
def q6_predicate():
    return ((col("l_shipdate") >= date("1994-01-01")) & (col("l_shipdate") < date("1995-01-01"))
            & col("l_discount").between(lit(0.06) - lit(0.01), lit(0.06) + lit(0.01))
            & (col("l_quantity") < lit(24)))


@functools.lru_cache(maxsize=None)
def q6_aggregates():
    return [sum_(col("l_extendedprice") * col("l_discount")).alias("revenue")]


def q6_partial(scan):
    filt = CoalesceBatchesExec(FilterExec(q6_predicate(), scan))
    return HashAggregateExec("Partial", [], q6_aggregates(), filt)


def q6_final(partial_states, input_schema):
    return HashAggregateExec("Final", [], q6_aggregates(), MergeExec(partial_states), input_schema=input_schema)


@functools.lru_cache(maxsize=64)
def _q3_predicates(segment, cutoff):
    return (col("c_mktsegment") == lit(segment), col("o_orderdate") < date(cutoff), col("l_shipdate") > date(cutoff))


def q3_join_tree(customer, orders, lineitem, segment="BUILDING", cutoff="1995-03-15"):
    """(customer ⋈ orders) ⋈ lineitem, left-deep, build = left, filters pushed below the joins."""
    pc, po, pl = _q3_predicates(segment, cutoff)
    c = CoalesceBatchesExec(FilterExec(pc, customer))
    o = CoalesceBatchesExec(FilterExec(po, orders))
    l = CoalesceBatchesExec(FilterExec(pl, lineitem))
    j1 = CoalesceBatchesExec(HashJoinExec(c, o, [("c_custkey", "o_custkey")]))
    j2 = CoalesceBatchesExec(HashJoinExec(j1, l, [("o_orderkey", "l_orderkey")]))
    return j2


def q3_distributed_join_tree(customer, orders, lineitem, nranks, segment="BUILDING", cutoff="1995-03-15"):
    """Q3 over row-range partitions (one per GPU): both inputs of each join are hash-repartitioned on the join key
    into one partition per rank (the ShuffleWriter -> ShuffleReaderExec hop, here an exchange over NVLink), so each
    rank joins only its own key range.  Filters sit directly below the repartitions and ride in the split kernels."""
    pc, po, pl = _q3_predicates(segment, cutoff)
    c = RepartitionExec(CoalesceBatchesExec(FilterExec(pc, customer)), "Hash", nranks, [col("c_custkey")])
    o = RepartitionExec(CoalesceBatchesExec(FilterExec(po, orders)), "Hash", nranks, [col("o_custkey")])
    j1 = CoalesceBatchesExec(HashJoinExec(c, o, [("c_custkey", "o_custkey")]))
    j1s = RepartitionExec(j1, "Hash", nranks, [col("o_orderkey")])
    l = RepartitionExec(CoalesceBatchesExec(FilterExec(pl, lineitem)), "Hash", nranks, [col("l_orderkey")])
    return CoalesceBatchesExec(HashJoinExec(j1s, l, [("o_orderkey", "l_orderkey")]))


def q3_distributed_partial(customer, orders, lineitem, nranks, **kw):
    """Partial aggregate over the partitioned join.  Groups are keyed by l_orderkey, the last shuffle key, so the
    states of different ranks hold disjoint groups and each rank finalises its own (no further exchange)."""
    return HashAggregateExec("Partial", q3_groups(), q3_aggregates(),
                             q3_distributed_join_tree(customer, orders, lineitem, nranks, **kw))


def q3_broadcast_stages(ctx, customer, orders, lineitem, nranks, segment="BUILDING", cutoff="1995-03-15"):
    """Q3 over row-range partitions with BROADCAST build sides: the reference's own join is collect-left
    (`HashJoinExec` builds over `MergeExec(left)`), which across ranks is an all-gather of the filtered
    build side; the probe side (orders, then lineitem: 98 % of the bytes) never leaves its GPU.  Run as Ballista runs
    stages, each stage's output materialised and handed to the next:
      1  Filter(customer) -> c_custkey                                   -> all-gather
      2  HashJoin(stage 1, Filter(orders)) -> order key / date / priority -> all-gather
      3  HashJoin(stage 2, Filter(lineitem)) -> Partial aggregate -> Repartition(Hash(group keys), N)
      4  Final aggregate of this rank's groups (the caller: `q3_final`)
    Returns (partial_states_op, input_schema_for_final): the states of this rank's partition of the groups."""
    pc, po, pl = _q3_predicates(segment, cutoff)
    rank = ctx.rank if nranks > 1 else 0

    def gather(plan):
        got = CoalesceBatchesExec(plan, (1 << 31) - 2).collect_device(0)
        if len(got) > 1:
            raise RuntimeError("a broadcast stage produced more than one batch")
        if not got:  # every rank joins the collective, with an empty batch if it has no rows
            got = [ctx.upload(pa.RecordBatch.from_pylist([], schema=plan.schema()))]
        if nranks > 1:
            got = [ctx.allgather(got[0])]
        return MemoryExec(ctx, plan.schema(), [got])
    s1 = gather(ProjectionExec([col("c_custkey")], CoalesceBatchesExec(FilterExec(pc, customer))))
    j1 = HashJoinExec(s1, CoalesceBatchesExec(FilterExec(po, orders)), [("c_custkey", "o_custkey")])
    s2 = gather(ProjectionExec([col("o_orderkey"), col("o_orderdate"), col("o_shippriority")], CoalesceBatchesExec(j1)))
    j2 = CoalesceBatchesExec(HashJoinExec(s2, CoalesceBatchesExec(FilterExec(pl, lineitem)), [("o_orderkey", "l_orderkey")]))
    partial = HashAggregateExec("Partial", q3_groups(), q3_aggregates(), j2)
    if nranks == 1:
        return partial, j2.schema(), 0
    # a group (one order) can straddle two lineitem row ranges: co-locate the states by group key before the Final
    names = [f.name for f in partial.schema()]
    return RepartitionExec(partial, "Hash", nranks, [col(names[0])]), j2.schema(), rank


@functools.lru_cache(maxsize=None)
def q3_aggregates():
    return [sum_(col("l_extendedprice") * (lit(1) - col("l_discount"))).alias("revenue")]


@functools.lru_cache(maxsize=None)
def q3_groups():
    return [col("l_orderkey"), col("o_orderdate"), col("o_shippriority")]


def q3_partial(customer, orders, lineitem, **kw):
    return HashAggregateExec("Partial", q3_groups(), q3_aggregates(), q3_join_tree(customer, orders, lineitem, **kw))


def q3_final(partial_states, input_schema):
    return HashAggregateExec("Final", q3_groups(), q3_aggregates(), MergeExec(partial_states),
                             input_schema=input_schema)


@functools.lru_cache(maxsize=16)
def _q5_predicates(region, date_lo, date_hi):
    return (col("r_name") == lit(region), (col("o_orderdate") >= date(date_lo)) & (col("o_orderdate") < date(date_hi)))


@functools.lru_cache(maxsize=None)
def q5_aggregates():
    return [sum_(col("l_extendedprice") * (lit(1) - col("l_discount"))).alias("revenue")]


def q5_join_tree(customer, orders, lineitem, supplier, nation, region, region_name="ASIA", date_lo="1994-01-01",
                 date_hi="1995-01-01"):
    """q5.sql as the reference's planner shapes it: comma joins in FROM order, left-deep, build = left,
    single-table predicates pushed below the joins; `l_suppkey = s_suppkey AND c_nationkey = s_nationkey` is ONE join
    with a two-column key."""
    pr, po = _q5_predicates(region_name, date_lo, date_hi)
    o = CoalesceBatchesExec(FilterExec(po, orders))
    r = CoalesceBatchesExec(FilterExec(pr, region))
    j1 = CoalesceBatchesExec(HashJoinExec(customer, o, [("c_custkey", "o_custkey")]))
    j2 = CoalesceBatchesExec(HashJoinExec(j1, lineitem, [("o_orderkey", "l_orderkey")]))
    j3 = CoalesceBatchesExec(HashJoinExec(j2, supplier, [("l_suppkey", "s_suppkey"), ("c_nationkey", "s_nationkey")]))
    j4 = CoalesceBatchesExec(HashJoinExec(j3, nation, [("s_nationkey", "n_nationkey")]))
    return CoalesceBatchesExec(HashJoinExec(j4, r, [("n_regionkey", "r_regionkey")]))


def q5_partial(customer, orders, lineitem, supplier, nation, region, **kw):
    return HashAggregateExec("Partial", [col("n_name")], q5_aggregates(),
                             q5_join_tree(customer, orders, lineitem, supplier, nation, region, **kw))


def q5_final(partial_states, input_schema):
    return HashAggregateExec("Final", [col("n_name")], q5_aggregates(), MergeExec(partial_states),
                             input_schema=input_schema)


def q5_sorted(final):
    """ORDER BY revenue DESC (q5.sql:24)."""
    return SortExec([(col("revenue"), False)], final)


Q5_TABLES = (("customer", Q5_CUSTOMER_COLUMNS), ("orders", Q5_ORDERS_COLUMNS), ("lineitem", Q5_LINEITEM_COLUMNS),
             ("supplier", Q5_SUPPLIER_COLUMNS), ("nation", Q5_NATION_COLUMNS), ("region", Q5_REGION_COLUMNS))


def q1_sorted(final):
    """ORDER BY l_returnflag, l_linestatus (q1.sql:17-21) over the Final aggregate."""
    return SortExec([col("l_returnflag"), col("l_linestatus")], final)


def q3_top(final, limit=10):
    """ORDER BY revenue DESC, o_orderdate LIMIT 10 (q3.sql:20-23) over the Final aggregate."""
    return GlobalLimitExec(SortExec([(col("revenue"), False), (col("o_orderdate"), True)], final), limit)


def scan(ctx, batches_per_partition, schema=None):
    """MemoryExec over a list (partitions) of lists of DeviceBatch / pyarrow RecordBatch."""
    if schema is None:
        first = next(b for p in batches_per_partition for b in p)
        schema = first.schema
    return MemoryExec(ctx, schema, batches_per_partition)


class TwoStage:
    """A Partial stage plan and its Final stage plan, built ONCE and executed per call - what an executor does when it
    keeps the decoded plan of a stage and runs one task per partition (b2_op_execute is the per-task call,
    rust/executor/src/flight_service.rs:117-121).  `run()` = Partial task on this GPU -> merge of the partial states
    (all-gather over NVLink when a communicator is attached, replacing the MergeExec stage hop) -> Final task."""

    def __init__(self, ctx, partial_plan, make_final, partition=0, merge="local"):
        self.ctx, self.partial, self.partition = ctx, partial_plan, partition
        self.gather = ctx.nranks > 1 and merge != "local"
        self.leaf = MemoryExec(ctx, partial_plan.schema(), [[]])
        self.final = make_final(self.leaf)

    def run(self):
        states = self.partial.collect_device(self.partition)
        if self.gather:
            states = [self.ctx.allgather(s) for s in states]
        self.leaf.clear()
        for s in states:
            self.leaf.add(0, s)
        return self.final.collect()


def run_two_stage(ctx, partial_plan, make_final, partition=0, merge="local"):
    """Run the Partial stage on this GPU, merge partial states (all-gather over NVLink when a communicator
    is attached, replacing the MergeExec stage hop), then the Final stage.  Returns a pyarrow Table."""
    states = partial_plan.collect_device(partition)
    if ctx.nranks > 1 and merge != "local":
        states = [ctx.allgather(s) for s in states]
    final = make_final(MemoryExec(ctx, partial_plan.schema(), [states]))
    return final.collect()
