"""Python face of the CPU oracle (oracle/liboracle.so + numpy).  TEST INFRASTRUCTURE: imported only by
tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs — never by ballista_b200.

PARITY UNPINNED (see oracle.cpp): restates DataFusion@46161d2 operator semantics on the
CPU.  Expressions are evaluated column-at-a-time with numpy float64 (IEEE, one rounding per operation — numpy
never fuses multiply-add), filters are numpy boolean compaction (arrow::compute::filter keeps order), and
the grouped accumulation / join / repartition loops are the C++ restatement in oracle.cpp.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

T_INT32, T_INT64, T_UINT64, T_FLOAT64, T_DATE32, T_UTF8 = 2, 3, 4, 5, 6, 7
BATCH = 32768  # from_proto.rs:100-102


class KeyCol(C.Structure):
    _fields_ = [("dtype", C.c_int32), ("pad", C.c_int32), ("data", C.c_void_p), ("chars", C.c_void_p)]


def build():
    subprocess.check_call(["make", "-s", "-C", _HERE])


def lib():
    global _LIB
    if _LIB is None:
        path = os.path.join(_HERE, "liboracle.so")
        if not os.path.exists(path):
            build()
        _LIB = C.CDLL(path)
        _LIB.oracle_agg_create.restype = C.c_void_p
        _LIB.oracle_agg_num_groups.restype = C.c_int64
        _LIB.oracle_hash_join.restype = C.c_int64
        _LIB.oracle_hash_join_run.restype = C.c_void_p
        _LIB.oracle_join_size.restype = C.c_int64
        _LIB.oracle_gen_customer.restype = C.c_int64
        _LIB.oracle_nation_name.restype = C.c_char_p
        _LIB.oracle_region_name.restype = C.c_char_p
        _LIB.oracle_nation_regionkey.restype = C.c_int
        _LIB.oracle_tpch_num_rows.restype = C.c_int64
        _LIB.oracle_q6.restype = C.c_double
        _LIB.oracle_q1.restype = C.c_int
    return _LIB


def _ptr(a):
    return C.c_void_p(a.ctypes.data) if a is not None else C.c_void_p(0)


class Utf8:
    """Arrow Utf8 column on the host: int32 offsets[n+1] + bytes."""

    def __init__(self, offsets, chars):
        self.offsets = np.ascontiguousarray(offsets, dtype=np.int32)
        self.chars = np.ascontiguousarray(chars, dtype=np.uint8)

    def __len__(self):
        return len(self.offsets) - 1

    @staticmethod
    def from_list(values):
        enc = [v.encode() for v in values]
        off = np.zeros(len(enc) + 1, dtype=np.int32)
        off[1:] = np.cumsum([len(e) for e in enc])
        return Utf8(off, np.frombuffer(b"".join(enc), dtype=np.uint8))

    def to_list(self):
        b = self.chars.tobytes()
        return [b[self.offsets[i]:self.offsets[i + 1]].decode() for i in range(len(self))]

    def take(self, idx):
        vals = self.to_list()
        return Utf8.from_list([vals[i] for i in idx])

    def equals_literal(self, s):
        lit = np.frombuffer(s.encode(), dtype=np.uint8)
        m = np.diff(self.offsets) == len(lit)
        if len(lit) and m.any():
            start = self.offsets[:-1][m].astype(np.int64)
            m[m] = (self.chars[start[:, None] + np.arange(len(lit))] == lit).all(axis=1)
        return m


def _keycols(cols, keep):
    arr = (KeyCol * max(len(cols), 1))()
    for k, c in enumerate(cols):
        if isinstance(c, Utf8):
            arr[k].dtype = T_UTF8
            arr[k].data = c.offsets.ctypes.data
            arr[k].chars = c.chars.ctypes.data
            keep.append(c)
        else:
            c = np.ascontiguousarray(c)
            keep.append(c)
            arr[k].dtype = {np.dtype(np.int32): T_INT32, np.dtype(np.int64): T_INT64, np.dtype(np.uint64): T_UINT64,
                            np.dtype(np.float64): T_FLOAT64}[c.dtype]
            arr[k].data = c.ctypes.data
    return arr


def take(col, idx):
    return col.take(idx) if isinstance(col, Utf8) else np.asarray(col)[idx]


def filter_columns(cols, mask):
    """A.1 FilterExec: compaction of every column, order kept."""
    idx = np.nonzero(mask)[0]
    return [take(c, idx) for c in cols]


def hash_aggregate(key_cols, val_cols, cnt_cols=None, batch_size=BATCH, partitions=None):
    """A.2: grouped SUM/COUNT state.  key_cols: list of numpy / Utf8 columns; val_cols: float64 columns summed;
    cnt_cols: None = count rows, else uint64 columns summed (Final merge of count states).
    `partitions`: optional list of (begin, end) row ranges fed in order (one Partial task each would be merged
    identically, as state addition is associative up to f64 rounding).
    Returns (rep_rows, sums[g, n_vals], counts[g, n_cnts or 1]); rep_rows index a representative input row."""
    L = lib()
    n = len(key_cols[0]) if key_cols else (len(val_cols[0]) if val_cols else len(cnt_cols[0]))
    keep = []
    keys = _keycols(key_cols, keep)
    vals = [np.ascontiguousarray(v, dtype=np.float64) for v in val_cols]
    cnts = [np.ascontiguousarray(c, dtype=np.uint64) for c in (cnt_cols or [])]
    vptr = (C.c_void_p * max(len(vals), 1))(*[v.ctypes.data for v in vals])
    cptr = (C.c_void_p * max(len(cnts), 1))(*[c.ctypes.data for c in cnts])
    h = C.c_void_p(L.oracle_agg_create(len(vals), len(cnts)))
    L.oracle_agg_update(h, C.c_int64(n), C.c_int64(batch_size), C.c_int64(0), len(key_cols), keys, vptr, cptr)
    g = L.oracle_agg_num_groups(h)
    rep = np.zeros(g, dtype=np.int64)
    sums = np.zeros((g, len(vals)), dtype=np.float64)
    counts = np.zeros((g, max(len(cnts), 1)), dtype=np.uint64)
    L.oracle_agg_result(h, _ptr(rep), _ptr(sums), _ptr(counts))
    L.oracle_agg_destroy(h)
    return rep, sums, counts


def hash_join(left_keys, right_keys):
    """A.3 inner join: (left_idx, right_idx) pairs in probe order."""
    L = lib()
    keep = []
    nl, nr = len(left_keys[0]), len(right_keys[0])
    lk, rk = _keycols(left_keys, keep), _keycols(right_keys, keep)
    h = C.c_void_p(L.oracle_hash_join_run(C.c_int64(nl), len(left_keys), lk, C.c_int64(nr), rk))
    m = L.oracle_join_size(h)
    li, ri = np.zeros(m, dtype=np.int64), np.zeros(m, dtype=np.int64)
    L.oracle_join_fetch(h, _ptr(li), _ptr(ri))
    return li, ri


def hash_join_outer(left_keys, right_keys, how, left_valid=None, right_valid=None):
    """A.3 Inner / Left / Right join over keys that may be NULL (`*_valid`: boolean array, False = some key part NULL).
    Null keys never match.  Returns (left_idx, right_idx) with -1 marking the null-extended side: the matched pairs in
    probe order, then (Left) the build rows nothing matched, or (Right) each probe row without a match."""
    nl, nr = len(left_keys[0]), len(right_keys[0])
    lv = np.ones(nl, bool) if left_valid is None else np.asarray(left_valid, bool)
    rv = np.ones(nr, bool) if right_valid is None else np.asarray(right_valid, bool)
    lrows, rrows = np.nonzero(lv)[0], np.nonzero(rv)[0]
    li, ri = hash_join([take(k, lrows) for k in left_keys], [take(k, rrows) for k in right_keys])
    li, ri = lrows[li], rrows[ri]
    if how == "Left":
        un = np.setdiff1d(np.arange(nl), li)
        li, ri = np.concatenate([li, un]), np.concatenate([ri, np.full(len(un), -1)])
    elif how == "Right":
        un = np.setdiff1d(np.arange(nr), ri)
        li, ri = np.concatenate([li, np.full(len(un), -1)]), np.concatenate([ri, un])
    elif how != "Inner":
        raise ValueError(how)
    return li.astype(np.int64), ri.astype(np.int64)


def aggregate_skip_nulls(keys, values, valid):
    """A.2 with NULL inputs: per group (dict key -> (count_rows, count_valid, sum or None)): SUM / AVG / COUNT(x) skip
    NULLs, SUM of nothing is NULL.  Sequential left-to-right f64 sums, as DataFusion's accumulators."""
    out = {}
    for k, v, ok in zip(keys, values, valid):
        rows, cnt, tot = out.get(k, (0, 0, None))
        if ok:
            cnt, tot = cnt + 1, (v if tot is None else tot + v)
        out[k] = (rows + 1, cnt, tot)
    return out


def hash_partition(key_cols, n_out):
    """A.4: destination partition of every row (equal keys co-locate; which partition is unobservable)."""
    L = lib()
    keep = []
    n = len(key_cols[0])
    dest = np.zeros(n, dtype=np.int32)
    L.oracle_hash_partition(C.c_int64(n), len(key_cols), _keycols(key_cols, keep), n_out, _ptr(dest))
    return dest


# ---- workload on the host ----
def tpch_num_rows(table, sf):
    return lib().oracle_tpch_num_rows(table.encode(), C.c_double(sf))


def gen_lineitem(sf, r0, n, threads=1):
    cols = {
        "l_orderkey": np.zeros(n, np.int32), "l_suppkey": np.zeros(n, np.int32), "l_quantity": np.zeros(n),
        "l_extendedprice": np.zeros(n), "l_discount": np.zeros(n), "l_tax": np.zeros(n),
        "l_shipdate": np.zeros(n, np.int32), "l_commitdate": np.zeros(n, np.int32),
        "l_receiptdate": np.zeros(n, np.int32),
    }
    rf = Utf8(np.zeros(n + 1, np.int32), np.zeros(n, np.uint8))
    ls = Utf8(np.zeros(n + 1, np.int32), np.zeros(n, np.uint8))
    lib().oracle_gen_lineitem(C.c_double(sf), C.c_int64(r0), C.c_int64(n), _ptr(cols["l_orderkey"]),
                              _ptr(cols["l_suppkey"]), _ptr(cols["l_quantity"]), _ptr(cols["l_extendedprice"]),
                              _ptr(cols["l_discount"]), _ptr(cols["l_tax"]), _ptr(cols["l_shipdate"]),
                              _ptr(cols["l_commitdate"]), _ptr(cols["l_receiptdate"]), _ptr(rf.offsets),
                              _ptr(rf.chars), _ptr(ls.offsets), _ptr(ls.chars), threads)
    cols["l_returnflag"], cols["l_linestatus"] = rf, ls
    return cols


def gen_orders(sf, r0, n):
    cols = {k: np.zeros(n, np.int32) for k in ("o_orderkey", "o_custkey", "o_orderdate", "o_shippriority")}
    lib().oracle_gen_orders(C.c_double(sf), C.c_int64(r0), C.c_int64(n), _ptr(cols["o_orderkey"]),
                            _ptr(cols["o_custkey"]), _ptr(cols["o_orderdate"]), _ptr(cols["o_shippriority"]))
    return cols


def gen_customer(sf, r0, n):
    ck, nk = np.zeros(n, np.int32), np.zeros(n, np.int32)
    off, chr_ = np.zeros(n + 1, np.int32), np.zeros(10 * n + 1, np.uint8)
    total = lib().oracle_gen_customer(C.c_double(sf), C.c_int64(r0), C.c_int64(n), _ptr(ck), _ptr(nk), _ptr(off),
                                      _ptr(chr_))
    return {"c_custkey": ck, "c_nationkey": nk, "c_mktsegment": Utf8(off, chr_[:total])}


def gen_supplier(sf, r0, n):
    sk, nk = np.zeros(n, np.int32), np.zeros(n, np.int32)
    lib().oracle_gen_supplier(C.c_double(sf), C.c_int64(r0), C.c_int64(n), _ptr(sk), _ptr(nk))
    return {"s_suppkey": sk, "s_nationkey": nk}


def gen_nation():
    L = lib()
    return {"n_nationkey": np.arange(25, dtype=np.int32),
            "n_name": Utf8.from_list([L.oracle_nation_name(i).decode() for i in range(25)]),
            "n_regionkey": np.array([L.oracle_nation_regionkey(i) for i in range(25)], dtype=np.int32)}


def gen_region():
    L = lib()
    return {"r_regionkey": np.arange(5, dtype=np.int32),
            "r_name": Utf8.from_list([L.oracle_region_name(i).decode() for i in range(5)])}


def to_arrow(cols, names=None):
    """host columns -> pyarrow RecordBatch (zero-copy for numerics)."""
    import pyarrow as pa
    names = names or list(cols)
    arrays = []
    for nm in names:
        c = cols[nm]
        if isinstance(c, Utf8):
            arrays.append(pa.Array.from_buffers(pa.utf8(), len(c), [None, pa.py_buffer(c.offsets), pa.py_buffer(c.chars)]))
        elif "date" in nm:
            arrays.append(pa.Array.from_buffers(pa.date32(), len(c), [None, pa.py_buffer(np.ascontiguousarray(c))]))
        else:
            arrays.append(pa.array(c))
    return pa.RecordBatch.from_arrays(arrays, names=names)


# ---- whole-query pipelines (multi-threaded: one task per partition), used as the CPU baseline ----
DATE_1998_09_02, DATE_1994_01_01, DATE_1995_01_01, DATE_1995_03_15 = 10471, 8766, 9131, 9204


def q1(cols, n_partitions=1, date_lit=DATE_1998_09_02):
    """-> list of dict rows in the Q1 output schema."""
    n = len(cols["l_quantity"])
    rf, ls = cols["l_returnflag"], cols["l_linestatus"]
    out_rf, out_ls = C.create_string_buffer(64), C.create_string_buffer(64)
    out_f = np.zeros((64, 7))
    out_c = np.zeros(64, np.uint64)
    ng = lib().oracle_q1(C.c_int64(n), _ptr(cols["l_quantity"]), _ptr(cols["l_extendedprice"]),
                         _ptr(cols["l_discount"]), _ptr(cols["l_tax"]), _ptr(cols["l_shipdate"]), _ptr(rf.offsets),
                         _ptr(rf.chars), _ptr(ls.offsets), _ptr(ls.chars), C.c_int32(date_lit), n_partitions, 64,
                         out_rf, out_ls, _ptr(out_f), _ptr(out_c))
    names = ["sum_qty", "sum_base_price", "sum_disc_price", "sum_charge", "avg_qty", "avg_price", "avg_disc"]
    rows = []
    for g in range(ng):
        r = {"l_returnflag": out_rf.raw[g:g + 1].decode(), "l_linestatus": out_ls.raw[g:g + 1].decode()}
        r.update({nm: float(out_f[g, k]) for k, nm in enumerate(names)})
        r["count_order"] = int(out_c[g])
        rows.append(r)
    return sorted(rows, key=lambda r: (r["l_returnflag"], r["l_linestatus"]))


def q6(cols, n_partitions=1, date_lo=DATE_1994_01_01, date_hi=DATE_1995_01_01, quantity_lt=24.0):
    n = len(cols["l_quantity"])
    passing = C.c_uint64(0)
    lo, hi = 0.06 - 0.01, 0.06 + 0.01  # folded in IEEE f64 exactly as the reference does at run time
    rev = lib().oracle_q6(C.c_int64(n), _ptr(cols["l_shipdate"]), _ptr(cols["l_discount"]), _ptr(cols["l_quantity"]),
                          _ptr(cols["l_extendedprice"]), C.c_int32(date_lo), C.c_int32(date_hi),
                          C.c_double(lo), C.c_double(hi), C.c_double(quantity_lt), n_partitions, C.byref(passing))
    # an ungrouped SUM over no rows is NULL, not 0.0 (the accumulator starts as Float64(None))
    return (float(rev) if passing.value else None), int(passing.value)


# ---- TPC-H Q3 (rust/benchmarks/tpch/queries/q3.sql) as the reference's plan runs it: filters below the joins,
#      (customer |x| orders) |x| lineitem with build = left, group by (l_orderkey, o_orderdate, o_shippriority) ----
def q3_range(sf, o_lo, o_hi, building_custkeys, cutoff=DATE_1995_03_15):
    """Q3 over the orders rows [o_lo, o_hi) (multiples of 8) and their lineitem rows [4 o_lo, 4 o_hi); every order's
    lines lie inside its range, so ranges yield disjoint groups.
    -> (l_orderkey, o_orderdate, o_shippriority, revenue, join statistics)"""
    o = gen_orders(sf, o_lo, o_hi - o_lo)
    l = gen_lineitem(sf, 4 * o_lo, 4 * (o_hi - o_lo))
    om = o["o_orderdate"] < cutoff
    of = {k: v[om] for k, v in o.items()}
    _, ri = hash_join([building_custkeys], [of["o_custkey"]])
    j1 = {k: v[ri] for k, v in of.items()}
    lm = l["l_shipdate"] > cutoff
    lf = {k: l[k][lm] for k in ("l_orderkey", "l_extendedprice", "l_discount")}
    # build rows / probes of the two joins (SURVEY 8d counts one 32-byte sector of table traffic for each)
    stats = {"probe1": int(om.sum()), "build2": int(len(ri)), "probe2": int(lm.sum())}
    empty = (np.zeros(0, np.int32),) * 3 + (np.zeros(0), stats)
    if len(ri) == 0 or not lm.any():
        return empty
    li2, ri2 = hash_join([j1["o_orderkey"]], [lf["l_orderkey"]])
    stats["joined"] = int(len(ri2))
    if len(ri2) == 0:
        return empty
    rev = lf["l_extendedprice"][ri2] * (1.0 - lf["l_discount"][ri2])
    keys = [lf["l_orderkey"][ri2], j1["o_orderdate"][li2], j1["o_shippriority"][li2]]
    rep, sums, _ = hash_aggregate(keys, [rev])
    return keys[0][rep], keys[1][rep], keys[2][rep], sums[:, 0].copy(), stats


def q3(sf, segment="BUILDING", cutoff=DATE_1995_03_15, threads=1, range_orders=1 << 21):
    """The whole query, one task per orders range on `threads` host threads (the C++ primitives release the GIL).
    -> dict of arrays sorted by l_orderkey."""
    from concurrent.futures import ThreadPoolExecutor
    nc, no = tpch_num_rows("customer", sf), tpch_num_rows("orders", sf)
    c = gen_customer(sf, 0, nc)
    ck = np.ascontiguousarray(c["c_custkey"][c["c_mktsegment"].equals_literal(segment)])
    step = max(8, min(range_orders, -(-no // max(threads, 1))) // 8 * 8)
    ranges = [(a, min(a + step, no)) for a in range(0, no, step)]
    with ThreadPoolExecutor(max_workers=max(threads, 1)) as ex:
        parts = list(ex.map(lambda r: q3_range(sf, r[0], r[1], ck, cutoff), ranges))
    out = [np.concatenate([p[k] for p in parts]) for k in range(4)]
    order = np.argsort(out[0], kind="stable")
    stats = {"build1": int(len(ck))}
    for p in parts:
        for k, v in p[4].items():
            stats[k] = stats.get(k, 0) + v
    return {"l_orderkey": out[0][order], "o_orderdate": out[1][order], "o_shippriority": out[2][order],
            "revenue": out[3][order], "stats": stats}


# ---- TPC-H Q5 (rust/benchmarks/tpch/queries/q5.sql): customer |x| orders |x| lineitem |x| supplier |x| nation |x|
#      region, left-deep in FROM order with build = left, the two-column key l_suppkey = s_suppkey AND
#      c_nationkey = s_nationkey, r_name = 'ASIA', o_orderdate in [1994-01-01, 1995-01-01), group by n_name ----
def q5_range(sf, o_lo, o_hi, cust, supp, nation, region_keys, date_lo=DATE_1994_01_01, date_hi=DATE_1995_01_01):
    """Q5 over the orders rows [o_lo, o_hi) and their lineitem rows; -> {n_name: revenue} of this range"""
    o = gen_orders(sf, o_lo, o_hi - o_lo)
    l = gen_lineitem(sf, 4 * o_lo, 4 * (o_hi - o_lo))
    om = (o["o_orderdate"] >= date_lo) & (o["o_orderdate"] < date_hi)
    of = {k: v[om] for k, v in o.items()}
    if not om.any():
        return {}
    li1, ri1 = hash_join([cust["c_custkey"]], [of["o_custkey"]])                      # customer |x| orders
    j1 = {"c_nationkey": cust["c_nationkey"][li1], "o_orderkey": of["o_orderkey"][ri1]}
    li2, ri2 = hash_join([j1["o_orderkey"]], [l["l_orderkey"]])                        # ... |x| lineitem
    j2 = {"c_nationkey": j1["c_nationkey"][li2], "l_suppkey": l["l_suppkey"][ri2],
          "rev": l["l_extendedprice"][ri2] * (1.0 - l["l_discount"][ri2])}
    if len(ri2) == 0:
        return {}
    li3, ri3 = hash_join([j2["l_suppkey"], j2["c_nationkey"]], [supp["s_suppkey"], supp["s_nationkey"]])  # ... |x| supplier
    j3 = {"s_nationkey": supp["s_nationkey"][ri3], "rev": j2["rev"][li3]}
    if len(ri3) == 0:
        return {}
    li4, ri4 = hash_join([j3["s_nationkey"]], [nation["n_nationkey"]])                 # ... |x| nation
    j4 = {"n_regionkey": nation["n_regionkey"][ri4], "n_name_id": ri4, "rev": j3["rev"][li4]}
    li5, ri5 = hash_join([j4["n_regionkey"]], [region_keys])                           # ... |x| region (r_name = 'ASIA')
    if len(li5) == 0:
        return {}
    names = nation["n_name"].to_list()
    ids, rev = j4["n_name_id"][li5], j4["rev"][li5]
    rep, sums, _ = hash_aggregate([ids.astype(np.int32)], [rev])
    return {names[int(ids[r])]: float(sums[g, 0]) for g, r in enumerate(rep)}


def q5(sf, region="ASIA", threads=1, range_orders=1 << 21, orders_range=None):
    """-> {n_name: revenue}; one task per orders range on `threads` host threads.  `orders_range` = (lo, hi) restricts
    the query to those orders rows and their lineitem rows (one rank's partition; partial sums add up over ranks)."""
    from concurrent.futures import ThreadPoolExecutor
    nc, no, ns = tpch_num_rows("customer", sf), tpch_num_rows("orders", sf), tpch_num_rows("supplier", sf)
    cust, supp, nation, reg = gen_customer(sf, 0, nc), gen_supplier(sf, 0, ns), gen_nation(), gen_region()
    rk = np.ascontiguousarray(reg["r_regionkey"][reg["r_name"].equals_literal(region)])
    o_lo, o_hi = orders_range if orders_range is not None else (0, no)
    step = max(8, min(range_orders, -(-(o_hi - o_lo) // max(threads, 1))) // 8 * 8)
    ranges = [(a, min(a + step, o_hi)) for a in range(o_lo, o_hi, step)]
    with ThreadPoolExecutor(max_workers=max(threads, 1)) as ex:
        parts = list(ex.map(lambda r: q5_range(sf, r[0], r[1], cust, supp, nation, rk), ranges))
    out = {}
    for p in parts:
        for k, v in p.items():
            out[k] = out.get(k, 0.0) + v
    return out
