#!/usr/bin/env python
"""bench.py — TPC-H Q1 (scan -> filter -> 4-group partial hash aggregate -> merge -> final) throughput.

    python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path (one process per GPU)
    python bench.py --impl reference --gpus N --steps K ...   # the reference's CPU algorithm on the host cores

Metric (BASELINE.json): input lineitem rows/s of TPC-H Q1 at SF100.  One "step" = one pass of the whole hot
path over the rank's lineitem partition: HashAggregateExec(Partial) with the FilterExec predicate fused into
its scan (ONE kernel), all-gather of the <=4 partial-state rows over NCCL when N > 1 (replaces the MergeExec
stage hop), HashAggregateExec(Final), and the device->host read of the 4 result rows.

  value    : rows of all ranks / max-over-ranks step time, inputs resident in HBM
  e2e      : same plan through the C ABI fed with HOST (pinned) Arrow RecordBatches, H2D inside the timed region
  roofline : the fused scan kernel (agg_tma_kernel<Q1>) — 46 algorithmic bytes per row over its
             mean launch duration, CUDA events on the launching stream, vs the measured HBM copy bandwidth
  cpu_baseline : oracle/ (C++ restatement of the DataFusion plan, one thread per partition) on a bounded sample;
             cpu_baseline_b = pyarrow Acero on the same sample (a production CPU engine as a sanity anchor)
  parity   : the result of the timed plan over ALL its rows against the oracle over the same rows (counts bit-exact,
             Float64 sums <= 1e-9 relative); every other config carries its own verdict; a mismatch fails the run

Weak scaling: every rank holds one SF100-sized partition (`--sf` x 6M rows) of an SF(100*N) lineitem;
`--scaling strong` instead splits one SF100 table across the N ranks (BASELINE.json config 3, literally).
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "tpch_q1_sf100_input_rows_per_s"
UNIT = "rows/s"
Q1_BYTES_PER_ROW = 46  # 4 x f64 + date32 + 2 x (int32 offset + 1 char)


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--sf", type=float, default=100.0, help="scale factor of the per-GPU partition (weak) / table (strong)")
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"])
    ap.add_argument("--e2e-batch-rows", type=int, default=1 << 22, help="rows per host RecordBatch in the e2e leg")
    ap.add_argument("--e2e-steps", type=int, default=3)
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-other", action="store_true", help="skip the other BASELINE configs (Q6 / Q3 at SF10, shuffle)")
    ap.add_argument("--other-sf", type=float, default=10.0)
    ap.add_argument("--q3-sf", type=float, default=100.0, help="Q3 at the metric's scale: one GPU's own copy at N=1, "
                    "row-range partitions of one database of this scale (strong scaling) at N>1")
    ap.add_argument("--cpu-seconds", type=float, default=3.0, help="wall-clock target of the cpu_baseline sample")
    ap.add_argument("--no-parity", action="store_true", help="skip the full-size oracle comparisons (untimed)")
    ap.add_argument("--parity-chunk-rows", type=int, default=1 << 26)
    return ap.parse_args()


# ---------------------------------------------------------------------------------------------------------
# clocks: sample nvidia-smi while the timed region runs
# ---------------------------------------------------------------------------------------------------------
class ClockSampler(threading.Thread):
    """SM clock + throttle reasons DURING the timed region: NVML every 5 ms (nvidia-smi as a fallback)."""
    BAD = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown")

    def __init__(self, gpu_index):
        super().__init__(daemon=True)
        self.idx, self.stop_flag, self.sm, self.reasons, self.sm_max = gpu_index, False, [], set(), None
        self.nvml = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nvml = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(self._physical_index(gpu_index))
            self.sm_max = float(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
        except Exception:
            self.nvml = None

    @staticmethod
    def _physical_index(i):
        vis = os.environ.get("CUDA_VISIBLE_DEVICES")
        if vis:
            try:
                return int(vis.split(",")[i])
            except Exception:
                pass
        return i

    def _smi(self, fields):
        out = subprocess.run(["nvidia-smi", f"--id={self.idx}", f"--query-gpu={fields}", "--format=csv,noheader,nounits"],
                             capture_output=True, text=True, timeout=10).stdout.strip()
        return [x.strip() for x in out.split(",")]

    def run(self):
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        if self.nvml is not None:
            n = self.nvml
            bits = {"hw_slowdown": n.nvmlClocksThrottleReasonHwSlowdown,
                    "hw_thermal_slowdown": n.nvmlClocksThrottleReasonHwThermalSlowdown,
                    "sw_thermal_slowdown": n.nvmlClocksThrottleReasonSwThermalSlowdown,
                    "sw_power_cap": n.nvmlClocksThrottleReasonSwPowerCap}
            while not self.stop_flag:
                try:
                    self.sm.append(float(n.nvmlDeviceGetClockInfo(self.h, n.NVML_CLOCK_SM)))
                    r = n.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                    for k, bit in bits.items():
                        if r & bit:
                            self.reasons.add(k)
                except Exception:
                    pass
                time.sleep(0.005)
            return
        fields = "clocks.sm," + ",".join("clocks_event_reasons." + x for x in names)
        try:
            self.sm_max = float(self._smi("clocks.max.sm")[0])
        except Exception:
            pass
        while not self.stop_flag:
            try:
                v = self._smi(fields)
                self.sm.append(float(v[0]))
                for k, flag in zip(names, v[1:]):
                    if flag.lower().startswith("active"):
                        self.reasons.add(k)
            except Exception:
                pass
            time.sleep(0.05)

    def result(self):
        self.stop_flag = True
        self.join(timeout=15)
        return {"sm_mhz": statistics.median(self.sm) if self.sm else None, "sm_max_mhz": self.sm_max,
                "reasons": sorted(self.reasons), "samples": len(self.sm),
                "source": "nvml" if self.nvml is not None else "nvidia-smi"}


def measured_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return json.load(f)
    except Exception:
        return {}


# ---------------------------------------------------------------------------------------------------------
# host-side helpers
# ---------------------------------------------------------------------------------------------------------
def env_rank():
    from ballista_b200 import cluster
    return cluster.env_rank()


def mem_available_bytes():
    try:
        with open("/proc/meminfo") as f:
            for line in f:
                if line.startswith("MemAvailable:"):
                    return int(line.split()[1]) * 1024
    except Exception:
        pass
    return 0


def pin_record_batch(torch, rb):
    """cudaHostRegister every buffer of a host RecordBatch so H2D copies run at PCIe speed."""
    rt = torch.cuda.cudart()
    pinned = []
    for col in rb.columns:
        for buf in col.buffers():
            if buf is not None and buf.size > 0:
                err = rt.cudaHostRegister(buf.address, buf.size, 0)
                if int(err) == 0:
                    pinned.append(buf.address)
    return pinned


def unpin(torch, addrs):
    rt = torch.cuda.cudart()
    for a in addrs:
        rt.cudaHostUnregister(a)




def check_q1_rows(rows, n_rows_total):
    """Cheap invariants of the Q1 result (the parity block compares every value with the oracle)."""
    assert 1 <= len(rows) <= 4, rows
    keys = sorted((r["l_returnflag"], r["l_linestatus"]) for r in rows)
    assert keys == sorted(set(keys)), keys
    total = sum(r["count_order"] for r in rows)
    assert 0 < total <= n_rows_total, (total, n_rows_total)
    for r in rows:
        assert abs(r["avg_qty"] - r["sum_qty"] / r["count_order"]) <= 1e-9 * abs(r["avg_qty"])
    return total


def workload_config(sf, scaling, world, rows_per_gpu, rows_total):
    """`config` of the JSON line: the SAME object in this repo's arm and in the reference arm of one run."""
    return {"workload": f"TPC-H Q1 over synthetic SF{sf:g} lineitem: filter l_shipdate <= 1998-09-02, group by "
                        "(l_returnflag, l_linestatus), 4 x SUM + 3 x AVG + COUNT (rust/benchmarks/tpch/queries/q1.sql)",
            "query": "q1", "sf": sf, "scaling": scaling, "n_gpus": world, "rows_per_gpu": int(rows_per_gpu),
            "rows_total": int(rows_total), "algorithmic_bytes_per_row": Q1_BYTES_PER_ROW,
            "partitioning": ("one SF%g partition per GPU (weak: an SF%g table over %d GPUs)" % (sf, sf * world, world))
            if scaling == "weak" else ("one SF%g table in %d row-range partitions" % (sf, world)),
            "l2": "inputs larger than L2 (no flush needed)" if rows_per_gpu * Q1_BYTES_PER_ROW > 512e6 else
                  "inputs smaller than L2"}


# ---------------------------------------------------------------------------------------------------------
# parity: the oracle over the SAME rows the timed plan read (untimed).  Rows are regenerated chunk by chunk by the
# library's counter-based generator, copied to the host and aggregated by oracle/; the first chunk is also compared
# bit for bit with the oracle's own host generator, so both sides of the comparison see one table.
# ---------------------------------------------------------------------------------------------------------
def arrow_to_oracle_cols(O, rb):
    import numpy as np
    import pyarrow as pa
    cols = {}
    for name, col in zip(rb.schema.names, rb.columns):
        n = len(col)
        bufs = col.buffers()
        if pa.types.is_string(col.type):
            cols[name] = O.Utf8(np.frombuffer(bufs[1], dtype=np.int32, count=n + 1 + col.offset)[col.offset:],
                                np.frombuffer(bufs[2], dtype=np.uint8) if bufs[2] is not None else np.zeros(0, np.uint8))
        elif pa.types.is_date32(col.type) or pa.types.is_int32(col.type):
            cols[name] = np.frombuffer(bufs[1], dtype=np.int32, count=n + col.offset)[col.offset:]
        else:
            cols[name] = col.to_numpy(zero_copy_only=True)
    return cols


def check_generator(O, sf, r0, cols, limit=1 << 20):
    """the first `limit` rows of a device-generated chunk == the oracle's host generator, bit for bit"""
    import numpy as np
    n = min(limit, len(next(iter(cols.values()))))
    want = O.gen_lineitem(sf, r0, n)
    for name, c in cols.items():
        w = want[name]
        if isinstance(c, O.Utf8):
            assert np.array_equal(c.offsets[:n + 1] - c.offsets[0], w.offsets) and \
                np.array_equal(c.chars[c.offsets[0]:c.offsets[n]], w.chars[:w.offsets[n]]), name
        else:
            assert np.array_equal(c[:n], w), name
    return n


def oracle_scan_chunks(ctx, O, table_sf, r0, r1, columns, chunk_rows, fn):
    """fn(oracle columns of one chunk) for every chunk of lineitem rows [r0, r1); -> rows the generators agreed on"""
    checked = 0
    for b in range(r0, r1, chunk_rows):
        e = min(b + chunk_rows, r1)
        db = ctx.tpch("lineitem", table_sf, b, e, columns=columns)
        rb = db.to_arrow()
        del db
        cols = arrow_to_oracle_cols(O, rb)
        if b == r0:
            checked = check_generator(O, table_sf, b, cols)
        fn(cols)
        del cols, rb
    return checked


def oracle_q1_partition(ctx, O, table_sf, r0, r1, threads, chunk_rows, columns, date_lit=None):
    """-> ({(flag, status): [sum_qty, sum_base, sum_disc_price, sum_charge, sum_disc, count]}, generator rows checked)"""
    acc = {}
    kw = {} if date_lit is None else {"date_lit": date_lit}

    def one(cols):
        for r in O.q1(cols, n_partitions=threads, **kw):
            a = acc.setdefault((r["l_returnflag"], r["l_linestatus"]), [0.0, 0.0, 0.0, 0.0, 0.0, 0])
            for k, nm in enumerate(("sum_qty", "sum_base_price", "sum_disc_price", "sum_charge")):
                a[k] += r[nm]
            a[4] += r["avg_disc"] * r["count_order"]
            a[5] += r["count_order"]
    checked = oracle_scan_chunks(ctx, O, table_sf, r0, r1, columns, chunk_rows, one)
    return acc, checked


def merge_q1_acc(accs):
    out = {}
    for acc in accs:
        for k, v in acc.items():
            a = out.setdefault(k, [0.0, 0.0, 0.0, 0.0, 0.0, 0])
            for i in range(6):
                a[i] += v[i]
    return out


def parity_q1(name, gpu_rows, acc, rows, checked):
    """GPU result rows vs the merged oracle accumulators: group keys and counts bit-exact, sums / averages <= 1e-9."""
    got = {(r["l_returnflag"], r["l_linestatus"]): r for r in gpu_rows}
    keys_equal = sorted(got) == sorted(acc)
    counts_equal = keys_equal and all(got[k]["count_order"] == acc[k][5] for k in acc)
    worst = 0.0
    if keys_equal:
        for k, a in acc.items():
            g, n = got[k], a[5]
            want = {"sum_qty": a[0], "sum_base_price": a[1], "sum_disc_price": a[2], "sum_charge": a[3],
                    "avg_qty": a[0] / n, "avg_price": a[1] / n, "avg_disc": a[4] / n}
            for nm, w in want.items():
                worst = max(worst, abs(g[nm] - w) / abs(w))
    ok = bool(keys_equal and counts_equal and worst <= 1e-9)
    return {"config": name, "rows": int(rows), "groups": len(got), "keys_equal": bool(keys_equal),
            "counts_equal": bool(counts_equal), "max_rel_err": worst, "tolerance": 1e-9, "ok": ok,
            "checker": "oracle/ Q1 over the same rows (chunks regenerated by the library's generator, D2H, aggregated on "
                       "the host)", "generator_rows_cross_checked": int(checked)}


def q3_table_arrays(table):
    """pa.Table of the Q3 Final -> (orderkey, orderdate days, shippriority, revenue) sorted by orderkey"""
    import numpy as np
    import pyarrow as pa
    ok = table["l_orderkey"].combine_chunks().to_numpy()
    od = table["o_orderdate"].combine_chunks().cast(pa.int32()).to_numpy()
    sp = table["o_shippriority"].combine_chunks().to_numpy()
    rv = table["revenue"].combine_chunks().to_numpy()
    o = np.argsort(ok, kind="stable")
    return ok[o], od[o], sp[o], rv[o]


def q3_checksum(arrs):
    """order-independent digest of a set of Q3 groups: (count, 64-bit key checksum, revenue sum)"""
    import math
    import numpy as np
    ok, od, sp, rv = arrs
    with np.errstate(over="ignore"):
        h = (ok.astype(np.uint64) * np.uint64(0x9E3779B97F4A7C15) + od.astype(np.uint64) * np.uint64(0xC2B2AE3D27D4EB4F)
             + sp.astype(np.uint64) * np.uint64(0x165667B19E3779F9))
        h ^= h >> np.uint64(29)
        return int(len(ok)), int(h.sum(dtype=np.uint64)), math.fsum(rv.tolist())


def parity_q3(name, arrs, want, rows):
    """full comparison with the oracle's Q3: group keys bit-exact, revenue <= 1e-9 relative"""
    import numpy as np
    ok, od, sp, rv = arrs
    keys_equal = (len(ok) == len(want["l_orderkey"]) and np.array_equal(ok, want["l_orderkey"]) and
                  np.array_equal(od, want["o_orderdate"]) and np.array_equal(sp, want["o_shippriority"]))
    worst = float(np.max(np.abs(rv - want["revenue"]) / np.abs(want["revenue"]))) if keys_equal and len(ok) else 0.0
    return {"config": name, "input_rows": int(rows), "groups": int(len(ok)), "keys_equal": bool(keys_equal),
            "counts_equal": bool(keys_equal), "max_rel_err": worst, "tolerance": 1e-9,
            "ok": bool(keys_equal and worst <= 1e-9),
            "checker": "oracle/ Q3 (filters, two collect-left hash joins, group-by) over the oracle's own host tables"}


# ---------------------------------------------------------------------------------------------------------
# this repo's arm
# ---------------------------------------------------------------------------------------------------------
class Env:
    pass


def setup_env(args):
    import torch
    import torch.distributed as dist
    import ballista_b200 as b2
    from ballista_b200 import cluster
    env = Env()
    env.rank, env.local_rank, env.world = env_rank()
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (ballista_b200 has no CPU fallback)")
    torch.cuda.set_device(env.local_rank)
    if env.world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", env.local_rank))
    env.torch, env.dist, env.cluster = torch, dist, cluster
    env.ctx = b2.Context(env.local_rank)
    cluster.attach_communicator(env.ctx, dist, env.rank, env.world)
    env.stream = torch.cuda.ExternalStream(env.ctx.cuda_stream)
    env.local_world = int(os.environ.get("LOCAL_WORLD_SIZE", env.world))
    env.host_threads = max(1, (os.cpu_count() or 1) // max(env.local_world, 1))
    return env


def barrier(env):
    if env.world > 1:
        env.dist.barrier()
    env.torch.cuda.synchronize()
    env.ctx.synchronize()


def time_steps(env, fn, steps, warmup, profile=False, sample_clocks=False):
    """W untimed + K timed calls of fn(), one CUDA event per step on the library's stream, barrier + synchronize on
    both sides.  -> dict(ms_total (max over ranks), per-step ms list (this rank), launches, kernel (ms, n), clocks, last)"""
    torch, ctx = env.torch, env.ctx
    last = None
    for _ in range(warmup):
        last = fn()
    barrier(env)
    sampler = None
    if sample_clocks:
        sampler = ClockSampler(env.local_rank)
        sampler.start()
    if profile:
        ctx.set_profiling(True)
        ctx.profile()
    l0 = ctx.kernel_launches()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(steps + 1)]
    ev[0].record(env.stream)
    for k in range(steps):
        last = fn()
        ev[k + 1].record(env.stream)
    barrier(env)
    ms = ev[0].elapsed_time(ev[steps])
    per_step = [ev[k].elapsed_time(ev[k + 1]) for k in range(steps)]
    launches = ctx.kernel_launches() - l0
    kern = ctx.profile() if profile else (0.0, 0)
    if profile:
        ctx.set_profiling(False)
    clocks = sampler.result() if sampler else None
    return {"ms_total": env.cluster.max_over_ranks(env.dist, torch, ms, env.world), "per_step": per_step,
            "launches": launches, "kernel": kern, "clocks": clocks, "last": last}


def step_stats(per_step):
    return {"median_ms": statistics.median(per_step), "best_ms": min(per_step), "worst_ms": max(per_step)}


def gather_objects(env, obj):
    if env.world == 1:
        return [obj]
    out = [None] * env.world
    env.dist.all_gather_object(out, obj)
    return out


def run_b200(args, env):
    from ballista_b200 import tpch, _capi
    ctx, rank, world, cluster = env.ctx, env.rank, env.world, env.cluster
    # ---- workload: this rank's lineitem partition, generated in HBM (setup, untimed) ----
    if args.scaling == "weak":
        table_sf = args.sf * world
        rows_local = _capi.lib().b2_tpch_num_rows(b"lineitem", float(args.sf))
        r0 = rank * rows_local
        rows_total = rows_local * world
    else:
        table_sf = args.sf
        rows_total = _capi.lib().b2_tpch_num_rows(b"lineitem", float(args.sf))
        r0, r1 = cluster.partition_rows(rows_total, world, rank, align=32)
        rows_local = r1 - r0
    rows_gpu0 = rows_local if args.scaling == "weak" else cluster.partition_rows(rows_total, world, 0, align=32)[1]
    max_batch = 1 << 30  # batches of at most 2^30 rows (Utf8 offsets are int32)
    dev_batches = [ctx.tpch("lineitem", table_sf, b, min(b + max_batch, r0 + rows_local), columns=tpch.Q1_COLUMNS)
                   for b in range(r0, r0 + rows_local, max_batch)]
    ctx.synchronize()

    def make_step(source_batches):
        # the two stage plans are built once (an executor keeps the decoded plan of a stage); a step executes them
        scan = tpch.scan(ctx, [source_batches])
        partial = tpch.q1_partial(scan)
        query = tpch.TwoStage(ctx, partial, lambda s: tpch.q1_final(s, scan.schema()), merge="allgather")

        def step():
            return query.run(), partial
        return step

    # ---- value: inputs resident in HBM ----
    t = time_steps(env, make_step(dev_batches), args.steps, args.warmup, profile=True, sample_clocks=True)
    table, partial = t["last"]
    rows = table.to_pylist()
    passed = check_q1_rows(rows, rows_total)
    assert partial.last_path().startswith("fused:") and "[tma" in partial.last_path(), partial.last_path()
    ms, (kern_ms, kern_n) = t["ms_total"], t["kernel"]
    value = rows_total * args.steps / (ms / 1e3)
    peaks = measured_peaks()
    peak = peaks.get("hbm_gbs")
    peak_src = "measured (MEASURED_PEAKS.json hbm_gbs, copy read+write, burst: the kernel is timed alone by its own events)"
    if not peak:
        peak, peak_src = 6650.0, "fallback (B200_PROFILING.md)"
    rows_per_launch = rows_local * args.steps / max(kern_n, 1)
    achieved = Q1_BYTES_PER_ROW * rows_per_launch / (kern_ms / max(kern_n, 1) / 1e3) / 1e9 if kern_n else None
    roofline = {"bound": "hbm", "kernel": "agg_tma_kernel<FusedSrc<Q1>, narrow> (TMA-fed scan + filter + expressions + partial aggregate)",
                "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": (achieved / peak) if achieved else None,
                "traffic": None, "peak_source": peak_src, "launches": kern_n,
                "avg_launch_ms": kern_ms / kern_n if kern_n else None,
                "algorithmic_bytes_per_row": Q1_BYTES_PER_ROW, "rows_per_launch": rows_per_launch,
                "kernel_share_of_step": kern_ms / ms if ms else None,
                "step_frac_of_peak": Q1_BYTES_PER_ROW * rows_local * args.steps / (ms / 1e3) / 1e9 / peak}
    traffic_file = os.path.join(ROOT, "profiles", "q1_traffic.json")
    if os.path.exists(traffic_file):
        try:
            with open(traffic_file) as f:
                tj = json.load(f)
            if tj.get("rows_per_launch") == rows_per_launch:
                roofline["traffic"] = tj.get("dram_bytes_per_launch")
                roofline["traffic_source"] = "recorded, not measured in this run: " + tj.get("source", "profiles/q1_traffic.json")
        except Exception:
            pass
    out = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
           "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None,
           "dtype": "f64", "data": "synthetic (counter-based TPC-H-shaped generator, seed 42, generated in HBM)",
           "config": workload_config(args.sf, args.scaling, world, rows_gpu0, rows_total),
           "gpu_launches": t["launches"], "clocks": t["clocks"], "roofline": roofline,
           "GB_per_s": value * Q1_BYTES_PER_ROW / 1e9, "rows_passing_filter": passed,
           "step_ms": dict(step_stats(t["per_step"]), mean_ms=ms / args.steps,
                           note="per-step CUDA-event times of rank 0; value uses the max-over-ranks total"),
           "plan_path": partial.last_path()}
    # ---- parity at full size: every rank runs the oracle over ITS rows, rank 0 merges and compares ----
    if not args.no_parity:
        from oracle import oracle_py as O
        O.lib()
        acc, checked = oracle_q1_partition(ctx, O, table_sf, r0, r0 + rows_local, env.host_threads, args.parity_chunk_rows,
                                           tpch.Q1_COLUMNS)
        accs = gather_objects(env, acc)
        out["parity"] = parity_q1(f"q1_sf{args.sf:g}_x{world}_{args.scaling}", rows, merge_q1_acc(accs), rows_total, checked)
    env.q1 = (dev_batches, make_step, table_sf, r0, rows_local)
    return out


def run_e2e(args, env):
    """Same plan fed with HOST RecordBatches (pinned): H2D of every input column inside the timed region."""
    dev_batches, make_step, _, _, rows_local = env.q1
    torch, world = env.torch, env.world
    need = rows_local * Q1_BYTES_PER_ROW
    budget = mem_available_bytes() // max(env.local_world, 1) // 3
    frac = 1.0
    if need > budget:
        frac = max(budget / need, 0.01)
    host_batches, pinned, rows_e2e = [], [], 0
    want_rows = int(rows_local * frac)
    for db in dev_batches:
        if rows_e2e >= want_rows:
            break
        rb = db.to_arrow()  # device -> host (setup)
        if rows_e2e + rb.num_rows > want_rows:
            rb = rb.slice(0, want_rows - rows_e2e)
        pinned += pin_record_batch(torch, rb)
        for o in range(0, rb.num_rows, args.e2e_batch_rows):
            host_batches.append(rb.slice(o, args.e2e_batch_rows))
        rows_e2e += rb.num_rows
    h2d = rows_e2e * Q1_BYTES_PER_ROW + 8 * len(host_batches)  # Utf8 offsets carry one extra int32 per batch
    # the ceiling of this leg: plain pinned host -> device copies issued by ALL ranks at the same time (same PCIe
    # links, same host memory), no kernels
    ceiling = None
    try:
        blk = torch.empty(256 << 20, dtype=torch.uint8).pin_memory()
        dst = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
        reps = 16
        for _ in range(2):
            dst.copy_(blk, non_blocking=True)
        barrier(env)
        c0, c1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        c0.record()
        for _ in range(reps):
            dst.copy_(blk, non_blocking=True)
        c1.record()
        barrier(env)
        ceiling = reps * blk.numel() / (env.cluster.max_over_ranks(env.dist, torch, c0.elapsed_time(c1), world) / 1e3) / 1e9
        del blk, dst
    except Exception:
        pass
    t = time_steps(env, make_step(host_batches), args.e2e_steps, 1)
    table, _ = t["last"]
    ms = t["ms_total"]
    d2h = sum(c.nbytes for c in table.columns)
    check_q1_rows(table.to_pylist(), rows_e2e * world)
    unpin(torch, pinned)
    return {"value": rows_e2e * world * args.e2e_steps / (ms / 1e3), "unit": UNIT, "h2d_bytes_per_step": int(h2d),
            "d2h_bytes_per_step": int(d2h), "rows_per_gpu": rows_e2e, "steps": args.e2e_steps,
            "ms_per_step": ms / args.e2e_steps, "host_batch_rows": args.e2e_batch_rows,
            "h2d_GB_per_s_per_gpu": h2d / (ms / args.e2e_steps / 1e3) / 1e9,
            "h2d_ceiling_GB_per_s_per_gpu": ceiling,
            "frac_of_h2d_ceiling": (h2d / (ms / args.e2e_steps / 1e3) / 1e9 / ceiling) if ceiling else None,
            "ceiling_note": "pinned 256 MiB host -> device copies issued by all ranks at once (CUDA events, max over ranks): "
                            "what the PCIe links and the host memory deliver with no kernel running",
            "note": "host Arrow RecordBatches in cudaHostRegister'ed memory pushed through b2_memory_exec_add_batch; "
                    "H2D of all 7 columns + result D2H inside the timed region"
                    + ("" if frac == 1.0 else f"; bounded to {frac:.2f} of the partition by host RAM")}


# ---------------------------------------------------------------------------------------------------------
# the reference's CPU algorithm (oracle/: C++ restatement of the DataFusion plan; the Rust reference cannot be
# built here — no cargo, DataFusion not vendored), and pyarrow Acero as a second, production-grade CPU engine
# ---------------------------------------------------------------------------------------------------------
def cpu_q1(sf, seconds, threads, steps=1, warmup=0, with_acero=False):
    """-> dict(rows/s of the oracle port, sample rows, s per pass, per-pass list, acero rows/s or None).  A bounded
    sample of the SF`sf` lineitem: the first `n` rows, n calibrated so one pass takes about `seconds` on `threads` threads."""
    from oracle import oracle_py as O
    O.lib()
    total = O.tpch_num_rows("lineitem", sf)
    probe = min(total, 250_000 * threads)
    cols = O.gen_lineitem(sf, 0, probe, threads=threads)
    t = time.perf_counter()
    O.q1(cols, n_partitions=threads)
    dt = time.perf_counter() - t
    n = int(min(total, max(probe, probe * seconds / max(dt, 1e-6))))
    budget = mem_available_bytes() // 4
    n = int(min(n, max(budget // 64, probe)))
    cols = O.gen_lineitem(sf, 0, n, threads=threads)
    for _ in range(warmup):
        O.q1(cols, n_partitions=threads)
    passes = []
    for _ in range(steps):
        t = time.perf_counter()
        rows = O.q1(cols, n_partitions=threads)
        passes.append(time.perf_counter() - t)
    check_q1_rows(rows, n)
    out = {"value": n * steps / sum(passes), "rows": n, "s_per_pass": sum(passes) / steps, "passes": passes, "acero": None}
    if with_acero:
        try:
            out["acero"] = acero_q1(O, cols, rows, n)
        except Exception as e:  # a sanity anchor, never fatal
            out["acero"] = {"error": repr(e)}
    return out


def acero_q1(O, cols, oracle_rows, n):
    """TPC-H Q1 with pyarrow's Acero engine (filter + computed columns + group_by().aggregate) on the same host sample;
    -> rows/s (best of 3) after checking its counts against the oracle's."""
    import pyarrow as pa
    import pyarrow.compute as pc
    t = pa.Table.from_batches([O.to_arrow(cols, tpch_q1_columns())])

    def run():
        f = t.filter(pc.less_equal(t["l_shipdate"], pa.scalar(O.DATE_1998_09_02, pa.int32()).cast(pa.date32())))
        one = pa.scalar(1.0)
        disc = pc.multiply(f["l_extendedprice"], pc.subtract(one, f["l_discount"]))
        f = f.append_column("disc_price", disc).append_column("charge", pc.multiply(disc, pc.add(one, f["l_tax"])))
        return f.group_by(["l_returnflag", "l_linestatus"]).aggregate(
            [("l_quantity", "sum"), ("l_extendedprice", "sum"), ("disc_price", "sum"), ("charge", "sum"),
             ("l_quantity", "mean"), ("l_extendedprice", "mean"), ("l_discount", "mean"), ("l_quantity", "count")])
    best, res = None, None
    for _ in range(3):
        t0 = time.perf_counter()
        res = run()
        dt = time.perf_counter() - t0
        best = dt if best is None else min(best, dt)
    got = {(r["l_returnflag"], r["l_linestatus"]): r["l_quantity_count"] for r in res.to_pylist()}
    want = {(r["l_returnflag"], r["l_linestatus"]): r["count_order"] for r in oracle_rows}
    assert got == want, (got, want)
    return {"value": n / best, "unit": UNIT, "cores": pa.cpu_count(), "kind": "pyarrow Acero " + pa.__version__,
            "sample": f"the same {n} rows, best of 3 passes ({best:.2f} s); group counts equal the oracle's"}


def tpch_q1_columns():
    return ["l_quantity", "l_extendedprice", "l_discount", "l_tax", "l_returnflag", "l_linestatus", "l_shipdate"]


def run_reference(args):
    rank, _, world = env_rank()
    if rank != 0:
        return None
    from oracle import oracle_py as O
    from ballista_b200 import cluster
    O.lib()
    threads = os.cpu_count() or 1
    # each step = one pass over a bounded sample; K + W passes must end within minutes
    per_step = min(20.0, 150.0 / max(args.steps + args.warmup, 1))
    r = cpu_q1(args.sf, per_step, threads, steps=args.steps, warmup=args.warmup)
    v, n, sec = r["value"], r["rows"], r["s_per_pass"]
    rows_sf = O.tpch_num_rows("lineitem", args.sf)
    if args.scaling == "weak":
        rows_gpu0, rows_total = rows_sf, rows_sf * world
    else:
        rows_gpu0, rows_total = cluster.partition_rows(rows_sf, world, 0, align=32)[1], rows_sf
    sample = (f"first {n} rows of the SF{args.sf:g} synthetic lineitem per step, {threads} partitions (one thread each); "
              "rows/s of the sample stands for the whole table (the plan is a streaming scan: cost is linear in rows)")
    return {"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": sec * 1e3, "higher_is_better": True, "scaling": args.scaling,
            "vs_baseline": None, "dtype": "f64", "data": "synthetic (same generator, host copy)",
            "config": workload_config(args.sf, args.scaling, world, rows_gpu0, rows_total),
            "cpu_baseline": {"value": v, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample},
            "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0, "step_ms": {"median_ms": statistics.median(r["passes"]) * 1e3, "best_ms": min(r["passes"]) * 1e3},
            "note": "oracle/ C++ restatement of DataFusion@46161d2's Q1 plan (32,768-row batches, FilterExec copy, "
                    "no-CSE temporaries, byte-key hash aggregate); the Rust reference itself cannot be built here"}


# ---------------------------------------------------------------------------------------------------------
# the other BASELINE.json configs: each timed the same way and each with its own parity verdict
# ---------------------------------------------------------------------------------------------------------
def oracle_q6_partition(ctx, O, table_sf, r0, r1, threads, chunk_rows, columns):
    tot = [0.0, 0]

    def one(cols):
        rev, passing = O.q6(cols, n_partitions=threads)
        tot[0] += rev or 0.0
        tot[1] += passing
    checked = oracle_scan_chunks(ctx, O, table_sf, r0, r1, columns, chunk_rows, one)
    return tot, checked


def q3_roofline(scan_bytes, stats, ms, peak):
    """scan-only GB/s and scan + table GB/s (one 32-byte sector per build insert and per probe)"""
    out = {"bound": "hbm", "unit": "GB/s", "peak": peak, "scan_bytes": int(scan_bytes),
           "scan_only": scan_bytes / (ms / 1e3) / 1e9, "frac_scan_only": scan_bytes / (ms / 1e3) / 1e9 / peak,
           "note": "whole-step figures (all kernels of the plan), not a single kernel's"}
    if stats:
        table_bytes = 32 * (stats["build1"] + stats["build2"] + stats["probe1"] + stats["probe2"])
        out.update({"table_bytes": int(table_bytes), "scan_plus_table": (scan_bytes + table_bytes) / (ms / 1e3) / 1e9,
                    "frac_scan_plus_table": (scan_bytes + table_bytes) / (ms / 1e3) / 1e9 / peak, "join_rows": stats})
    return out


def run_other_configs(args, env):
    from ballista_b200 import MemoryExec, RepartitionExec, _capi, col, tpch
    import numpy as np
    ctx, rank, world, cluster, dist, torch = env.ctx, env.rank, env.world, env.cluster, env.dist, env.torch
    sf = args.other_sf
    peak = measured_peaks().get("hbm_gbs") or 6650.0
    O = None
    if not args.no_parity:
        from oracle import oracle_py as O
        O.lib()
    out = {}
    nrows = lambda t, s=None: _capi.lib().b2_tpch_num_rows(t.encode(), float(sf if s is None else s))

    def timeit(fn, steps=10, warmup=3, profile=False):
        t = time_steps(env, fn, steps, warmup, profile=profile)
        return t["ms_total"] / steps, t["launches"] / steps, t["last"], t

    # ---- Q6 @ SF10: filter + global SUM over this rank's lineitem (28 B/row) ----
    n6 = nrows("lineitem")
    li6 = ctx.tpch("lineitem", sf * world, rank * n6, (rank + 1) * n6, columns=tpch.Q6_COLUMNS)

    scan6 = tpch.scan(ctx, [[li6]])
    part6 = tpch.q6_partial(scan6)
    query6 = tpch.TwoStage(ctx, part6, lambda s: tpch.q6_final(s, scan6.schema()), merge="allgather")

    def q6():
        return query6.run(), part6
    ms, launches, (tab, part), t = timeit(q6, steps=20, profile=True)
    k_ms, k_n = t["kernel"]
    n = li6.num_rows
    revenue = tab.to_pylist()[0]["revenue"]
    out["q6"] = {"workload": f"TPC-H Q6, SF{sf:g} lineitem per GPU ({n} rows x 28 B)", "ms_per_step": ms,
                 "step_ms": step_stats(t["per_step"]),
                 "rows_per_s": n * world / (ms / 1e3), "GB_per_s_per_gpu": n * 28 / (ms / 1e3) / 1e9,
                 "frac_of_hbm_peak": n * 28 / (ms / 1e3) / 1e9 / peak, "launches_per_step": launches,
                 "scan_kernel": {"avg_launch_ms": k_ms / k_n if k_n else None,
                                 "GB_per_s": n * 28 / (k_ms / k_n / 1e3) / 1e9 if k_n else None,
                                 "frac_of_hbm_peak": n * 28 / (k_ms / k_n / 1e3) / 1e9 / peak if k_n else None,
                                 "note": "the fused scan kernel alone; the rest of the step is fixed per-step overhead"},
                 "path": part.last_path(), "revenue": revenue}
    if O:
        tot, checked = oracle_q6_partition(ctx, O, sf * world, rank * n6, (rank + 1) * n6, env.host_threads,
                                           args.parity_chunk_rows, tpch.Q6_COLUMNS)
        tots = gather_objects(env, tot)
        want, passing = sum(x[0] for x in tots), sum(x[1] for x in tots)
        err = abs(revenue - want) / abs(want) if passing and revenue is not None else (0.0 if revenue is None and not passing else 1.0)
        # counts: COUNT(*) under the same filter, bit-exact
        from ballista_b200 import CoalesceBatchesExec, FilterExec, HashAggregateExec, count
        cnt = HashAggregateExec("Partial", [], [count().alias("c")],
                                CoalesceBatchesExec(FilterExec(tpch.q6_predicate(), tpch.scan(ctx, [[li6]]))))
        mine = cnt.collect().to_pylist()[0]["c[count]"]
        got_passing = sum(gather_objects(env, mine))
        out["q6"]["parity"] = {"config": f"q6_sf{sf:g}_x{world}", "rows": int(n * world), "counts_equal": bool(got_passing == passing),
                               "rows_passing": int(passing), "max_rel_err": err, "tolerance": 1e-9,
                               "ok": bool(got_passing == passing and err <= 1e-9), "generator_rows_cross_checked": checked,
                               "checker": "oracle/ Q6 over the same rows"}
    del li6, scan6, part6, query6

    # ---- Q3 on one GPU's own copy of all three tables: SF10, and (N = 1) the metric's SF100 ----
    def q3_local(qsf, name):
        cu = ctx.tpch("customer", qsf, columns=tpch.Q3_CUSTOMER_COLUMNS)
        od = ctx.tpch("orders", qsf, columns=tpch.Q3_ORDERS_COLUMNS)
        li = ctx.tpch("lineitem", qsf, columns=tpch.Q3_LINEITEM_COLUMNS)

        c, o, l = (tpch.scan(ctx, [[x]]) for x in (cu, od, li))
        partial = tpch.q3_partial(c, o, l)
        leaf = MemoryExec(ctx, partial.schema(), [[]])
        final = tpch.q3_final(leaf, partial.children()[0].schema())

        def plan():  # Partial task, then the Final task over its states (plans built once, executed per step)
            # re-feeding the leaves moves their version: the joins rebuild their hash tables in every step, as every
            # Ballista task does (a build table kept from the previous step would be work skipped in the timed region)
            for leaf_scan, batch in ((c, cu), (o, od), (l, li)):
                leaf_scan.clear()
                leaf_scan.add(0, batch)
            states = partial.collect_device(0)
            leaf.clear()
            for s_ in states:
                leaf.add(0, s_)
            return final

        def q3():
            return sum(b.num_rows for b in plan().collect_device(0))
        ms, launches, groups, t = timeit(q3)
        rows = cu.num_rows + od.num_rows + li.num_rows
        scan_bytes = cu.num_bytes + od.num_bytes + li.num_bytes
        res = {"workload": f"TPC-H Q3, SF{qsf:g} on each GPU (customer {cu.num_rows} + orders {od.num_rows} + "
                           f"lineitem {li.num_rows} rows; {scan_bytes / 1e9:.3f} GB of scanned columns)",
               "ms_per_step": ms, "step_ms": step_stats(t["per_step"]), "input_rows_per_s": rows * world / (ms / 1e3),
               "scan_GB_per_s_per_gpu": scan_bytes / (ms / 1e3) / 1e9, "groups": groups, "launches_per_step": launches,
               "note": "late materialisation reads only the key / predicate columns of non-matching rows, so "
                       "scan GB/s is algorithmic bytes / time, not DRAM traffic"}
        arrs = q3_table_arrays(plan().collect())
        stats = None
        if O:
            want = O.q3(qsf, threads=env.host_threads)
            stats = want["stats"]
            res["parity"] = parity_q3(name, arrs, want, rows)
        res["roofline"] = q3_roofline(scan_bytes, stats, ms, peak)
        res["digest"] = q3_checksum(arrs)
        return res
    out["q3"] = q3_local(sf, f"q3_sf{sf:g}")
    if world == 1 and args.q3_sf > sf:
        out[f"q3_sf{args.q3_sf:g}"] = q3_local(args.q3_sf, f"q3_sf{args.q3_sf:g}")
    # ---- plan shapes the library has no pre-built kernel for: compiled at run time (NVRTC) into the same TMA-fed scan
    # kernel.  (a) Q1 with `l_shipdate < d` instead of `<=`; (b) a disjunctive predicate, one key, other aggregates. ----
    if world == 1:
        from ballista_b200 import CoalesceBatchesExec, FilterExec, HashAggregateExec, avg, count, date, lit
        from ballista_b200 import sum as sum_
        n1 = nrows("lineitem", args.sf)
        li1 = ctx.tpch("lineitem", args.sf, columns=tpch.Q1_COLUMNS)

        def unseen(name, partial, make_final, bytes_per_row, check):
            query = tpch.TwoStage(ctx, partial, make_final)
            ms, launches, tab, t = timeit(query.run, steps=10, profile=True)
            k_ms, k_n = t["kernel"]
            path = partial.last_path()
            assert path.startswith("fused:") and "[tma" in path and "{jit" in path, path
            res = {"workload": name, "ms_per_step": ms, "step_ms": step_stats(t["per_step"]),
                   "rows_per_s": n1 / (ms / 1e3), "frac_of_hbm_peak": n1 * bytes_per_row / (ms / 1e3) / 1e9 / peak,
                   "scan_kernel": {"avg_launch_ms": k_ms / k_n, "GB_per_s": n1 * bytes_per_row / (k_ms / k_n / 1e3) / 1e9,
                                   "frac_of_hbm_peak": n1 * bytes_per_row / (k_ms / k_n / 1e3) / 1e9 / peak},
                   "algorithmic_bytes_per_row": bytes_per_row, "launches_per_step": launches, "path": path}
            if O:
                res["parity"] = check(tab.to_pylist())
            return res
        scan_a = tpch.scan(ctx, [[li1]])
        part_a = HashAggregateExec("Partial", tpch._q1_groups(), tpch.q1_aggregates(),
                                   CoalesceBatchesExec(FilterExec(col("l_shipdate") < date("1998-09-02"), scan_a)))

        def check_a(rows_):
            acc_, chk_ = oracle_q1_partition(ctx, O, args.sf, 0, n1, env.host_threads, args.parity_chunk_rows,
                                             tpch.Q1_COLUMNS, date_lit=O.DATE_1998_09_02 - 1)  # x < d  ==  x <= d - 1
            return parity_q1(f"q1_lt_sf{args.sf:g}", rows_, acc_, n1, chk_)
        out["q1_lt_jit"] = unseen(f"TPC-H Q1 with l_shipdate < date (not <=), SF{args.sf:g}: a shape compiled at run time",
                                  part_a, lambda s_: tpch.q1_final(s_, scan_a.schema()), Q1_BYTES_PER_ROW, check_a)
        pred_b = col("l_discount").between(lit(0.02), lit(0.06)) | (col("l_quantity") < lit(10))
        aggs_b = [sum_(col("l_quantity") * col("l_extendedprice")).alias("s"), avg(col("l_tax")).alias("t"), count().alias("n")]
        scan_b = tpch.scan(ctx, [[li1]])
        part_b = HashAggregateExec("Partial", [col("l_linestatus")], aggs_b, CoalesceBatchesExec(FilterExec(pred_b, scan_b)))
        mk_b = lambda s_: HashAggregateExec("Final", [col("l_linestatus")], aggs_b, tpch.MergeExec(s_), input_schema=scan_b.schema())

        def check_b(rows_):  # the column-at-a-time path (eval.cu + materialised source) over the same rows
            os.environ["B2_DISABLE_FUSION"] = "1"
            try:
                p2 = HashAggregateExec("Partial", [col("l_linestatus")], aggs_b, CoalesceBatchesExec(FilterExec(pred_b, tpch.scan(ctx, [[li1]]))))
                ref = tpch.run_two_stage(ctx, p2, mk_b).to_pylist()
            finally:
                del os.environ["B2_DISABLE_FUSION"]
            got_ = {r_["l_linestatus"]: r_ for r_ in rows_}
            want_ = {r_["l_linestatus"]: r_ for r_ in ref}
            keys_eq = sorted(got_) == sorted(want_)
            counts_eq = keys_eq and all(got_[k_]["n"] == want_[k_]["n"] for k_ in want_)
            err_ = max(abs(got_[k_][c_] - want_[k_][c_]) / abs(want_[k_][c_]) for k_ in want_ for c_ in ("s", "t")) if keys_eq else 1.0
            return {"config": f"unseen_shape_b_sf{args.sf:g}", "rows": int(n1), "keys_equal": bool(keys_eq),
                    "counts_equal": bool(counts_eq), "max_rel_err": err_, "tolerance": 1e-9,
                    "ok": bool(keys_eq and counts_eq and err_ <= 1e-9),
                    "checker": "the library's own column-at-a-time path (B2_DISABLE_FUSION) over the same rows; the shape's "
                               "semantics are checked against numpy in tests/test_gpu_jit.py"}
        out["unseen_shape_b_jit"] = unseen("disjunctive predicate (l_discount BETWEEN .02 AND .06 OR l_quantity < 10), group by "
                                           f"l_linestatus, SUM(qty*price) AVG(tax) COUNT(*), SF{args.sf:g}: compiled at run time",
                                           part_b, mk_b, 37, check_b)
        del li1, scan_a, scan_b, part_a, part_b

    # ---- Q5: six tables, five joins (one on a two-column key), group by n_name.  Orders / lineitem are row-range
    # partitions of ONE database (one per GPU); customer, supplier, nation and region are replicated, as the reference's
    # collect-left join reads its whole build side in every task.  An order's lines lie in its own partition, so every
    # join is local and only the 5-row partial states cross NVLink (all-gather). ----
    def q5_run(qsf, name):
        no = nrows("orders", qsf)
        o0, o1 = cluster.partition_rows(no, world, rank, align=8)
        batches = []
        for table, cols_ in tpch.Q5_TABLES:
            if table == "orders":
                batches.append(ctx.tpch(table, qsf, o0, o1, columns=cols_))
            elif table == "lineitem":
                batches.append(ctx.tpch(table, qsf, 4 * o0, 4 * o1, columns=cols_))
            else:
                batches.append(ctx.tpch(table, qsf, columns=cols_))
        scans = [tpch.scan(ctx, [[b_]]) for b_ in batches]
        partial = tpch.q5_partial(*scans)
        query = tpch.TwoStage(ctx, partial, lambda s_: tpch.q5_final(s_, partial.children()[0].schema()), merge="allgather")

        def q5():
            for leaf_scan, batch in zip(scans, batches):  # the joins rebuild their tables in every step (see Q3)
                leaf_scan.clear()
                leaf_scan.add(0, batch)
            return query.run()
        ms, launches, tab, t = timeit(q5)
        rows_all = sum(nrows(tb, qsf) for tb, _ in tpch.Q5_TABLES)
        scan_bytes = sum(b_.num_bytes for b_ in batches)
        res = {"workload": f"TPC-H Q5 over an SF{qsf:g} database: orders / lineitem in {world} row-range partition(s), "
                           "customer / supplier / nation / region replicated", "ms_per_step": ms,
               "step_ms": step_stats(t["per_step"]), "input_rows_per_s": rows_all / (ms / 1e3),
               "scan_GB_per_s_per_gpu": scan_bytes / (ms / 1e3) / 1e9, "launches_per_step": launches,
               "revenue": {r_["n_name"]: r_["revenue"] for r_ in tab.to_pylist()}}
        if O:
            mine = O.q5(qsf, threads=env.host_threads, orders_range=(o0, o1))
            want = {}
            for part_ in gather_objects(env, mine):
                for k_, v_ in part_.items():
                    want[k_] = want.get(k_, 0.0) + v_
            got = res["revenue"]
            keys_equal = sorted(got) == sorted(want)
            err = max(abs(got[k_] - want[k_]) / abs(want[k_]) for k_ in want) if keys_equal and want else 0.0
            res["parity"] = {"config": f"{name}_x{world}", "input_rows": int(rows_all), "groups": len(got),
                             "keys_equal": bool(keys_equal), "counts_equal": bool(keys_equal), "max_rel_err": err,
                             "tolerance": 1e-9, "ok": bool(keys_equal and err <= 1e-9),
                             "checker": "oracle/ Q5 (five collect-left hash joins in FROM order + group-by) over the oracle's own host tables"}
        return res
    out["q5"] = q5_run(sf, f"q5_sf{sf:g}")
    if args.q3_sf > sf:
        out[f"q5_sf{args.q3_sf:g}"] = q5_run(args.q3_sf, f"q5_sf{args.q3_sf:g}")
    if world == 1:
        out["parquet_ingest"] = run_parquet_ingest(env, check=not args.no_parity)
    if world > 1:
        out.update(run_multi_gpu_configs(args, env, O, timeit))
    return out


def run_parquet_ingest(env, sf=1.0, check=True):
    """the ParquetExec leaf's column decode on the GPU.  A lineitem SF1 file written the way the reference's
    `convert` command writes it (snappy pages, dictionary coding, the benchmark's REQUIRED columns; main.rs:233-252),
    read into page-locked memory, decoded by b2_parquet_decode; pyarrow's reader on the same bytes is both the parity
    check (every column bit-exact) and the CPU baseline."""
    import io
    import pyarrow as pa
    import pyarrow.parquet as pq
    import ballista_b200 as b2
    ctx = env.ctx
    table = pa.Table.from_batches([ctx.tpch("lineitem", sf).to_arrow()])
    bio = io.BytesIO()
    pq.write_table(table, bio, compression="snappy", use_dictionary=True)
    data = bio.getvalue()
    f = b2.ParquetFile(data, pinned_by=ctx)
    times = []
    for _ in range(6):
        ctx.synchronize()
        t0 = time.perf_counter()
        dev = f.read(ctx)
        ctx.synchronize()
        times.append((time.perf_counter() - t0) * 1e3)
        out_bytes = dev.num_bytes
        got = dev.to_arrow() if check and len(times) == 1 else None
        if got is not None:
            first = got
        del dev
    times = sorted(times[1:])
    cpu = {}
    for threads in (False, True):
        best = None
        for _ in range(2):
            t0 = time.perf_counter()
            ref = pq.read_table(io.BytesIO(data), use_threads=threads)
            dt = (time.perf_counter() - t0) * 1e3
            best = dt if best is None or dt < best else best
        cpu[threads] = best
    res = {"workload": f"lineitem SF{sf:g} ({table.num_rows} rows x {table.num_columns} columns) as one Parquet file: snappy pages, "
                       "dictionary coding, REQUIRED columns; file bytes in page-locked host memory -> columns in HBM",
           "file_MB": len(data) / 1e6, "decoded_MB": out_bytes / 1e6, "ms_best": times[0], "ms_median": times[len(times) // 2],
           "decoded_GB_per_s": out_bytes / (times[0] / 1e3) / 1e9,
           "cpu_reader": {"what": "pyarrow.parquet.read_table on the same bytes (Arrow C++ reader)",
                          "ms_1_thread": cpu[False], "ms_all_threads": cpu[True], "threads": os.cpu_count()}}
    if check:
        ref = ref.combine_chunks()
        same = [first.column(i).equals(ref.column(i).chunk(0) if ref.column(i).num_chunks else pa.array([], ref.schema.field(i).type))
                for i in range(ref.num_columns)]
        res["parity"] = {"config": f"parquet_ingest_sf{sf:g}", "input_rows": int(table.num_rows), "columns": len(same),
                         "keys_equal": bool(all(same)), "counts_equal": bool(first.num_rows == ref.num_rows), "max_rel_err": 0.0,
                         "tolerance": 0.0, "ok": bool(all(same) and first.num_rows == ref.num_rows),
                         "checker": "pyarrow.parquet reader on the same file bytes, every column Array.equals (bit-exact)"}
    return res


def run_multi_gpu_configs(args, env, O, timeit):
    """N > 1: the shuffle, Q3 over ONE SF100 database in N row-range partitions (both plan shapes) and the strong-scaled
    Q1.  Every leg checks itself: digests reduced over the ranks against the single-GPU plan over the same database."""
    from ballista_b200 import MemoryExec, RepartitionExec, _capi, col, tpch
    import numpy as np
    ctx, rank, world, cluster = env.ctx, env.rank, env.world, env.cluster
    sf = args.other_sf
    out = {}
    nrows = lambda t, s: _capi.lib().b2_tpch_num_rows(t.encode(), float(s))

    def row_digest(ok, price, disc):
        with np.errstate(over="ignore"):
            h = (ok.astype(np.uint64) * np.uint64(0x9E3779B97F4A7C15) ^ price.view(np.uint64) * np.uint64(0xC2B2AE3D27D4EB4F)
                 ^ disc.view(np.uint64) * np.uint64(0x165667B19E3779F9))
            return int(len(ok)), int(h.sum(dtype=np.uint64))

    # ---- shuffle: RepartitionExec(Hash(l_orderkey), N) with the communicator attached ----
    cols = ["l_orderkey", "l_extendedprice", "l_discount"]
    n = nrows("lineitem", sf)
    part = ctx.tpch("lineitem", sf * world, rank * n, (rank + 1) * n, columns=cols)
    path = [None]

    def exchange():
        rep = RepartitionExec(tpch.scan(ctx, [[part]]), "Hash", world, [col("l_orderkey")])
        got = rep.collect_device(rank)
        path[0] = rep.last_path()
        return got

    def shuffle():
        return sum(b.num_rows for b in exchange())
    sent0 = ctx.comm_bytes_sent()
    ms, launches, _, t = timeit(shuffle, steps=10, warmup=3, profile=True)
    pull_ms, pull_n = t["kernel"]  # CUDA events around the exchange kernel of every timed exchange
    pull_ms = cluster.max_over_ranks(env.dist, env.torch, pull_ms / max(pull_n, 1), world)
    sent = (ctx.comm_bytes_sent() - sent0) / 13.0
    out["shuffle"] = {"workload": f"hash repartition of lineitem SF{sf:g} per GPU ({n} rows x 20 B) into {world} "
                                  "partitions, exchanged over NVLink", "path": path[0], "ms_per_step": ms,
                      "step_ms": step_stats(t["per_step"]),
                      "rows_per_s": n * world / (ms / 1e3), "egress_bytes_per_gpu": sent,
                      "egress_GB_per_s_per_gpu": sent / (ms / 1e3) / 1e9,
                      "frac_of_nvlink_900GBps": sent / (ms / 1e3) / 1e9 / 900.0, "launches_per_step": launches,
                      "exchange_step": {"kernel": "the exchange kernel alone (CUDA events on the library's stream)", "ms": pull_ms,
                                        "egress_GB_per_s_per_gpu": sent / (pull_ms / 1e3) / 1e9 if pull_ms else None,
                                        "frac_of_nvlink_900GBps": sent / (pull_ms / 1e3) / 1e9 / 900.0 if pull_ms else None},
                      "note": "ms_per_step is the whole operator (split + metadata + exchange)"}
    # parity: nothing lost or invented (row count + 64-bit digest of whole rows, reduced over ranks) and every key
    # on exactly one rank (distinct keys per rank add up to the table's orders: every order has a line)
    src = part.to_arrow()
    d_in = row_digest(*(src[c].to_numpy() for c in cols))
    got = [b.to_arrow() for b in exchange()]
    oks = np.concatenate([g["l_orderkey"].to_numpy() for g in got]) if got else np.zeros(0, np.int32)
    d_out = row_digest(oks, np.concatenate([g["l_extendedprice"].to_numpy() for g in got]),
                       np.concatenate([g["l_discount"].to_numpy() for g in got]))
    mine = (d_in, d_out, int(len(np.unique(oks))))
    alls = gather_objects(env, mine)
    rows_in, rows_out = sum(a[0][0] for a in alls), sum(a[1][0] for a in alls)
    dig_in = sum(a[0][1] for a in alls) % (1 << 64)
    dig_out = sum(a[1][1] for a in alls) % (1 << 64)
    distinct = sum(a[2] for a in alls)
    orders = nrows("orders", sf * world)
    out["shuffle"]["parity"] = {"config": f"shuffle_sf{sf:g}_x{world}", "rows": rows_in, "counts_equal": bool(rows_in == rows_out),
                                "row_digest_equal": bool(dig_in == dig_out),
                                "keys_colocated": bool(distinct == orders), "distinct_keys": distinct,
                                "ok": bool(rows_in == rows_out and dig_in == dig_out and distinct == orders),
                                "checker": "row count + order-independent 64-bit digest of whole rows before / after the "
                                           "exchange, reduced over ranks; distinct keys per rank sum to the table's orders"}
    del part, src, got, oks

    # ---- Q3 over ONE SF100 database in N row-range partitions ----
    dsf = args.q3_sf
    def tpart(table, columns, align=1):
        total = nrows(table, dsf)
        r0, r1 = cluster.partition_rows(total, world, rank, align=align)
        return ctx.tpch(table, dsf, r0, r1, columns=columns)
    # reference digest: the single-GPU plan over the whole database (every rank computes it: no communication)
    fc, fo, fl = (ctx.tpch(t, dsf, columns=c) for t, c in (("customer", tpch.Q3_CUSTOMER_COLUMNS),
                                                           ("orders", tpch.Q3_ORDERS_COLUMNS),
                                                           ("lineitem", tpch.Q3_LINEITEM_COLUMNS)))
    scans = [tpch.scan(ctx, [[x]]) for x in (fc, fo, fl)]
    partial = tpch.q3_partial(*scans)
    states = partial.collect_device(0)
    ref = q3_checksum(q3_table_arrays(tpch.q3_final(MemoryExec(ctx, partial.schema(), [states]),
                                                    partial.children()[0].schema()).collect()))
    del fc, fo, fl, scans, partial, states
    dcu, dod = tpart("customer", tpch.Q3_CUSTOMER_COLUMNS), tpart("orders", tpch.Q3_ORDERS_COLUMNS)
    dli = tpart("lineitem", tpch.Q3_LINEITEM_COLUMNS, align=32)
    rows_all = sum(nrows(t, dsf) for t in ("customer", "orders", "lineitem"))

    def q3d_plan():
        scans = [tpch.scan(ctx, [[x]]) for x in (dcu, dod, dli)]
        partial = tpch.q3_distributed_partial(*scans, world)
        states = partial.collect_device(rank)
        return tpch.q3_final(MemoryExec(ctx, partial.schema(), [states]), partial.children()[0].schema())

    def q3b_plan():
        scans = [tpch.scan(ctx, [[x]]) for x in (dcu, dod, dli)]
        partial, in_schema, p = tpch.q3_broadcast_stages(ctx, *scans, world)
        states = partial.collect_device(p)
        return tpch.q3_final(MemoryExec(ctx, partial.schema(), [states]), in_schema)

    def digest_parity(name, plan):
        d = q3_checksum(q3_table_arrays(plan().collect()))
        alls = gather_objects(env, d)
        cnt, dig = sum(a[0] for a in alls), sum(a[1] for a in alls) % (1 << 64)
        rev = math_fsum([a[2] for a in alls])
        err = abs(rev - ref[2]) / abs(ref[2]) if ref[2] else 0.0
        return {"config": name, "input_rows": int(rows_all), "groups": cnt, "counts_equal": bool(cnt == ref[0]),
                "keys_equal": bool(dig == ref[1]), "max_rel_err": err, "tolerance": 1e-9,
                "ok": bool(cnt == ref[0] and dig == ref[1] and err <= 1e-9),
                "checker": "group count, 64-bit digest of the group keys and total revenue, reduced over ranks, against the "
                           "single-GPU plan over the same database (itself compared with oracle/ in the N = 1 run)"}
    for name, plan, what in (("q3_distributed", q3d_plan, "both join inputs hash-repartitioned over NVLink (filters fused into the split)"),
                             ("q3_broadcast", q3b_plan, "build sides all-gathered (collect-left), probe sides local, partial states exchanged by group key")):
        sent0 = ctx.comm_bytes_sent()
        ms, launches, groups, t = timeit(lambda: sum(b.num_rows for b in plan().collect_device(0)), steps=10, warmup=3)
        sent = (ctx.comm_bytes_sent() - sent0) / 13.0
        out[name] = {"workload": f"TPC-H Q3 over an SF{dsf:g} database in {world} row-range partitions, " + what,
                     "ms_per_step": ms, "step_ms": step_stats(t["per_step"]), "input_rows_per_s": rows_all / (ms / 1e3),
                     "groups_rank0": groups, "egress_bytes_per_gpu": sent, "launches_per_step": launches,
                     "parity": digest_parity(f"{name}_sf{dsf:g}_x{world}", plan)}
    del dcu, dod, dli

    # ---- Q1 strong scaling: ONE SF100 lineitem in N row-range partitions (BASELINE config 3, literally) ----
    total = nrows("lineitem", args.sf)
    r0, r1 = cluster.partition_rows(total, world, rank, align=32)
    li = ctx.tpch("lineitem", args.sf, r0, r1, columns=tpch.Q1_COLUMNS)

    scan1 = tpch.scan(ctx, [[li]])
    query1 = tpch.TwoStage(ctx, tpch.q1_partial(scan1), lambda s: tpch.q1_final(s, scan1.schema()), merge="allgather")
    ms, launches, tab, t = timeit(query1.run, steps=20, warmup=3, profile=True)
    k_ms, k_n = t["kernel"]
    peak = measured_peaks().get("hbm_gbs") or 6650.0
    out["q1_strong"] = {"workload": f"TPC-H Q1, one SF{args.sf:g} lineitem ({total} rows) in {world} row-range partitions",
                        "ms_per_step": ms, "step_ms": step_stats(t["per_step"]), "rows_per_s": total / (ms / 1e3),
                        "GB_per_s_per_gpu": (r1 - r0) * Q1_BYTES_PER_ROW / (ms / 1e3) / 1e9,
                        "frac_of_hbm_peak_per_gpu": (r1 - r0) * Q1_BYTES_PER_ROW / (ms / 1e3) / 1e9 / peak,
                        "scan_kernel_ms": k_ms / k_n if k_n else None, "launches_per_step": launches}
    if O:
        acc, checked = oracle_q1_partition(ctx, O, args.sf, r0, r1, env.host_threads, args.parity_chunk_rows, tpch.Q1_COLUMNS)
        out["q1_strong"]["parity"] = parity_q1(f"q1_strong_sf{args.sf:g}_x{world}", tab.to_pylist(),
                                               merge_q1_acc(gather_objects(env, acc)), total, checked)
    del li, scan1, query1
    return out


def math_fsum(xs):
    import math
    return math.fsum(xs)


def emit(real_stdout, obj):
    os.write(real_stdout, (json.dumps(obj) + "\n").encode())


def collect_parity(out):
    """every parity verdict of the line -> (all ok, list of failing configs)"""
    bad = []
    def walk(o):
        if isinstance(o, dict):
            p = o.get("parity")
            if isinstance(p, dict) and not p.get("ok", False):
                bad.append(p.get("config"))
            for v in o.values():
                walk(v)
    walk(out)
    return not bad, bad


def main():
    args = parse_args()
    # stdout carries exactly ONE JSON line: anything libraries print (e.g. NCCL's version banner) goes to stderr
    sys.stdout.flush()
    real_stdout = os.dup(1)
    os.dup2(2, 1)
    if args.impl == "reference":
        out = run_reference(args)
        if out is not None:
            emit(real_stdout, out)
        return
    env = setup_env(args)
    out = run_b200(args, env)
    rank, world = env.rank, env.world
    if not args.no_e2e:
        out["e2e"] = run_e2e(args, env)
    env.q1 = None
    if not args.no_other:
        out["other_configs"] = run_other_configs(args, env)
    if rank == 0 and world == 1 and not args.no_cpu:
        threads = os.cpu_count() or 1
        r = cpu_q1(args.sf, args.cpu_seconds, threads, with_acero=True)
        out["cpu_baseline"] = {"value": r["value"], "unit": UNIT, "cores": threads, "kind": "port",
                               "sample": f"first {r['rows']} rows of the SF{args.sf:g} lineitem, one pass of {r['s_per_pass']:.2f} s on "
                                         f"{threads} threads (oracle/: C++ restatement of the DataFusion plan)"}
        out["cpu_baseline_b"] = r["acero"]
    ok, bad = collect_parity(out)
    out["parity_all_ok"] = ok
    if rank == 0:
        emit(real_stdout, out)
    if world > 1:
        env.dist.barrier()
        env.dist.destroy_process_group()
    if not ok:
        sys.stderr.write(f"bench.py: PARITY FAILED for {bad}\n")
        sys.exit(3)


if __name__ == "__main__":
    main()
