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
  cpu_baseline : oracle/ (C++ restatement of the DataFusion plan, one thread per partition) on a bounded sample

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
    """Cheap invariants of the Q1 result (full parity is tests/; this guards the timed path)."""
    assert 1 <= len(rows) <= 4, rows
    keys = sorted((r["l_returnflag"], r["l_linestatus"]) for r in rows)
    assert keys == sorted(set(keys)), keys
    total = sum(r["count_order"] for r in rows)
    assert 0 < total <= n_rows_total, (total, n_rows_total)
    for r in rows:
        assert abs(r["avg_qty"] - r["sum_qty"] / r["count_order"]) <= 1e-9 * abs(r["avg_qty"])
    return total


# ---------------------------------------------------------------------------------------------------------
# this repo's arm
# ---------------------------------------------------------------------------------------------------------
def run_b200(args):
    import torch
    import torch.distributed as dist
    import ballista_b200 as b2
    from ballista_b200 import tpch, _capi

    rank, local_rank, world = env_rank()
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (ballista_b200 has no CPU fallback)")
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    ctx = b2.Context(local_rank)
    from ballista_b200 import cluster
    cluster.attach_communicator(ctx, dist, rank, world)
    stream = torch.cuda.ExternalStream(ctx.cuda_stream)

    # ---- workload: this rank's lineitem partition, generated in HBM (setup, untimed) ----
    if args.scaling == "weak":
        table_sf = args.sf * world
        rows_local = _capi.lib().b2_tpch_num_rows(b"lineitem", float(args.sf))
        r0 = rank * rows_local
    else:
        table_sf = args.sf
        total = _capi.lib().b2_tpch_num_rows(b"lineitem", float(args.sf))
        r0, r1 = cluster.partition_rows(total, world, rank, align=32)
        rows_local = r1 - r0
    rows_total = rows_local * world if args.scaling == "weak" else _capi.lib().b2_tpch_num_rows(b"lineitem", float(args.sf))
    # batches of at most 2^30 rows (Utf8 offsets are int32)
    max_batch = 1 << 30
    dev_batches = [ctx.tpch("lineitem", table_sf, b, min(b + max_batch, r0 + rows_local), columns=tpch.Q1_COLUMNS)
                   for b in range(r0, r0 + rows_local, max_batch)]
    ctx.synchronize()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        ctx.synchronize()

    def step(source_batches):
        scan = tpch.scan(ctx, [source_batches])
        partial = tpch.q1_partial(scan)
        table = tpch.run_two_stage(ctx, partial, lambda s: tpch.q1_final(s, scan.schema()), merge="allgather")
        return table, partial

    def timed(source_batches, steps, warmup, profile):
        for _ in range(warmup):
            step(source_batches)
        barrier()
        sampler = ClockSampler(local_rank)
        sampler.start()
        if profile:
            ctx.set_profiling(True)
            ctx.profile()
        l0 = ctx.kernel_launches()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(steps):
            table, partial = step(source_batches)
        e1.record(stream)
        barrier()
        ms = e0.elapsed_time(e1)
        launches = ctx.kernel_launches() - l0
        kern = ctx.profile() if profile else (0.0, 0)
        if profile:
            ctx.set_profiling(False)
        clocks = sampler.result()
        return cluster.max_over_ranks(dist, torch, ms, world), launches, kern, clocks, table, partial

    # ---- value: inputs resident in HBM ----
    ms, launches, (kern_ms, kern_n), clocks, table, partial = timed(dev_batches, args.steps, args.warmup, True)
    rows = table.to_pylist()
    passed = check_q1_rows(rows, rows_total)
    assert partial.last_path().startswith("fused:") and "[tma" in partial.last_path(), partial.last_path()
    value = rows_total * args.steps / (ms / 1e3)
    peaks = measured_peaks()
    peak = peaks.get("hbm_gbs")
    peak_src = "measured (MEASURED_PEAKS.json hbm_gbs, copy read+write)"
    if not peak:
        peak, peak_src = 6650.0, "fallback (B200_PROFILING.md)"
    achieved = (Q1_BYTES_PER_ROW * rows_local * args.steps / max(kern_n, 1) * 1.0) / (kern_ms / max(kern_n, 1) / 1e3) / 1e9 if kern_n else None
    roofline = {"bound": "hbm", "kernel": "agg_tma_kernel<FusedSrc<Q1>, narrow> (TMA-fed scan + filter + expressions + partial aggregate)",
                "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": (achieved / peak) if achieved else None,
                "traffic": None, "peak_source": peak_src, "launches": kern_n,
                "avg_launch_ms": kern_ms / kern_n if kern_n else None,
                "algorithmic_bytes_per_row": Q1_BYTES_PER_ROW, "rows_per_launch": rows_local * args.steps / max(kern_n, 1),
                "kernel_share_of_step": kern_ms / ms if ms else None}
    traffic_file = os.path.join(ROOT, "profiles", "q1_traffic.json")
    if os.path.exists(traffic_file):
        try:
            with open(traffic_file) as f:
                tj = json.load(f)
            if tj.get("rows_per_launch") == roofline["rows_per_launch"]:
                roofline["traffic"] = tj.get("dram_bytes_per_launch")
        except Exception:
            pass

    out = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
           "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None,
           "dtype": "f64", "data": "synthetic (counter-based TPC-H-shaped generator, seed 42, generated in HBM)",
           "config": {"workload": f"TPC-H Q1, SF{args.sf:g} lineitem ({rows_local} rows x 46 B = "
                                  f"{rows_local * 46 / 1e9:.2f} GB) per GPU" if args.scaling == "weak" else
                                  f"TPC-H Q1, one SF{args.sf:g} lineitem split over {world} GPUs",
                      "rows_per_gpu": rows_local, "rows_total": rows_total, "rows_passing_filter": passed,
                      "l2": "inputs larger than L2 (no flush needed)" if rows_local * 46 > 512e6 else "inputs smaller than L2!",
                      "parallelism": f"partition-per-GPU x{world}, partial states all-gathered over NCCL" if world > 1 else "1 partition"},
           "gpu_launches": launches, "clocks": clocks, "roofline": roofline,
           "GB_per_s": value * Q1_BYTES_PER_ROW / 1e9}
    return ctx, out, dev_batches, timed, (rank, world, torch)


def run_e2e(args, ctx, dev_batches, timed, env):
    """Same plan fed with HOST RecordBatches (pinned): H2D of every input column inside the timed region."""
    rank, world, torch = env
    rows_local = sum(b.num_rows for b in dev_batches)
    need = rows_local * Q1_BYTES_PER_ROW
    local_world = int(os.environ.get("LOCAL_WORLD_SIZE", world))
    budget = mem_available_bytes() // max(local_world, 1) // 3
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
    ms, launches, _, clocks, table, _ = timed(host_batches, args.e2e_steps, 1, False)
    d2h = sum(c.nbytes for c in table.columns)
    check_q1_rows(table.to_pylist(), rows_e2e * world)
    unpin(torch, pinned)
    return {"value": rows_e2e * world * args.e2e_steps / (ms / 1e3), "unit": UNIT, "h2d_bytes_per_step": int(h2d),
            "d2h_bytes_per_step": int(d2h), "rows_per_gpu": rows_e2e, "steps": args.e2e_steps,
            "ms_per_step": ms / args.e2e_steps, "host_batch_rows": args.e2e_batch_rows,
            "note": "host Arrow RecordBatches in cudaHostRegister'ed memory pushed through b2_memory_exec_add_batch; "
                    "H2D of all 7 columns + result D2H inside the timed region"
                    + ("" if frac == 1.0 else f"; bounded to {frac:.2f} of the partition by host RAM")}


# ---------------------------------------------------------------------------------------------------------
# the reference's CPU algorithm (oracle/: C++ restatement of the DataFusion plan; the Rust reference cannot be
# built here — no cargo, DataFusion not vendored)
# ---------------------------------------------------------------------------------------------------------
def cpu_q1(sf, seconds, threads, steps=1, warmup=0):
    """-> (rows/s, sample description).  A bounded sample of the SF`sf` lineitem: the first `n` rows, n
    calibrated so that one pass takes about `seconds` of wall clock on `threads` threads."""
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
    t = time.perf_counter()
    for _ in range(steps):
        rows = O.q1(cols, n_partitions=threads)
    dt = time.perf_counter() - t
    check_q1_rows(rows, n)
    return n * steps / dt, n, dt / steps


def run_reference(args):
    rank, _, world = env_rank()
    if rank != 0:
        return None
    threads = os.cpu_count() or 1
    # each step = one pass over a bounded sample; K + W passes must end within minutes
    per_step = min(20.0, 150.0 / max(args.steps + args.warmup, 1))
    v, n, sec = cpu_q1(args.sf, per_step, threads, steps=args.steps, warmup=args.warmup)
    sample = f"first {n} rows of the SF{args.sf:g} synthetic lineitem per step, {threads} partitions (one thread each)"
    return {"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": sec * 1e3, "higher_is_better": True, "scaling": args.scaling,
            "vs_baseline": None, "dtype": "f64", "data": "synthetic (same generator, host copy)",
            "config": {"workload": f"TPC-H Q1, SF{args.sf:g} lineitem, bounded sample of {n} rows per step"},
            "cpu_baseline": {"value": v, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample},
            "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0,
            "note": "oracle/ C++ restatement of DataFusion@46161d2's Q1 plan (32,768-row batches, FilterExec copy, "
                    "no-CSE temporaries, byte-key hash aggregate); the Rust reference itself cannot be built here"}


def run_other_configs(args, ctx, env):
    """The other BASELINE.json configs, timed the same way (CUDA events on the library's stream, max over ranks)
    and reported under "other_configs": Q6 and Q3 at SF10 on one GPU's partition, Q3 at SF100 (one GPU's own copy at N = 1), and at N > 1 the hash-repartition
    shuffle of the Q3 lineitem side (l_orderkey, l_extendedprice, l_discount = 20 B/row) over NVLink."""
    import torch.distributed as dist
    from ballista_b200 import MemoryExec, RepartitionExec, _capi, cluster, col, tpch
    rank, world, torch = env
    stream = torch.cuda.ExternalStream(ctx.cuda_stream)
    sf = args.other_sf
    peak = measured_peaks().get("hbm_gbs") or 6650.0

    def timeit(fn, steps=10, warmup=3):
        for _ in range(warmup):
            fn()
        if world > 1:
            dist.barrier()
        ctx.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        l0 = ctx.kernel_launches()
        e0.record(stream)
        for _ in range(steps):
            r = fn()
        e1.record(stream)
        ctx.synchronize()
        torch.cuda.synchronize()
        return cluster.max_over_ranks(dist, torch, e0.elapsed_time(e1) / steps, world), (ctx.kernel_launches() - l0) / steps, r

    out = {}
    nrows = lambda t: _capi.lib().b2_tpch_num_rows(t.encode(), float(sf))
    # ---- Q6 @ SF10: filter + global SUM over this rank's lineitem (28 B/row) ----
    li6 = ctx.tpch("lineitem", sf * world, rank * nrows("lineitem"), (rank + 1) * nrows("lineitem"), columns=tpch.Q6_COLUMNS)

    def q6():
        scan = tpch.scan(ctx, [[li6]])
        part = tpch.q6_partial(scan)
        return tpch.run_two_stage(ctx, part, lambda s: tpch.q6_final(s, scan.schema()), merge="allgather"), part
    ctx.set_profiling(True)
    ctx.profile()
    ms, launches, (tab, part) = timeit(q6)
    k_ms, k_n = ctx.profile()  # CUDA events around every fused-scan launch since the reset (warm-up included)
    ctx.set_profiling(False)
    n = li6.num_rows
    out["q6"] = {"workload": f"TPC-H Q6, SF{sf:g} lineitem per GPU ({n} rows x 28 B)", "ms_per_step": ms,
                 "rows_per_s": n * world / (ms / 1e3), "GB_per_s_per_gpu": n * 28 / (ms / 1e3) / 1e9,
                 "frac_of_hbm_peak": n * 28 / (ms / 1e3) / 1e9 / peak, "launches_per_step": launches,
                 "scan_kernel": {"avg_launch_ms": k_ms / k_n if k_n else None,
                                 "GB_per_s": n * 28 / (k_ms / k_n / 1e3) / 1e9 if k_n else None,
                                 "frac_of_hbm_peak": n * 28 / (k_ms / k_n / 1e3) / 1e9 / peak if k_n else None,
                                 "note": "the fused scan kernel alone; the rest of the step is fixed per-step overhead"},
                 "path": part.last_path(), "revenue": tab.to_pylist()[0]["revenue"]}
    del li6
    # ---- Q3 @ SF10 (and, on one GPU, the metric's SF100): customer |x| orders |x| lineitem + group-by over this
    # rank's own copy of all three tables ----
    def q3_local(qsf):
        cu = ctx.tpch("customer", qsf, columns=tpch.Q3_CUSTOMER_COLUMNS)
        od = ctx.tpch("orders", qsf, columns=tpch.Q3_ORDERS_COLUMNS)
        li = ctx.tpch("lineitem", qsf, columns=tpch.Q3_LINEITEM_COLUMNS)

        def q3():
            c, o, l = (tpch.scan(ctx, [[x]]) for x in (cu, od, li))
            partial = tpch.q3_partial(c, o, l)
            states = partial.collect_device(0)
            final = tpch.q3_final(MemoryExec(ctx, partial.schema(), [states]), partial.children()[0].schema())
            return sum(b.num_rows for b in final.collect_device(0))
        ms, launches, groups = timeit(q3)
        rows = cu.num_rows + od.num_rows + li.num_rows
        scan_bytes = cu.num_bytes + od.num_bytes + li.num_bytes
        return {"workload": f"TPC-H Q3, SF{qsf:g} on each GPU (customer {cu.num_rows} + orders {od.num_rows} + "
                            f"lineitem {li.num_rows} rows; {scan_bytes / 1e9:.3f} GB of scanned columns)",
                "ms_per_step": ms, "input_rows_per_s": rows * world / (ms / 1e3),
                "scan_GB_per_s_per_gpu": scan_bytes / (ms / 1e3) / 1e9, "groups": groups, "launches_per_step": launches,
                "note": "late materialisation reads only the key / predicate columns of non-matching rows, so "
                        "scan GB/s is algorithmic bytes / time, not DRAM traffic"}
    out["q3"] = q3_local(sf)
    if world == 1 and args.q3_sf > sf:
        out[f"q3_sf{args.q3_sf:g}"] = q3_local(args.q3_sf)
    # ---- shuffle: RepartitionExec(Hash(l_orderkey), N) with the communicator attached (N > 1) ----
    if world > 1:
        cols = ["l_orderkey", "l_extendedprice", "l_discount"]
        n = nrows("lineitem")
        part = ctx.tpch("lineitem", sf * world, rank * n, (rank + 1) * n, columns=cols)
        path = [None]

        def shuffle():
            rep = RepartitionExec(tpch.scan(ctx, [[part]]), "Hash", world, [col("l_orderkey")])
            got = rep.collect_device(rank)
            path[0] = rep.last_path()
            return sum(b.num_rows for b in got)
        sent0 = ctx.comm_bytes_sent()
        ctx.set_profiling(True)
        ctx.profile()
        ms, launches, _ = timeit(shuffle, steps=10, warmup=3)
        pull_ms, pull_n = ctx.profile()  # CUDA events around the pull kernel of every exchange (warm-up included)
        ctx.set_profiling(False)
        pull_ms = cluster.max_over_ranks(dist, torch, pull_ms / max(pull_n, 1), world)
        sent = (ctx.comm_bytes_sent() - sent0) / 13.0
        out["shuffle"] = {"workload": f"hash repartition of lineitem SF{sf:g} per GPU ({n} rows x 20 B) into {world} "
                                      "partitions, exchanged over NVLink", "path": path[0], "ms_per_step": ms,
                          "rows_per_s": n * world / (ms / 1e3), "egress_bytes_per_gpu": sent,
                          "egress_GB_per_s_per_gpu": sent / (ms / 1e3) / 1e9,
                          "frac_of_nvlink_900GBps": sent / (ms / 1e3) / 1e9 / 900.0, "launches_per_step": launches,
                          "exchange_step": {"kernel": "pull_kernel (every rank reads its slices from all peer arenas; "
                                                      "also copies the local slice)", "ms": pull_ms,
                                            "egress_GB_per_s_per_gpu": sent / (pull_ms / 1e3) / 1e9 if pull_ms else None,
                                            "frac_of_nvlink_900GBps": sent / (pull_ms / 1e3) / 1e9 / 900.0 if pull_ms else None,
                                            "frac_of_measured_peer_copy_770GBps": sent / (pull_ms / 1e3) / 1e9 / 770.0 if pull_ms else None},
                          "note": "ms_per_step is the whole operator (split + metadata + move into the arena + exchange)"}
        del part
        # ---- Q3 with the shuffle: row-range partitions of the metric's SF100 database, both join inputs exchanged ----
        dsf = args.q3_sf
        def tpart(table, columns, align=1):
            total = _capi.lib().b2_tpch_num_rows(table.encode(), float(dsf))
            r0, r1 = cluster.partition_rows(total, world, rank, align=align)
            return ctx.tpch(table, dsf, r0, r1, columns=columns)
        dcu, dod = tpart("customer", tpch.Q3_CUSTOMER_COLUMNS), tpart("orders", tpch.Q3_ORDERS_COLUMNS)
        dli = tpart("lineitem", tpch.Q3_LINEITEM_COLUMNS, align=32)
        paths = [None]

        def q3d():
            scans = [tpch.scan(ctx, [[x]]) for x in (dcu, dod, dli)]
            partial = tpch.q3_distributed_partial(*scans, world)
            states = partial.collect_device(rank)
            final = tpch.q3_final(MemoryExec(ctx, partial.schema(), [states]), partial.children()[0].schema())
            return sum(b.num_rows for b in final.collect_device(0))
        sent0 = ctx.comm_bytes_sent()
        ms, launches, groups = timeit(q3d, steps=10, warmup=3)
        sent = (ctx.comm_bytes_sent() - sent0) / 13.0
        rows_all = sum(_capi.lib().b2_tpch_num_rows(t.encode(), float(dsf)) for t in ("customer", "orders", "lineitem"))
        out["q3_distributed"] = {"workload": f"TPC-H Q3 over an SF{dsf:g} database in {world} row-range partitions, both join "
                                             "inputs hash-repartitioned over NVLink (filters fused into the split)",
                                 "ms_per_step": ms, "input_rows_per_s": rows_all / (ms / 1e3), "groups_rank0": groups,
                                 "egress_bytes_per_gpu": sent, "launches_per_step": launches}
        # ---- the same query and partitions with BROADCAST build sides (the reference's collect-left join across
        # ranks): filtered customers and the first join's output are all-gathered, lineitem / orders never move,
        # partial states are exchanged by group key ----
        def q3b():
            scans = [tpch.scan(ctx, [[x]]) for x in (dcu, dod, dli)]
            partial, in_schema, p = tpch.q3_broadcast_stages(ctx, *scans, world)
            states = partial.collect_device(p)
            final = tpch.q3_final(MemoryExec(ctx, partial.schema(), [states]), in_schema)
            return sum(b.num_rows for b in final.collect_device(0))
        sent0 = ctx.comm_bytes_sent()
        ms, launches, groups = timeit(q3b, steps=10, warmup=3)
        sent = (ctx.comm_bytes_sent() - sent0) / 13.0
        out["q3_broadcast"] = {"workload": f"TPC-H Q3 over an SF{dsf:g} database in {world} row-range partitions, build sides "
                                           "all-gathered (collect-left), probe sides local, partial states exchanged by group key",
                               "ms_per_step": ms, "input_rows_per_s": rows_all / (ms / 1e3), "groups_rank0": groups,
                               "egress_bytes_per_gpu": sent, "launches_per_step": launches}
        del dcu, dod, dli
    return out



def emit(real_stdout, obj):
    os.write(real_stdout, (json.dumps(obj) + "\n").encode())


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
    ctx, out, dev_batches, timed, env = run_b200(args)
    rank, world, torch = env
    if not args.no_e2e:
        out["e2e"] = run_e2e(args, ctx, dev_batches, timed, env)
    if not args.no_other:
        out["other_configs"] = run_other_configs(args, ctx, env)
    if rank == 0 and world == 1 and not args.no_cpu:
        threads = os.cpu_count() or 1
        v, n, sec = cpu_q1(args.sf, args.cpu_seconds, threads)
        out["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": threads, "kind": "port",
                               "sample": f"first {n} rows of the SF{args.sf:g} lineitem, one pass of {sec:.2f} s on "
                                         f"{threads} threads (oracle/: C++ restatement of the DataFusion plan)"}
    if rank == 0:
        emit(real_stdout, out)
    if world > 1:
        import torch.distributed as dist
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
