"""Compact view of a bench.py JSON line: python profiles/tools/show_bench.py gpurun_out/bench_n1.json"""
import json, sys
d = json.load(open(sys.argv[1]))
print(f"N={d['n_gpus']} value={d['value']:.4g} {d['unit']}  ms/step={d['ms_per_step']:.3f} median={d['step_ms']['median_ms']:.3f} best={d['step_ms']['best_ms']:.3f} launches/step={d['gpu_launches']/d['steps']:.1f}")
r = d["roofline"]; print(f"  roofline: kernel {r['avg_launch_ms']:.3f} ms  {r['achieved']:.0f} GB/s frac={r['frac']:.3f}  step_frac={r['step_frac_of_peak']:.3f} share={r['kernel_share_of_step']:.3f}")
print("  parity:", {k: d.get("parity", {}).get(k) for k in ("ok", "counts_equal", "max_rel_err")}, "all_ok:", d.get("parity_all_ok"))
if "e2e" in d: print(f"  e2e: {d['e2e']['value']:.4g} rows/s  {d['e2e']['ms_per_step']:.1f} ms/step")
for k, v in d.get("other_configs", {}).items():
    extra = ""
    if "frac_of_hbm_peak" in v: extra += f" frac_hbm={v['frac_of_hbm_peak']:.3f}"
    if "frac_of_nvlink_900GBps" in v: extra += f" frac_nvlink={v['frac_of_nvlink_900GBps']:.3f} xchg={v['exchange_step']['frac_of_nvlink_900GBps']:.3f}"
    if "roofline" in v: extra += f" scan_frac={v['roofline']['frac_scan_only']:.3f}" + (f" scan+table={v['roofline'].get('frac_scan_plus_table', 0):.3f}" if 'frac_scan_plus_table' in v['roofline'] else "")
    if "frac_of_hbm_peak_per_gpu" in v: extra += f" frac_hbm={v['frac_of_hbm_peak_per_gpu']:.3f}"
    p = v.get("parity", {})
    if not isinstance(v, dict) or "ms_per_step" not in v:  # legs with their own shape (parquet_ingest)
        print(f"  {k:16s} {json.dumps(v)[:200]}")
        continue
    print(f"  {k:16s} {v['ms_per_step']:.3f} ms (median {v['step_ms']['median_ms']:.3f}) launches={v.get('launches_per_step')} {extra} parity_ok={p.get('ok')} err={p.get('max_rel_err')}")
for k in ("cpu_baseline", "cpu_baseline_b"):
    if d.get(k): print(f"  {k}: {d[k].get('value')}")
