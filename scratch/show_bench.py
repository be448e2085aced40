import json, sys
d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
o = d.pop('other_configs', {})
keep = {k: d[k] for k in ('metric', 'value', 'n_gpus', 'ms_per_step', 'gpu_launches') if k in d}
keep['roofline_frac'] = d.get('roofline', {}).get('frac'); keep['kernel_ms'] = d.get('roofline', {}).get('avg_launch_ms')
keep['e2e'] = d.get('e2e', {}).get('value'); keep['clocks'] = d.get('clocks')
print(json.dumps(keep))
for k, v in o.items():
    print(' ', k, json.dumps({a: (round(b, 4) if isinstance(b, float) else b) for a, b in v.items() if a not in ('note', 'workload', 'path')})[:600])
