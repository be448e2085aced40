"""per-launch time + DRAM bytes of the last N launches of an ncu csv: python profiles/tools/launch_table.py file.csv N"""
import csv, sys
lines = [l for l in open(sys.argv[1]) if not l.startswith('==')]
by = {}
for x in csv.DictReader(lines):
    by.setdefault(x['ID'], {'k': x['Kernel Name'], 'g': x['Grid Size']})[x['Metric Name']] = (x['Metric Value'].replace(',', ''), x['Metric Unit'])
ids = sorted(by, key=int)[-int(sys.argv[2]):]
tot = 0
for i in ids:
    d = by[i]; v, u = d['gpu__time_duration.sum']; t = float(v) * {'ns': 1e-6, 'us': 1e-3, 'ms': 1, 's': 1e3}[u]
    def b(m):
        if m not in d: return 0.0
        v, u = d[m]; return float(v) * {'byte': 1, 'Kbyte': 1e3, 'Mbyte': 1e6, 'Gbyte': 1e9}[u]
    rd, wr = b('dram__bytes_read.sum'), b('dram__bytes_write.sum')
    tot += t
    print(f"{t:8.3f} ms rd {rd/1e6:9.1f} MB wr {wr/1e6:9.1f} MB {(rd+wr)/t/1e6 if t else 0:7.0f} GB/s {d['g']:>14} {d['k'][:60]}")
print('total', tot)
