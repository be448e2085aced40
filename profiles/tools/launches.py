"""print the last N launches of an ncu gpu__time_duration launch list: python profiles/tools/launches.py file.csv N"""
import csv, sys
lines = [l for l in open(sys.argv[1]) if not l.startswith('==')]
L = [x for x in csv.DictReader(lines) if x.get('Metric Name') == 'gpu__time_duration.sum']
def t(x):
    return float(x['Metric Value'].replace(',', '')) * {'ns': 1e-6, 'us': 1e-3, 'ms': 1, 's': 1e3}[x['Metric Unit']]
tail = L[-int(sys.argv[2]):]
print('total ms', sum(t(x) for x in tail))
for x in tail:
    print(f"{t(x):8.3f}  {x['Kernel Name'][:100]}")
