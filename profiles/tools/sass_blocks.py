"""group the SASS page of an ncu report into runs of equal execution count: python profiles/tools/sass_blocks.py rep.ncu-rep [kernel_index]"""
import sys, csv, subprocess
from collections import Counter
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
# split per kernel: a row starting with "Kernel Name"
starts = [i for i, r in enumerate(rows) if r and r[0] == "Kernel Name"]
which = int(sys.argv[2]) if len(sys.argv) > 2 else 0
lo = starts[which]; hi = starts[which + 1] if which + 1 < len(starts) else len(rows)
hdr = rows[lo + 1]; data = [r for r in rows[lo + 2:hi] if len(r) == len(hdr)]
ia, ie, is_ = hdr.index('Source'), hdr.index('Instructions Executed'), hdr.index('# Samples')
op = lambda t: (t.split()[1] if t.strip().startswith('@') else t.split()[0]).split('.')[0]
tot = sum(int(r[ie]) for r in data); ts = sum(int(r[is_]) for r in data)
print('total warp instr', tot, 'samples', ts, 'static', len(data))
blocks = []; cur = None
for idx, r in enumerate(data):
    e = int(r[ie]); s = int(r[is_])
    if cur and cur['e'] == e:
        cur['n'] += 1; cur['s'] += s; cur['ops'].append(op(r[ia]))
    else:
        cur = {'i': idx, 'e': e, 'n': 1, 's': s, 'ops': [op(r[ia])]}; blocks.append(cur)
for b in blocks:
    if b['e'] * b['n'] > tot * 0.01:
        print(f"@{b['i']:5d} n={b['n']:4d} exec={b['e']:9d} share={b['e']*b['n']/tot*100:5.1f}% samp={b['s']/ts*100:5.1f}% ", dict(Counter(b['ops']).most_common(7)))
