#!/usr/bin/env python
"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: per-kernel totals for the LAST step.
usage: summarize_launches.py launches.csv [n_steps_in_run]   (splits by the flush kernel's fill launches)"""
import csv, sys, re, collections
rows = []
with open(sys.argv[1]) as f:
    lines = [l for l in f if l.startswith('"')]
for r in csv.DictReader(lines):
    rows.append((int(r['ID']), r['Kernel Name'], float(r['Metric Value']), r['Grid Size'], r['Block Size']))
def short(n):
    n = re.sub(r'\(.*', '', n)
    n = re.sub(r'<.*', '', n)
    return n.replace('void ', '').replace('stgcn::', '').replace('at::native::', '')[:60]
# step boundary: the 160 MB flush = a FillFunctor launch with a huge grid
marks = [i for i, r in enumerate(rows) if 'FillFunctor' in r[1] and 'unsigned char' in r[1]]
print('launches', len(rows), 'flush marks at', marks)
lo = marks[0] if len(marks) >= 2 else 0
hi = marks[1] if len(marks) >= 2 else len(rows)
if len(sys.argv) > 2 and sys.argv[2] == 'all': lo, hi = 0, len(rows)
step = rows[lo + 1:hi]
agg = collections.OrderedDict()
for _, n, ns, g, b in step:
    k = short(n)
    a = agg.setdefault(k, [0, 0.0])
    a[0] += 1; a[1] += ns
tot = sum(a[1] for a in agg.values())
print('one step: %d launches, %.3f ms summed kernel time' % (len(step), tot / 1e6))
for k, a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print('%-62s %4d %9.1f us %5.1f%%' % (k, a[0], a[1] / 1e3, 100 * a[1] / tot))
