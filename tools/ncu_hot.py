#!/usr/bin/env python
"""Hot SASS instructions (warp stall samples) of one launch in an .ncu-rep: ncu_hot.py report.ncu-rep [launch-index] [min-samples]"""
import csv, subprocess, sys, io
rep = sys.argv[1]; kid = sys.argv[2] if len(sys.argv) > 2 else '1'; thr = int(sys.argv[3]) if len(sys.argv) > 3 else 40
raw = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv', '--print-source', 'sass', '--kernel-id', ':::' + kid],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr = next(i for i, r in enumerate(rows) if len(r) > 3 and r[1] == 'Source')
h = rows[hdr]; sa = h.index('# Samples'); ie = h.index('Instructions Executed')
data = []
for r in rows[hdr + 1:]:
    if len(r) <= max(sa, ie): continue
    if r[1] == 'Source': break            # next launch
    data.append(r)
tot = sum(int(r[sa] or 0) for r in data)
print('total samples', tot, 'instructions', len(data))
for n, r in enumerate(data):
    s = int(r[sa] or 0)
    if s >= thr:
        print('%5d %6d %5.1f%% exec=%-8s %s' % (n, s, 100.0 * s / tot, r[ie], r[1][:100]))
if len(sys.argv) > 4:
    w = int(sys.argv[4])
    for a in range(0, len(data), w):
        s = sum(int(r[sa] or 0) for r in data[a:a + w])
        if s: print('[%5d..%5d) %6d %5.1f%% exec(first)=%s' % (a, a + w, s, 100.0 * s / tot, data[a][ie]))
