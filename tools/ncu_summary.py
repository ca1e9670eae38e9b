#!/usr/bin/env python
"""Compact per-launch summary of an .ncu-rep (run where ncu is installed): ncu_summary.py report.ncu-rep > profiles/x.txt"""
import csv, subprocess, sys, io
raw = subprocess.run(['ncu', '-i', sys.argv[1], '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units, data = rows[0], rows[1], rows[2:]
idx = {h: i for i, h in enumerate(hdr)}
cols = [('Kernel Name', 'kernel'), ('Grid Size', 'grid'), ('gpu__time_duration.sum', 'time'),
        ('dram__bytes_read.sum', 'dram_rd'), ('dram__bytes_write.sum', 'dram_wr'),
        ('gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'dram_%'),
        ('sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active', 'tensor_%'),
        ('l1tex__m_xbar2l1tex_read_bytes.sum', 'l2->sm_rd'), ('launch__registers_per_thread', 'regs'),
        ('launch__shared_mem_per_block_dynamic', 'dyn_smem'), ('sm__warps_active.avg.pct_of_peak_sustained_active', 'warps_%')]
print('# ' + sys.argv[1])
print(' | '.join('%s[%s]' % (n, units[idx[c]]) if c in idx else n for c, n in cols))
for d in data:
    out = []
    for c, n in cols:
        v = d[idx[c]] if c in idx else '-'
        if c == 'Kernel Name':
            v = v.split('(')[0].replace('void ', '').replace('stgcn::', '')[:40]
        out.append(v)
    print(' | '.join(out))
