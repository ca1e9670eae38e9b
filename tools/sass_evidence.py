#!/usr/bin/env python
"""Per-kernel counts of the SASS mnemonics that prove a Blackwell-native path (tcgen05 -> UTC*MMA, tcgen05.ld -> LDTM, TMA ->
UTMALDG / UTMASTG, legacy mma.sync -> HMMA) in the shipped libstgcn_b200.so.  Writes profiles/sass_tcgen05.txt.

    python tools/sass_evidence.py
"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, 'stgcn_b200', 'libstgcn_b200.so')
PAT = {'UTC*MMA (tcgen05.mma)': re.compile(r'\bUTC\w*MMA'), 'LDTM (tcgen05.ld)': re.compile(r'\bLDTM'),
       'STTM (tcgen05.st)': re.compile(r'\bSTTM'), 'UTMALDG (TMA load)': re.compile(r'\bUTMALDG'),
       'UTMASTG (TMA store)': re.compile(r'\bUTMASTG'), 'UTCBAR (tcgen05.commit)': re.compile(r'\bUTCBAR'),
       'SYNCS (mbarrier)': re.compile(r'\bSYNCS'), 'HMMA (mma.sync)': re.compile(r'\bHMMA'), 'RED/ATOM global': re.compile(r'\b(RED|ATOMG)\b')}


def main():
    sass = subprocess.run(['cuobjdump', '-sass', LIB], capture_output=True, text=True, check=True).stdout
    counts = collections.OrderedDict()
    cur = None
    for line in sass.splitlines():
        m = re.search(r'Function : (\S+)', line)
        if m:
            cur = subprocess.run(['c++filt', m.group(1)], capture_output=True, text=True).stdout.strip()
            cur = re.sub(r'\(.*', '', cur)
            counts.setdefault(cur, collections.Counter())
            continue
        if cur:
            for name, pat in PAT.items():
                if pat.search(line):
                    counts[cur][name] += 1
    out = ['# cuobjdump -sass stgcn_b200/libstgcn_b200.so: SASS mnemonic counts per kernel (tools/sass_evidence.py)',
           '# arch: ' + ', '.join(sorted(set(re.findall(r'arch = (sm_\w+)', sass))))]
    tot = collections.Counter()
    for k, c in counts.items():
        tot.update(c)
        if c['UTC*MMA (tcgen05.mma)'] or c['UTMALDG (TMA load)'] or c['HMMA (mma.sync)']:
            out.append('%-90s %s' % (k[:90], '  '.join('%s=%d' % (n.split(' ')[0], c[n]) for n in PAT if c[n])))
    out.append('TOTAL over %d kernels: %s' % (len(counts), '  '.join('%s=%d' % (n, tot[n]) for n in PAT)))
    path = os.path.join(ROOT, 'profiles', 'sass_tcgen05.txt')
    open(path, 'w').write('\n'.join(out) + '\n')
    print('\n'.join(out[-12:]))


if __name__ == '__main__':
    sys.exit(main())
