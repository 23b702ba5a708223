#!/usr/bin/env python
"""Per-kernel counts of the SASS mnemonics that prove tcgen05 / TMEM / TMA use (B200_PROFILING.md), from
`cuobjdump -sass` of the built library.   python tools/sass_counts.py > profiles/r02_sass_counts.txt"""
import collections
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, 'thesis-clip-nerf_b200', 'lib', 'libclipnerf_b200.so')
WANT = ['UTCHMMA', 'UTCQMMA', 'UTCBAR', 'LDTM', 'STTM', 'UTCATOMSWS', 'UBLKCP', 'UBLKRED', 'UBLKPF', 'SYNCS', 'REDG', 'RED.', 'ATOMG', 'HMMA', 'MUFU']
out = subprocess.run(['cuobjdump', '-sass', LIB], capture_output=True, text=True, check=True).stdout
counts, name = collections.OrderedDict(), None
for line in out.splitlines():
    m = re.match(r'\s*Function : (\S+)', line)
    if m:
        name = subprocess.run(['c++filt', m.group(1)], capture_output=True, text=True).stdout.strip().split('(')[0]
        counts[name] = collections.Counter()
        continue
    if name is None:
        continue
    m = re.match(r'\s*/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)', line)
    if m:
        op = m.group(1)
        counts[name]['_total'] += 1
        for w in WANT:
            if op.startswith(w):
                counts[name][w] += 1
print(f'# cuobjdump -sass {os.path.relpath(LIB, ROOT)} (sm_100a): instruction counts per kernel')
print('# UTCHMMA = tcgen05.mma kind::f16, LDTM / STTM = tcgen05.ld / st, UTCBAR = tcgen05.commit, UBLKCP = cp.async.bulk (TMA),')
print('# UBLKRED = cp.reduce.async.bulk, UBLKPF = bulk prefetch, SYNCS = mbarrier ops, HMMA = legacy mma.sync (none expected)')
hdr = ['kernel', 'instrs'] + WANT
print(' | '.join(hdr))
tot = collections.Counter()
for k, c in counts.items():
    print(' | '.join([k.replace('cn::', ''), str(c['_total'])] + [str(c[w]) for w in WANT]))
    tot.update(c)
print(' | '.join(['TOTAL', str(tot['_total'])] + [str(tot[w]) for w in WANT]))
