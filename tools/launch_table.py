"""per-kernel totals of an ncu launch list (`--metrics gpu__time_duration.sum --csv`); .csv or .csv.gz"""
import collections
import csv
import gzip
import sys

path = sys.argv[1]
f = gzip.open(path, 'rt', errors='replace') if path.endswith('.gz') else open(path, errors='replace')
rows = [r for r in csv.reader(f) if len(r) > 10]
hdr = rows[0]
ik, iv, iu = hdr.index('Kernel Name'), hdr.index('Metric Value'), hdr.index('Metric Unit')
tot, cnt = collections.defaultdict(float), collections.Counter()
for r in rows[1:]:
    v, u = float(r[iv].replace(',', '')), r[iu]
    us = v / 1e3 if u.startswith('n') else v if u.startswith('u') else v * 1e3
    k = r[ik].replace('void ', '')[:72]
    tot[k] += us
    cnt[k] += 1
s = sum(tot.values())
for k, v in sorted(tot.items(), key=lambda kv: -kv[1])[:30]:
    print(f'{v:10.1f} us {100 * v / s:5.1f} %  x{cnt[k]:5d}  {k}')
print(f'{s:10.1f} us total, {sum(cnt.values())} launches')
