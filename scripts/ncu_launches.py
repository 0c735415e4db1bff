#!/usr/bin/env python
"""Aggregate an ncu launch list (gpu__time_duration.sum per launch) by kernel name."""
import collections
import csv
import re
import sys

lines = [l for l in open(sys.argv[1]) if not l.startswith('==')]
agg = collections.defaultdict(lambda: [0, 0.0])
tot = 0.0
for row in csv.DictReader(lines):
    if row.get('Metric Name') != 'gpu__time_duration.sum':
        continue
    name = re.sub(r'\(.*', '', row['Kernel Name'])
    name = re.sub(r'^void ', '', name)[:80]
    v = float(row['Metric Value'].replace(',', ''))
    u = row['Metric Unit']
    v *= {'ns': 1e-6, 'us': 1e-3, 'ms': 1.0, 's': 1e3}[u]
    agg[name][0] += 1
    agg[name][1] += v
    tot += v
print(f'total {tot:.3f} ms over {sum(c for c, _ in agg.values())} launches (cold-cache, serialised: compare shares)')
for k, (c, ms) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f'{ms:9.3f} ms {100 * ms / tot:5.1f}% {c:5d}  {k}')
