#!/usr/bin/env python
"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: per-kernel count, total
and average device time and share.  usage: ncu_launch_summary.py launches.csv"""
import collections
import csv
import io
import sys

lines = [l for l in open(sys.argv[1]) if not l.startswith('==')]
agg = collections.defaultdict(lambda: [0, 0.0])
for row in csv.DictReader(io.StringIO(''.join(lines))):
    if row.get('Metric Name') != 'gpu__time_duration.sum':
        continue
    scale = {'ns': 1e-6, 'us': 1e-3, 'ms': 1.0, 's': 1e3}.get(row['Metric Unit'], 1e-6)
    a = agg[row['Kernel Name'][:96]]
    a[0] += 1
    a[1] += float(row['Metric Value'].replace(',', '')) * scale
tot = sum(v[1] for v in agg.values())
print(f'total device time {tot:.2f} ms over {sum(v[0] for v in agg.values())} launches')
print(f'{"ms":>10} {"share":>6} {"n":>5} {"avg ms":>8}  kernel')
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1])[:15]:
    print(f'{v[1]:10.2f} {100 * v[1] / tot:5.1f}% {v[0]:5d} {v[1] / v[0]:8.3f}  {k}')
