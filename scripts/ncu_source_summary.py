#!/usr/bin/env python
"""Summarise `ncu --page source --csv` output: stall-reason totals, instruction mix and the
hottest SASS lines of the first kernel in the file.  usage: ncu_source_summary.py file.csv [top]"""
import csv
import collections
import sys

path = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 25
rows = list(csv.reader(open(path)))
hdr_i = next(i for i, r in enumerate(rows) if r and r[0] == 'Address')
hdr = rows[hdr_i]
col = {h: i for i, h in enumerate(hdr)}
body = []
for r in rows[hdr_i + 1:]:
    if len(r) != len(hdr) or r[0] == 'Address':
        break
    body.append(r)
stalls = [h for h in hdr if h.startswith('stall_') and 'Not Issued' not in h]
tot = collections.Counter()
mix = collections.Counter()
samples = 0
for r in body:
    for s in stalls:
        tot[s] += int(r[col[s]] or 0)
    samples += int(r[col['# Samples']] or 0)
    op = r[col['Source']].split()
    op = [o for o in op if not o.startswith('@')]
    if op:
        mix[op[0].split('.')[0]] += int(r[col['Instructions Executed']] or 0)
print(f'{len(body)} SASS lines, {samples} samples')
print('stalls:', ', '.join(f'{k[6:]} {100 * v / max(1, samples):.1f}%' for k, v in tot.most_common(8)))
n_inst = sum(mix.values())
print('warp-instruction mix:', ', '.join(f'{k} {100 * v / n_inst:.1f}%' for k, v in mix.most_common(14)), f'(total {n_inst})')
print('hottest lines:')
for r in sorted(body, key=lambda r: -int(r[col['# Samples']] or 0))[:top]:
    why = max(stalls, key=lambda s: int(r[col[s]] or 0))
    print(f"  {int(r[col['# Samples']]):7d} {why[6:]:12s} {r[col['Source']].strip()[:90]}")
