#!/usr/bin/env python
"""Per CUDA source line: warp instructions executed and stall samples, from
`ncu -i rep --page source --print-source cuda,sass --csv`.  usage: ncu_cuda_lines.py file.csv [top]"""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 30
out = []
fname, hdr, cur = None, None, None
for r in rows:
    if not r:
        continue
    if r[0] == 'File Path':
        fname = r[1].split('/')[-1]
        continue
    if r[0] == 'Function Name':
        continue
    if r[0] == 'Line No':
        hdr = {h: i for i, h in enumerate(r)}
        # two 'Source' columns: first is CUDA, second SASS
        continue
    if hdr is None:
        continue
    if r[0] != '':
        cur = [fname, r[0], r[1].strip(), 0, 0]
        out.append(cur)
    elif cur is not None:
        try:
            cur[3] += int(r[hdr['Instructions Executed']] or 0)
            cur[4] += int(r[hdr['# Samples']] or 0)
        except (ValueError, IndexError):
            pass
ti = sum(o[3] for o in out) or 1
ts = sum(o[4] for o in out) or 1
print(f'total warp instructions {ti}, samples {ts}')
for o in sorted(out, key=lambda o: -o[3])[:top]:
    print(f'{100 * o[3] / ti:5.1f}% inst {100 * o[4] / ts:5.1f}% smp  {o[0]}:{o[1]}  {o[2][:90]}')
