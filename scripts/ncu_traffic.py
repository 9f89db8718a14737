#!/usr/bin/env python
"""Writes the per-launch ncu counters of one kernel into profiles/traffic.json, keyed by the C-ABI entry point and tagged
with the hash of the kernel sources, so that bench.py reports `roofline.traffic` only for the build it measures.

    ncu -i gpurun_out/<rep>.ncu-rep --page raw --csv > raw.csv
    python scripts/ncu_traffic.py raw.csv <kernel-name regex> <entry point> "<where the capture came from>"
"""
import csv
import json
import os
import re
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from torchimize_b200 import _lib                                   # noqa: E402

UNIT = {'byte': 1, 'Kbyte': 1e3, 'Mbyte': 1e6, 'Gbyte': 1e9, 'Tbyte': 1e12, 'ns': 1e-6, 'us': 1e-3, 'ms': 1, 's': 1e3}


def main():
    raw, pattern, entry, source = sys.argv[1:5]
    rows = list(csv.reader(open(raw)))
    hdr, units = rows[0], rows[1]
    col = {h: i for i, h in enumerate(hdr)}
    sel = [r for r in rows[2:] if re.search(pattern, r[col['Kernel Name']])]
    assert sel, f'no launch matches {pattern}'

    def avg(name, scale=True):
        i = col[name]
        f = UNIT.get(units[i], 1) if scale else 1
        return sum(float(r[i]) * f for r in sel) / len(sel)

    rec = {'kernel': sel[0][col['Kernel Name']][:160], 'launches': len(sel),
           'dram_bytes': avg('dram__bytes_read.sum') + avg('dram__bytes_write.sum'),
           'dram_bytes_read': avg('dram__bytes_read.sum'), 'dram_bytes_write': avg('dram__bytes_write.sum'),
           'duration_ms': avg('gpu__time_duration.sum'),
           'fp64_pipe_pct': avg('sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_elapsed', False),
           'issue_active_pct': avg('sm__inst_executed.avg.pct_of_peak_sustained_elapsed', False) if 'sm__inst_executed.avg.pct_of_peak_sustained_elapsed' in col else None,
           'warp_instructions': avg('smsp__inst_executed.sum', False),
           'registers_per_thread': avg('launch__registers_per_thread', False),
           'source_hash': _lib.source_hash(), 'source': source}
    path = os.path.join(ROOT, 'profiles', 'traffic.json')
    try:
        tr = json.load(open(path))
    except Exception:                                              # noqa: BLE001
        tr = {}
    tr['_comment'] = ('per-launch ncu counters (--set full --clock-control none) of the dominant kernels, written by scripts/ncu_traffic.py; '
                      'bench.py copies dram_bytes into roofline.traffic only when source_hash matches the benched build')
    tr[entry] = rec
    json.dump(tr, open(path, 'w'), indent=1)
    print(entry, json.dumps(rec, indent=1))


if __name__ == '__main__':
    main()
