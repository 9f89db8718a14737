#!/usr/bin/env python
"""Pretty-print gpurun_out/configs.log (output of scripts/bench_configs.py)."""
import json
import sys
for l in open(sys.argv[1] if len(sys.argv) > 1 else 'gpurun_out/configs.log'):
    if not l.startswith('{'):
        print(l[:300].rstrip())
        continue
    d = json.loads(l)
    if 'error' in d or 'variant' in d:
        print(d)
        continue
    print(d['case'], 'B', d['B'], 'ms/solve', round(d.get('ms_per_solve', 0), 2), 'iters', d.get('iterations'), 'prob/s',
          round(d.get('problems_per_s', 0)))
    for k, v in d['kernels'].items():
        print('    %-26s avg %.3f ms  %s' % (k, v['avg_ms'], ('%.0f GB/s %.1f%%' % (v['GBps'], 100 * v['frac_of_measured_hbm']))
                                             if 'GBps' in v else ''))
