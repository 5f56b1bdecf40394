#!/usr/bin/env python
"""Summarise an ncu launch list (gpu__time_duration.sum CSV) into per-kernel totals for ONE training step.

  python scripts/summarise_ncu.py gpurun_out/launches_r01.csv [--steps N] > profiles/r01_launches.md

The bench command profiled runs N identical steps (3 warm-up + 2 enqueue-timing + 2 timed + 2 e2e warm-up + 2 e2e + 2 profiled
= 13 by default); totals are divided by N.  The spin kernels bench.py puts ahead of its profiled steps are left out.  ncu times are cold-cache and serialised: read SHARES.
"""
import csv
import re
import sys
from collections import defaultdict


def main():
    path = sys.argv[1]
    steps = 13
    if '--steps' in sys.argv:
        steps = int(sys.argv[sys.argv.index('--steps') + 1])
    rows = []
    with open(path, newline='') as f:
        lines = [l for l in f if not l.startswith('==')]
    rd = csv.DictReader(lines)
    for r in rd:
        if r.get('Metric Name') != 'gpu__time_duration.sum':
            continue
        v = float(r['Metric Value'].replace(',', ''))
        unit = r['Metric Unit']
        us = v / 1e3 if unit in ('ns', 'nsecond') else v * {'us': 1, 'usecond': 1, 'ms': 1e3, 'msecond': 1e3}.get(unit, 1)
        if 'spin_kernel' in r['Kernel Name']:
            continue
        rows.append((r['Kernel Name'], us))
    agg = defaultdict(lambda: [0.0, 0])
    for name, us in rows:
        if 'b200seg' in name:
            m = re.search(r'(\w+_kernel)', name)
            short = 'b200seg:' + (m.group(1) if m else name[:40])
        else:
            short = re.sub(r'\(.*$', '', name)
            short = re.sub(r'<.*$', '', short)
        agg[short][0] += us
        agg[short][1] += 1
    total = sum(v[0] for v in agg.values())
    ours = sum(v[0] for k, v in agg.items() if k.startswith('b200seg:'))
    print('# ncu launch list summary: %s' % path)
    print()
    print('%d launches over %d steps; per step: %.2f ms of kernel time (cold-cache, serialised), '
          '%.2f ms (%.1f %%) in libb200seg kernels, %.0f launches' % (len(rows), steps, total / steps / 1e3,
                                                                     ours / steps / 1e3, 100 * ours / total,
                                                                     len(rows) / steps))
    print()
    print('| kernel | launches/step | us/step | share | avg us |')
    print('|---|---|---|---|---|')
    for k, (us, n) in sorted(agg.items(), key=lambda kv: -kv[1][0])[:45]:
        print('| %s | %.1f | %.1f | %.1f %% | %.1f |' % (k[:90], n / steps, us / steps, 100 * us / total, us / n))


OURS = set()
if __name__ == '__main__':
    main()
