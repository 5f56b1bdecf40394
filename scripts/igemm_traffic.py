#!/usr/bin/env python
"""profiles/igemm_traffic.json from an `ncu --set full` summary of the igemm_kernel launches of one training step
(scripts/summarise_full.py table): DRAM bytes per launch (read by bench.py as roofline.traffic), time-weighted tensor-pipe
activity.   python scripts/igemm_traffic.py profiles/r02_full_igemm.md > profiles/igemm_traffic.json"""
import json
import re
import sys

rows = []
for ln in open(sys.argv[1]):
    m = re.match(r'\| (\d+) \| [^|]*\| \([\d, ]+\) \| ([\d.]+) \| ([\d.]+) \| ([\d.]+) \| ([\d.]+) \|', ln)
    if m:
        rows.append(tuple(float(m.group(i)) for i in (2, 3, 4, 5)))
n = len(rows)
us = sum(r[0] for r in rows)
print(json.dumps({
    'dram_bytes_per_launch': sum(r[1] + r[2] for r in rows) * 1e6 / n,
    'dram_read_bytes_per_launch': sum(r[1] for r in rows) * 1e6 / n,
    'dram_write_bytes_per_launch': sum(r[2] for r in rows) * 1e6 / n,
    'launches': n,
    'us_total_under_ncu': round(us, 1),
    'tensor_pipe_active_pct_time_weighted': sum(r[0] * r[3] for r in rows) / us,
    'source': '%s (round 2 final state, ncu --set full --clock-control none, the %d igemm_kernel launches of ONE eager '
              'training step, scripts/gpu_profile_r02.sh)' % (sys.argv[1], n)}, indent=1))
