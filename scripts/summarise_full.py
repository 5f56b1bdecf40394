#!/usr/bin/env python
"""Summarise an `ncu --set full` capture (.ncu-rep) per launch: duration, DRAM bytes, tensor-pipe activity.

  python scripts/summarise_full.py gpurun_out/convs_r01c.ncu-rep > profiles/r01_convs_full.md
  python scripts/summarise_full.py gpurun_out/head_r02.ncu-rep --tags gpurun_out/head_tags.json > profiles/r02_full_aspp.md

``--tags`` (scripts/profile_head.py): the ordered list of conv launches inside the captured range; adds the launch's
name, its nominal dense FLOPs and TFLOP/s (nominal / ncu duration — ncu durations are cold-cache and at the profiler's
clocks, so read them as evidence of the pipe utilisation, not as bench numbers).
"""
import json
import csv
import io
import subprocess
import sys

COLS = ['Kernel Name', 'Grid Size', 'gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active',
        'TPC.TriageCompute.sm__pipe_tensor_cycles_active_realtime.avg.pct_of_peak_sustained_elapsed',
        'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'launch__registers_per_thread',
        'sm__warps_active.avg.pct_of_peak_sustained_active', 'lts__t_bytes.sum',
        'l1tex__m_xbar2l1tex_read_bytes.sum']


def main():
    path = sys.argv[1]
    tags = json.load(open(sys.argv[sys.argv.index('--tags') + 1])) if '--tags' in sys.argv else None
    out = subprocess.check_output(['ncu', '-i', path, '--page', 'raw', '--csv'], stderr=subprocess.DEVNULL).decode()
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    idx = {c: hdr.index(c) for c in COLS if c in hdr}
    print('# ncu --set full capture: %s (%d launches)\n' % (path, len(data)))
    if tags is not None and len(tags) != len(data):
        print('(tag list has %d entries, capture %d launches: tags dropped)\n' % (len(tags), len(data)))
        tags = None
    print('| # | kernel | grid | us | dram rd MB | dram wr MB | tensor pipe % (active) | tensor realtime % (elapsed) | '
          'dram % | L2 bytes MB |' + (' launch | nominal GFLOP | nominal TFLOP/s |' if tags else ''))
    print('|---|---|---|---|---|---|---|---|---|---|' + ('---|---|---|' if tags else ''))

    def get(r, c, scale=1.0):
        if c not in idx:
            return ''
        v = r[idx[c]].replace(',', '')
        try:
            return '%.1f' % (float(v) * scale)
        except ValueError:
            return v

    def unit_scale(c, want):
        if c not in idx:
            return 1.0
        u = units[idx[c]]
        return {'byte': 1e-6, 'Kbyte': 1e-3, 'Mbyte': 1.0, 'Gbyte': 1e3, 'ns': 1e-3, 'us': 1.0, 'ms': 1e3}.get(u, 1.0)

    for i, r in enumerate(data):
        name = r[idx['Kernel Name']]
        name = name.split('::')[-1].split('(')[0]
        extra = ''
        if tags:
            try:
                us = float(get(r, 'gpu__time_duration.sum', unit_scale('gpu__time_duration.sum', 'us')))
                extra = ' %s | %.1f | %.0f |' % (tags[i]['tag'], tags[i]['flops'] / 1e9, tags[i]['flops'] / us / 1e6)
            except ValueError:
                extra = ' %s | | |' % tags[i]['tag']
        print('| %d | %s | %s | %s | %s | %s | %s | %s | %s | %s |%s' % (
            i, name, r[idx['Grid Size']].replace(', 1, 1', ''),
            get(r, 'gpu__time_duration.sum', unit_scale('gpu__time_duration.sum', 'us')),
            get(r, 'dram__bytes_read.sum', unit_scale('dram__bytes_read.sum', 'MB')),
            get(r, 'dram__bytes_write.sum', unit_scale('dram__bytes_write.sum', 'MB')),
            get(r, 'sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active'),
            get(r, 'TPC.TriageCompute.sm__pipe_tensor_cycles_active_realtime.avg.pct_of_peak_sustained_elapsed'),
            get(r, 'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed'),
            get(r, 'lts__t_bytes.sum', unit_scale('lts__t_bytes.sum', 'MB')), extra))


if __name__ == '__main__':
    main()
