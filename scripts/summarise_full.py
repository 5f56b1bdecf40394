#!/usr/bin/env python
"""Summarise an `ncu --set full` capture (.ncu-rep) per launch: duration, DRAM bytes, tensor-pipe activity.

  python scripts/summarise_full.py gpurun_out/convs_r01c.ncu-rep > profiles/r01_convs_full.md
"""
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
    out = subprocess.check_output(['ncu', '-i', path, '--page', 'raw', '--csv'], stderr=subprocess.DEVNULL).decode()
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    idx = {c: hdr.index(c) for c in COLS if c in hdr}
    print('# ncu --set full capture: %s (%d launches)\n' % (path, len(data)))
    print('| # | kernel | grid | us | dram rd MB | dram wr MB | tensor pipe % (active) | tensor realtime % (elapsed) | '
          'dram % | L2 bytes MB |')
    print('|---|---|---|---|---|---|---|---|---|---|')

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
        print('| %d | %s | %s | %s | %s | %s | %s | %s | %s | %s |' % (
            i, name, r[idx['Grid Size']].replace(', 1, 1', ''),
            get(r, 'gpu__time_duration.sum', unit_scale('gpu__time_duration.sum', 'us')),
            get(r, 'dram__bytes_read.sum', unit_scale('dram__bytes_read.sum', 'MB')),
            get(r, 'dram__bytes_write.sum', unit_scale('dram__bytes_write.sum', 'MB')),
            get(r, 'sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active'),
            get(r, 'TPC.TriageCompute.sm__pipe_tensor_cycles_active_realtime.avg.pct_of_peak_sustained_elapsed'),
            get(r, 'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed'),
            get(r, 'lts__t_bytes.sum', unit_scale('lts__t_bytes.sum', 'MB'))))


if __name__ == '__main__':
    main()
