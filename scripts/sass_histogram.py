#!/usr/bin/env python
"""Per-kernel opcode histogram of libb200seg.so from `cuobjdump -sass` — the evidence that the convolution kernels are
Blackwell-native: UTC*MMA = tcgen05.mma, LDTM / STTM = tcgen05.ld / .st, UTMALDG / UTMASTG = TMA tensor loads / stores,
SYNCS = mbarrier traffic; HMMA (the legacy mma.sync path) must not appear.

  python scripts/sass_histogram.py > profiles/sass_r02.txt
"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, 'pytorch_segmentation_b200', 'libb200seg.so')
KEY = ('UTCHMMA', 'UTCQMMA', 'UTCBAR', 'UTCCP', 'LDTM', 'STTM', 'UTMALDG', 'UTMASTG', 'UTMAPF', 'UBLKCP', 'SYNCS', 'HMMA',
       'HGMMA', 'LDGSTS', 'ATOMG', 'REDG', 'RED', 'ATOMS', 'LDG', 'STG', 'LDS', 'STS', 'SHFL', 'FFMA', 'MUFU', 'ACQBULK')


def main():
    out = subprocess.check_output(['cuobjdump', '-sass', LIB]).decode()
    kernels, cur = collections.OrderedDict(), None
    for line in out.splitlines():
        m = re.search(r'Function : (\S+)', line)
        if m:
            name = subprocess.check_output(['c++filt', m.group(1)]).decode().strip()
            name = re.sub(r'\(anonymous namespace\)::', '', name)
            name = re.sub(r'\(.*$', '', name).replace('b200seg::', '').replace('void ', '')
            cur = kernels.setdefault(name, collections.Counter())
            continue
        m = re.match(r'\s+/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)', line)
        if m and cur is not None:
            cur[m.group(1).split('.')[0]] += 1
            cur['_total'] += 1
    print('# SASS opcode histogram of libb200seg.so (cuobjdump -sass, sm_100a), %d kernels' % len(kernels))
    print('# columns: kernel | instructions | Blackwell / memory opcodes that appear')
    tot = collections.Counter()
    for name, c in kernels.items():
        keys = ['%s %d' % (k, c[k]) for k in KEY if c.get(k)]
        print('%-72s %6d  %s' % (name[:72], c['_total'], ', '.join(keys)))
        tot.update({k: v for k, v in c.items() if k in KEY})
    print()
    print('# whole library: ' + ', '.join('%s %d' % (k, tot[k]) for k in KEY if tot.get(k)))
    if tot.get('HMMA') or tot.get('HGMMA'):
        print('# WARNING: legacy tensor-core opcodes present')
        sys.exit(1)


if __name__ == '__main__':
    main()
