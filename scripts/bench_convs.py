#!/usr/bin/env python
"""Per-shape microbenchmark of the tcgen05 conv kernels (BASELINE.json configs[2]: the ASPP block isolated, plus
the decoder / cls / backbone shapes).  fprop (with fused BN statistics), dgrad and wgrad are timed separately
with CUDA events on the launching stream; a 256 MB read between iterations flushes the 126 MB L2.

  python scripts/bench_convs.py [--iters 20] [--set head|backbone|all] [--json out.json]

TFLOP/s are nominal dense (2*N*P*Q*K*C*R*S, padding taps counted), fractions are of the measured sustained
bf16 peak in MEASURED_PEAKS.json; GB/s are algorithmic bytes (input + output + weights once) of the measured
copy bandwidth.
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

HEAD = [
    # name, n, h, w, c, k, ksize, dil, stride
    ('aspp.3x3.d6', 16, 32, 32, 2048, 256, 3, 6, 1),
    ('aspp.3x3.d12', 16, 32, 32, 2048, 256, 3, 12, 1),
    ('aspp.3x3.d18', 16, 32, 32, 2048, 256, 3, 18, 1),
    ('aspp.1x1', 16, 32, 32, 2048, 256, 1, 1, 1),
    ('aspp.project', 16, 32, 32, 1280, 256, 1, 1, 1),
    ('dec.project', 16, 128, 128, 256, 128, 1, 1, 1),
    ('cls_conv(24)', 16, 128, 128, 384, 24, 3, 1, 1),
]
BACKBONE = [
    ('l1.conv1', 16, 128, 128, 256, 64, 1, 1, 1),
    ('l1.conv2', 16, 128, 128, 64, 64, 3, 1, 1),
    ('l1.conv3', 16, 128, 128, 64, 256, 1, 1, 1),
    ('l2.conv1', 16, 64, 64, 512, 128, 1, 1, 1),
    ('l2.conv2', 16, 64, 64, 128, 128, 3, 1, 1),
    ('l2.conv3', 16, 64, 64, 128, 512, 1, 1, 1),
    ('l3.conv1', 16, 32, 32, 1024, 256, 1, 1, 1),
    ('l3.conv2', 16, 32, 32, 256, 256, 3, 1, 1),
    ('l3.conv3', 16, 32, 32, 256, 1024, 1, 1, 1),
    ('l4.conv1', 16, 32, 32, 2048, 512, 1, 1, 1),
    ('l4.conv2.d2', 16, 32, 32, 512, 512, 3, 2, 1),
    ('l4.conv3', 16, 32, 32, 512, 2048, 1, 1, 1),
    ('l2.0.conv2.s2', 16, 128, 128, 128, 128, 3, 1, 2),
    ('l2.0.down.s2', 16, 128, 128, 256, 512, 1, 1, 2),
    ('l3.0.conv2.s2', 16, 64, 64, 256, 256, 3, 1, 2),
    ('l3.0.down.s2', 16, 64, 64, 512, 1024, 1, 1, 2),
]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--iters', type=int, default=20)
    ap.add_argument('--set', default='head')
    ap.add_argument('--json', default=None)
    args = ap.parse_args()
    from pytorch_segmentation_b200 import engine as E
    pk = json.load(open(os.path.join(ROOT, 'MEASURED_PEAKS.json'))) if os.path.exists(
        os.path.join(ROOT, 'MEASURED_PEAKS.json')) else {'bf16_tflops_sustained': 1400.0, 'hbm_gbs': 6650.0}
    peak_tf, peak_gb = pk['bf16_tflops_sustained'], pk['hbm_gbs']
    flush = torch.zeros(64 << 20, dtype=torch.float32, device='cuda')      # 256 MB, read (not written): clean lines
    cases = {'head': HEAD, 'backbone': BACKBONE, 'all': HEAD + BACKBONE}[args.set]

    def timed(fn):
        for _ in range(3):
            fn()
        tot = 0.0
        for _ in range(args.iters):
            flush.sum()
            torch.cuda._sleep(1500000)      # ~0.8 ms of GPU spin: host-side launch preparation is not timed
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            fn()
            e1.record()
            torch.cuda.synchronize()
            tot += e0.elapsed_time(e1)
        return tot / args.iters * 1e3      # us

    # calibration: plain library write / read / copy of 134 MB (what a 16x128x128x256 bf16 tensor costs to stream)
    cal = torch.empty(16 * 128 * 128 * 256, dtype=torch.bfloat16, device='cuda')
    cal2 = torch.empty_like(cal)
    for nm, fn, nbytes in (('write 134MB', lambda: cal.zero_(), cal.numel() * 2),
                           ('read 134MB', lambda: cal.view(torch.float32).sum(), cal.numel() * 2),
                           ('copy 134MB', lambda: cal2.copy_(cal), cal.numel() * 4)):
        us = timed(fn)
        print('calibration %-12s %8.1f us %8.0f GB/s' % (nm, us, nbytes / us / 1e3))
    del cal, cal2
    out = []
    print('%-16s %-6s %9s %9s %6s %8s %6s' % ('conv', 'pass', 'us', 'TFLOP/s', 'frac', 'GB/s', 'frac'))
    for name, n, h, w, c, k, ks, dil, stride in cases:
        oh, ow = h // stride, w // stride
        x = E.Act(torch.randn(n, h, w, c, device='cuda').to(torch.bfloat16))
        wt = torch.randn(k, c, ks, ks, device='cuda') * (2.0 / (c * ks * ks)) ** 0.5
        pw = E.PackedWeight().update(wt, need_dgrad=True)
        y = E.Act.empty(n, oh, ow, k, 'cuda')
        dy = E.Act(torch.randn(n, oh, ow, k, device='cuda').to(torch.bfloat16))
        dx = E.Act.empty(n, h, w, c, 'cuda')
        stats = torch.zeros(2 * k, device='cuda')
        flops = 2.0 * n * oh * ow * k * c * ks * ks
        byts = {'fprop0': 2.0 * (x.t.numel() + y.t.numel() + wt.numel()), 'fprop': 2.0 * (x.t.numel() + y.t.numel() + wt.numel()),
                'dgrad': 2.0 * (x.t.numel() + y.t.numel() + wt.numel()),
                'wgrad': 2.0 * (x.t.numel() + y.t.numel()) + 4.0 * wt.numel()}
        fns = {'fprop0': lambda: E.conv_fprop([(x, pw, dil, stride)], y),
               'fprop': lambda: E.conv_fprop([(x, pw, dil, stride)], y, stats=stats),
               'dgrad': lambda: E.conv_dgrad([(dy, pw, dil, stride)], dx),
               'wgrad': lambda: E.conv_wgrad(dy, x, k, ks, ks, dil, stride=stride)}
        for what, fn in fns.items():
            us = timed(fn)
            tf, gb = flops / us / 1e6, byts[what] / us / 1e3
            print('%-16s %-6s %9.1f %9.1f %6.3f %8.0f %6.3f' % (name, what, us, tf, tf / peak_tf, gb, gb / peak_gb))
            out.append({'conv': name, 'pass': what, 'shape': [n, h, w, c, k, ks, dil, stride], 'us': us, 'tflops': tf,
                        'frac_tensor': tf / peak_tf, 'gbs': gb, 'frac_hbm': gb / peak_gb})
    if args.json:
        json.dump({'peaks': pk, 'rows': out}, open(args.json, 'w'), indent=1)


if __name__ == '__main__':
    main()
