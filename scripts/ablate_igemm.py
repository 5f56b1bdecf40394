#!/usr/bin/env python
"""Timing ablations of igemm_kernel on the small-channel backbone shapes (results are garbage, only the time matters):
B200SEG_IGEMM_ABLATE bitmask — 1 no A loads, 2 no B loads, 4 no MMAs, 8 epilogue does no work, 16 no BN statistics.
One process per mask (the switch is read once); prints us per launch of fprop with fused statistics.
The switches are NOT in the shipped kernel: apply profiles/dropped/r02_igemm_ablation_switches.patch and rebuild first
(results: profiles/r02_igemm_ablation.md).

  python scripts/ablate_igemm.py            # driver: runs itself once per mask
"""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
SHAPES = [('l1.conv1 256>64 1x1', 16, 128, 128, 256, 64, 1), ('l1.conv2 64>64 3x3', 16, 128, 128, 64, 64, 3),
          ('l1.conv3 64>256 1x1', 16, 128, 128, 64, 256, 1), ('l2.conv2 128>128 3x3', 16, 64, 64, 128, 128, 3),
          ('l3.conv2 256>256 3x3', 16, 32, 32, 256, 256, 3)]
MASKS = [int(m) for m in os.environ.get('ABLATE_MASKS', '0 16 8 24 4 1 2 3 7 15 31').split()]


def child():
    import torch
    from pytorch_segmentation_b200 import engine as E
    flush = torch.zeros(64 << 20, dtype=torch.float32, device='cuda')
    row = []
    for name, n, h, w, c, k, ks in SHAPES:
        x = E.Act(torch.randn(n, h, w, c, device='cuda').to(torch.bfloat16))
        wt = torch.randn(k, c, ks, ks, device='cuda') * 0.05
        pw = E.PackedWeight().update(wt, need_dgrad=False)
        y = E.Act.empty(n, h, w, k, 'cuda')
        stats = torch.zeros(2 * k, device='cuda')
        fn = lambda: E.conv_fprop([(x, pw, 1, 1)], y, stats=stats)
        for _ in range(3):
            fn()
        tot = 0.0
        for _ in range(10):
            flush.sum()
            torch.cuda._sleep(1500000)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            fn()
            e1.record()
            torch.cuda.synchronize()
            tot += e0.elapsed_time(e1)
        row.append(tot / 10 * 1e3)
    print('ROW %s ' % os.environ.get('B200SEG_IGEMM_ABLATE', '0') + ' '.join('%.1f' % v for v in row), flush=True)


def main():
    if len(sys.argv) > 1 and sys.argv[1] == 'child':
        return child()
    print('mask | ' + ' | '.join(s[0] for s in SHAPES))
    for m in MASKS:
        env = dict(os.environ, B200SEG_IGEMM_ABLATE=str(m))
        out = subprocess.run([sys.executable, os.path.abspath(__file__), 'child'], env=env, capture_output=True, text=True, timeout=300)
        rows = [ln for ln in out.stdout.splitlines() if ln.startswith('ROW ')]
        print(rows[0] if rows else 'mask %d FAILED: %s' % (m, out.stderr[-500:]))


if __name__ == '__main__':
    main()
