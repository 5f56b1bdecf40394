#!/usr/bin/env python
"""Stand-alone timing of the head / stem bandwidth kernels at the BASELINE.json training shapes (batch 16, 512x512):

  upsample_fwd   16x32x32x256 -> channels [0,256) of the 16x128x128x304 decoder concat buffer, align_corners=True
  upsample_bwd   its adjoint
  maxpool_fwd    16x256x256x64 -> 16x128x128x64 (+ window positions)
  maxpool_bwd    its adjoint
  pack_multi     every conv weight of DeepLabV3+ (39 M fp32 -> bf16 fprop + dgrad packings), one multi-tensor call

CUDA events on the launching stream, an L2 flush (256 MB write) before every timed call, median of --reps.  ``--lib``
loads another build of the library (same-box A/B of a kernel change).  Prints one JSON line.
"""
import argparse
import json
import os
import statistics
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--lib', default=None)
    ap.add_argument('--reps', type=int, default=15)
    ap.add_argument('--batch', type=int, default=16)
    args = ap.parse_args()
    from pytorch_segmentation_b200 import _lib
    if args.lib:
        _lib.LIB_PATH = os.path.abspath(args.lib)
    import pytorch_segmentation_b200 as P
    from pytorch_segmentation_b200 import engine as E
    dev = 'cuda'
    n = args.batch
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)

    def timed(fn, nbytes):
        ts = []
        for _ in range(args.reps + 2):
            flush.fill_(1)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            fn()
            e1.record()
            torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1) * 1e3)
        us = statistics.median(ts[2:])
        return {'us': round(us, 1), 'algorithmic_MB': round(nbytes / 1e6, 1), 'GB_s': round(nbytes / us / 1e3, 0)}

    out = {}
    bf = torch.bfloat16
    # decoder upsample into the concat buffer
    x = E.Act(torch.randn(n, 32, 32, 256, device=dev).to(bf))
    cat = torch.zeros(n, 128, 128, 304, device=dev, dtype=bf)
    y = E.Act(cat, 0, 256)
    out['upsample_fwd'] = timed(lambda: E.upsample_fwd(x, y, align_corners=True), x.t.numel() * 2 + n * 128 * 128 * 256 * 2)
    dcat = torch.randn(n, 128, 128, 304, device=dev).to(bf)
    dy = E.Act(dcat, 0, 256)
    dx = E.Act.empty(n, 32, 32, 256, dev)
    out['upsample_bwd'] = timed(lambda: E.upsample_bwd(dy, dx, align_corners=True), x.t.numel() * 2 + n * 128 * 128 * 256 * 2)
    # UNet-like x2 (align_corners=True) and HRNet-like x2 / x4 (align_corners=False), full tensors
    for tag, (h, c, s, ac) in {'x2_ac': (64, 96, 2, True), 'x4_noac': (32, 32, 4, False)}.items():
        xs = E.Act(torch.randn(n, h, h, c, device=dev).to(bf))
        ys = E.Act.empty(n, h * s, h * s, c, dev)
        nb = xs.t.numel() * 2 + ys.t.numel() * 2
        out['upsample_fwd_' + tag] = timed(lambda: E.upsample_fwd(xs, ys, align_corners=ac), nb)
        out['upsample_bwd_' + tag] = timed(lambda: E.upsample_bwd(ys, xs, align_corners=ac), nb)
    # stem max pool
    a = E.Act(torch.randn(n, 256, 256, 64, device=dev).to(bf))
    res = {}

    def fwd():
        res['y'], res['idx'] = E.maxpool_fwd(a)
    nb = a.t.numel() * 2 + a.t.numel() // 4 * 3
    out['maxpool_fwd'] = timed(fwd, nb)
    g = E.Act(torch.randn(n, 128, 128, 64, device=dev).to(bf))
    out['maxpool_bwd'] = timed(lambda: E.maxpool_bwd(g, res['idx'], 256, 256), nb)
    # weight packing of the whole model
    model = P.DeepLabV3Plus(21).cuda().train()
    entries = E.packed_entries(model)
    E.pack_all(entries)
    numel = sum(w.numel() for _, w in entries)

    def pack():
        E.WEIGHT_EPOCH += 1
        E.pack_all(entries)
    out['pack_multi'] = timed(pack, numel * 4 + numel * 2 * 2)
    print(json.dumps({'lib': args.lib or 'default', 'kernels': out}))


if __name__ == '__main__':
    main()
