#!/usr/bin/env python
"""Microbenchmark of the HBM-bound kernels: achieved GB/s (algorithmic bytes / CUDA-event time) against the measured
copy bandwidth in MEASURED_PEAKS.json.  L2 is flushed by a 256 MB read between iterations.

  python scripts/bench_bandwidth.py [--iters 10] [--json out.json]
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--iters', type=int, default=10)
    ap.add_argument('--json', default=None)
    args = ap.parse_args()
    from pytorch_segmentation_b200 import engine as E
    pk = json.load(open(os.path.join(ROOT, 'MEASURED_PEAKS.json'))) if os.path.exists(
        os.path.join(ROOT, 'MEASURED_PEAKS.json')) else {'hbm_gbs': 6650.0}
    peak = pk['hbm_gbs']
    flush = torch.zeros(64 << 20, dtype=torch.float32, device='cuda')

    def timed(fn):
        for _ in range(2):
            fn()
        tot = 0.0
        for _ in range(args.iters):
            flush.sum()
            torch.cuda._sleep(1500000)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            fn()
            e1.record()
            torch.cuda.synchronize()
            tot += e0.elapsed_time(e1)
        return tot / args.iters * 1e3

    def act(n, h, w, c):
        return E.Act(torch.randn(n, h, w, c, device='cuda').to(torch.bfloat16))

    rows = []

    def report(name, us, nbytes):
        gbs = nbytes / us / 1e3
        print('%-44s %9.1f us %8.0f GB/s  %5.3f of copy peak' % (name, us, gbs, gbs / peak))
        rows.append({'kernel': name, 'us': us, 'gbs': gbs, 'frac_hbm': gbs / peak, 'bytes': nbytes})

    for (n, h, w, c) in [(16, 128, 128, 256), (16, 128, 128, 64), (16, 64, 64, 512), (16, 32, 32, 1024), (16, 32, 32, 256)]:
        tag = '%dx%dx%dx%d' % (n, h, w, c)
        x, y, dy, dx, res = act(n, h, w, c), E.Act.empty(n, h, w, c, 'cuda'), act(n, h, w, c), E.Act.empty(n, h, w, c, 'cuda'), act(n, h, w, c)
        vec = torch.rand(4, c, device='cuda') + 0.5
        b = x.t.numel() * 2
        report('bn_apply+relu ' + tag, timed(lambda: E.bn_apply(x, y, vec, True)), 2 * b)
        report('bn_apply+res+relu ' + tag, timed(lambda: E.bn_apply(x, y, vec, True, res=res)), 3 * b)
        report('bn_backward (reduce+apply) ' + tag, timed(lambda: E.bn_backward(dy, x, dx, vec, True, True)), 5 * b)
        report('bn_backward masked ' + tag, timed(lambda: E.bn_backward(dy, x, dx, vec, True, True, mask=y)), 7 * b)
        report('relu_bwd_add ' + tag, timed(lambda: E.relu_bwd_add(dx, dy, y)), 4 * b)
        report('bn_stats ' + tag, timed(lambda: E.bn_stats(x)), b)

    # decoder / loss kernels at the DeepLabV3+ 512x512 batch-16 shapes
    n = 16
    a32, dec = act(n, 32, 32, 256), E.Act.empty(n, 128, 128, 384, 'cuda')
    report('upsample_fwd x4 256ch 32->128 into concat', timed(lambda: E.upsample_fwd(a32, dec.slice(0, 256))),
           a32.t.numel() * 2 + n * 128 * 128 * 256 * 2)
    ddec, da = act(n, 128, 128, 384), E.Act.empty(n, 32, 32, 256, 'cuda')
    report('upsample_bwd x4 256ch 128->32', timed(lambda: E.upsample_bwd(ddec.slice(0, 256), da)),
           n * 128 * 128 * 256 * 2 + da.t.numel() * 2)
    lr = torch.randn(n, 128, 128, 24, device='cuda')
    report('logits_upsample_fwd 24ch(21) 128->512 NHWC f32 -> NCHW f32', timed(lambda: E.logits_upsample_fwd(lr, 21, 512, 512)),
           lr.numel() * 4 + n * 21 * 512 * 512 * 4)
    logits = torch.randn(n, 21, 512, 512, device='cuda')
    tgt = torch.randint(0, 21, (n, 512, 512), device='cuda')
    report('softmax_ce fwd+bwd 21 classes 512x512', timed(lambda: E.softmax_ce(logits, tgt)),
           logits.numel() * 8 + tgt.numel() * 8)
    report('logits_upsample_bwd 512->128 NCHW f32 -> NHWC bf16', timed(lambda: E.logits_upsample_bwd(logits, 128, 128, 24)),
           logits.numel() * 4 + n * 128 * 128 * 24 * 2)
    report('argmax_confusion 21 classes 512x512', timed(lambda: E.argmax_confusion(logits, tgt)),
           logits.numel() * 4 + tgt.numel() * 8)
    a0 = act(n, 256, 256, 64)
    p0, idx = E.maxpool_fwd(a0)
    report('maxpool3x3s2 fwd 64ch 256->128', timed(lambda: E.maxpool_fwd(a0)), a0.t.numel() * 2 + p0.t.numel() * 3)
    dp = act(n, 128, 128, 64)
    report('maxpool3x3s2 bwd', timed(lambda: E.maxpool_bwd(dp, idx, 256, 256)), a0.t.numel() * 2 + p0.t.numel() * 3)
    hi = act(n, 32, 32, 2048)
    report('gap_fwd 2048ch 32x32', timed(lambda: E.gap_fwd(hi)), hi.t.numel() * 2)
    if args.json:
        json.dump({'peak_hbm_gbs': peak, 'rows': rows}, open(args.json, 'w'), indent=1)


if __name__ == '__main__':
    main()
