#!/usr/bin/env python
"""Throughput of the other BASELINE.json configurations on one B200 (bench.py is the headline configs[3] line):

  configs[0]  UNet 2-class 320x320 batch 4 fwd+bwd            --model unet --mode train --batch 4 --size 320 320
  configs[1]  DeepLabV3+ 21-class 512x512 inference batch 16   --model deeplab --mode infer --batch 16 --size 512 512
  configs[4]  HRNet 19-class 1024x512 training                 --model hrnet --mode train --batch 8 --size 512 1024

Synthetic inputs, random-init weights; CUDA-event timing after warm-up; prints one JSON line per run.
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
    ap.add_argument('--model', default='hrnet', choices=['hrnet', 'unet', 'deeplab'])
    ap.add_argument('--mode', default='train', choices=['train', 'infer'])
    ap.add_argument('--batch', type=int, default=8)
    ap.add_argument('--size', type=int, nargs=2, default=[512, 1024], help='H W')
    ap.add_argument('--steps', type=int, default=20)
    ap.add_argument('--warmup', type=int, default=5)
    ap.add_argument('--graph', default='off', choices=['on', 'off'])
    args = ap.parse_args()
    import pytorch_segmentation_b200 as P
    from pytorch_segmentation_b200 import _lib
    nc = {'hrnet': 19, 'unet': 2, 'deeplab': 21}[args.model]
    cls = {'hrnet': P.HRNet, 'unet': P.UNet, 'deeplab': P.DeepLabV3Plus}[args.model]
    torch.manual_seed(0)
    model = cls(nc).cuda()
    h, w = args.size
    g = torch.Generator().manual_seed(1)
    x = torch.randn(args.batch, 3, h, w, generator=g).cuda()
    t = torch.randint(0, nc, (args.batch, h, w), generator=g).cuda()
    if args.mode == 'train':
        model.train()
        opt = P.optim.SGD(model.parameters(), lr=1e-3, momentum=0.9)
        step = P.TrainStep(model, opt, graph=args.graph == 'on')
        fn = lambda: step(x, t)
        extra = 4 if args.graph == 'on' else 0
    else:
        model.eval()
        cm = torch.zeros((nc, nc), dtype=torch.int64, device='cuda')
        from pytorch_segmentation_b200 import engine as E

        def fn():
            with torch.no_grad():
                y = model(x)
                E.argmax_confusion(y, t, cm)          # test.py:28-46: forward, argmax, per-class counts
        extra = 0
    for _ in range(args.warmup + extra):
        fn()
    torch.cuda.synchronize()
    l0 = _lib.launches()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / args.steps
    print(json.dumps({'model': args.model, 'mode': args.mode, 'batch': args.batch, 'size': [h, w], 'graph': args.graph,
                      'ms_per_step': ms, 'images_per_sec': args.batch / ms * 1e3,
                      'libb200seg_launches_per_step': (_lib.launches() - l0) / args.steps,
                      'peak_mem_gb': torch.cuda.max_memory_allocated() / 2 ** 30}), flush=True)


if __name__ == '__main__':
    main()
