#!/usr/bin/env python
"""Where the HOST time of one eager training step goes (the reference's external Trainer calls model / compute_loss /
backward / optimizer eagerly; the CUDA-graph step removes this cost, the eager step pays it): cProfile over a few eager
DeepLabV3+ steps at the BASELINE.json shape, top functions by own time and by cumulative time."""
import argparse
import cProfile
import io
import os
import pstats
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--batch', type=int, default=16)
    ap.add_argument('--steps', type=int, default=5)
    ap.add_argument('--top', type=int, default=45)
    args = ap.parse_args()
    import pytorch_segmentation_b200 as P
    torch.manual_seed(0)
    model = P.DeepLabV3Plus(21).cuda().train()
    opt = P.optim.SGD(model.parameters(), lr=1e-3, momentum=0.9, weight_decay=1e-4)
    step = P.TrainStep(model, opt, None, graph=False)
    x = torch.randint(0, 256, (args.batch, 3, 512, 512), dtype=torch.uint8, device='cuda')
    t = torch.randint(0, 21, (args.batch, 512, 512), dtype=torch.uint8, device='cuda')
    for _ in range(3):
        step(x, t)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step(x, t)
    host = (time.perf_counter() - t0) / args.steps * 1e3
    torch.cuda.synchronize()
    wall = (time.perf_counter() - t0) / args.steps * 1e3
    print('eager step: host enqueue %.2f ms, wall %.2f ms per step (no profiler)' % (host, wall))
    pr = cProfile.Profile()
    pr.enable()
    for _ in range(args.steps):
        step(x, t)
    pr.disable()
    torch.cuda.synchronize()
    for key in ('tottime', 'cumtime'):
        s = io.StringIO()
        pstats.Stats(pr, stream=s).sort_stats(key).print_stats(args.top)
        print(s.getvalue()[:9000])


if __name__ == '__main__':
    main()
