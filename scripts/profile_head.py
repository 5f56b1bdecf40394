#!/usr/bin/env python
"""One training step of the DeepLabV3+ HEAD alone (ASPP + decoder + cls_conv + class-map loss, forward and backward) at
the BASELINE.json shapes (batch 16, 512x512 input => features[1] 16x256x128x128, features[-1] 16x2048x32x32) — the convs
the metric names ("conv tensor-pipe util vs peak" on the ASPP / decoder convs), isolated so an `ncu --set full` capture
holds exactly them:

  python scripts/profile_head.py --tags gpurun_out/head_tags.json                      # plain run (must exit 0 first)
  ncu --set full --clock-control none --import-source on --profile-from-start off \
      -k regex:'igemm_kernel|wgrad_kernel' -o gpurun_out/head_r02 -f python scripts/profile_head.py
  python scripts/summarise_full.py gpurun_out/head_r02.ncu-rep --tags gpurun_out/head_tags.json > profiles/r02_full_aspp.md

Warm-up iterations run outside the profiler range (cudaProfilerStart/Stop bracket ONE step).  ``--tags`` writes the
ordered list of tcgen05 conv launches of that step (kind, geometry, nominal FLOPs) so the summary can name each row.
The weight-gradient lane is off here so launches are serialised in host order (ncu serialises them anyway).
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
    ap.add_argument('--batch', type=int, default=16)
    ap.add_argument('--size', type=int, default=512)
    ap.add_argument('--warmup', type=int, default=3)
    ap.add_argument('--tags', default=None)
    args = ap.parse_args()
    import pytorch_segmentation_b200 as P
    from pytorch_segmentation_b200 import engine as E
    E._LANE_ON = False
    nc, n, h, w = 21, args.batch, args.size, args.size
    torch.manual_seed(0)
    model = P.DeepLabV3Plus(nc).cuda().train()
    low = torch.randn(n, h // 4, w // 4, 256, device='cuda').relu().to(torch.bfloat16).permute(0, 3, 1, 2).requires_grad_(True)
    high = torch.randn(n, h // 16, w // 16, 2048, device='cuda').relu().to(torch.bfloat16).permute(0, 3, 1, 2).requires_grad_(True)
    tgt = torch.randint(0, nc, (n, h, w), device='cuda')

    def step():
        low.grad = high.grad = None
        model.zero_grad(set_to_none=True)
        E.pack_all(E.packed_entries(model))
        loss = P.compute_loss(model.head(low, high), tgt, model)
        loss.backward()
        return loss

    for _ in range(args.warmup):
        step()
    torch.cuda.synchronize()
    E.PROFILE = []
    torch.cuda.cudart().cudaProfilerStart()
    loss = step()
    torch.cuda.synchronize()
    torch.cuda.cudart().cudaProfilerStop()
    rec, E.PROFILE = E.PROFILE, None
    rows = [{'tag': r[0], 'flops': r[1], 'us_events': r[2].elapsed_time(r[3]) * 1e3, 'run_share': r[4]['run_share'],
             'bytes': r[4]['bytes']} for r in rec]
    if args.tags:
        json.dump(rows, open(args.tags, 'w'), indent=1)
    print('head step: loss %.5f, %d tcgen05 conv launches' % (float(loss), len(rows)))
    for r in rows:
        print('  %-64s %8.1f us  %7.1f TFLOP/s nominal' % (r['tag'], r['us_events'], r['flops'] / r['us_events'] / 1e6))


if __name__ == '__main__':
    main()
