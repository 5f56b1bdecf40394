#!/usr/bin/env python
"""Attribute epilogue time of the igemm kernel: globaltimer stamps of CTA 0 per staged slab (diagnostics)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from pytorch_segmentation_b200 import _lib, engine as E
lib = _lib.load()
names = ['ld_issue->ld_done', 'bar1', 'cvt+sts', 'fence', 'bar2', 'tma+wait', 'stats', 'next']
for (n, h, w, c, k, with_stats) in [(16, 128, 128, 64, 256, 1), (16, 32, 32, 256, 1024, 1), (16, 32, 32, 1024, 256, 1), (16, 64, 64, 128, 512, 1)]:
    x = E.Act(torch.randn(n, h, w, c, device='cuda').to(torch.bfloat16))
    pw = E.PackedWeight().update(torch.randn(k, c, 1, 1, device='cuda') * 0.05, need_dgrad=False)
    y = E.Act.empty(n, h, w, k, 'cuda')
    stats = torch.zeros(2 * k, device='cuda') if with_stats else None
    slots = 64
    buf = torch.zeros(slots * 8, dtype=torch.int64, device='cuda')
    for _ in range(2):
        E.conv_fprop([(x, pw, 1)], y, stats=stats)
    lib.b200seg_debug_timing(buf.data_ptr(), slots)
    E.conv_fprop([(x, pw, 1)], y, stats=stats)
    lib.b200seg_debug_timing(None, 0)
    torch.cuda.synchronize()
    t = buf.view(slots, 8).cpu()
    used = int((t[:, 0] > 0).sum())
    t = t[:used]
    print('shape', (n, h, w, c, k), 'stats', with_stats, 'slabs by CTA0:', used, 'span us %.1f' % ((t[-1, 7] - t[0, 0]) / 1e3))
    d = (t[:, 1:] - t[:, :-1]).float()
    for i in range(7):
        print('   %-18s mean %7.0f ns  max %7.0f' % (names[i], d[:, i].mean(), d[:, i].max()))
    gap = (t[1:, 0] - t[:-1, 7]).float()
    print('   %-18s mean %7.0f ns  max %7.0f' % ('slab->next slab', gap.mean(), gap.max()))
