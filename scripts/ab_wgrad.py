"""A/B of two builds of the library on isolated wgrad launches: python scripts/ab_wgrad.py [path/to/lib.so]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from pytorch_segmentation_b200 import _lib
if len(sys.argv) > 1:
    _lib.LIB_PATH = os.path.abspath(sys.argv[1])
from pytorch_segmentation_b200 import engine as E

SHAPES = [  # n, h, w, k, c, r, dil
    (16, 32, 32, 256, 2048, 3, 6), (16, 32, 32, 1024, 256, 1, 1), (16, 128, 128, 256, 64, 1, 1),
    (16, 32, 32, 512, 512, 3, 1), (16, 128, 128, 64, 64, 3, 1), (16, 64, 64, 128, 128, 3, 1), (16, 32, 32, 256, 256, 3, 1),
]
flush = torch.empty(256 << 20, dtype=torch.uint8, device='cuda')
for n, h, w, k, c, r, dil in SHAPES:
    dy = E.Act(torch.randn(n, h, w, k, device='cuda').to(torch.bfloat16))
    x = E.Act(torch.randn(n, h, w, c, device='cuda').to(torch.bfloat16))
    for _ in range(3):
        E._conv_wgrad(dy, x, k, r, r, dil)
    ts = []
    for _ in range(10):
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        E._conv_wgrad(dy, x, k, r, r, dil)
        b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b) * 1e3)
    ts.sort()
    print('wgrad %dx%dx%d k%d c%d r%d d%d: median %.1f us (min %.1f)' % (n, h, w, k, c, r, dil, ts[len(ts) // 2], ts[0]))
