#!/usr/bin/env python
"""Split-K workspace bytes b200seg_conv_wgrad_workspace() reports for a list of layer shapes (pointers are not
dereferenced by the query)."""
import ctypes
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402,F401
from pytorch_segmentation_b200 import _lib as L  # noqa: E402

lib = L.load()
torch.zeros(1, device='cuda')
for n, h, w, c, k, r in [(32, 128, 128, 256, 256, 3), (16, 256, 256, 256, 256, 3), (16, 128, 128, 256, 320, 3), (16, 128, 128, 384, 256, 3),
                         (64, 128, 128, 256, 256, 3), (16, 128, 128, 320, 320, 3), (16, 128, 128, 256, 512, 3), (8, 128, 128, 256, 256, 3),
                         (4, 64, 64, 64, 256, 1), (2, 32, 32, 256, 256, 3), (4, 64, 64, 128, 128, 3), (16, 128, 128, 128, 128, 3), (16, 128, 128, 304, 256, 3)]:
    d = L.WgradDesc()
    d.n, d.h, d.w = n, h, w
    d.dy, d.dy_c_total, d.dy_c_offset, d.k = 256, k, 0, k
    d.x, d.x_c_total, d.x_c_offset, d.c = 256, c, 0, c
    d.r, d.s, d.dil, d.stride = r, r, 1, 1
    d.dw, d.k_out, d.accumulate = 256, k, 0
    print((n, h, w, c, k, r), lib.b200seg_conv_wgrad_workspace(ctypes.byref(d)), L.last_error())
