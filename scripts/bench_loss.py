"""Per-kernel timings of the loss path at the headline shape (16 x 21 x 512 x 512 from a 128 x 128 class map)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from pytorch_segmentation_b200 import engine as E


def timeit(f, reps=20):
    for _ in range(3):
        f()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    a.record()
    for _ in range(reps):
        f()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / reps * 1e3


n, h, w, nc, ct, s = 16, 128, 128, 21, 24, 4
lr = torch.randn(n, h, w, ct, device='cuda')
t = torch.randint(0, nc, (n, h * s, w * s), device='cuda')
g = torch.tensor(1.0, device='cuda')
acc = E.lowres_ce_fwd(lr, nc, t, True)
cnt = acc[1:2].view(torch.int64)
logits = E.logits_upsample_fwd(lr, nc, h * s, w * s, True)
dl = E.softmax_ce_bwd(logits, t, g, cnt)
print('lowres_ce_fwd      %7.1f us' % timeit(lambda: E.lowres_ce_fwd(lr, nc, t, True)))
print('lowres_ce_bwd      %7.1f us' % timeit(lambda: E.lowres_ce_bwd(lr, nc, t, g, cnt, True)))
print('softmax_ce fwd     %7.1f us' % timeit(lambda: E.softmax_ce(logits, t, want_grad=False)))
print('softmax_ce_bwd     %7.1f us' % timeit(lambda: E.softmax_ce_bwd(logits, t, g, cnt)))
print('logits_up_fwd      %7.1f us' % timeit(lambda: E.logits_upsample_fwd(lr, nc, h * s, w * s, True)))
print('logits_up_bwd      %7.1f us' % timeit(lambda: E.logits_upsample_bwd(dl, h, w, ct, True)))
