"""How far is a *library* bf16 pipeline (PyTorch autocast, cuDNN) from the fp32 oracle on the same ASPP case?"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from oracle import golden as G, restatement as R
import pytorch_segmentation_b200 as P

def errs(a, b):
    a, b = a.detach().float().cpu(), b.detach().float().cpu()
    return float((a - b).abs().max()) / float(b.abs().max()), float((a - b).norm() / b.norm())

for name in ['aspp_small', 'aspp_full']:
    c = G.CASES[name]
    g = torch.Generator().manual_seed(c['seed'] + 1)
    x = torch.randn(c['n'], c['cin'], c['h'], c['w'], generator=g)
    gy = torch.randn(c['n'], c['cout'], c['h'], c['w'], generator=g)
    def run(mode):
        sd = {('a.' + k): v.clone().cuda() for k, v in G.small_aspp_state(c).items()}
        for k, v in sd.items():
            if v.is_floating_point() and 'running' not in k:
                v.requires_grad_(True)
        xx = x.cuda().requires_grad_(True)
        if mode == 'bf16':
            with torch.autocast('cuda', dtype=torch.bfloat16):
                y = R.aspp(sd, 'a', xx, c['rates'], True)
        else:
            y = R.aspp(sd, 'a', xx, c['rates'], True)
        y.backward(gy.cuda().to(y.dtype))
        return y, xx.grad, {k[2:]: v.grad for k, v in sd.items() if v.requires_grad}
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    y32, gx32, gp32 = run('fp32')
    y16, gx16, gp16 = run('bf16')
    m = P.ASPP(c['cin'], c['cout'], list(c['rates']))
    m.load_state_dict(G.small_aspp_state(c))
    m.cuda().train()
    xg = x.cuda().requires_grad_(True)
    y = m(xg)
    y.backward(gy.cuda().to(y.dtype))
    print(name, 'out   ours %s  autocast %s' % (errs(y, y32), errs(y16, y32)))
    print(name, 'gradx ours %s  autocast %s' % (errs(xg.grad, gx32), errs(gx16, gx32)))
    for k, p in m.named_parameters():
        print(name, k, 'ours (max %.4f l2 %.4f)  autocast (max %.4f l2 %.4f)' % (errs(p.grad, gp32[k]) + errs(gp16[k], gp32[k])))
