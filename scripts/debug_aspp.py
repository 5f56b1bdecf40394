import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from oracle import golden as G
import pytorch_segmentation_b200 as P

for name in ['aspp_small', 'aspp_full']:
    c = G.CASES[name]
    gold = G.load(name)
    m = P.ASPP(c['cin'], c['cout'], list(c['rates']))
    m.load_state_dict(G.small_aspp_state(c))
    m.cuda()
    g = torch.Generator().manual_seed(c['seed'] + 1)
    x = torch.randn(c['n'], c['cin'], c['h'], c['w'], generator=g)
    gy = torch.randn(c['n'], c['cout'], c['h'], c['w'], generator=g)
    m.eval()
    with torch.no_grad():
        y = m(x.cuda())
    print(name, 'eval_out', G.summary_error(gold, 'eval_out', y))
    m.train()
    xg = x.cuda().requires_grad_(True)
    y = m(xg)
    y.backward(gy.cuda().to(y.dtype))
    print(name, 'train_out', G.summary_error(gold, 'train_out', y))
    print(name, 'grad_in', G.summary_error(gold, 'grad_in', xg.grad))
    for k, p in m.named_parameters():
        print(name, k, G.summary_error(gold, 'grad:' + k, p.grad))
    for k, v in m.state_dict().items():
        if 'running_' in k:
            ref = torch.from_numpy(gold['stat:' + k])
            print(name, k, float((v.cpu() - ref).abs().max()) / float(ref.abs().max()))
