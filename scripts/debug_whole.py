import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from oracle import golden as G, restatement as R
import pytorch_segmentation_b200 as P
torch.backends.cudnn.allow_tf32 = False
torch.backends.cuda.matmul.allow_tf32 = False
def errs(a, b):
    a, b = a.detach().float().cpu(), b.detach().float().cpu()
    return float((a - b).abs().max()) / float(b.abs().max()), float((a - b).norm() / b.norm())
for name in ['deeplab_small', 'unet_320']:
    c = G.CASES[name]
    x, t = G.make_inputs(c)
    ref = R.OracleModel(c['kind'], c['nc'], seed=c['seed']).to('cuda')
    flo = R.OracleModel(c['kind'], c['nc'], seed=c['seed']).to('cuda')
    m = (P.DeepLabV3Plus if c['kind'] == 'deeplabv3plus' else P.UNet)(c['nc'])
    m.load_state_dict(ref.state_dict()); m.cuda()
    for training in (False, True):
        ref.train(training); flo.train(training); m.train(training)
        yr = ref(x.cuda()); lr = R.compute_loss(yr, t.cuda())
        with torch.autocast('cuda', dtype=torch.bfloat16):
            yf = flo(x.cuda())
        lf = R.compute_loss(yf.float(), t.cuda())
        y = m(x.cuda()); l = P.compute_loss(y, t.cuda(), m)
        print(name, 'train' if training else 'eval', 'logits ours', errs(y, yr), 'floor', errs(yf, yr), 'loss', float(l), float(lf), float(lr))
        if training:
            lr.backward(); lf.backward(); l.backward()
            params = dict(m.named_parameters())
            for k, v in ref.sd.items():
                if v.requires_grad:
                    print('   ', k, 'ours max %.4f l2 %.4f | floor max %.4f l2 %.4f' % (errs(params[k].grad, v.grad) + errs(flo.sd[k].grad, v.grad)))
