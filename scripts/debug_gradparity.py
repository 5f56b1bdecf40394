#!/usr/bin/env python
"""Per-tensor table (max-norm and L2 relative error, in network order) of the mask-matched gradient comparison of the
ResNet-50 backbone (tests/test_gpu_gradparity.py) — shows where along the depth an error enters."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
from oracle import restatement as R  # noqa: E402
import pytorch_segmentation_b200 as P  # noqa: E402
from pytorch_segmentation_b200 import engine as E  # noqa: E402

n, h, w, seed = 2, 128, 128, 5
oracle = R.round_conv_weights_to_bf16_(R.OracleModel('deeplabv3plus', 21, seed=seed))
ours = P.DeepLabV3Plus(21)
ours.load_state_dict(oracle.state_dict())
ours.cuda().train()
g = torch.Generator().manual_seed(seed)
x = torch.randn(n, 3, h, w, generator=g)
gl = torch.randn(n, 256, h // 4, w // 4, generator=g).to(torch.bfloat16).float()
gh = torch.randn(n, 2048, h // 16, w // 16, generator=g).to(torch.bfloat16).float()
if len(sys.argv) > 1 and sys.argv[1] == 'high-only':
    gl = None
names = {id(p): k for k, p in ours.backbone.named_parameters()}
E.CAPTURE = {}
feats = ours.backbone(x.cuda())
outs, gs = ([feats[4]], [gh.cuda().to(feats[4].dtype)]) if gl is None else (
    [feats[1], feats[4]], [gl.cuda().to(feats[1].dtype), gh.cuda().to(feats[4].dtype)])
torch.autograd.backward(outs, gs)
vals = {names[pid] + '.' + kind: v.cpu() for (pid, kind), v in E.CAPTURE.items()}
E.CAPTURE = None
oracle.train()
handles = R.hook_backbone(oracle.backbone, vals, prefix='')
fo = oracle.backbone(x)
torch.autograd.backward([fo[4]] if gl is None else [fo[1], fo[4]], [gh] if gl is None else [gl, gh])
ref = dict(oracle.backbone.named_parameters())
for k, p in ours.backbone.named_parameters():
    a, b = p.grad.detach().float().cpu(), ref[k].grad
    print('%-34s max %.4f  l2 %.4f   |ref|max %.3e' % (k, float((a - b).abs().max()) / float(b.abs().max()),
                                                      float((a - b).norm()) / float(b.norm()), float(b.abs().max())))
