#!/usr/bin/env python
"""Run-to-run spread of the train-mode loss of the small whole-model golden case (tests/test_gpu_models.py), in one
process, before and after other work (a CUDA-graph TrainStep) has run in it."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from oracle import golden as G  # noqa: E402
from oracle import restatement as R  # noqa: E402
import pytorch_segmentation_b200 as P  # noqa: E402


def once(name):
    c = G.CASES[name]
    gold = G.load(name)
    oracle = R.OracleModel(c['kind'], c['nc'], seed=c['seed'])
    m = (P.DeepLabV3Plus if c['kind'] == 'deeplabv3plus' else P.UNet)(c['nc'])
    m.load_state_dict(oracle.state_dict())
    m.cuda().train()
    x, t = G.make_inputs(c)
    y = m(x.cuda())
    loss = P.compute_loss(y, t.cuda(), m)
    loss.backward()
    return float(loss.detach()), float(gold['train_loss'])


def main():
    for name in ('deeplab_small',):
        print(name, 'fresh:', ['%.4f' % once(name)[0] for _ in range(6)], 'gold %.4f' % once(name)[1])
    model = P.DeepLabV3Plus(21).cuda().train()
    opt = P.optim.SGD(model.parameters(), lr=1e-3, momentum=0.9)
    step = P.TrainStep(model, opt, None, graph=True)
    x = torch.randint(0, 256, (2, 3, 128, 128), dtype=torch.uint8, device='cuda')
    t = torch.randint(0, 21, (2, 128, 128), dtype=torch.uint8, device='cuda')
    for _ in range(6):
        step(x, t)
    torch.cuda.synchronize()
    for name in ('deeplab_small',):
        print(name, 'after a graph TrainStep:', ['%.4f' % once(name)[0] for _ in range(6)])


if __name__ == '__main__':
    main()
