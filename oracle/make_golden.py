"""Generate tests/golden/*.npz by running the UNMODIFIED reference (build container only).

TEST INFRASTRUCTURE.  Usage:  python -m oracle.make_golden   (from the repo root; needs
/root/reference).  The vectors pin ``oracle/restatement.py`` — and through it the CUDA path — to what
the reference's own models/aspp.py, models/deeplabv3plus.py, models/unet.py, utils/utils.py and
test.py:test compute on seeded synthetic inputs with seeded random-init weights.

Weights: ``OracleModel(kind, nc, seed)`` draws them deterministically (CPU RNG); its state dict is
loaded into the real reference module, so the GPU box can regenerate identical weights without
the reference tree.  Large tensors are stored as (norm, mean, strided sample) summaries so the
fixtures stay small; ``oracle/golden.py`` holds the shared summary/compare helpers.
"""
import os
import sys
import tempfile

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from oracle import reference_loader, restatement as R            # noqa: E402
from oracle.golden import GOLDEN_DIR, CASES, make_inputs, summarize, small_aspp_state  # noqa: E402


def _save(name, arrays):
    os.makedirs(GOLDEN_DIR, exist_ok=True)
    path = os.path.join(GOLDEN_DIR, name + '.npz')
    np.savez_compressed(path, **{k: np.asarray(v) for k, v in arrays.items()})
    print('wrote', path, os.path.getsize(path), 'bytes')


def model_case(ref, name):
    c = CASES[name]
    oracle = R.OracleModel(c['kind'], c['nc'], seed=c['seed'])
    cls = ref.DeepLabV3Plus if c['kind'] == 'deeplabv3plus' else ref.UNet
    torch.manual_seed(1234)
    model = cls(c['nc'])
    model.load_state_dict(oracle.state_dict())
    x, t = make_inputs(c)
    out = {}
    # eval forward (test.py:16,28)
    model.eval()
    with torch.no_grad():
        y = model(x)
    out.update(summarize('eval_logits', y))
    out['eval_loss'] = float(ref.compute_loss(y, t, model))
    out['eval_argmax_sample'] = y.max(1)[1].flatten()[::997].numpy()
    # train forward + backward (train.py:63 via Trainer.step)
    model.train()
    y = model(x)
    loss = ref.compute_loss(y, t, model)
    loss.backward()
    out.update(summarize('train_logits', y.detach()))
    out['train_loss'] = float(loss)
    sd = model.state_dict()
    for k, p in model.named_parameters():
        if not k.startswith('backbone.'):
            out.update(summarize('grad:' + k, p.grad))
    for k, v in sd.items():
        if not k.startswith('backbone.') and 'running_' in k:
            out['stat:' + k] = v.numpy()
    # a couple of backbone grads prove dgrad into the features is right
    bk = [k for k, _ in model.named_parameters() if k.startswith('backbone.')]
    for k in (bk[0], bk[len(bk) // 2], bk[-1]):
        out.update(summarize('grad:' + k, dict(model.named_parameters())[k].grad))
    _save(name, out)


def hrnet_case(ref, name):
    """The UNMODIFIED reference HRNet (models/hrnet.py) on seeded inputs with the seeded state of
    oracle/hrnet_restatement.make_state: eval logits, train logits / loss, gradients of a spread of parameters, and the
    BN running statistics after one training step."""
    from oracle import hrnet_restatement as H
    c = CASES[name]
    model = ref.HRNet(c['nc'])
    model.load_state_dict(H.make_state(c['nc'], seed=c['seed']))
    x, t = make_inputs(c)
    out = {}
    model.eval()
    with torch.no_grad():
        y = model(x)
    out.update(summarize('eval_logits', y))
    out['eval_loss'] = float(ref.compute_loss(y, t, model))
    out['eval_argmax_sample'] = y.max(1)[1].flatten()[::97].numpy()
    model.train()
    y = model(x)
    loss = ref.compute_loss(y, t, model)
    loss.backward()
    out.update(summarize('train_logits', y.detach()))
    out['train_loss'] = float(loss)
    names = [k for k, _ in model.named_parameters()]
    for k in names[::max(1, len(names) // 24)] + ['final_layer.weight', 'final_layer.bias', 'stem.0.0.weight']:
        out.update(summarize('grad:' + k, dict(model.named_parameters())[k].grad))
    sd = model.state_dict()
    for k in [k for k in sd if 'running_' in k][::16]:
        out['stat:' + k] = sd[k].numpy()
    _save(name, out)


def aspp_case(ref, name):
    """ASPP standalone: SURVEY §8c fixture (i) and a small all-values variant."""
    c = CASES[name]
    torch.manual_seed(c['seed'])
    m = ref.ASPP(c['cin'], c['cout'], list(c['rates']))
    m.load_state_dict(small_aspp_state(c))
    g = torch.Generator().manual_seed(c['seed'] + 1)
    x = torch.randn(c['n'], c['cin'], c['h'], c['w'], generator=g).requires_grad_(True)
    gy = torch.randn(c['n'], c['cout'], c['h'], c['w'], generator=g)
    out = {}
    m.eval()
    with torch.no_grad():
        y = m(x)
    out.update(summarize('eval_out', y))
    m.train()
    y = m(x)
    y.backward(gy)
    out.update(summarize('train_out', y.detach()))
    out.update(summarize('grad_in', x.grad))
    for k, p in m.named_parameters():
        out.update(summarize('grad:' + k, p.grad))
    for k, v in m.state_dict().items():
        if 'running_' in k:
            out['stat:' + k] = v.numpy()
    _save(name, out)


def loss_case(ref):
    g = torch.Generator().manual_seed(7)
    out = {}
    for i, (n, nc, h, w, th, tw) in enumerate([(2, 21, 32, 32, 32, 32), (2, 21, 16, 16, 64, 64),
                                               (3, 2, 20, 24, 40, 48), (1, 19, 8, 16, 29, 37)]):
        logits = torch.randn(n, nc, h, w, generator=g) * 3
        logits.requires_grad_(True)
        tgt = torch.randint(0, nc, (n, th, tw), generator=g)
        loss = ref.compute_loss(logits, tgt, None)
        loss.backward()
        out['shape%d' % i] = np.array([n, nc, h, w, th, tw])
        out['logits%d' % i] = logits.detach().numpy()
        out['targets%d' % i] = tgt.numpy()
        out['loss%d' % i] = float(loss)
        out['dlogits%d' % i] = logits.grad.numpy()
    _save('compute_loss', out)


class _FakeFetcher:
    """test.py:18 reads fetcher.loader.dataset.classes; iteration yields (inputs, targets)."""

    def __init__(self, batches, classes):
        self._b = batches
        self.loader = type('L', (), {})()
        self.loader.dataset = type('D', (), {})()
        self.loader.dataset.classes = classes

    def __iter__(self):
        return iter(self._b)

    def __len__(self):
        return len(self._b)


class _FixedLogits(torch.nn.Module):
    """A 'model' that replays prepared logits so test.py:test runs its real loop on known predictions."""

    def __init__(self, logits):
        super().__init__()
        self._l, self._i = logits, 0

    def forward(self, x):
        y = self._l[self._i % len(self._l)]
        self._i += 1
        return y


def metrics_case(ref):
    g = torch.Generator().manual_seed(11)
    out = {}
    cases = []
    for nc in (2, 21, 19):
        pred = torch.randint(0, nc, (3, 2, 24, 24), generator=g)
        tgt = torch.randint(0, nc, (3, 2, 24, 24), generator=g)
        cases.append((nc, pred, tgt))
    nc = 5   # class 4 absent from both -> union 0 -> guard; all-correct; all-wrong
    t = torch.randint(0, 4, (1, 2, 24, 24), generator=g)
    cases.append((nc, t.clone(), t.clone()))
    cases.append((nc, (t + 1) % 4, t.clone()))
    cwd = os.getcwd()
    with tempfile.TemporaryDirectory() as tmp:
        os.chdir(tmp)             # test.py:32-33 writes batch.png into the CWD
        try:
            for i, (nc, pred, tgt) in enumerate(cases):
                logits = [torch.nn.functional.one_hot(p, nc).permute(0, 3, 1, 2).float() * 4 for p in pred]
                imgs = torch.zeros(pred.shape[1], 3, pred.shape[2], pred.shape[3])
                fetcher = _FakeFetcher([(imgs.clone(), tg) for tg in tgt], ['c%d' % j for j in range(nc)])
                miou = ref.test(_FixedLogits(logits), fetcher)
                tp, fn, fp = R.tp_fn_fp_loop(pred, tgt, nc)
                T, P, Rr, iou, F1 = ref.compute_metrics(tp.float(), fn.float(), fp.float())
                out['nc%d' % i] = nc
                out['pred%d' % i] = pred.numpy()
                out['tgt%d' % i] = tgt.numpy()
                out['miou%d' % i] = np.float32(miou)
                out['tp%d' % i], out['fn%d' % i], out['fp%d' % i] = tp.numpy(), fn.numpy(), fp.numpy()
                out['iou%d' % i], out['P%d' % i], out['R%d' % i], out['F1%d' % i] = \
                    iou.numpy(), P.numpy(), Rr.numpy(), F1.numpy()
        finally:
            os.chdir(cwd)
    out['ncases'] = len(cases)
    _save('metrics', out)


def main():
    ref = reference_loader.load()
    torch.set_num_threads(os.cpu_count())
    loss_case(ref)
    metrics_case(ref)
    aspp_case(ref, 'aspp_small')
    aspp_case(ref, 'aspp_full')
    model_case(ref, 'unet_320')
    model_case(ref, 'deeplab_small')
    model_case(ref, 'deeplab_512')
    hrnet_case(ref, 'hrnet_small')


if __name__ == '__main__':
    main()
