"""Model-level parity of the drop-in modules (ASPP, DeepLabV3+ head, UNet head, compute_loss, test()) against

  (a) the golden vectors generated from the UNMODIFIED reference (tests/golden/, oracle/make_golden.py), and
  (b) the CPU oracle (oracle/restatement.py) run here on the same seeded inputs and weights.

Tolerance (BASELINE.json north_star): logits and gradients max|d|/max|ref| <= 2e-2 per tensor, argmax
agreement >= 99.9 %, loss relative error <= 1e-2; confusion matrix / mIoU bit-exact for identical
predictions.  Head tests feed both sides the same backbone features so only the hot path is compared;
whole-model tests (backbone in bf16 autocast on torchvision/cuDNN, outside the hot path) use 5e-2.
"""
import numpy as np
import pytest
import torch

from oracle import golden as G
from oracle import restatement as R

pytestmark = pytest.mark.gpu

TOL = 2e-2
TOL_WHOLE = 5e-2


def rel_err(got, ref):
    got, ref = got.detach().float().cpu(), ref.detach().float().cpu()
    return float((got - ref).abs().max()) / max(float(ref.abs().max()), 1e-30)


@pytest.fixture(scope='module')
def P(lib):
    import pytorch_segmentation_b200 as pkg
    return pkg


# ---------------------------------------------------------------------------------------------------
# ASPP standalone vs golden
# ---------------------------------------------------------------------------------------------------
@pytest.mark.parametrize('name', ['aspp_small', 'aspp_full'])
def test_aspp_matches_reference_golden(P, name):
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
    assert G.summary_error(gold, 'eval_out', y)[0] < TOL
    m.train()
    xg = x.cuda().requires_grad_(True)
    y = m(xg)
    y.backward(gy.cuda().to(y.dtype))
    assert G.summary_error(gold, 'train_out', y)[0] < TOL
    assert G.summary_error(gold, 'grad_in', xg.grad)[0] < TOL
    worst = 0.0
    for k, p in m.named_parameters():
        e = G.summary_error(gold, 'grad:' + k, p.grad)[0]
        worst = max(worst, e)
        assert e < TOL, (k, e)
    for k, v in m.state_dict().items():
        if 'running_' in k:
            ref = torch.from_numpy(gold['stat:' + k])
            assert rel_err(v, ref) < 5e-3, k
        if k.endswith('num_batches_tracked'):
            assert int(v) == 1


# ---------------------------------------------------------------------------------------------------
# heads vs the CPU oracle on identical features
# ---------------------------------------------------------------------------------------------------
def _head_case(P, kind, nc, n, h, w, seed):
    oracle = R.OracleModel(kind, nc, seed=seed)
    ours = (P.DeepLabV3Plus if kind == 'deeplabv3plus' else P.UNet)(nc)
    ours.load_state_dict(oracle.state_dict())
    ours.cuda()
    x, t = G.make_inputs(dict(n=n, nc=nc, h=h, w=w, seed=seed))
    oracle.eval()
    with torch.no_grad():
        feats = [f.detach() for f in oracle.features(x)]
    return oracle, ours, feats, t


def _run_heads(oracle, ours, feats, t, training, kind):
    oracle.train(training)
    ours.train(training)
    oracle.zero_grad()
    ours.zero_grad()
    f_ref = [f.clone().requires_grad_(True) for f in feats]
    f_dev = [f.cuda().to(torch.bfloat16).contiguous(memory_format=torch.channels_last).requires_grad_(True)
             for f in feats]
    y_ref = oracle.head(f_ref)
    loss_ref = R.compute_loss(y_ref, t)
    from pytorch_segmentation_b200 import compute_loss
    if kind == 'deeplabv3plus':
        y = ours.head(f_dev[1], f_dev[-1])
    else:
        y = ours.head(f_dev[1], f_dev[2], f_dev[3], f_dev[4])
    loss = compute_loss(y, t.cuda(), ours)
    if training:
        loss_ref.backward()
        loss.backward()
    return y_ref, loss_ref, f_ref, y, loss, f_dev


@pytest.mark.parametrize('kind,nc,n,h,w,seed', [('deeplabv3plus', 21, 2, 128, 160, 4), ('unet', 2, 4, 320, 320, 2),
                                                 ('deeplabv3plus', 21, 2, 512, 512, 3), ('unet', 5, 2, 96, 128, 8)])
def test_head_matches_oracle(P, kind, nc, n, h, w, seed):
    oracle, ours, feats, t = _head_case(P, kind, nc, n, h, w, seed)
    # eval mode: Conv+BN+ReLU fused in the epilogue
    with torch.no_grad():
        y_ref, loss_ref, _, y, loss, _ = _run_heads(oracle, ours, feats, t, False, kind)
    assert y.shape == y_ref.shape and y.dtype == torch.float32
    assert rel_err(y, y_ref) < TOL
    assert abs(float(loss) - float(loss_ref)) <= 1e-2 * abs(float(loss_ref))
    # train mode: batch statistics, backward through every kernel
    y_ref, loss_ref, f_ref, y, loss, f_dev = _run_heads(oracle, ours, feats, t, True, kind)
    assert rel_err(y, y_ref) < TOL
    assert abs(float(loss) - float(loss_ref)) <= 1e-2 * abs(float(loss_ref))
    used = (1, 4) if kind == 'deeplabv3plus' else (1, 2, 3, 4)
    for i in used:
        assert rel_err(f_dev[i].grad, f_ref[i].grad) < TOL, 'feature %d grad' % i
    params = dict(ours.named_parameters())
    for k, v in oracle.sd.items():
        if v.requires_grad:
            assert rel_err(params[k].grad, v.grad) < TOL, k
    sd = ours.state_dict()
    for k, v in oracle.sd.items():
        if 'running_' in k:
            assert rel_err(sd[k], v) < 5e-3, k


def test_argmax_agreement_on_separated_logits(P):
    """argmax agreement >= 99.9 % — measured where classes are separated by more than bf16 noise: the
    cls_conv bias is set so every pixel has a clear winner that varies over the image."""
    kind, nc, n, h, w, seed = 'deeplabv3plus', 21, 2, 128, 128, 11
    oracle, ours, feats, t = _head_case(P, kind, nc, n, h, w, seed)
    with torch.no_grad():
        wt = oracle.sd['cls_conv.weight']
        wt.mul_(8.0)
        ours.load_state_dict(oracle.state_dict())
        ours.cuda()
        y_ref, _, _, y, _, _ = _run_heads(oracle, ours, feats, t, False, kind)
    p_ref, p = y_ref.max(1)[1], y.max(1)[1].cpu()
    top2 = y_ref.topk(2, dim=1)[0]
    margin = (top2[:, 0] - top2[:, 1])
    clear = margin > 2e-2 * float(y_ref.abs().max())
    agree_clear = float((p_ref == p)[clear].float().mean())
    agree_all = float((p_ref == p).float().mean())
    assert agree_clear >= 0.999, (agree_clear, agree_all, float(clear.float().mean()))
    assert agree_all >= 0.98


# ---------------------------------------------------------------------------------------------------
# whole model (backbone included) vs the reference's golden vectors
# ---------------------------------------------------------------------------------------------------
@pytest.mark.parametrize('name', ['deeplab_small', 'unet_320', 'deeplab_512'])
def test_whole_model_matches_reference_golden(P, name):
    c = G.CASES[name]
    gold = G.load(name)
    oracle = R.OracleModel(c['kind'], c['nc'], seed=c['seed'])
    m = (P.DeepLabV3Plus if c['kind'] == 'deeplabv3plus' else P.UNet)(c['nc'])
    m.load_state_dict(oracle.state_dict())
    m.cuda()
    x, t = G.make_inputs(c)
    x, t = x.cuda(), t.cuda()
    m.eval()
    with torch.no_grad():
        y = m(x)
        loss = P.compute_loss(y, t, m)
    assert list(y.shape) == [c['n'], c['nc'], c['h'], c['w']]
    assert G.summary_error(gold, 'eval_logits', y)[0] < TOL_WHOLE
    assert abs(float(loss) - float(gold['eval_loss'])) <= 2e-2 * abs(float(gold['eval_loss']))
    m.train()
    y = m(x)
    loss = P.compute_loss(y, t, m)
    loss.backward()
    assert G.summary_error(gold, 'train_logits', y)[0] < TOL_WHOLE
    assert abs(float(loss) - float(gold['train_loss'])) <= 2e-2 * abs(float(gold['train_loss']))
    for k, p in m.named_parameters():
        if not k.startswith('backbone.'):
            assert G.summary_error(gold, 'grad:' + k, p.grad)[0] < 2 * TOL_WHOLE, k


# ---------------------------------------------------------------------------------------------------
# loss and metrics vs golden
# ---------------------------------------------------------------------------------------------------
def test_compute_loss_matches_reference_golden(P):
    gold = G.load('compute_loss')
    i = 0
    while 'shape%d' % i in gold:
        logits = torch.from_numpy(gold['logits%d' % i]).cuda().requires_grad_(True)
        tgt = torch.from_numpy(gold['targets%d' % i]).cuda()
        loss = P.compute_loss(logits, tgt, None)
        loss.backward()
        assert loss.dim() == 0
        assert abs(float(loss) - float(gold['loss%d' % i])) < 1e-5 * max(1.0, abs(float(gold['loss%d' % i])))
        assert rel_err(logits.grad, torch.from_numpy(gold['dlogits%d' % i])) < 1e-4
        i += 1
    assert i == 4


class _Fetcher:
    def __init__(self, batches, classes):
        self._b = batches
        self.loader = type('L', (), {})()
        self.loader.dataset = type('D', (), {})()
        self.loader.dataset.classes = classes

    def __iter__(self):
        return iter(self._b)


class _Replay(torch.nn.Module):
    def __init__(self, logits):
        super().__init__()
        self._l, self._i = logits, 0

    def forward(self, x):
        y = self._l[self._i % len(self._l)]
        self._i += 1
        return y


def test_test_loop_matches_reference_golden_bit_exact(P):
    """test(model, fetcher) on the predictions the reference's own test.py:test saw: mIoU bit-exact."""
    gold = G.load('metrics')
    for i in range(int(gold['ncases'])):
        nc = int(gold['nc%d' % i])
        pred, tgt = torch.from_numpy(gold['pred%d' % i]), torch.from_numpy(gold['tgt%d' % i])
        logits = [(torch.nn.functional.one_hot(p, nc).permute(0, 3, 1, 2).float() * 4).cuda() for p in pred]
        imgs = torch.zeros(pred.shape[1], 3, pred.shape[2], pred.shape[3], device='cuda')
        fetcher = _Fetcher([(imgs, tg.cuda()) for tg in tgt], ['c%d' % j for j in range(nc)])
        miou = P.test(_Replay(logits), fetcher, verbose=False)
        assert np.float32(miou) == gold['miou%d' % i], (i, miou, gold['miou%d' % i])
        from pytorch_segmentation_b200 import engine as E
        cm = E.confusion(pred.reshape(-1).cuda(), tgt.reshape(-1).cuda(), nc)
        from pytorch_segmentation_b200.utils.utils import counts_from_confusion
        tp, fn, fp = (v.cpu() for v in counts_from_confusion(cm))
        assert np.array_equal(tp.numpy(), gold['tp%d' % i])
        assert np.array_equal(fn.numpy(), gold['fn%d' % i])
        assert np.array_equal(fp.numpy(), gold['fp%d' % i])


def test_full_size_properties(P):
    """BASELINE full size (N=16, 512x512, 21 classes) through size-independent properties:
    confusion total == pixel count, row sums == target histogram, loss of uniform logits == log(nc),
    gradient of the mean loss sums to zero per pixel."""
    from pytorch_segmentation_b200 import engine as E
    n, nc, h, w = 16, 21, 512, 512
    g = torch.Generator(device='cuda').manual_seed(0)
    logits = torch.randn(n, nc, h, w, device='cuda', generator=g)
    tgt = torch.randint(0, nc, (n, h, w), device='cuda', generator=g)
    cm = E.argmax_confusion(logits, tgt)
    assert int(cm.sum()) == n * h * w
    assert torch.equal(cm.sum(1), torch.bincount(tgt.reshape(-1), minlength=nc))
    assert torch.equal(cm.sum(0), torch.bincount(E.argmax(logits).reshape(-1), minlength=nc))
    flat = torch.zeros(n, nc, h, w, device='cuda').requires_grad_(True)
    loss = P.compute_loss(flat, tgt, None)
    loss.backward()
    assert abs(float(loss) - float(np.log(nc))) < 1e-5
    assert float(flat.grad.sum(1).abs().max()) < 1e-9
