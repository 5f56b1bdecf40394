"""The oracle (oracle/restatement.py) against the golden vectors generated from the UNMODIFIED reference
(oracle/make_golden.py -> tests/golden/*.npz), and — when /root/reference is present (build container) — against
the reference itself, live.  CPU only."""
import numpy as np
import pytest
import torch

from oracle import golden as G
from oracle import reference_loader
from oracle import restatement as R

FP32_TOL = 2e-4     # same library calls as the reference; differences are thread-count / summation order only


def test_compute_loss_golden():
    gold = G.load('compute_loss')
    i = 0
    while 'shape%d' % i in gold:
        logits = torch.from_numpy(gold['logits%d' % i]).requires_grad_(True)
        loss = R.compute_loss(logits, torch.from_numpy(gold['targets%d' % i]))
        loss.backward()
        assert abs(float(loss) - float(gold['loss%d' % i])) < 1e-6
        assert np.allclose(logits.grad.numpy(), gold['dlogits%d' % i], atol=1e-8, rtol=1e-5)
        i += 1
    assert i == 4


def test_metrics_golden_bit_exact():
    gold = G.load('metrics')
    for i in range(int(gold['ncases'])):
        nc = int(gold['nc%d' % i])
        pred, tgt = torch.from_numpy(gold['pred%d' % i]), torch.from_numpy(gold['tgt%d' % i])
        tp, fn, fp = R.tp_fn_fp_loop(pred, tgt, nc)
        assert np.array_equal(tp.numpy(), gold['tp%d' % i])
        assert np.array_equal(fn.numpy(), gold['fn%d' % i])
        assert np.array_equal(fp.numpy(), gold['fp%d' % i])
        # the confusion matrix is a superset of the loop
        tp2, fn2, fp2 = R.counts_from_confusion(R.confusion_matrix(pred, tgt, nc))
        assert torch.equal(tp2, tp) and torch.equal(fn2, fn) and torch.equal(fp2, fp)
        T, P, Rr, iou, F1 = R.compute_metrics(tp.float(), fn.float(), fp.float())
        for name, v in (('iou', iou), ('P', P), ('R', Rr), ('F1', F1)):
            assert np.array_equal(v.numpy(), gold['%s%d' % (name, i)]), name
        assert np.float32(iou.mean().item()) == gold['miou%d' % i]


def test_evaluate_replays_reference_test_loop():
    gold = G.load('metrics')

    class Replay:
        def __init__(self, logits):
            self.l, self.i = logits, 0

        def eval(self):
            return self

        def __call__(self, x):
            y = self.l[self.i % len(self.l)]
            self.i += 1
            return y

    for i in range(int(gold['ncases'])):
        nc = int(gold['nc%d' % i])
        pred, tgt = torch.from_numpy(gold['pred%d' % i]), torch.from_numpy(gold['tgt%d' % i])
        logits = [torch.nn.functional.one_hot(p, nc).permute(0, 3, 1, 2).float() * 4 for p in pred]
        miou, _, _ = R.evaluate(Replay(logits), [(None, tg) for tg in tgt], nc)
        assert np.float32(miou) == gold['miou%d' % i]


@pytest.mark.parametrize('name', ['aspp_small', 'aspp_full'])
def test_aspp_golden(name):
    c = G.CASES[name]
    gold = G.load(name)
    sd = {('a.' + k): v.clone() for k, v in G.small_aspp_state(c).items()}
    for k, v in sd.items():
        if v.is_floating_point() and 'running_' not in k:
            v.requires_grad_(True)
    g = torch.Generator().manual_seed(c['seed'] + 1)
    x = torch.randn(c['n'], c['cin'], c['h'], c['w'], generator=g).requires_grad_(True)
    gy = torch.randn(c['n'], c['cout'], c['h'], c['w'], generator=g)
    with torch.no_grad():
        y = R.aspp(sd, 'a', x, c['rates'], False)
    assert G.summary_error(gold, 'eval_out', y)[0] < FP32_TOL
    y = R.aspp(sd, 'a', x, c['rates'], True)
    y.backward(gy)
    assert G.summary_error(gold, 'train_out', y)[0] < FP32_TOL
    assert G.summary_error(gold, 'grad_in', x.grad)[0] < FP32_TOL
    for k, v in sd.items():
        if v.requires_grad:
            assert G.summary_error(gold, 'grad:' + k[2:], v.grad)[0] < FP32_TOL, k
        if 'running_' in k:
            assert np.allclose(v.numpy(), gold['stat:' + k[2:]], rtol=1e-4, atol=1e-6), k


@pytest.mark.parametrize('name', ['unet_320', 'deeplab_small', 'deeplab_512'])
def test_model_golden(name):
    c = G.CASES[name]
    gold = G.load(name)
    m = R.OracleModel(c['kind'], c['nc'], seed=c['seed'])
    x, t = G.make_inputs(c)
    m.eval()
    with torch.no_grad():
        y = m(x)
    assert G.summary_error(gold, 'eval_logits', y)[0] < FP32_TOL
    assert abs(float(R.compute_loss(y, t)) - float(gold['eval_loss'])) < 1e-4 * abs(float(gold['eval_loss']))
    assert np.array_equal(y.max(1)[1].flatten()[::997].numpy(), gold['eval_argmax_sample'])
    m.train()
    y = m(x)
    loss = R.compute_loss(y, t)
    loss.backward()
    assert G.summary_error(gold, 'train_logits', y)[0] < 5e-4
    assert abs(float(loss.detach()) - float(gold['train_loss'])) < 1e-4 * abs(float(gold['train_loss']))
    for k, v in m.sd.items():
        if v.requires_grad:
            assert G.summary_error(gold, 'grad:' + k, v.grad)[0] < 2e-3, k
        if 'running_' in k:
            assert np.allclose(v.numpy(), gold['stat:' + k], rtol=1e-3, atol=1e-5), k


@pytest.mark.skipif(not reference_loader.available(), reason='/root/reference not present (GPU box)')
def test_restatement_equals_reference_live():
    """The unmodified reference modules, imported from /root/reference, against the restatement on a fresh case."""
    ref = reference_loader.load()
    oracle = R.OracleModel('unet', 3, seed=21)
    torch.manual_seed(0)
    model = ref.UNet(3)
    model.load_state_dict(oracle.state_dict())
    x, t = G.make_inputs(dict(n=2, nc=3, h=64, w=96, seed=21))
    model.train()
    oracle.train()
    y_ref = model(x)
    loss_ref = ref.compute_loss(y_ref, t, model)
    loss_ref.backward()
    y = oracle(x)
    loss = R.compute_loss(y, t)
    loss.backward()
    assert torch.allclose(y, y_ref, rtol=1e-4, atol=1e-5)
    assert abs(float(loss.detach()) - float(loss_ref.detach())) < 1e-6
    grads = dict(model.named_parameters())
    for k, v in oracle.sd.items():
        if v.requires_grad:
            assert torch.allclose(v.grad, grads[k].grad, rtol=1e-3, atol=1e-6), k
    tp = torch.tensor([5., 0., 7.])
    fn = torch.tensor([1., 0., 0.])
    fp = torch.tensor([2., 0., 3.])
    for a, b in zip(ref.compute_metrics(tp.clone(), fn.clone(), fp.clone()), R.compute_metrics(tp, fn, fp)):
        assert torch.equal(a, b)


# ---------------------------------------------------------------------------------------------------
# HRNet restatement (oracle/hrnet_restatement.py) vs the golden vectors from the unmodified reference and vs the live
# reference
# ---------------------------------------------------------------------------------------------------
def _hrnet_run(training):
    from oracle import hrnet_restatement as H
    c = G.CASES['hrnet_small']
    sd = H.make_state(c['nc'], seed=c['seed'])
    for k, v in sd.items():
        if v.is_floating_point() and 'running_' not in k:
            v.requires_grad_(True)
    x, t = G.make_inputs(c)
    if not training:
        with torch.no_grad():
            y = H.hrnet(sd, x, training=False)
        return sd, y, R.compute_loss(y, t)
    y = H.hrnet(sd, x, training=True)
    loss = R.compute_loss(y, t)
    loss.backward()
    return sd, y, loss


def test_hrnet_restatement_golden():
    gold = G.load('hrnet_small')
    sd, y, loss = _hrnet_run(False)
    assert G.summary_error(gold, 'eval_logits', y)[0] < FP32_TOL
    assert abs(float(loss) - float(gold['eval_loss'])) < FP32_TOL * abs(float(gold['eval_loss']))
    assert np.array_equal(y.max(1)[1].flatten()[::97].numpy(), gold['eval_argmax_sample'])
    sd, y, loss = _hrnet_run(True)
    assert G.summary_error(gold, 'train_logits', y.detach())[0] < FP32_TOL
    assert abs(float(loss) - float(gold['train_loss'])) < FP32_TOL * abs(float(gold['train_loss']))
    checked = 0
    for k in gold:
        if k.startswith('grad:') and k.endswith('/shape'):
            name = k[len('grad:'):-len('/shape')]
            assert G.summary_error(gold, 'grad:' + name, sd[name].grad)[0] < 5e-3, name
            checked += 1
        elif k.startswith('stat:'):
            assert np.allclose(sd[k[len('stat:'):]].numpy(), gold[k], rtol=1e-4, atol=1e-6), k
    assert checked >= 20


@pytest.mark.skipif(not reference_loader.available(), reason='reference tree not present')
def test_hrnet_restatement_equals_reference_live():
    from oracle import hrnet_restatement as H
    ref = reference_loader.load()
    c = G.CASES['hrnet_small']
    state = H.make_state(c['nc'], seed=c['seed'])
    model = ref.HRNet(c['nc'])
    model.load_state_dict(state)          # strict: the enumerated keys ARE the reference's
    x = torch.randn(1, 3, 64, 64, generator=torch.Generator().manual_seed(1))
    model.eval()
    with torch.no_grad():
        assert torch.allclose(model(x), H.hrnet({k: v.clone() for k, v in state.items()}, x), rtol=1e-4, atol=1e-4)
    # the drop-in module exposes the same keys and shapes, so reference checkpoints load into it unchanged
    from pytorch_segmentation_b200.models.hrnet import HRNet
    mine = HRNet(c['nc'])
    ours, theirs = mine.state_dict(), model.state_dict()
    assert list(ours.keys()) == list(theirs.keys())
    assert all(ours[k].shape == theirs[k].shape for k in ours)


def test_mask_matched_substitution_machinery():
    """oracle/restatement.py ``substituted`` / ``hook_backbone`` (used by tests/test_gpu_gradparity.py): recording the
    oracle's own boundary tensors and substituting them back is the identity (same logits, same gradients, bit for
    bit); substituting bf16-rounded values moves the logits by rounding noise only and keeps the fp32 autograd graph
    intact (every parameter still receives a gradient)."""
    n, nc, h, w, seed = 2, 5, 64, 64, 3
    x, t = G.make_inputs(dict(n=n, nc=nc, h=h, w=w, seed=seed))

    def run(values, record=False):
        oracle = R.OracleModel('deeplabv3plus', nc, seed=seed).train()
        handles = R.hook_backbone(oracle.backbone, values, prefix='backbone.', record=record)
        with R.substituted(values, record=record):
            y = oracle(x)
            loss = R.compute_loss(y, t)
        loss.backward()
        for hd in handles:
            hd.remove()
        grads = {'backbone.' + k: p.grad for k, p in oracle.backbone.named_parameters()}
        grads.update({k: v.grad for k, v in oracle.sd.items() if v.requires_grad})
        return y.detach(), grads

    vals = {}
    y0, g0 = run(vals, record=True)
    assert sum(k.endswith('.raw') for k in vals) == 53 + 7            # ResNet-50 convs + project, 5 ASPP branches, ASPP project
    assert 'backbone.layer4.2.conv3.weight.out' in vals and 'backbone.conv1.weight.out' in vals
    y1, g1 = run(vals)
    assert torch.equal(y0, y1)
    for k in g0:
        assert torch.equal(g0[k], g1[k]), k
    rounded = {k: v.to(torch.bfloat16).float() for k, v in vals.items()}
    y2, g2 = run(rounded)
    err = float((y2 - y0).abs().max()) / float(y0.abs().max())
    assert 0 < err < 5e-2, err
    assert all(g is not None and bool(torch.isfinite(g).all()) for g in g2.values())


def test_hrnet_mask_matched_machinery():
    """oracle/hrnet_restatement.py under ``restatement.substituted``: recording the oracle's own boundary tensors and
    gradients, then substituting them back with the backward re-synchronisation on, is the identity — same logits, same
    parameter gradients, and every logged (own, stored) gradient pair equal; the number of boundaries is the number of
    convs in front of a BatchNorm."""
    from oracle import hrnet_restatement as H
    nc, seed = 5, 2
    g = torch.Generator().manual_seed(seed)
    x, t = torch.randn(2, 3, 64, 64, generator=g), torch.randint(0, nc, (2, 64, 64), generator=g)

    def run(values, record=False, grad_log=None):
        sd = H.make_state(nc, seed=seed)
        for v in sd.values():
            if v.is_floating_point():
                v.requires_grad_(True)
        for k in sd:
            if 'running_' in k:
                sd[k] = sd[k].detach()
        ctx = R.substituted(values, record=record, grad_log=grad_log) if values is not None else None
        if ctx is not None:
            with ctx:
                y = H.hrnet(sd, x, training=True)
                loss = R.compute_loss(y, t)
                loss.backward()
        else:
            y = H.hrnet(sd, x, training=True)
            R.compute_loss(y, t).backward()
        return y.detach(), {k: v.grad for k, v in sd.items() if v.requires_grad}

    y0, g0 = run(None)
    values = {}
    y1, g1 = run(values, record=True)
    assert torch.equal(y0, y1)
    convs = [k for k in g0 if k.endswith('.weight') and g0[k].dim() == 4 and k != 'final_layer.weight']
    assert len([k for k in values if k.endswith('.raw')]) == len(convs)
    assert len([k for k in values if k.endswith('.graw')]) == len(convs)
    assert len([k for k in values if k.startswith('sum_relu.')]) == 2 + 3 + 1
    log = {}
    y2, g2 = run(values, grad_log=log)
    assert len(log) == len(convs)
    assert torch.allclose(y0, y2, rtol=1e-5, atol=1e-6)
    for k, (own, stored) in log.items():
        assert torch.allclose(own, stored, rtol=1e-4, atol=1e-7 + 1e-5 * float(stored.abs().max())), k
    for k in g0:
        assert g2[k] is not None, k
        assert torch.allclose(g0[k], g2[k], rtol=1e-3, atol=1e-5 * float(g0[k].abs().max()) + 1e-9), k
