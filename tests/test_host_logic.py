"""Host-side logic of the drop-in package that needs no GPU: metric arithmetic, module surface / state-dict
keys, loud failures, bucket construction."""
import numpy as np
import pytest
import torch

import pytorch_segmentation_b200 as P
from oracle import golden as G
from oracle import restatement as R
from pytorch_segmentation_b200 import _lib
from pytorch_segmentation_b200.parallel import GradBucketer
from pytorch_segmentation_b200.utils.utils import counts_from_confusion


def test_compute_metrics_bit_exact_vs_reference_golden():
    gold = G.load('metrics')
    for i in range(int(gold['ncases'])):
        tp, fn, fp = (torch.from_numpy(gold['%s%d' % (n, i)]).float() for n in ('tp', 'fn', 'fp'))
        keep = (tp.clone(), fn.clone(), fp.clone())
        T, Pp, Rr, iou, F1 = P.compute_metrics(tp, fn, fp)
        assert np.array_equal(iou.numpy(), gold['iou%d' % i]) and np.array_equal(Pp.numpy(), gold['P%d' % i])
        assert np.array_equal(Rr.numpy(), gold['R%d' % i]) and np.array_equal(F1.numpy(), gold['F1%d' % i])
        assert np.float32(iou.mean().item()) == gold['miou%d' % i]
        assert all(torch.equal(a, b) for a, b in zip((tp, fn, fp), keep))       # inputs untouched


def test_zero_denominator_guards():
    z = torch.zeros(4)
    T, Pp, Rr, iou, F1 = P.compute_metrics(z.clone(), z.clone(), z.clone())
    assert torch.equal(iou, z) and torch.equal(Pp, z) and torch.equal(Rr, z) and torch.equal(F1, z)


def test_counts_from_confusion_matches_loop():
    g = torch.Generator().manual_seed(0)
    for nc in (2, 7, 21):
        pred, tgt = torch.randint(0, nc, (5000,), generator=g), torch.randint(0, nc, (5000,), generator=g)
        cm = R.confusion_matrix(pred, tgt, nc)
        tp, fn, fp = counts_from_confusion(cm)
        a, b, c = R.tp_fn_fp_loop(pred, tgt, nc)
        assert torch.equal(tp, a) and torch.equal(fn, b) and torch.equal(fp, c)


@pytest.mark.parametrize('kind,nc', [('deeplabv3plus', 21), ('unet', 2)])
def test_state_dict_keys_and_shapes_match_reference_layout(kind, nc):
    ours = (P.DeepLabV3Plus if kind == 'deeplabv3plus' else P.UNet)(nc)
    ref_sd = R.OracleModel(kind, nc, seed=0).state_dict()
    sd = ours.state_dict()
    assert set(sd) == set(ref_sd)
    for k in sd:
        assert tuple(sd[k].shape) == tuple(ref_sd[k].shape), k
    ours.load_state_dict(ref_sd)        # strict


def test_aspp_constructor_surface():
    m = P.ASPP(64, 32)                  # default rates [12, 24, 36] (models/aspp.py:23)
    assert [b.dilation for b in list(m.blocks)[2:]] == [12, 24, 36]
    assert m.project[0].in_channels == 32 * 5
    assert 'blocks.0.gap.1.0.weight' in m.state_dict() and 'blocks.0.gap.1.1.running_var' in m.state_dict()


def test_no_cpu_fallback():
    m = P.ASPP(64, 32, [2])
    with pytest.raises(RuntimeError, match='no CPU fallback'):
        m(torch.zeros(2, 64, 8, 8))
    with pytest.raises(RuntimeError, match='no CPU fallback'):
        P.compute_loss(torch.zeros(1, 3, 4, 4), torch.zeros(1, 4, 4, dtype=torch.int64))


def test_unsupported_convnormact_is_an_error_not_a_fallback():
    with pytest.raises(NotImplementedError):
        P.ConvNormAct(8, 8, 3, 3)                      # stride 3
    with pytest.raises(NotImplementedError):
        P.ConvNormAct(8, 8, 3, 2, dilation=2)          # strided and dilated
    with pytest.raises(NotImplementedError):
        P.ConvNormAct(8, 8, 3, 1, groups=8)            # depthwise
    with pytest.raises(NotImplementedError):
        P.ConvNormAct(8, 8, 5)
    assert P.ConvNormAct(8, 8, 3, 2)[0].stride == (2, 2)          # HRNet transitions / fuse layers (models/hrnet.py:213-224)
    assert P.ConvNormAct(3, 64, 3, 2, activate=None)[0].in_channels == 3   # HRNet stem (models/hrnet.py:260)
    with pytest.raises(NotImplementedError):
        P.ConvNormAct(8, 8, activate=torch.nn.Sigmoid())
    P.ConvNormAct(8, 8, 1, activate=None)


def test_missing_library_fails_loudly(monkeypatch, tmp_path):
    monkeypatch.setattr(_lib, '_lib', None)
    monkeypatch.setattr(_lib, 'LIB_PATH', str(tmp_path / 'nope.so'))
    with pytest.raises(RuntimeError, match='libb200seg.so is missing'):
        _lib.load()


def test_bucketer_layout_reverse_order_and_sizes():
    ps = [torch.nn.Parameter(torch.zeros(n)) for n in (10, 300000, 200000, 7000000, 5)]
    b = GradBucketer(ps, bucket_mb=8.0, first_bucket_mb=1.0)
    # reverse registration order: first (small) bucket holds the last parameters
    assert b.bucket_sizes()[0] == 5
    assert sum(b.bucket_sizes()) == sum(p.numel() for p in ps)
    assert b.num_buckets >= 3
    for p in ps:
        p.grad = torch.ones_like(p)
        b._on_grad(p)
    b.finish()
    for flat, slots in b._buckets:
        assert float(flat.sum()) == flat.numel()
        for p, off, n in slots:
            assert p.grad.data_ptr() == flat[off:off + n].data_ptr()


def test_bucketer_early_delivery_feeds_buckets_before_autograd():
    from pytorch_segmentation_b200 import engine as E
    ps = [torch.nn.Parameter(torch.zeros(n)) for n in (4, 6, 300000, 8)]
    b = GradBucketer(ps, bucket_mb=1.0, first_bucket_mb=1.0).attach()
    try:
        assert E.GRAD_SINK is not None
        launched = []
        b._launch = lambda bi, _orig=b._launch: (launched.append(bi), _orig(bi))[1]
        # a hand-sequenced node finished the stage that owns ps[2] and ps[3] (reverse order: they fill bucket 0 / 1 first)
        grads = {ps[3]: torch.full((8,), 2.0), ps[2]: torch.full((300000,), 3.0), ps[1]: None}
        E.deliver_grads(grads, [ps[3], ps[2], ps[1]])
        assert grads[ps[3]] is None and grads[ps[2]] is None          # the node returns None for them
        assert ps[3].grad is not None and float(ps[3].grad.sum()) == 16.0 and float(ps[2].grad[5]) == 3.0
        assert len(launched) >= 1                                       # a complete bucket went out immediately
        b._on_grad(ps[3])                                               # autograd still runs the hook: a no-op
        # the rest arrives through autograd's hooks as usual
        for p in ps[:2]:
            p.grad = torch.ones_like(p)
            b._on_grad(p)
        b.finish()
        assert sorted(set(launched)) == list(range(b.num_buckets))
        for flat, slots in b._buckets:
            for p, off, n in slots:
                assert p.grad.data_ptr() == flat[off:off + n].data_ptr()
        # a parameter that already holds a gradient (accumulation) is not delivered early
        grads = {ps[0]: torch.ones(4)}
        E.deliver_grads(grads, [ps[0]])
        assert grads[ps[0]] is not None
    finally:
        b.remove()
    assert E.GRAD_SINK is None


def test_class_map_loss_predicate_covers_the_heads_and_rejects_the_rest():
    from pytorch_segmentation_b200 import engine as E
    ok = E.lowres_ce_supported
    assert ok((16, 128, 128, 24), 21, 512, 512, True)            # DeepLabV3+ x4, align_corners (deeplabv3plus.py:40-43)
    assert ok((4, 160, 160, 8), 2, 320, 320, True)               # UNet x2 (unet.py:52-55)
    assert ok((8, 128, 256, 24), 19, 512, 1024, False)           # HRNet x4, half-pixel centres (hrnet.py:400-404)
    assert not ok((1, 64, 64, 24), 21, 64, 64, True)             # no upsample: the logits path
    assert not ok((1, 16, 16, 24), 21, 128, 128, True)           # x8: more than four patch windows per cell
    assert not ok((1, 64, 64, 40), 33, 256, 256, True)           # more than 32 classes
    assert not ok((1, 1, 64, 24), 21, 4, 256, True)              # degenerate class map
