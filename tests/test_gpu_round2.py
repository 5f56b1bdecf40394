"""Round-2 rows of SURVEY.md §8: the input side (f4: uint8 images / masks straight into the kernels), the fused evaluation
from the class map (a9 / a13), Adam / AdamW and gradient accumulation (f3), and the graph step's hyper-parameter plumbing.

Bit-exact where the arithmetic is integer or a fixed fp32 op sequence (normalisation, space-to-depth, argmax, confusion),
tolerances stated where fp32 accumulation order differs (optimizer updates, losses)."""
import copy

import pytest
import torch

pytestmark = pytest.mark.gpu

MEAN = (123.675, 116.28, 103.53)      # utils/datasets.py:201-205
STD = (58.395, 57.12, 57.375)


@pytest.fixture(scope='module')
def P(lib):
    import pytorch_segmentation_b200 as pkg
    return pkg


@pytest.fixture(scope='module')
def E(lib):
    from pytorch_segmentation_b200 import engine
    return engine


def _post_fetch_reference(imgs_u8):
    """utils/datasets.py:199-205, the reference's three fp32 tensor ops (on the CPU)."""
    imgs = imgs_u8.cpu().float()
    imgs -= torch.FloatTensor(list(MEAN)).reshape(1, 3, 1, 1)
    imgs /= torch.FloatTensor(list(STD)).reshape(1, 3, 1, 1)
    return imgs


# ---------------------------------------------------------------------------------------------------
# f4: input side
# ---------------------------------------------------------------------------------------------------
@pytest.mark.parametrize('n,h,w', [(2, 64, 96), (3, 37, 53), (16, 512, 512)])
def test_normalize_u8_bit_exact(E, n, h, w):
    g = torch.Generator().manual_seed(n * h + w)
    x = torch.randint(0, 256, (n, 3, h, w), generator=g, dtype=torch.uint8)
    want = _post_fetch_reference(x)
    got = E.normalize_u8(x.cuda())
    assert got.dtype == torch.float32 and torch.equal(got.cpu(), want)
    # channels_last uint8 (an NHWC buffer, what cv2 decodes to) is read in place
    xl = x.cuda().contiguous(memory_format=torch.channels_last)
    assert torch.equal(E.normalize_u8(xl).cpu(), want)


@pytest.mark.parametrize('n,h,w,oh,ow', [(2, 64, 96, 96, 64), (3, 37, 53, 64, 33), (2, 224, 224, 160, 320),
                                          (16, 512, 512, 736, 384)])
def test_resize_nearest_u8_matches_interpolate(E, n, h, w, oh, ow):
    """multi_scale of post_fetch_fn (utils/datasets.py:206-212): F.interpolate(imgs, (h, w)) — nearest — after the
    normalisation == the uint8 nearest resize before it, bit for bit."""
    g = torch.Generator().manual_seed(h * w + oh)
    x = torch.randint(0, 256, (n, 3, h, w), generator=g, dtype=torch.uint8)
    want = torch.nn.functional.interpolate(_post_fetch_reference(x), (oh, ow))
    y = E.resize_nearest_u8(x.cuda(), oh, ow)
    assert y.dtype == torch.uint8 and list(y.shape) == [n, 3, oh, ow]
    assert torch.equal(E.normalize_u8(y).cpu(), want)
    m = torch.randint(0, 21, (n, h, w), generator=g, dtype=torch.uint8)
    want_m = torch.nn.functional.interpolate(m[:, None].float(), (oh, ow))[:, 0].to(torch.uint8)
    assert torch.equal(E.resize_nearest_u8(m.cuda(), oh, ow).cpu(), want_m)


def test_post_fetch_fn_multi_scale_follows_the_reference_draw(P):
    """Same ``random`` state => the same scale draw and output size as the reference's post_fetch_fn, same pixels."""
    import random
    from pytorch_segmentation_b200.utils.utils import model_input, post_fetch_fn
    g = torch.Generator().manual_seed(5)
    x = torch.randint(0, 256, (2, 3, 224, 224), generator=g, dtype=torch.uint8)
    t = torch.randint(0, 21, (2, 224, 224), generator=g, dtype=torch.uint8)
    for seed in (0, 1, 2, 3):
        random.seed(seed)
        scale = random.uniform(0.7, 1.5)                              # utils/datasets.py:209-211
        h = w = int(224 * scale / 32) * 32
        want = torch.nn.functional.interpolate(_post_fetch_reference(x), (h, w))
        random.seed(seed)
        imgs, segs = post_fetch_fn((x.cuda(), t.cuda()), multi_scale=True)
        assert imgs.dtype == torch.uint8 and list(imgs.shape) == [2, 3, h, w]
        assert list(segs.shape) == [2, 224, 224]                      # the reference leaves the masks alone
        assert torch.equal(model_input(imgs).cpu(), want)
        random.seed(seed)
        imgs, segs = post_fetch_fn((x.cuda(), t.cuda()), multi_scale=True, resize_masks=True)
        assert list(segs.shape) == [2, h, w]


@pytest.mark.parametrize('n,h,w,r,pad', [(2, 64, 96, 7, 3), (2, 62, 70, 3, 1), (16, 512, 512, 7, 3)])
def test_space_to_depth_u8_equals_normalise_then_fp32_path(E, n, h, w, r, pad):
    """One pass over 1-byte pixels == post_fetch_fn followed by the fp32 space-to-depth pass, bit for bit (bf16)."""
    g = torch.Generator().manual_seed(h + w)
    x = torch.randint(0, 256, (n, 3, h, w), generator=g, dtype=torch.uint8)
    ref, oh, ow = E.space_to_depth(_post_fetch_reference(x).cuda(), r, r, pad)
    for xin in (x.cuda(), x.cuda().contiguous(memory_format=torch.channels_last)):
        got, oh2, ow2 = E.space_to_depth_u8(xin, r, r, pad)
        assert (oh2, ow2) == (oh, ow) and got.c == ref.c
        assert torch.equal(got.t.view(torch.int16), ref.t.view(torch.int16))


def test_models_take_uint8_batches(P):
    """post_fetch_fn drop-in: uint8 images + uint8 masks give the logits / loss of the float + int64 path exactly
    (eval mode: no batch-statistics atomics, so the forward is bit-reproducible)."""
    from pytorch_segmentation_b200.utils.utils import post_fetch_fn
    g = torch.Generator().manual_seed(5)
    x8 = torch.randint(0, 256, (2, 3, 128, 160), generator=g, dtype=torch.uint8)
    t8 = torch.randint(0, 7, (2, 128, 160), generator=g, dtype=torch.uint8)
    xf = _post_fetch_reference(x8).cuda()
    for ctor in (P.DeepLabV3Plus, P.UNet, lambda nc: P.HRNet(nc)):
        torch.manual_seed(3)
        m = ctor(7).cuda().eval()
        imgs, segs = post_fetch_fn((x8.cuda(), t8.cuda()))
        with torch.no_grad():
            y8 = m(imgs)
            l8 = P.compute_loss(y8, segs, m)
            yf = m(xf)
            lf = P.compute_loss(yf, t8.cuda().long(), m)
        assert torch.equal(y8, yf), ctor
        assert abs(float(l8) - float(lf)) <= 1e-6 * abs(float(lf))


@pytest.mark.parametrize('dtype', [torch.uint8, torch.int32])
def test_narrow_targets_equal_int64(P, E, dtype):
    g = torch.Generator().manual_seed(11)
    n, nc, h, w = 2, 21, 96, 128
    logits = torch.randn(n, nc, h, w, generator=g).cuda()
    t64 = torch.randint(0, nc, (n, h, w), generator=g).cuda()
    tn = t64.to(dtype)
    # logits path
    la, lb = logits.clone().requires_grad_(True), logits.clone().requires_grad_(True)
    a, b = P.compute_loss(la, t64, None), P.compute_loss(lb, tn, None)
    a.backward()
    b.backward()
    assert abs(float(a) - float(b)) <= 1e-6 * abs(float(a)) and torch.equal(la.grad, lb.grad)
    assert torch.equal(E.argmax_confusion(logits, t64), E.argmax_confusion(logits, tn))
    pred = E.argmax(logits)
    assert torch.equal(E.confusion(pred.reshape(-1), t64.reshape(-1), nc), E.confusion(pred.reshape(-1), tn.reshape(-1), nc))
    # class-map path
    lr = torch.randn(n, h // 4, w // 4, 24, generator=g).cuda()
    acc64, accn = E.lowres_ce_fwd(lr, nc, t64), E.lowres_ce_fwd(lr, nc, tn)
    assert abs(float(acc64[0]) - float(accn[0])) <= 1e-9 * abs(float(acc64[0]))
    assert torch.equal(acc64[1:2].view(torch.int64), accn[1:2].view(torch.int64))
    with pytest.raises(RuntimeError):
        P.compute_loss(logits, t64.to(torch.int16), None)


# ---------------------------------------------------------------------------------------------------
# a9 / a13: evaluation from the class map
# ---------------------------------------------------------------------------------------------------
@pytest.mark.parametrize('n,h,w,up,nc,ct', [(2, 37, 41, 4, 21, 24), (3, 40, 24, 2, 2, 8), (16, 128, 128, 4, 21, 24)])
def test_class_map_confusion_bit_exact(E, n, h, w, up, nc, ct):
    g = torch.Generator().manual_seed(n + h)
    lr = torch.randn(n, h, w, ct, generator=g).cuda()
    tgt = torch.randint(0, nc, (n, h * up, w * up), generator=g).cuda()
    logits = E.logits_upsample_fwd(lr, nc, h * up, w * up)
    want_cm = E.argmax_confusion(logits, tgt)
    cm, pred = E.class_map_confusion(lr, nc, tgt, True, want_pred=True)
    assert torch.equal(pred, E.argmax(logits))
    assert torch.equal(pred, logits.max(1)[1])               # outputs.max(1)[1] of test.py:31
    assert torch.equal(cm, want_cm) and int(cm.sum()) == tgt.numel()
    # accumulates into a caller's matrix; wrong-sized matrices are refused
    cm2, _ = E.class_map_confusion(lr, nc, tgt.to(torch.uint8), True, cm=cm.clone())
    assert torch.equal(cm2, 2 * want_cm)
    with pytest.raises(ValueError):
        E.class_map_confusion(lr, nc, tgt, True, cm=torch.zeros((nc + 1, nc + 1), dtype=torch.int64, device='cuda'))


def test_class_map_confusion_tie_rule(E):
    """Ties go to the LOWEST class index (the first maximum, like ``outputs.max(1)[1]`` on the written logits): integer
    class maps interpolate to exact ties at every cell centre and along every edge between equal cells."""
    n, h, w, nc, ct = 2, 9, 11, 5, 8
    g = torch.Generator().manual_seed(0)
    lr = torch.randint(0, 2, (n, h, w, ct), generator=g).float().cuda()       # values in {0, 1}: ties everywhere
    tgt = torch.randint(0, nc, (n, h * 4, w * 4), generator=g).cuda()
    logits = E.logits_upsample_fwd(lr, nc, h * 4, w * 4)
    cm, pred = E.class_map_confusion(lr, nc, tgt, True, want_pred=True)
    assert torch.equal(pred, logits.max(1)[1]) and torch.equal(pred, E.argmax(logits))
    ties = (logits == logits.max(1, keepdim=True)[0]).sum(1) > 1
    assert float(ties.float().mean()) > 0.2                      # the case really is exercised
    first = (logits == logits.max(1, keepdim=True)[0]).float().argmax(1)
    assert torch.equal(pred, first)
    zero = torch.zeros(n, h, w, ct, device='cuda')
    _, p0 = E.class_map_confusion(zero, nc, tgt, True, want_pred=True)
    assert int(p0.abs().sum()) == 0                              # all classes tie -> class 0


def test_test_loop_uses_the_class_map(P, E, monkeypatch):
    """test(model, fetcher) (test.py:14-73) on a DeepLabV3+ head: the prediction comes from the class map (no pass over
    the logits) and mIoU equals the logits path's bit for bit."""
    torch.manual_seed(2)
    nc = 6
    m = P.DeepLabV3Plus(nc).cuda().eval()
    g = torch.Generator().manual_seed(4)
    batches = [(torch.randn(2, 3, 128, 128, generator=g).cuda(), torch.randint(0, nc, (2, 128, 128), generator=g).cuda())
               for _ in range(2)]

    class F:
        def __init__(self, b):
            self.b = b
            self.loader = type('L', (), {})()
            self.loader.dataset = type('D', (), {'classes': ['c%d' % i for i in range(nc)]})()

        def __iter__(self):
            return iter(self.b)

    calls = {'cm': 0, 'logits': 0}
    real_cm, real_lg = E.class_map_confusion, E.argmax_confusion
    monkeypatch.setattr(E, 'class_map_confusion', lambda *a, **k: (calls.__setitem__('cm', calls['cm'] + 1), real_cm(*a, **k))[1])
    monkeypatch.setattr(E, 'argmax_confusion', lambda *a, **k: (calls.__setitem__('logits', calls['logits'] + 1), real_lg(*a, **k))[1])
    miou_fused = P.test(m, F(batches), verbose=False)
    assert calls == {'cm': 2, 'logits': 0}
    monkeypatch.setattr(E, 'LOWRES_CE', False)

    class Untag(torch.nn.Module):
        def __init__(self, inner):
            super().__init__()
            self.inner = inner

        def forward(self, x):
            return self.inner(x).clone()                      # a fresh tensor carries no class-map tag

    miou_logits = P.test(Untag(m), F(batches), verbose=False)
    assert calls['logits'] == 2
    assert miou_fused == miou_logits
    with pytest.raises(ValueError):
        bad = F(batches)
        bad.loader.dataset.classes = ['a', 'b']
        P.test(m, bad, verbose=False)


# ---------------------------------------------------------------------------------------------------
# f3: Adam / AdamW, hyper-parameters under the graph, accumulation
# ---------------------------------------------------------------------------------------------------
@pytest.mark.parametrize('name,kw', [('AdamW', dict(lr=3e-3, weight_decay=0.05)), ('Adam', dict(lr=3e-3, weight_decay=0.01)),
                                     ('AdamW', dict(lr=1e-2, betas=(0.8, 0.95), eps=1e-6, weight_decay=0.0))])
def test_adam_multi_matches_torch(P, name, kw):
    g = torch.Generator().manual_seed(1)
    shapes = [(64, 3, 7, 7), (256,), (33, 17), (1,), (2048, 256, 1, 1), (5, 8, 3, 3)]
    ps = [torch.randn(s, generator=g).cuda() for s in shapes]
    ps[0] = ps[0].contiguous(memory_format=torch.channels_last)
    a = [p.clone().requires_grad_(True) for p in ps]
    b = [p.clone().requires_grad_(True) for p in ps]
    oa = getattr(P.optim, name)(a, **kw)
    ob = getattr(torch.optim, name)(b, **kw)
    for step in range(5):
        if step == 3:                                  # a schedule moving the learning rate
            for o in (oa, ob):
                o.param_groups[0]['lr'] *= 0.5
        for x, y in zip(a, b):
            gr = torch.randn(x.shape, generator=g).cuda()
            x.grad = gr.clone().contiguous(memory_format=torch.channels_last) if x.dim() == 4 and x.shape == (64, 3, 7, 7) else gr.clone()
            y.grad = x.grad.clone()
        oa.step()
        ob.step()
        for x, y in zip(a, b):
            assert float((x - y).abs().max()) <= 2e-6 * max(1.0, float(y.abs().max())), (name, step, tuple(x.shape))
    sa, sb = oa.state_dict()['state'], ob.state_dict()['state']
    assert set(sa[0]) >= {'step', 'exp_avg', 'exp_avg_sq'} and float(sa[0]['step']) == float(sb[0]['step']) == 5
    assert float((sa[4]['exp_avg_sq'] - sb[4]['exp_avg_sq']).abs().max()) <= 1e-6 * float(sb[4]['exp_avg_sq'].abs().max())


def _params(m):
    return torch.cat([p.detach().flatten() for p in m.parameters()]).clone()


def test_graph_replay_follows_param_groups(P):
    """The captured step reads lr / momentum / weight decay from the device at replay time: lr = 0 freezes the
    parameters exactly, restoring it resumes training, and the momentum path is the steady-state one (a step captured
    on the first-step path would reset the momentum buffer every replay)."""
    torch.manual_seed(5)
    m = P.DeepLabV3Plus(4).cuda().train()
    opt = P.optim.SGD(m.parameters(), lr=1e-2, momentum=0.9)
    s = P.TrainStep(m, opt, graph=True, warmup=1)
    with pytest.raises(ValueError):
        P.TrainStep(m, opt, graph=True, warmup=0)
    x, t = torch.randn(2, 3, 64, 64).cuda(), torch.randint(0, 4, (2, 64, 64)).cuda()
    s(x, t)                     # eager warm-up
    s(x, t)                     # capture + replay
    p0 = _params(m)
    s(x, t)
    p1 = _params(m)
    assert float((p1 - p0).abs().max()) > 0
    opt.param_groups[0]['lr'] = 0.0
    s(x, t)
    p2 = _params(m)
    assert torch.equal(p2, p1)                          # the replay saw lr = 0
    buf = opt.state[m.cls_conv.weight]['momentum_buffer'].clone()
    s(x, t)
    buf2 = opt.state[m.cls_conv.weight]['momentum_buffer']
    assert not torch.equal(buf, buf2)                   # momentum keeps integrating: m = 0.9 m + g, not m = g
    grad = m.cls_conv.weight.grad
    assert float((buf2 - (0.9 * buf + grad)).abs().max()) <= 1e-5 * max(1.0, float(buf2.abs().max()))
    opt.param_groups[0]['lr'] = 1e-2
    s(x, t)
    assert float((_params(m) - p2).abs().max()) > 0
    # eval after the replays reads the CURRENT weights (packed bf16 copies refreshed): a fresh model given the same
    # state dict must produce the same logits
    m.eval()
    fresh = P.DeepLabV3Plus(4).cuda().eval()
    fresh.load_state_dict(m.state_dict())
    with torch.no_grad():
        assert torch.equal(m(x), fresh(x))


def _floor_check(ua, ub, uc, what):
    """A bf16 Conv-BN-ReLU stack trained on random data is chaotic (fp32 atomics reorder -> bf16 roundings flip -> ReLU
    masks flip; tests/test_gpu_trainstep.py): two IDENTICAL eager runs A and B differ, and that distance is the
    yardstick for the run under test C."""
    floor = float((ua - ub).norm())
    assert float(ua.norm()) > 0
    assert float((ua - uc).norm()) <= max(3.0 * floor, 0.05 * float(ua.norm())), \
        (what, float((ua - uc).norm()), floor, float(ua.norm()))


def test_adamw_under_graph_advances_its_step_count(P):
    """Three copies, same batches: A and B step AdamW eagerly, C through the captured graph, whose replays must see the
    advancing step count (bias corrections) through the device-resident hyper-parameters."""
    torch.manual_seed(6)
    nc, n, h, w = 5, 4, 128, 128
    m_a = P.DeepLabV3Plus(nc).cuda().train()
    m_b, m_c = copy.deepcopy(m_a), copy.deepcopy(m_a)
    kw = dict(lr=1e-4, weight_decay=0.01)
    steps = [P.TrainStep(m, P.optim.AdamW(m.parameters(), **kw), graph=gr, warmup=1)
             for m, gr in ((m_a, False), (m_b, False), (m_c, True))]
    g = torch.Generator().manual_seed(2)
    batches = [(torch.randn(n, 3, h, w, generator=g).cuda(), torch.randint(0, nc, (n, h, w), generator=g).cuda())
               for _ in range(5)]
    init = _params(m_a)
    for x, t in batches:
        for s in steps:
            s(x, t)
    for s in steps:
        assert int(s.optimizer.param_groups[0]['_b200seg_step']) == 5
    p0 = next(iter(steps[2].optimizer.state.values()))
    assert float(p0['step']) == 5
    _floor_check(_params(m_a) - init, _params(m_b) - init, _params(m_c) - init, 'AdamW graph vs eager')
    # a replay that kept the capture-time bias corrections (step 2 forever) would take steps ~1.6x too large by step 5:
    # compare the size of the LAST update between eager and graph
    before = [_params(m) for m in (m_a, m_c)]
    x, t = batches[0]
    steps[0](x, t)
    steps[2](x, t)
    da, dc = float((_params(m_a) - before[0]).norm()), float((_params(m_c) - before[1]).norm())
    assert abs(da - dc) <= 0.1 * da, (da, dc)


def test_gradient_accumulation(P):
    """accumulate=2 (train.py:94's default): the optimizer steps on every second call with the mean of the two
    micro-batches' gradients.  A and B: the sequence written out by hand (zero_grad; two half-weighted backward passes;
    step) on two copies — their distance is the run-to-run floor; C: TrainStep(accumulate=2) eager; D: the same under a
    CUDA graph (one captured graph over both micro-steps)."""
    torch.manual_seed(8)
    nc, n, h, w = 5, 4, 128, 128
    m_a = P.DeepLabV3Plus(nc).cuda().train()
    m_b, m_c, m_d = copy.deepcopy(m_a), copy.deepcopy(m_a), copy.deepcopy(m_a)
    mk = lambda m: P.optim.SGD(m.parameters(), lr=1e-3, momentum=0.9)
    o_a, o_b = mk(m_a), mk(m_b)
    s_c = P.TrainStep(m_c, mk(m_c), graph=False, accumulate=2)
    s_d = P.TrainStep(m_d, mk(m_d), graph=True, warmup=1, accumulate=2)
    g = torch.Generator().manual_seed(1)
    data = [(torch.randn(n, 3, h, w, generator=g).cuda(), torch.randint(0, nc, (n, h, w), generator=g).cuda())
            for _ in range(8)]
    init = _params(m_a)
    for i in range(0, 8, 2):
        for m, o in ((m_a, o_a), (m_b, o_b)):
            o.zero_grad(set_to_none=True)
            for x, t in data[i:i + 2]:
                (P.compute_loss(m(x), t, m) * 0.5).backward()
            o.step()
        for s, m in ((s_c, m_c), (s_d, m_d)):
            before = _params(m)
            r1 = s(*data[i])
            assert torch.equal(_params(m), before), 'no update on the first micro-batch'
            r2 = s(*data[i + 1])
            assert float((_params(m) - before).abs().max()) > 0
            assert r2 is not None and bool(torch.isfinite(r2))
            if s is s_d and i >= 2:
                assert r1 is None                              # staging call of the captured two-micro-step graph
    ua = _params(m_a) - init
    _floor_check(ua, _params(m_b) - init, _params(m_c) - init, 'eager accumulate')
    _floor_check(ua, _params(m_b) - init, _params(m_d) - init, 'graph accumulate')
    assert s_d.launches_per_step is not None and s_d.launches_per_step > 800      # two forward/backward passes in one graph
    for m in (m_a, m_c, m_d):
        assert int(m.aspp.project[1].num_batches_tracked) == 8


def test_trainstep_defers_the_logits_and_loss_is_unchanged(P, E):
    """TrainStep lets the head skip writing the full-resolution logits (its loss comes from the class map); the loss is
    the same as the plain model(x) -> compute_loss sequence, and a loss that cannot use the class map gets them
    materialised first."""
    torch.manual_seed(4)
    m = P.DeepLabV3Plus(5).cuda().eval()            # eval: deterministic forward
    x, t = torch.randn(2, 3, 128, 128).cuda(), torch.randint(0, 5, (2, 128, 128)).cuda()
    with torch.no_grad():
        want = float(P.compute_loss(m(x), t, m))
        E.DEFER_LOGITS = True
        try:
            y = m(x)
        finally:
            E.DEFER_LOGITS = False
        assert y._b200seg_class_map[4] is True      # allocated, not written
        got = float(P.compute_loss(y, t, m))
        assert abs(got - want) <= 1e-6 * abs(want)
        # half-size targets: compute_loss resizes the logits (utils/utils.py:18-20) -> needs them -> fills them in
        E.DEFER_LOGITS = True
        try:
            y2 = m(x)
        finally:
            E.DEFER_LOGITS = False
        t_half = t[:, ::2, ::2].contiguous()
        got2 = float(P.compute_loss(y2, t_half, m))
        want2 = float(P.compute_loss(m(x), t_half, m))
        assert y2._b200seg_class_map[4] is False and torch.equal(y2, m(x))
        assert abs(got2 - want2) <= 1e-6 * abs(want2)
    s = P.TrainStep(m.train(), P.optim.SGD(m.parameters(), lr=0.0), graph=False)
    before = _lib_launches()
    s(x, t)
    assert E.DEFER_LOGITS is False


def _lib_launches():
    from pytorch_segmentation_b200 import _lib
    return _lib.launches()


@pytest.mark.parametrize('n,nc,h,w,up,dtype', [(2, 21, 32, 40, 4, torch.int64), (3, 2, 80, 80, 2, torch.uint8),
                                               (16, 21, 128, 128, 4, torch.uint8)])
def test_fused_class_map_loss_and_gradient(P, E, monkeypatch, n, nc, h, w, up, dtype):
    """Training takes loss and class-map gradient from one launch (gradient unscaled, then * g / count in place): the
    same loss bit for bit and the same gradient to fp32 round-off as the two-launch form, ignored pixels included."""
    from pytorch_segmentation_b200.utils.utils import _ClassMapCrossEntropyFn
    g = torch.Generator().manual_seed(n * h + nc)
    ct = (nc + 7) // 8 * 8
    lr = torch.randn(n, h, w, ct, generator=g).cuda()
    t = torch.randint(0, nc, (n, h * up, w * up), generator=g)
    if dtype == torch.int64:
        t[0, :7, :] = -100
    t = t.to(dtype).cuda()
    assert E.lowres_ce_supported(lr.shape, nc, h * up, w * up, True)
    out = {}
    for fused in (False, True):
        monkeypatch.setattr(E, 'FUSED_CE', fused)
        x = lr.clone().requires_grad_(True)
        loss = _ClassMapCrossEntropyFn.apply(x, t, nc, True)
        (loss * 3.0).backward()
        out[fused] = (loss.detach().clone(), x.grad.clone())
    assert torch.equal(out[True][0], out[False][0])
    ref = out[False][1]
    assert float((out[True][1] - ref).abs().max()) <= 2e-6 * float(ref.abs().max())
    assert float(out[True][1][..., nc:].abs().max() if ct > nc else 0.0) == 0.0
