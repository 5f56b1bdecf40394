"""Per-kernel parity of the C-ABI entry points against the CPU oracle (PyTorch fp32 on the host — the same
library calls the reference makes).  Tolerances: bf16 kernels max|d|/max|ref| <= 2e-2 (BASELINE.json
north_star), in practice ~4e-3; integer work bit-exact."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

TOL = 2e-2


def rel_err(got, ref):
    got, ref = got.detach().float().cpu(), ref.detach().float().cpu()
    return float((got - ref).abs().max()) / max(float(ref.abs().max()), 1e-30)


def bf16_round(t):
    return t.to(torch.bfloat16).float()


@pytest.fixture(scope='module')
def E(lib):
    from pytorch_segmentation_b200 import engine
    return engine


def nhwc(x):
    """NCHW fp32 CPU -> Act on cuda"""
    from pytorch_segmentation_b200.engine import Act
    return Act(x.permute(0, 2, 3, 1).contiguous().to('cuda', torch.bfloat16))


def from_act(a):
    return a.nchw().float().cpu()


CONV_CASES = [
    # n, h, w, c, k, ksize, dil
    (1, 8, 16, 64, 32, 1, 1),
    (2, 32, 32, 128, 256, 3, 6),
    (2, 32, 32, 192, 256, 3, 12),
    (2, 32, 32, 64, 64, 3, 18),
    (3, 10, 10, 160, 64, 3, 1),
    (2, 20, 20, 352, 128, 3, 1),
    (4, 1, 1, 256, 256, 1, 1),
    (1, 40, 40, 128, 352, 3, 2),
    (2, 17, 23, 88, 8, 3, 1),
    # <= 64 output channels over enough pixels: the operand-role swap of igemm_kernel (256-pixel patches, transposing
    # epilogue), with a width that leaves border patches and a 32-channel output
    (4, 128, 128, 64, 64, 3, 1),
    (3, 128, 120, 256, 64, 1, 1),
    (5, 100, 84, 64, 32, 3, 2),
    # dilated, one CTA per (tap, tile) already fills the GPU (the ASPP weight-gradient regime)
    (2, 32, 32, 2048, 256, 3, 12),
    (1, 32, 32, 1152, 128, 3, 18),
]


@pytest.mark.parametrize('n,h,w,c,k,ks,dil', CONV_CASES)
def test_conv_fprop_dgrad_wgrad(E, n, h, w, c, k, ks, dil):
    g = torch.Generator().manual_seed(n * 1000 + h * 10 + c + k + dil)
    x = bf16_round(torch.randn(n, c, h, w, generator=g))
    wt = bf16_round(torch.randn(k, c, ks, ks, generator=g) * (2.0 / (c * ks * ks)) ** 0.5)
    dy = bf16_round(torch.randn(n, k, h, w, generator=g))
    pad = (ks // 2) * dil
    x.requires_grad_(True)
    wt.requires_grad_(True)
    y_ref = F.conv2d(x, wt, None, 1, pad, dil)
    y_ref.backward(dy)

    pw = E.PackedWeight().update(wt.detach().cuda(), need_dgrad=True)
    xa = nhwc(x.detach())
    ya = E.Act.empty(n, h, w, k, 'cuda')
    stats = torch.zeros(2 * k, device='cuda')
    E.conv_fprop([(xa, pw, dil)], ya, stats=stats)
    assert rel_err(from_act(ya), y_ref) < TOL
    # fused BN statistics: sum and sum of squares of the conv output as stored (bf16-rounded)
    y_st = from_act(ya)
    assert rel_err(stats[:k], y_st.sum((0, 2, 3))) < 1e-4 + 1e-3 * (n * h * w > 64)
    assert rel_err(stats[k:], (y_st ** 2).sum((0, 2, 3))) < 1e-4

    dya = nhwc(dy)
    dxa = E.Act.empty(n, h, w, c, 'cuda')
    E.conv_dgrad([(dya, pw, dil)], dxa)
    assert rel_err(from_act(dxa), x.grad) < TOL

    dw = E.conv_wgrad(dya, xa, k, ks, ks, dil)
    assert rel_err(dw, wt.grad) < 5e-3


STRIDE2_CASES = [
    # n, h_out, w_out, c, k, ksize
    (2, 16, 16, 128, 128, 3),
    (1, 8, 24, 64, 96, 3),
    (2, 16, 16, 256, 512, 1),
    (3, 4, 4, 72, 40, 3),
]


@pytest.mark.parametrize('n,h,w,c,k,ks,dil', [(8, 128, 128, 256, 256, 3, 1), (16, 32, 32, 2048, 256, 3, 12), (4, 64, 64, 128, 128, 3, 1),
                                              (16, 128, 128, 64, 64, 3, 1), (16, 32, 32, 1024, 256, 1, 1)])
def test_conv_wgrad_is_bit_identical_run_to_run(E, n, h, w, c, k, ks, dil):
    """Split-K partials meet in a fixed order — inside a cluster over DSMEM, then in wgrad_reduce_kernel — never through
    floating-point atomics: the same operands give the same bits, also when the gradient accumulates."""
    g = torch.Generator().manual_seed(c + k + dil)
    xa = E.Act(torch.randn(n, h, w, c, generator=g).to(torch.bfloat16).cuda())
    dya = E.Act(torch.randn(n, h, w, k, generator=g).to(torch.bfloat16).cuda())
    first = E.conv_wgrad(dya, xa, k, ks, ks, dil).clone()
    for _ in range(3):
        assert torch.equal(E.conv_wgrad(dya, xa, k, ks, ks, dil), first)
    base = torch.randn(k, c, ks, ks, generator=g).cuda()
    acc = [E.conv_wgrad(dya, xa, k, ks, ks, dil, dw=base.clone()).clone() for _ in range(2)]
    assert torch.equal(acc[0], acc[1])
    assert float((acc[0] - (base + first)).abs().max()) <= 1e-5 * float(first.abs().max()) + 1e-6


@pytest.mark.parametrize('n,h,w,c,k,ks', STRIDE2_CASES)
def test_conv_stride2_fprop_dgrad_wgrad(E, n, h, w, c, k, ks):
    """Stride-2 convs of the ResNet backbone (3x3 conv2 and 1x1 downsample of layer2.0 / layer3.0): the input is
    read through its four parity sub-grid views, dgrad writes the four parity sub-grids of dX."""
    g = torch.Generator().manual_seed(n * 1000 + h * 10 + c + k)
    x = bf16_round(torch.randn(n, c, 2 * h, 2 * w, generator=g))
    wt = bf16_round(torch.randn(k, c, ks, ks, generator=g) * (2.0 / (c * ks * ks)) ** 0.5)
    dy = bf16_round(torch.randn(n, k, h, w, generator=g))
    x.requires_grad_(True)
    wt.requires_grad_(True)
    y_ref = F.conv2d(x, wt, None, 2, ks // 2, 1)
    assert y_ref.shape == dy.shape
    y_ref.backward(dy)

    pw = E.PackedWeight().update(wt.detach().cuda(), need_dgrad=True)
    xa = nhwc(x.detach())
    ya = E.Act.empty(n, h, w, k, 'cuda')
    stats = torch.zeros(2 * k, device='cuda')
    E.conv_fprop([(xa, pw, 1, 2)], ya, stats=stats)
    assert rel_err(from_act(ya), y_ref) < TOL
    y_st = from_act(ya)
    assert rel_err(stats[:k], y_st.sum((0, 2, 3))) < 2e-3
    assert rel_err(stats[k:], (y_st ** 2).sum((0, 2, 3))) < 1e-4

    dya = nhwc(dy)
    dw = E.conv_wgrad(dya, xa, k, ks, ks, 1, stride=2)
    assert rel_err(dw, wt.grad) < 5e-3

    if ks == 3:
        dxa = E.Act(torch.full((n, 2 * h, 2 * w, c), 7.0, dtype=torch.bfloat16, device='cuda'))
        E.conv_dgrad([(dya, pw, 1, 2)], dxa)
        assert rel_err(from_act(dxa), x.grad) < TOL
    else:
        # a stride-2 1x1 alone leaves three of four dX pixels without gradient: refused (no silent garbage)
        dxa = E.Act.empty(n, 2 * h, 2 * w, c, 'cuda')
        with pytest.raises(RuntimeError):
            E.conv_dgrad([(dya, pw, 1, 2)], dxa)


def test_conv_dgrad_mixed_stride_sources(E):
    """Block-input gradient of a ResNet downsampling bottleneck: dX = dgrad(conv1: 1x1 s1) + dgrad(downsample: 1x1 s2)
    in the same launches."""
    g = torch.Generator().manual_seed(11)
    n, h, w, c, k1, k2 = 2, 16, 16, 128, 64, 256
    x = bf16_round(torch.randn(n, c, h, w, generator=g)).requires_grad_(True)
    w1 = bf16_round(torch.randn(k1, c, 1, 1, generator=g) * 0.1).requires_grad_(True)
    w2 = bf16_round(torch.randn(k2, c, 1, 1, generator=g) * 0.1).requires_grad_(True)
    dy1 = bf16_round(torch.randn(n, k1, h, w, generator=g))
    dy2 = bf16_round(torch.randn(n, k2, h // 2, w // 2, generator=g))
    (F.conv2d(x, w1) * dy1).sum().backward()
    (F.conv2d(x, w2, None, 2) * dy2).sum().backward()
    p1 = E.PackedWeight().update(w1.detach().cuda(), True)
    p2 = E.PackedWeight().update(w2.detach().cuda(), True)
    dxa = E.Act.empty(n, h, w, c, 'cuda')
    E.conv_dgrad([(nhwc(dy1), p1, 1, 1), (nhwc(dy2), p2, 1, 2)], dxa)
    assert rel_err(from_act(dxa), x.grad) < TOL


def test_conv_dgrad_mixed_3x3_s1_and_s2(E):
    g = torch.Generator().manual_seed(12)
    n, h, w, c = 1, 12, 20, 64
    x = bf16_round(torch.randn(n, c, h, w, generator=g)).requires_grad_(True)
    w1 = bf16_round(torch.randn(32, c, 3, 3, generator=g) * 0.1).requires_grad_(True)
    w2 = bf16_round(torch.randn(48, c, 3, 3, generator=g) * 0.1).requires_grad_(True)
    dy1 = bf16_round(torch.randn(n, 32, h, w, generator=g))
    dy2 = bf16_round(torch.randn(n, 48, h // 2, w // 2, generator=g))
    (F.conv2d(x, w1, None, 1, 2, 2) * dy1).sum().backward()
    (F.conv2d(x, w2, None, 2, 1) * dy2).sum().backward()
    p1 = E.PackedWeight().update(w1.detach().cuda(), True)
    p2 = E.PackedWeight().update(w2.detach().cuda(), True)
    dxa = E.Act.empty(n, h, w, c, 'cuda')
    E.conv_dgrad([(nhwc(dy1), p1, 2, 1), (nhwc(dy2), p2, 1, 2)], dxa)
    assert rel_err(from_act(dxa), x.grad) < TOL


@pytest.mark.parametrize('n,c,h,w,k,r,stride,pad', [(2, 3, 64, 96, 64, 7, 2, 3), (1, 3, 50, 38, 64, 3, 2, 1),
                                                     (2, 3, 33, 200, 32, 7, 2, 3), (1, 4, 16, 16, 8, 3, 1, 1)])
def test_patch_gather_stem_conv(E, n, c, h, w, k, r, stride, pad):
    # few-channel stem: im2col gather (bit-exact vs unfold of the bf16-rounded image) + 1x1 GEMM fprop + wgrad
    g = torch.Generator().manual_seed(n * 100 + h)
    x = torch.randn(n, c, h, w, generator=g)
    wt = torch.randn(k, c, r, r, generator=g) * 0.1
    col = E.im2col(x.cuda(), r, r, stride, pad)
    oh, ow = (h + 2 * pad - r) // stride + 1, (w + 2 * pad - r) // stride + 1
    assert (col.n, col.h, col.w, col.c) == (n, oh, ow, (r * r * c + 7) // 8 * 8)
    unf = F.unfold(bf16_round(x), r, padding=pad, stride=stride)            # [n, c*r*r, oh*ow], order (c, r, s)
    unf = unf.view(n, c, r * r, oh, ow).permute(0, 3, 4, 2, 1).reshape(n, oh, ow, r * r * c)   # -> (tap, c)
    got = col.t.float().cpu()
    assert torch.equal(got[..., :r * r * c], unf)
    assert float(got[..., r * r * c:].abs().max() if col.c > r * r * c else 0.0) == 0.0
    pw = E.PackedIm2colWeight().update(wt.cuda())
    y = E.Act.empty(n, oh, ow, k, 'cuda')
    E.conv_fprop([(col, pw, 1)], y)
    xr = bf16_round(x).requires_grad_(True)
    wr = bf16_round(wt).requires_grad_(True)
    ref = F.conv2d(xr, wr, None, stride, pad)
    assert rel_err(from_act(y), ref) < TOL
    gy = bf16_round(torch.randn(n, k, oh, ow, generator=g))
    ref.backward(gy)
    dwcol = E.conv_wgrad(nhwc(gy), col, k, 1, 1, 1)
    dw = E.unpack_wgrad_im2col(dwcol, pw)
    assert dw.shape == wt.shape
    assert rel_err(dw, wr.grad) < TOL


def test_pack_all_matches_single_tensor_packing(E):
    g = torch.Generator().manual_seed(5)
    shapes = [(64, 64, 1, 1), (256, 64, 3, 3), (21, 384, 3, 3), (128, 2048, 3, 3), (8, 8, 1, 1)]
    ws = [torch.randn(*sh, generator=g).cuda() for sh in shapes]
    stem = torch.randn(64, 3, 7, 7, generator=g).cuda()
    single = [E.PackedWeight().update(w, need_dgrad=True) for w in ws]
    single_stem = E.PackedIm2colWeight().update(stem)
    multi = [E.PackedWeight() for _ in ws]
    multi_stem = E.PackedIm2colWeight()
    n = E.pack_all(list(zip(multi, ws)) + [(multi_stem, stem)], need_dgrad=True)
    assert n == len(ws) + 1
    for a, b in zip(single, multi):
        assert torch.equal(a.wf, b.wf) and torch.equal(a.wd, b.wd)
        assert (a.k, a.c, a.r, a.s, a.k_rows, a.c_rows, a.k_cols) == (b.k, b.c, b.r, b.s, b.k_rows, b.c_rows, b.k_cols)
    assert torch.equal(single_stem.wf, multi_stem.wf)
    # nothing stale: no launch; a version bump makes exactly that tensor stale again
    assert E.pack_all(list(zip(multi, ws)) + [(multi_stem, stem)], need_dgrad=True) == 0
    ws[1].mul_(2.0)
    assert E.pack_all(list(zip(multi, ws)), need_dgrad=True) == 1
    assert torch.equal(multi[1].wf, E.PackedWeight().update(ws[1]).wf)


@pytest.mark.parametrize('momentum,nesterov,wd', [(0.9, False, 0.0), (0.9, True, 5e-4), (0.0, False, 1e-3)])
def test_sgd_multi_matches_torch_sgd(E, momentum, nesterov, wd):
    from pytorch_segmentation_b200 import optim
    g = torch.Generator().manual_seed(11)
    shapes = [(64, 3, 7, 7), (21,), (256, 64, 3, 3), (1,), (2048, 513), (7, 5, 3)] + [(16,)] * 120
    ref = [torch.nn.Parameter(torch.randn(*sh, generator=g).cuda()) for sh in shapes]
    # a channels_last conv weight (torchvision backbones under .to(memory_format=channels_last))
    ref.append(torch.nn.Parameter(torch.randn(32, 16, 3, 3, generator=g).cuda().contiguous(memory_format=torch.channels_last)))
    shapes.append((32, 16, 3, 3))
    ours = [torch.nn.Parameter(p.detach().clone()) for p in ref]
    o_ref = torch.optim.SGD(ref, lr=0.05, momentum=momentum, nesterov=nesterov, weight_decay=wd)
    o_ours = optim.SGD(ours, lr=0.05, momentum=momentum, nesterov=nesterov, weight_decay=wd)
    for step in range(3):
        for a, b in zip(ref, ours):
            if step == 1 and a.numel() == 1:
                a.grad = b.grad = None                   # a parameter without a gradient is skipped
                continue
            gr = torch.randn(a.shape, generator=g).cuda()
            a.grad, b.grad = gr.clone(), gr.clone()
        v0 = ours[2]._version
        o_ref.step()
        o_ours.step()
        assert ours[2]._version > v0
        for a, b in zip(ref, ours):
            assert torch.allclose(a, b, rtol=1e-6, atol=1e-6)
    if momentum:
        assert torch.allclose(o_ref.state[ref[4]]['momentum_buffer'], o_ours.state[ours[4]]['momentum_buffer'],
                              rtol=1e-6, atol=1e-6)


def test_grad_add_masked(E):
    g = torch.Generator().manual_seed(2)
    n, h, w, c = 2, 9, 7, 24
    dx, dy, y, xin = (bf16_round(torch.randn(n, c, h, w, generator=g)) for _ in range(4))
    a = nhwc(dx)
    E.grad_add_masked(a, dy=nhwc(dy), y=nhwc(y), xin=nhwc(xin))
    assert rel_err(from_act(a), bf16_round((dx + dy * (y > 0)) * (xin > 0))) < 1e-6
    a = nhwc(dx)
    E.grad_add_masked(a, dy=nhwc(dy), xin=nhwc(xin))
    assert rel_err(from_act(a), bf16_round((dx + dy) * (xin > 0))) < 1e-6
    a = nhwc(dx)
    E.grad_add_masked(a, xin=nhwc(xin))
    assert torch.equal(from_act(a), dx * (xin > 0))
    a = nhwc(dx)
    E.grad_add_masked(a, dy=nhwc(dy), y=nhwc(y))
    assert rel_err(from_act(a), bf16_round(dx + dy * (y > 0))) < 1e-6


@pytest.mark.parametrize('n,h,w,c,k', [(2, 24, 40, 384, 21), (1, 17, 23, 88, 2), (2, 16, 16, 128, 19)])
def test_narrow_output_conv_taps_form(E, n, h, w, c, k):
    # cls_conv through the taps-as-channels form: Z GEMM + scatter-add forward; tap gather + 1x1 dgrad / wgrad backward
    from pytorch_segmentation_b200.models.deeplabv3plus import cls_conv_backward, cls_conv_forward, make_cls_packed
    g = torch.Generator().manual_seed(c + k)
    conv = torch.nn.Conv2d(c, k, 3, padding=1).cuda()
    with torch.no_grad():
        conv.weight.copy_(bf16_round(torch.randn(k, c, 3, 3, generator=g) * 0.05))
        conv.bias.copy_(torch.randn(k, generator=g))
    packed = make_cls_packed(conv)
    assert isinstance(packed, E.PackedTapWeight)
    x = bf16_round(torch.randn(n, c, h, w, generator=g))
    xa = nhwc(x)
    y = cls_conv_forward(conv, packed, xa, True)                      # fp32 NHWC, padded to 8 channels
    ncp = (k + 7) // 8 * 8
    assert tuple(y.shape) == (n, h, w, ncp)
    xr = x.clone().requires_grad_(True)
    wr = conv.weight.detach().cpu().clone().requires_grad_(True)
    ref = F.conv2d(xr, wr, conv.bias.detach().cpu(), padding=1)
    assert rel_err(y[..., :k].permute(0, 3, 1, 2), ref) < TOL
    assert float(y[..., k:].abs().max()) == 0.0 if ncp > k else True
    gy = bf16_round(torch.randn(n, k, h, w, generator=g))
    ref.backward(gy)
    gyp = torch.zeros(n, ncp, h, w)
    gyp[:, :k] = gy
    dx = E.Act.empty(n, h, w, c, 'cuda')
    dw, db = cls_conv_backward(conv, packed, xa, nhwc(gyp), dx)
    assert rel_err(from_act(dx), xr.grad) < TOL
    assert dw.shape == conv.weight.shape and rel_err(dw, wr.grad) < TOL
    assert rel_err(db, gy.sum((0, 2, 3))) < TOL


def test_conv_epilogue_scale_bias_relu_into_slice(E):
    g = torch.Generator().manual_seed(5)
    n, h, w, c, k = 2, 16, 16, 64, 128
    x = bf16_round(torch.randn(n, c, h, w, generator=g))
    wt = bf16_round(torch.randn(k, c, 1, 1, generator=g) * 0.2)
    scale, bias = torch.rand(k, generator=g) + 0.5, torch.randn(k, generator=g)
    ref = F.relu(F.conv2d(x, wt) * scale.view(1, -1, 1, 1) + bias.view(1, -1, 1, 1))
    # input is a slice [64:128) of a 192-channel buffer, output a slice [256:384) of a 384-channel buffer
    xb = torch.zeros(n, h, w, 192, dtype=torch.bfloat16, device='cuda')
    xb[..., 64:128] = x.permute(0, 2, 3, 1).cuda()
    yb = torch.zeros(n, h, w, 384, dtype=torch.bfloat16, device='cuda')
    pw = E.PackedWeight().update(wt.cuda(), need_dgrad=False)
    E.conv_fprop([(E.Act(xb, 64, 64), pw, 1)], E.Act(yb, 256, 128), scale=scale.cuda(), bias=bias.cuda(), relu=True)
    assert rel_err(yb[..., 256:384].permute(0, 3, 1, 2), ref) < TOL
    assert float(yb[..., :256].abs().max()) == 0.0      # neighbours untouched


def test_conv_multi_source_sum(E):
    """Two sources accumulate in TMEM: conv over a concat == sum of convs over the parts."""
    g = torch.Generator().manual_seed(6)
    n, h, w = 2, 12, 12
    x1 = bf16_round(torch.randn(n, 64, h, w, generator=g))
    x2 = bf16_round(torch.randn(n, 24, h, w, generator=g))
    w1 = bf16_round(torch.randn(32, 64, 3, 3, generator=g) * 0.1)
    w2 = bf16_round(torch.randn(32, 24, 1, 1, generator=g) * 0.1)
    ref = F.conv2d(x1, w1, None, 1, 2, 2) + F.conv2d(x2, w2)
    p1 = E.PackedWeight().update(w1.cuda(), False)
    p2 = E.PackedWeight().update(w2.cuda(), False)
    ya = E.Act.empty(n, h, w, 32, 'cuda')
    E.conv_fprop([(nhwc(x1), p1, 2), (nhwc(x2), p2, 1)], ya)
    assert rel_err(from_act(ya), ref) < TOL


def test_cls_conv_fp32_narrow(E):
    """21 classes padded to 24: fp32 NHWC class map with bias; wgrad/dgrad through the padded gradient."""
    g = torch.Generator().manual_seed(7)
    n, h, w, c, nc = 2, 24, 24, 384, 21
    x = bf16_round(torch.randn(n, c, h, w, generator=g)).requires_grad_(True)
    wt = bf16_round(torch.randn(nc, c, 3, 3, generator=g) * 0.05).requires_grad_(True)
    b = torch.randn(nc, generator=g)
    dy = bf16_round(torch.randn(n, nc, h, w, generator=g))
    ref = F.conv2d(x, wt, b, 1, 1)
    ref.backward(dy)
    pw = E.PackedWeight().update(wt.detach().cuda(), True)
    bias = torch.zeros(24, device='cuda')
    bias[:nc] = b.cuda()
    out = torch.empty(n, h, w, 24, device='cuda')
    xa = nhwc(x.detach())
    E.conv_fprop_f32([(xa, pw, 1)], out, 24, bias=bias)
    assert rel_err(out[..., :nc].permute(0, 3, 1, 2), ref) < 5e-3
    dyp = torch.zeros(n, h, w, 24, dtype=torch.bfloat16, device='cuda')
    dyp[..., :nc] = dy.permute(0, 2, 3, 1).cuda()
    dya = E.Act(dyp)
    dxa = E.Act.empty(n, h, w, c, 'cuda')
    E.conv_dgrad([(dya, pw, 1)], dxa)
    assert rel_err(from_act(dxa), x.grad) < TOL
    dw = E.conv_wgrad(dya, xa, nc, 3, 3, 1)
    assert rel_err(dw, wt.grad) < 5e-3
    assert rel_err(E.channel_sum(dya)[:nc], dy.sum((0, 2, 3))) < 1e-3


@pytest.mark.parametrize('n,h,w,c,ct,off', [(2, 16, 16, 64, 64, 0), (3, 7, 9, 24, 88, 64), (16, 1, 1, 256, 256, 0),
                                            (2, 32, 32, 2048, 2048, 0)])
def test_batchnorm_train_forward_backward(E, n, h, w, c, ct, off):
    g = torch.Generator().manual_seed(c + h)
    x = bf16_round(torch.randn(n, c, h, w, generator=g) * 1.5 + 0.3).requires_grad_(True)
    gamma = (1 + 0.2 * torch.randn(c, generator=g)).requires_grad_(True)
    beta = (0.1 * torch.randn(c, generator=g)).requires_grad_(True)
    rm, rv = 0.1 * torch.randn(c, generator=g), 1 + 0.1 * torch.rand(c, generator=g)
    dz = bf16_round(torch.randn(n, c, h, w, generator=g))
    rm_ref, rv_ref = rm.clone(), rv.clone()
    z_ref = F.relu(F.batch_norm(x, rm_ref, rv_ref, gamma, beta, True, 0.1, 1e-5))
    z_ref.backward(dz)

    xb = torch.zeros(n, h, w, ct, dtype=torch.bfloat16, device='cuda')
    xb[..., off:off + c] = x.detach().permute(0, 2, 3, 1).cuda()
    xa = E.Act(xb, off, c)
    stats = E.bn_stats(xa)
    rm_d, rv_d = rm.cuda(), rv.cuda()
    vec = E.bn_finalize(stats, c, n * h * w, gamma.detach().cuda(), beta.detach().cuda(), 1e-5, 0.1, rm_d, rv_d)
    za = E.Act.empty(n, h, w, c, 'cuda')
    E.bn_apply(xa, za, vec, True)
    assert rel_err(from_act(za), z_ref) < TOL
    assert rel_err(rm_d, rm_ref) < 1e-3 and rel_err(rv_d, rv_ref) < 1e-3
    dza = nhwc(dz)
    dxa = E.Act.empty(n, h, w, c, 'cuda')
    dgamma, dbeta = E.bn_backward(dza, xa, dxa, vec, True, True)
    assert rel_err(from_act(dxa), x.grad) < TOL
    assert rel_err(dgamma, gamma.grad) < 5e-3 and rel_err(dbeta, beta.grad) < 5e-3


def test_batchnorm_eval_affine_and_backward(E):
    g = torch.Generator().manual_seed(3)
    n, h, w, c = 2, 8, 8, 32
    x = bf16_round(torch.randn(n, c, h, w, generator=g)).requires_grad_(True)
    gamma = (1 + 0.2 * torch.randn(c, generator=g)).requires_grad_(True)
    beta = (0.1 * torch.randn(c, generator=g)).requires_grad_(True)
    rm, rv = 0.1 * torch.randn(c, generator=g), 1 + 0.1 * torch.rand(c, generator=g)
    dz = bf16_round(torch.randn(n, c, h, w, generator=g))
    z_ref = F.relu(F.batch_norm(x, rm, rv, gamma, beta, False, 0.1, 1e-5))
    z_ref.backward(dz)
    vec = E.bn_eval_affine(c, gamma.detach().cuda(), beta.detach().cuda(), rm.cuda(), rv.cuda(), 1e-5)
    xa, za = nhwc(x.detach()), E.Act.empty(n, h, w, c, 'cuda')
    E.bn_apply(xa, za, vec, True)
    assert rel_err(from_act(za), z_ref) < TOL
    dxa = E.Act.empty(n, h, w, c, 'cuda')
    dgamma, dbeta = E.bn_backward(nhwc(dz), xa, dxa, vec, True, False)
    assert rel_err(from_act(dxa), x.grad) < TOL
    assert rel_err(dgamma, gamma.grad) < 5e-3 and rel_err(dbeta, beta.grad) < 5e-3


@pytest.mark.parametrize('h,w,oh,ow,ac', [(8, 8, 32, 32, True), (10, 10, 20, 20, True), (5, 7, 20, 28, True),
                                           (8, 8, 8, 8, True), (6, 9, 13, 31, False), (1, 1, 16, 16, False),
                                           (16, 16, 7, 5, True)])
def test_bilinear_upsample_and_adjoint(E, h, w, oh, ow, ac):
    g = torch.Generator().manual_seed(h * w + oh)
    n, c = 2, 24
    x = bf16_round(torch.randn(n, c, h, w, generator=g)).requires_grad_(True)
    dy = bf16_round(torch.randn(n, c, oh, ow, generator=g))
    ref = F.interpolate(x, size=(oh, ow), mode='bilinear', align_corners=ac)
    ref.backward(dy)
    # into a slice of a wider buffer
    yb = torch.zeros(n, oh, ow, 40, dtype=torch.bfloat16, device='cuda')
    E.upsample_fwd(nhwc(x.detach()), E.Act(yb, 8, c), align_corners=ac)
    assert rel_err(yb[..., 8:8 + c].permute(0, 3, 1, 2), ref) < 1e-2
    assert float(yb[..., :8].abs().max()) == 0.0 and float(yb[..., 32:].abs().max()) == 0.0
    dxa = E.Act.empty(n, h, w, c, 'cuda')
    E.upsample_bwd(nhwc(dy), dxa, align_corners=ac)
    assert rel_err(from_act(dxa), x.grad) < 1e-2


def test_gap_broadcast_sum(E):
    g = torch.Generator().manual_seed(9)
    n, c, h, w = 3, 2048, 16, 16
    x = bf16_round(torch.randn(n, c, h, w, generator=g))
    xa = nhwc(x)
    ga = E.gap_fwd(xa)
    assert ga.dtype == torch.float32 and rel_err(ga, x.mean((2, 3))) < 1e-5
    assert rel_err(E.sum_hw(xa), x.sum((2, 3))) < 1e-5
    yb = torch.zeros(n, h, w, c + 8, dtype=torch.bfloat16, device='cuda')
    E.broadcast_hw(ga, E.Act(yb, 8, c))
    assert torch.equal(yb[..., 8:], ga.to(torch.bfloat16).reshape(n, 1, 1, c).expand(n, h, w, c))
    assert float(yb[..., :8].abs().max()) == 0.0
    base = bf16_round(torch.randn(n, c, h, w, generator=g))
    ba = nhwc(base)
    E.broadcast_hw(ga, ba, scale=0.5, accumulate=True)
    ref = base + 0.5 * ga.reshape(n, c, 1, 1).cpu()
    assert rel_err(from_act(ba), ref) < 1e-2


@pytest.mark.parametrize('n,c,k,training', [(16, 2048, 256, True), (3, 128, 64, True), (2, 2048, 256, True),
                                            (20, 96, 40, False)])
def test_pooled_branch_fp32_kernels(E, n, c, k, training):
    """1x1 ConvNormAct on the N x C x 1 x 1 pooled map (models/aspp.py:12) in fp32: forward, BN over the N rows,
    and every gradient against F.conv2d / F.batch_norm."""
    g = torch.Generator().manual_seed(n + c)
    x = (torch.randn(n, c, generator=g) * 0.3 + 2.0).requires_grad_(True)     # mean >> spread, like pooled features
    wt = (torch.randn(k, c, 1, 1, generator=g) * 0.05).requires_grad_(True)
    bn = torch.nn.BatchNorm2d(k)
    with torch.no_grad():
        bn.weight.copy_(1 + 0.2 * torch.randn(k, generator=g))
        bn.bias.copy_(0.1 * torch.randn(k, generator=g))
        bn.running_mean.copy_(0.1 * torch.randn(k, generator=g))
        bn.running_var.copy_(1 + 0.1 * torch.rand(k, generator=g))
    bn.train(training)
    import copy
    bn_dev = copy.deepcopy(bn).cuda()
    dz = torch.randn(n, k, generator=g)
    raw_ref = F.conv2d(x.view(n, c, 1, 1), wt)
    z_ref = F.relu(bn(raw_ref))
    z_ref.backward(dz.view(n, k, 1, 1))
    xd, wd = x.detach().cuda(), wt.detach().cuda()
    raw = E.linear_f32_fwd(xd, wd)
    assert rel_err(raw, raw_ref.view(n, k)) < 1e-5
    z, save = E.bn_rows_fwd(raw, bn_dev, training, True)
    assert rel_err(z, z_ref.view(n, k)) < 1e-4
    if training:
        assert rel_err(bn_dev.running_mean, bn.running_mean) < 1e-5
        assert rel_err(bn_dev.running_var, bn.running_var) < 1e-4
    draw, dgamma, dbeta = E.bn_rows_bwd(dz.cuda(), raw, bn_dev, save, training, True)
    dx, dw = E.linear_f32_bwd(draw, xd, wd)
    tol = 2e-3 if n > 2 else 5e-2          # n == 2: the BN input gradient is ~0 by cancellation
    assert rel_err(dgamma, bn.weight.grad) < 1e-4 and rel_err(dbeta, bn.bias.grad) < 1e-4
    assert rel_err(dw, wt.grad) < tol
    assert rel_err(dx, x.grad) < tol


@pytest.mark.parametrize('nc,h,w,oh,ow', [(21, 8, 8, 32, 32), (2, 10, 12, 20, 24), (19, 6, 5, 24, 20)])
def test_logits_upsample_and_adjoint(E, nc, h, w, oh, ow):
    g = torch.Generator().manual_seed(nc)
    n = 2
    ncp = (nc + 7) // 8 * 8
    x = torch.randn(n, nc, h, w, generator=g).requires_grad_(True)
    dy = torch.randn(n, nc, oh, ow, generator=g)
    ref = F.interpolate(x, size=(oh, ow), mode='bilinear', align_corners=True)
    ref.backward(dy)
    xp = torch.zeros(n, h, w, ncp)
    xp[..., :nc] = x.detach().permute(0, 2, 3, 1)
    y = E.logits_upsample_fwd(xp.cuda(), nc, oh, ow, True)
    assert rel_err(y, ref) < 1e-5
    dxa = E.logits_upsample_bwd(dy.cuda(), h, w, ncp, True)
    got = dxa.t.float().cpu()
    assert rel_err(got[..., :nc].permute(0, 3, 1, 2), x.grad) < 1e-2
    assert float(got[..., nc:].abs().max()) == 0.0 if ncp > nc else True


def test_resize_nchw_and_adjoint(E):
    g = torch.Generator().manual_seed(4)
    x = torch.randn(2, 5, 8, 16, generator=g).requires_grad_(True)
    dy = torch.randn(2, 5, 29, 37, generator=g)
    ref = F.interpolate(x, size=(29, 37), mode='bilinear', align_corners=True)
    ref.backward(dy)
    assert rel_err(E.resize_nchw_fwd(x.detach().cuda(), 29, 37, True), ref) < 1e-5
    assert rel_err(E.resize_nchw_bwd(dy.cuda(), 8, 16, True), x.grad) < 1e-5


@pytest.mark.parametrize('nc', [2, 21, 19, 40])
def test_softmax_cross_entropy(E, nc):
    g = torch.Generator().manual_seed(nc)
    n, h, w = 3, 17, 23
    logits = (torch.randn(n, nc, h, w, generator=g) * 3).requires_grad_(True)
    tgt = torch.randint(0, nc, (n, h, w), generator=g)
    tgt[0, 0, :5] = -100                                  # ignore_index pixels
    loss_ref = F.cross_entropy(logits, tgt)
    loss_ref.backward()
    acc, dl = E.softmax_ce(logits.detach().cuda(), tgt.cuda(), -100, True)
    count = int(acc[1:2].view(torch.int64))
    assert count == int((tgt != -100).sum())
    assert abs(float(acc[0]) / count - float(loss_ref.detach())) < 1e-5 * max(1.0, abs(float(loss_ref)))
    assert int(acc[2:3].view(torch.int32)[0]) == 0
    assert rel_err(dl / count, logits.grad) < 1e-5
    # out-of-range target sets the status flag
    bad = tgt.clone()
    bad[1, 2, 3] = nc
    acc, _ = E.softmax_ce(logits.detach().cuda(), bad.cuda(), -100, False)
    assert int(acc[2:3].view(torch.int32)[0]) == 1


@pytest.mark.parametrize('nc', [2, 21, 40])
def test_softmax_cross_entropy_backward_recompute(E, nc):
    # backward entry: dlogits = (softmax - onehot) * g / count with g and count read on the device
    g = torch.Generator().manual_seed(100 + nc)
    n, h, w = 2, 19, 31
    logits = (torch.randn(n, nc, h, w, generator=g) * 3).requires_grad_(True)
    tgt = torch.randint(0, nc, (n, h, w), generator=g)
    tgt[1, 3, 2:9] = -100
    (F.cross_entropy(logits, tgt) * 0.37).backward()
    count = torch.tensor([int((tgt != -100).sum())], dtype=torch.int64, device='cuda')
    gout = torch.tensor(0.37, dtype=torch.float32, device='cuda')
    dl = E.softmax_ce_bwd(logits.detach().cuda(), tgt.cuda(), gout, count, -100)
    assert rel_err(dl, logits.grad) < 1e-5
    assert float(dl[1, :, 3, 2:9].abs().max()) == 0.0


@pytest.mark.parametrize('nc', [2, 5, 21, 19, 64])
def test_argmax_and_confusion_bit_exact(E, nc):
    g = torch.Generator().manual_seed(100 + nc)
    n, h, w = 2, 37, 41
    logits = torch.randn(n, nc, h, w, generator=g)
    logits[0, :, 0, 0] = 1.0                               # full tie -> lowest index
    logits[1, nc - 1, 3, 3] = logits[1, 0, 3, 3] = 9.0     # two-way tie -> index 0
    tgt = torch.randint(0, nc, (n, h, w), generator=g)
    pred_ref = logits.max(1)[1]
    pred = E.argmax(logits.cuda())
    assert torch.equal(pred.cpu(), pred_ref)
    cm_ref = torch.bincount(tgt.reshape(-1) * nc + pred_ref.reshape(-1), minlength=nc * nc).reshape(nc, nc)
    cm = E.confusion(pred, tgt.cuda(), nc)
    assert torch.equal(cm.cpu(), cm_ref)
    cm2 = E.argmax_confusion(logits.cuda(), tgt.cuda())
    assert torch.equal(cm2.cpu(), cm_ref)
    # accumulation across calls and labels outside [0, nc) skipped
    tgt2 = tgt.clone()
    tgt2[0, 0, :7] = 255
    E.confusion(pred, tgt2.cuda(), nc, cm)
    keep = tgt2.reshape(-1) < nc
    cm_ref2 = cm_ref + torch.bincount((tgt2.reshape(-1) * nc + pred_ref.reshape(-1))[keep],
                                      minlength=nc * nc).reshape(nc, nc)
    assert torch.equal(cm.cpu(), cm_ref2)


def test_confusion_skewed_and_large(E):
    """Uniform regions (every lane of a warp hits one bin) and > 2**24 pixels: int64 counts stay exact
    where the reference's float32 accumulators (test.py:22-24) would not."""
    nc = 21
    n = 1 << 25
    tgt = torch.zeros(n, dtype=torch.int64, device='cuda')
    pred = torch.zeros(n, dtype=torch.int64, device='cuda')
    pred[::3] = 7
    cm = E.confusion(pred, tgt, nc)
    assert int(cm[0, 7]) == (n + 2) // 3 and int(cm[0, 0]) == n - (n + 2) // 3 and int(cm.sum()) == n


def test_layout_kernels_roundtrip(E, lib):
    g = torch.Generator().manual_seed(1)
    x = torch.randn(2, 24, 9, 11, generator=g)
    a = E.as_act(x.cuda())
    assert torch.equal(a.t.cpu(), x.to(torch.bfloat16).permute(0, 2, 3, 1))
    back = E.grad_like(x.cuda(), a)
    assert torch.equal(back.cpu(), bf16_round(x))
    # bf16 channels_last input is a zero-copy view
    xc = x.cuda().to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    assert E.as_act(xc).ptr == xc.data_ptr()


def test_errors_are_loud(E):
    with pytest.raises(RuntimeError, match='no CPU fallback'):
        E.as_act(torch.zeros(1, 8, 4, 4))
    with pytest.raises(ValueError, match='multiple of 8'):
        E.as_act(torch.zeros(1, 3, 4, 4, device='cuda'))
    from pytorch_segmentation_b200 import ConvNormAct
    with pytest.raises(NotImplementedError):
        ConvNormAct(8, 8, 3, 3)                   # stride 3 (stride 2 is the HRNet transition conv and is supported)


# ---- loss through the low-resolution class map (csrc/loss_metrics.cu lowres_ce_kernel) ----------------------------------
LOWRES_CASES = [
    # n, h, w, scale, nc, ct, align_corners
    (2, 32, 40, 4, 21, 24, True),       # DeepLabV3+ head (models/deeplabv3plus.py:40-43); 160 wide = 2.5 tiles
    (1, 23, 19, 4, 19, 24, False),      # HRNet head (models/hrnet.py:400-404), ragged
    (3, 40, 48, 2, 2, 8, True),         # UNet head (models/unet.py:52-55)
    (1, 9, 9, 4, 32, 32, True),         # widest class count
]


@pytest.mark.parametrize('n,h,w,scale,nc,ct,ac', LOWRES_CASES)
def test_lowres_ce_matches_interpolate_cross_entropy(E, n, h, w, scale, nc, ct, ac):
    g = torch.Generator().manual_seed(n * 100 + h)
    H, W = h * scale, w * scale
    lr = torch.zeros(n, h, w, ct)
    lr[..., :nc] = torch.randn(n, h, w, nc, generator=g) * 3
    t = torch.randint(0, nc, (n, H, W), generator=g)
    t[torch.rand(n, H, W, generator=g) < 0.1] = -100                    # ignored pixels
    ref_in = lr[..., :nc].permute(0, 3, 1, 2).contiguous().double().requires_grad_(True)
    ref = F.cross_entropy(F.interpolate(ref_in, size=(H, W), mode='bilinear', align_corners=ac), t)
    ref.backward()
    upstream = torch.tensor(0.7, device='cuda')
    lr_d, t_d = lr.cuda(), t.cuda()
    acc = E.lowres_ce_fwd(lr_d, nc, t_d, ac)
    loss = E.ce_finalize(acc)
    assert int(acc[1:2].view(torch.int64)) == int((t != -100).sum())
    assert abs(float(loss) - float(ref)) <= 2e-5 * abs(float(ref))
    dlr = E.lowres_ce_bwd(lr_d, nc, t_d, upstream, acc[1:2].view(torch.int64), ac)
    want = 0.7 * ref_in.grad.permute(0, 2, 3, 1).float()
    assert rel_err(dlr[..., :nc], want) < 1e-4
    assert float(dlr[..., nc:].abs().max()) == 0.0 if ct > nc else True
    # the same numbers as the unfused path the model falls back to (upsampled logits -> softmax_ce)
    logits = E.logits_upsample_fwd(lr_d, nc, H, W, align_corners=ac)
    acc2, _ = E.softmax_ce(logits, t_d, want_grad=False)
    assert abs(float(E.ce_finalize(acc2)) - float(loss)) <= 1e-5 * abs(float(loss))
    # bf16 Act for the classifier backward, with and without a second gradient through the logits
    a = E.class_map_grad(dlr)
    assert rel_err(a.t.float(), dlr) < 4e-3
    b = E.class_map_grad(dlr, a)
    assert rel_err(b.t.float(), 2 * dlr) < 8e-3


def test_lowres_ce_all_ignored_and_bad_target(E):
    lr = torch.randn(1, 8, 8, 8, device='cuda')
    t = torch.full((1, 32, 32), -100, dtype=torch.int64, device='cuda')
    acc = E.lowres_ce_fwd(lr, 5, t, True)
    assert torch.isnan(E.ce_finalize(acc))                             # 0 / 0, like nn.CrossEntropyLoss
    t[0, 3, 3] = 7                                                       # out of range -> NaN, no host sync
    assert torch.isnan(E.ce_finalize(E.lowres_ce_fwd(lr, 5, t, True)))
    with pytest.raises(RuntimeError, match='lowres_ce'):
        E.lowres_ce_fwd(torch.randn(1, 64, 64, 24, device='cuda'), 21, torch.zeros(1, 64, 64, dtype=torch.int64, device='cuda'))


@pytest.mark.parametrize('kind', ['deeplabv3plus', 'unet'])
def test_compute_loss_takes_the_class_map_shortcut_and_matches_the_logits_path(E, kind, monkeypatch):
    import pytorch_segmentation_b200 as P
    torch.manual_seed(0)
    m = (P.DeepLabV3Plus(21) if kind == 'deeplabv3plus' else P.UNet(2)).cuda().eval()
    nc = m.num_classes
    x = torch.randn(2, 3, 128, 192, device='cuda')
    t = torch.randint(0, nc, (2, 128, 192), device='cuda')
    res = {}
    for on in (True, False):
        monkeypatch.setattr(E, 'LOWRES_CE', on)
        m.zero_grad(set_to_none=True)
        out = m(x)
        assert hasattr(out, '_b200seg_class_map')
        before = E._lib.launches()
        loss = P.compute_loss(out, t)
        res[on] = dict(loss=float(loss), launches=E._lib.launches() - before)
        loss.backward()
        res[on]['grads'] = {k: p.grad.clone() for k, p in m.named_parameters() if p.grad is not None}
    assert res[True]['launches'] == 2 and res[False]['launches'] == 2
    assert abs(res[True]['loss'] - res[False]['loss']) <= 1e-5 * abs(res[False]['loss'])
    assert set(res[True]['grads']) == set(res[False]['grads'])
    worst = max(float((res[True]['grads'][k] - g).norm()) / max(float(g.norm()), 1e-12) for k, g in res[False]['grads'].items())
    assert worst < 2e-2, worst
    # an in-place edit of the logits turns the shortcut off; so does a size mismatch with the targets
    with torch.no_grad():
        out = m(x)
        out.mul_(2.0)
        assert abs(float(P.compute_loss(out, t)) - res[False]['loss']) > 1e-3
        small = P.compute_loss(m(x), t[:, ::2, ::2].contiguous())
        assert torch.isfinite(small)
    # a gradient through the logits AND the class-map loss: both reach the head
    m.zero_grad(set_to_none=True)
    monkeypatch.setattr(E, 'LOWRES_CE', True)
    out = m(x)
    (P.compute_loss(out, t) + 1e-3 * out.square().mean()).backward()
    assert all(torch.isfinite(p.grad).all() for p in m.parameters() if p.grad is not None)


# ---- space-to-depth stem: stride-2 few-channel conv as row taps over overlapping windows (csrc/im2col.cu) ---------------
S2D_CASES = [
    # n, c, h, w, k, r, pad
    (2, 3, 64, 96, 64, 7, 3),           # ResNet-50 conv1 (torchvision resnet.py via models/deeplabv3plus.py:17)
    (1, 3, 70, 54, 64, 7, 3),           # sizes that are not multiples of the tile
    (2, 3, 32, 64, 64, 3, 1),           # HRNet stem conv (models/hrnet.py:260)
    (1, 1, 33, 47, 16, 5, 2),           # odd image, single channel
]


@pytest.mark.parametrize('n,c,h,w,k,r,pad', S2D_CASES)
def test_space_to_depth_stem_conv(E, n, c, h, w, k, r, pad):
    g = torch.Generator().manual_seed(n * 100 + h + r)
    x = torch.randn(n, c, h, w, generator=g)
    wt = torch.randn(k, c, r, r, generator=g) * 0.1
    win, oh, ow = E.space_to_depth(x.cuda(), r, r, pad)
    assert (oh, ow) == ((h + 2 * pad - r) // 2 + 1, (w + 2 * pad - r) // 2 + 1)
    a = (r + 1) // 2
    cs = (4 * c + 7) // 8 * 8
    assert tuple(win.t.shape) == (n, oh + a - 1, ow + a - 1, cs) and win.c == a * cs
    # the transform is a bit-exact re-arrangement of the bf16-rounded padded image
    zh, zw = oh + a - 1, ow + a - 1
    xp = torch.zeros(n, c, 2 * zh, 2 * zw)
    hh, ww = min(h, 2 * zh - pad), min(w, 2 * zw - pad)
    xp[:, :, pad:pad + hh, pad:pad + ww] = bf16_round(x)[:, :, :hh, :ww]
    want = xp.view(n, c, zh, 2, zw, 2).permute(0, 2, 4, 3, 5, 1).reshape(n, zh, zw, 4 * c)      # (b, b', ch)
    got = win.t.float().cpu()
    assert torch.equal(got[..., :4 * c], want)
    assert cs == 4 * c or float(got[..., 4 * c:].abs().max()) == 0.0
    pw = E.PackedS2DWeight().update(wt.cuda())
    assert (pw.r, pw.s, pw.c) == (a, 1, a * cs)
    y = E.Act.empty(n, oh, ow, k, 'cuda')
    stats = torch.zeros(2 * k, device='cuda')
    E.conv_fprop([(win, pw, 1)], y, stats=stats)
    xr = bf16_round(x)
    wr = bf16_round(wt).requires_grad_(True)
    ref = F.conv2d(xr, wr, None, 2, pad)
    assert rel_err(from_act(y), ref) < TOL
    assert rel_err(stats[:k], from_act(y).sum((0, 2, 3))) < 1e-3
    gy = bf16_round(torch.randn(n, k, oh, ow, generator=g))
    ref.backward(gy)
    dwin = E.conv_wgrad(nhwc(gy), win, k, pw.r, 1, 1)
    dw = E.unpack_wgrad_s2d(dwin, pw)
    assert dw.shape == wt.shape
    assert rel_err(dw, wr.grad) < TOL


def test_anchored_window_source_is_checked(E):
    win, oh, ow = E.space_to_depth(torch.randn(1, 3, 32, 32, device='cuda'), 7, 7, 3)
    pw = E.PackedS2DWeight().update(torch.randn(64, 3, 7, 7, device='cuda'))
    with pytest.raises(RuntimeError, match='exceed'):
        E.conv_fprop([(win, pw, 1)], E.Act.empty(1, oh + 1, ow, 64, 'cuda'))      # one output row too many
    with pytest.raises(RuntimeError, match='forward-only'):
        E._conv(E._lib.load().b200seg_conv_dgrad, 'conv_dgrad', 1, oh, ow,
                [(win, pw.wf.data_ptr(), pw.k_rows, pw.r, pw.s, 1, 1)], E.Act.empty(1, oh, ow, 64, 'cuda').ptr, 64, 0, 64,
                False, None, None, False, None)


def test_bn_apply_train_equals_finalize_then_apply(E):
    # the fused launch (lazy finalize consumed by bn_apply) against the two-launch form, residual variants included
    g = torch.Generator().manual_seed(11)
    n, h, w, c = 2, 12, 20, 64
    x = nhwc(torch.randn(n, c, h, w, generator=g) * 2 + 0.5)
    res = nhwc(torch.randn(n, c, h, w, generator=g))
    gamma, beta = (1 + 0.1 * torch.randn(c, generator=g)).cuda(), (0.1 * torch.randn(c, generator=g)).cuda()
    stats = E.bn_stats(x)
    for use_res, res_affine in ((False, False), (True, False), (True, True)):
        rm1, rv1 = torch.zeros(c, device='cuda'), torch.ones(c, device='cuda')
        rm2, rv2 = rm1.clone(), rv1.clone()
        rvec = E.bn_finalize(E.bn_stats(res), c, res.pixels, gamma, beta, 1e-5, 0.1, None, None) if res_affine else None
        v1 = E.bn_finalize(stats, c, x.pixels, gamma, beta, 1e-5, 0.1, rm1, rv1)
        y1 = E.Act.empty(n, h, w, c, 'cuda')
        E.bn_apply(x, y1, v1, True, res=res if use_res else None, res_vec=rvec)
        v2 = E.bn_finalize_lazy(stats, c, x.pixels, gamma, beta, 1e-5, 0.1, rm2, rv2)
        assert getattr(v2, '_b200seg_pending', None) is not None
        before = E._lib.launches()
        y2 = E.Act.empty(n, h, w, c, 'cuda')
        E.bn_apply(x, y2, v2, True, res=res if use_res else None, res_vec=rvec)
        assert E._lib.launches() - before == 1 and v2._b200seg_pending is None
        assert torch.equal(y1.t, y2.t) and torch.equal(v1, v2) and torch.equal(rm1, rm2) and torch.equal(rv1, rv2)
    # a lazy vec nobody applied resolves to the same numbers on demand
    v3 = E.bn_resolve(E.bn_finalize_lazy(stats, c, x.pixels, gamma, beta, 1e-5, 0.1, None, None))
    assert torch.equal(v3[0], v1[0]) and torch.equal(v3[3], v1[3])


def test_relu_mask_bits_drive_the_gradient_join(E):
    # bn_apply(want_bits) leaves (y > 0) as one byte per 8 channels; grad_add_masked reads it instead of y
    g = torch.Generator().manual_seed(12)
    n, h, w, c = 2, 9, 14, 48
    x = nhwc(torch.randn(n, c, h, w, generator=g))
    res = nhwc(torch.randn(n, c, h, w, generator=g))
    gamma, beta = (1 + 0.1 * torch.randn(c, generator=g)).cuda(), (0.1 * torch.randn(c, generator=g)).cuda()
    vec = E.bn_finalize_lazy(E.bn_stats(x), c, x.pixels, gamma, beta, 1e-5, 0.1, None, None)
    y = E.Act.empty(n, h, w, c, 'cuda')
    E.bn_apply(x, y, vec, True, res=res, want_bits=True)
    assert y.bits is not None and tuple(y.bits.shape) == (n * h * w, c // 8)
    want = (y.t.float() > 0).view(-1, c // 8, 8).to(torch.int32)
    want = (want << torch.arange(8, device='cuda', dtype=torch.int32)).sum(-1).to(torch.uint8)
    assert torch.equal(y.bits, want)
    dx0 = nhwc(torch.randn(n, c, h, w, generator=g))
    dy = nhwc(torch.randn(n, c, h, w, generator=g))
    for use_dy in (True, False):
        a, b = E.Act(dx0.t.clone()), E.Act(dx0.t.clone())
        plain = E.Act(y.t)                                      # same tensor without the bits: reads y itself
        before = E._lib.launches()
        E.grad_add_masked(a, dy=dy if use_dy else None, xin=y)
        E.grad_add_masked(b, dy=dy if use_dy else None, xin=plain)
        assert E._lib.launches() - before == 2 and torch.equal(a.t, b.t)
