"""CPU fp32 restatement of the reference HRNet (models/hrnet.py) — functional, over a reference-keyed state dict.

TEST INFRASTRUCTURE — the checker for ``pytorch_segmentation_b200.models.HRNet`` on the GPU box (the reference tree
does not travel); never imported by the product.  ``tests/test_oracle.py`` checks it against the UNMODIFIED reference
HRNet run here and against ``tests/golden/hrnet_small.npz`` (written by ``oracle/make_golden.py`` from the reference).

The walk below follows the reference's module tree by state-dict prefix:
  stem.0 / stem.1                ConvNormAct(3,64,3,2,activate=None), ConvNormAct(64,64,3,2)      hrnet.py:260-261
  stem.2.<i>                     4 Bottlenecks, the first with a 1x1 shortcut                    hrnet.py:262,328-348
  transition<k>.<i>              3x3 ConvNormAct (width change) or stride-2 chain (new branch)   hrnet.py:304-326,375-398
  stage<k>.0.branches.<b>.<i>    4 BasicBlocks per branch                                        hrnet.py:146-177
  stage<k>.0.fuse_layers.<i>.<j> 1x1 ConvNormAct + bilinear up (j>i) / stride-2 3x3 chain (j<i)  hrnet.py:187-226
  final_layer                    1x1 conv with bias, then bilinear x4 (align_corners=False)      hrnet.py:302,400-404
"""
import torch
import torch.nn.functional as F

from . import restatement as _R

BN_EPS, BN_MOMENTUM = 1e-5, 0.1     # nn.BatchNorm2d defaults; hrnet.py:14 passes momentum=0.1 explicitly

# Mask-matched mode (tests/test_gpu_gradparity.py): under ``restatement.substituted(values, grad_log=...)`` every conv
# result, BatchNorm / ReLU / residual output and fused branch sum takes the VALUE (and ReLU mask) the CUDA run stored,
# and the backward is re-synchronised at every stored gradient of a raw conv output.  Keys are the reference's
# state-dict names of the conv weight + '.raw' / '.out' / '.graw'; the fused sums are 'sum_relu.<n>' in forward order.
# Outside that context the helpers below are the identity / a plain ReLU.
_SEQ = [0]


def _conv(sd, wkey, x, stride=1, pad=0):
    y = _R._subst(wkey + '.raw', F.conv2d(x, sd[wkey], None, stride, pad))
    return _R._resync(wkey + '.graw', y)


def _act(wkey, y, relu):
    if relu and _R._SUBST is not None:
        return _R._relu_adopt(y, wkey + '.out', _R._SUBST, _R._RECORD)
    return _R._subst(wkey + '.out', F.relu(y) if relu else y)


def _bn(sd, p, y, training):
    if training:
        sd[p + '.num_batches_tracked'] += 1
    return F.batch_norm(y, sd[p + '.running_mean'], sd[p + '.running_var'], sd[p + '.weight'], sd[p + '.bias'],
                        training, BN_MOMENTUM, BN_EPS)


def _cna(sd, p, x, ksize, stride, relu, training):
    """ConvNormAct under prefix p: conv = p.0, bn = p.1 (oracle/shim/pytorch_modules/nn)."""
    y = _conv(sd, p + '.0.weight', x, stride, (ksize - 1) // 2)
    return _act(p + '.0.weight', _bn(sd, p + '.1', y, training), relu)


def _shortcut(sd, p, x, training):
    if (p + '.downsample.0.weight') not in sd:
        return x
    k = p + '.downsample.0.weight'
    return _act(k, _bn(sd, p + '.downsample.1', _conv(sd, k, x), training), False)


def basic_block(sd, p, x, training):
    """hrnet.py:39-56"""
    y = _act(p + '.conv1.weight', _bn(sd, p + '.bn1', _conv(sd, p + '.conv1.weight', x, 1, 1), training), True)
    y = _bn(sd, p + '.bn2', _conv(sd, p + '.conv2.weight', y, 1, 1), training)
    return _act(p + '.conv2.weight', y + _shortcut(sd, p, x, training), True)


def bottleneck(sd, p, x, training):
    """hrnet.py:83-103"""
    y = _act(p + '.conv1.weight', _bn(sd, p + '.bn1', _conv(sd, p + '.conv1.weight', x), training), True)
    y = _act(p + '.conv2.weight', _bn(sd, p + '.bn2', _conv(sd, p + '.conv2.weight', y, 1, 1), training), True)
    y = _bn(sd, p + '.bn3', _conv(sd, p + '.conv3.weight', y), training)
    return _act(p + '.conv3.weight', y + _shortcut(sd, p, x, training), True)


def _count(sd, prefix):
    """Number of consecutively indexed children under ``prefix`` (``prefix.0``, ``prefix.1``, ...)."""
    n = 0
    while any(k.startswith('%s.%d.' % (prefix, n)) for k in sd):
        n += 1
    return n


def hr_module(sd, p, xs, multi_scale_output, training):
    """hrnet.py:228-251"""
    nb = len(xs)
    xs = list(xs)
    for b in range(nb):
        for i in range(_count(sd, '%s.branches.%d' % (p, b))):
            xs[b] = basic_block(sd, '%s.branches.%d.%d' % (p, b, i), xs[b], training)
    if nb == 1:
        return xs
    outs = []
    for i in range(nb if multi_scale_output else 1):
        y = None
        for j in range(nb):
            q = '%s.fuse_layers.%d.%d' % (p, i, j)
            if j == i:
                t = xs[j]
            elif j > i:
                t = _cna(sd, q + '.0', xs[j], 1, 1, True, training)
                t = F.interpolate(t, scale_factor=2 ** (j - i), mode='bilinear', align_corners=False)
            else:
                t = xs[j]
                for k in range(i - j):
                    last = k == i - j - 1
                    t = _cna(sd, '%s.%d' % (q, k) if last else '%s.%d.0' % (q, k), t, 3, 2, not last, training)
            y = t if y is None else y + t
        outs.append(_act('sum_relu.%d' % _SEQ[0], y, True))
        _SEQ[0] += 1
    return outs


def hrnet(sd, x, branches=(2, 3, 4), training=False):
    """hrnet.py:373-406.  ``sd`` is mutated in training mode (BN running statistics), like the module."""
    _SEQ[0] = 0
    x = _cna(sd, 'stem.0', x, 3, 2, False, training)
    x = _cna(sd, 'stem.1', x, 3, 2, True, training)
    for i in range(_count(sd, 'stem.2')):
        x = bottleneck(sd, 'stem.2.%d' % i, x, training)
    ys = [x]
    for stage, nb in zip((2, 3, 4), branches):
        xs = []
        for i in range(nb):
            p = 'transition%d.%d' % (stage - 1, i)
            has = any(k.startswith(p + '.') for k in sd)
            if not has:
                xs.append(ys[i])                                   # hrnet.py:388,397: None -> pass through
                continue
            src = ys[-1] if (stage > 2 or i >= len(ys)) else ys[i]  # hrnet.py:378 vs :386,:395
            if (p + '.0.weight') in sd:                             # a single ConvNormAct: width change, stride 1
                xs.append(_cna(sd, p, src, 3, 1, True, training))
            else:                                                   # Sequential of stride-2 ConvNormActs
                t = src
                for s in range(_count(sd, p)):
                    t = _cna(sd, '%s.%d' % (p, s), t, 3, 2, True, training)
                xs.append(t)
        ys = hr_module(sd, 'stage%d.0' % stage, xs, stage != 4, training)
    y = F.conv2d(ys[0], sd['final_layer.weight'], sd['final_layer.bias'])
    return F.interpolate(y, scale_factor=(4, 4), mode='bilinear', align_corners=False)


def make_state(num_classes, branches=(2, 3, 4), seed=0):
    """A seeded, reference-keyed HRNet state dict with non-trivial BN affine parameters and running statistics
    (so scale, shift and eval-mode statistics are all exercised).  Built WITHOUT the reference or the product: shapes
    are enumerated from the architecture (hrnet.py:255-302)."""
    g = torch.Generator().manual_seed(seed)
    sd = {}

    def conv(name, co, ci, k):
        sd[name] = torch.randn(co, ci, k, k, generator=g) * (1.0 / (ci * k * k)) ** 0.5

    def bn(p, c):
        sd[p + '.weight'] = 1.0 + 0.2 * (torch.rand(c, generator=g) - 0.5)
        sd[p + '.bias'] = 0.1 * torch.randn(c, generator=g)
        sd[p + '.running_mean'] = 0.1 * torch.randn(c, generator=g)
        sd[p + '.running_var'] = 1.0 + 0.2 * (torch.rand(c, generator=g) - 0.5)
        sd[p + '.num_batches_tracked'] = torch.tensor(0)

    def cna(p, ci, co, k):
        conv(p + '.0.weight', co, ci, k)
        bn(p + '.1', co)

    cna('stem.0', 3, 64, 3)
    cna('stem.1', 64, 64, 3)
    inpl = 64
    for i in range(4):
        p = 'stem.2.%d' % i
        conv(p + '.conv1.weight', 64, inpl, 1); bn(p + '.bn1', 64)
        conv(p + '.conv2.weight', 64, 64, 3); bn(p + '.bn2', 64)
        conv(p + '.conv3.weight', 256, 64, 1); bn(p + '.bn3', 256)
        if inpl != 256:
            conv(p + '.downsample.0.weight', 256, inpl, 1); bn(p + '.downsample.1', 256)
        inpl = 256
    pre = [256]
    for stage, nb in zip((2, 3, 4), branches):
        ch = [32 * 2 ** i for i in range(nb)]
        for i in range(nb):
            p = 'transition%d.%d' % (stage - 1, i)
            if i < len(pre):
                if ch[i] != pre[i]:
                    cna(p, pre[i], ch[i], 3)
            else:
                steps = i + 1 - len(pre)
                for s in range(steps):
                    cna('%s.%d' % (p, s), pre[-1], ch[i] if s == steps - 1 else pre[-1], 3)
        m = 'stage%d.0' % stage
        for b in range(nb):
            for i in range(4):
                p = '%s.branches.%d.%d' % (m, b, i)
                conv(p + '.conv1.weight', ch[b], ch[b], 3); bn(p + '.bn1', ch[b])
                conv(p + '.conv2.weight', ch[b], ch[b], 3); bn(p + '.bn2', ch[b])
        for i in range(nb if stage != 4 else 1):
            for j in range(nb):
                q = '%s.fuse_layers.%d.%d' % (m, i, j)
                if j > i:
                    cna(q + '.0', ch[j], ch[i], 1)
                elif j < i:
                    for k in range(i - j):
                        if k == i - j - 1:
                            cna('%s.%d' % (q, k), ch[j], ch[i], 3)
                        else:
                            cna('%s.%d.0' % (q, k), ch[j], ch[j], 3)
        pre = ch
    sd['final_layer.weight'] = torch.randn(num_classes, pre[0], 1, 1, generator=g) * (1.0 / pre[0]) ** 0.5
    sd['final_layer.bias'] = 0.1 * torch.randn(num_classes, generator=g)
    return sd
