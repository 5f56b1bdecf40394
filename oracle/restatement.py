"""CPU fp32 restatement of the reference hot path (functional, over a reference-keyed state dict).

TEST INFRASTRUCTURE — the checker for the CUDA path and the timed CPU baseline; never shipped, never
imported by ``pytorch_segmentation_b200``.  It travels to the GPU box (the reference tree does not).

Every function cites the reference lines it restates.  It is validated against the unmodified
reference (``oracle/reference_loader.py``) in ``tests/test_oracle.py`` and pinned by the vectors in
``tests/golden/`` produced by ``oracle/make_golden.py``.  Parity note: the reference has no tests or
golden vectors of its own; the pin is "outputs of the reference itself run here".

Arithmetic is whatever PyTorch-CPU fp32 does for conv2d / batch_norm / interpolate / cross_entropy —
the same library calls the reference makes — so agreement with the reference is to fp32 round-off.
"""
import os
import sys

import torch
import torch.nn.functional as F

_SHIM = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'shim')

BN_EPS = 1e-5        # nn.BatchNorm2d default, never overridden for ConvNormAct
BN_MOMENTUM = 0.1


def _shim_backbones():
    if _SHIM not in sys.path:
        sys.path.insert(0, _SHIM)
    import pytorch_modules.backbones as bb
    return bb


# ----------------------------------------------------------------------------------------------
# Mask-matched differentiation (tests/test_gpu_gradparity.py)
# ----------------------------------------------------------------------------------------------
# A bf16 pipeline stores every conv result rounded to bf16; BatchNorm statistics and ReLU masks downstream are
# functions of the STORED values.  Compared against an oracle that keeps fp32 intermediates, ~0.3 % of the ReLU masks
# differ and composed gradients differ by several per cent although every kernel is right — too loose a bound to catch
# a wiring error in a hand-written backward pass.  With ``substituted(values)`` active, ``conv_norm_act`` (and the
# backbone hooks below) replace the VALUE of each raw conv output / block output by the one the CUDA run stored while
# keeping the fp32 autograd graph (straight-through: y + (stored - y).detach()), so oracle and CUDA path differentiate
# the same piecewise-linear function and gradients must agree to the plain 2e-2 tolerance.
# ``values``: {'<reference key prefix>.raw' | '<prefix>.out': fp32 CPU tensor}.
_SUBST = None


_RECORD = False


_GRAD_LOG = None


class substituted:
    """``record=True`` fills ``values`` with the oracle's own tensors at the same keys instead (CPU self-tests).
    ``grad_log`` (a dict) switches on the backward re-synchronisation of ``_resync`` for the functional restatements."""

    def __init__(self, values, record=False, grad_log=None):
        self.values, self.record, self.grad_log = values, record, grad_log

    def __enter__(self):
        global _SUBST, _RECORD, _GRAD_LOG
        self._prev = (_SUBST, _RECORD, _GRAD_LOG)
        _SUBST, _RECORD, _GRAD_LOG = self.values, self.record, self.grad_log
        return self

    def __exit__(self, *exc):
        global _SUBST, _RECORD, _GRAD_LOG
        _SUBST, _RECORD, _GRAD_LOG = self._prev


def _resync(key, y):
    """Backward re-synchronisation at a tensor whose gradient the CUDA run stored (bf16) under ``key``: when backward
    reaches ``y`` the oracle's own gradient is logged beside the stored one (the caller compares the pair: one
    kernel-step of error) and REPLACED by the stored one, so storage noise does not accumulate from layer to layer and
    every layer's backward is checked on the gradient its CUDA counterpart actually received."""
    if _SUBST is None or not y.requires_grad:
        return y
    if _RECORD:                       # CPU self-tests: the oracle's own gradient becomes the "stored" one
        vals = _SUBST
        y = y.view_as(y)
        y.register_hook(lambda g: vals.__setitem__(key, g.detach().clone()))
        return y
    if _GRAD_LOG is None:
        return y
    g_stored = _SUBST.get(key)
    if g_stored is None:
        return y
    log = _GRAD_LOG

    def on_grad(g):
        gs = g_stored.to(g.dtype).reshape(g.shape)
        log[key] = (g.detach().clone(), gs)
        return gs
    y = y.view_as(y)
    y.register_hook(on_grad)
    return y


def _subst(key, y):
    if _SUBST is None:
        return y
    if _RECORD:
        _SUBST[key] = y.detach().clone()
        return y
    v = _SUBST.get(key)
    if v is None:
        return y
    assert v.numel() == y.numel(), (key, tuple(v.shape), tuple(y.shape))
    return y + (v.to(y.dtype).reshape(y.shape) - y).detach()


def _relu_adopt(pre, key, values, record=False):
    """ReLU whose VALUE and MASK are the stored ones: forward value = values[key] (the bf16 tensor the CUDA run wrote),
    backward d/d(pre) = [values[key] > 0].  An fp32 oracle's own mask differs from the kernels' on the handful of
    elements whose pre-activation is within fp32 round-off of zero (different operation order inside BatchNorm); on a
    12 x 12 map one such flip moves a weight-gradient entry by several per cent.  Without a stored value: plain ReLU."""
    if record:
        y = F.relu(pre)
        values[key] = y.detach().clone()
        return y
    v = values.get(key) if values is not None else None
    if v is None:
        return F.relu(pre)
    assert v.numel() == pre.numel(), (key, tuple(v.shape), tuple(pre.shape))
    v = v.to(pre.dtype).reshape(pre.shape)
    y = pre * (v > 0).to(pre.dtype)
    return y + (v - y).detach()


class _Restore:
    def __init__(self, fn):
        self.fn = fn

    def remove(self):
        self.fn()


def hook_backbone(backbone, values, prefix='backbone.', record=False, grad_log=None):
    """Makes a torchvision ResNet (through the shim) consume what the CUDA run stored: forward hooks substitute the conv
    results ('<prefix><conv name>.weight.raw'), and every ReLU takes value and mask from the stored activation
    ('<prefix><conv name>.weight.out': conv1 -> the stem output that feeds the max pool; block.conv1 / conv2 -> the ReLU
    outputs inside the block; block.conv3 -> the block output) — the Bottleneck forward is restated here for that
    (torchvision/models/resnet.py Bottleneck.forward: conv1-bn1-relu, conv2-bn2-relu, conv3-bn3, + identity or
    downsample(x), relu).  Returns handles with ``.remove()``.

    ``grad_log`` (a dict) switches on the backward re-synchronisation at block boundaries: where ``values`` holds
    '<prefix><block>.conv3.weight.gout' — the gradient the CUDA run stored (bf16) on entry to that block's backward,
    possibly pre-multiplied by the ReLU mask of the block output — the oracle's own gradient of the block output is
    logged (masked the same way) under that key and REPLACED by the stored one, so bf16 gradient-storage noise does
    not accumulate from block to block and every block's backward is checked on the gradient its CUDA counterpart
    actually received.  The logged gradients are compared with the stored ones by the caller, which closes the chain."""
    import torch.nn as nn
    handles = []

    def resync(key, out):
        g_stored = values.get(key) if grad_log is not None else None
        if g_stored is None or not out.requires_grad:
            return
        mask = (out.detach() > 0).to(out.dtype)

        def on_grad(g):
            grad_log[key] = (g.detach() * mask, g_stored.to(g.dtype) * mask)
            return g_stored.to(g.dtype)
        out.register_hook(on_grad)

    def sub(key):
        def hook(_m, _i, out):
            if record:
                values[key] = out.detach().clone()
                return None
            v = values.get(key)
            if v is None:
                return None
            assert v.shape == out.shape, (key, tuple(v.shape), tuple(out.shape))
            new = out + (v.to(out.dtype) - out).detach()
            # re-synchronisation inside the blocks as well: '<conv>.weight.graw' is the gradient of this raw conv output
            # the CUDA run stored (the BatchNorm backward's result, bf16) — logged against the oracle's and then used in
            # its place, so every dgrad / wgrad / BN backward is checked on exactly the operands its kernel had
            g_stored = values.get(key[:-len('.raw')] + '.graw') if grad_log is not None else None
            if g_stored is not None and new.requires_grad:
                def on_grad(g, gkey=key[:-len('.raw')] + '.graw'):
                    grad_log[gkey] = (g.detach(), g_stored.to(g.dtype))
                    return g_stored.to(g.dtype)
                new.register_hook(on_grad)
            return new
        return hook

    def patch_forward(mod, fn):
        had = 'forward' in mod.__dict__
        old = mod.__dict__.get('forward')
        mod.forward = fn

        def undo():
            if had:
                mod.forward = old
            else:
                del mod.forward
        handles.append(_Restore(undo))

    for name, m in backbone.named_modules():
        if isinstance(m, nn.Conv2d):
            handles.append(m.register_forward_hook(sub(prefix + name + '.weight.raw')))
        if hasattr(m, 'conv3') and hasattr(m, 'bn3'):          # Bottleneck
            def block_forward(x, blk=m, key=prefix + name):
                out = _relu_adopt(blk.bn1(blk.conv1(x)), key + '.conv1.weight.out', values, record)
                out = _relu_adopt(blk.bn2(blk.conv2(out)), key + '.conv2.weight.out', values, record)
                out = blk.bn3(blk.conv3(out))
                identity = x if blk.downsample is None else blk.downsample(x)
                out = _relu_adopt(out + identity, key + '.conv3.weight.out', values, record)
                resync(key + '.conv3.weight.gout', out)
                return out
            patch_forward(m, block_forward)
    if hasattr(backbone, 'relu') and hasattr(backbone, 'maxpool'):
        # the stem's ReLU (the shim computes relu(bn1(conv1 x)) with this module): stored value and mask
        old_relu = backbone.relu

        class _StemReLU(nn.Module):
            def forward(self, pre):
                return _relu_adopt(pre, prefix + 'conv1.weight.out', values, record)
        backbone.relu = _StemReLU()

        def undo_relu():
            backbone.relu = old_relu
        handles.append(_Restore(undo_relu))
    return handles


def round_conv_weights_to_bf16_(model):
    """In place: every conv weight of an ``OracleModel`` (4-D tensors of ``sd`` and of the backbone) becomes its nearest
    bf16-representable fp32 value.  The CUDA path multiplies bf16 packings of the fp32 master weights; with
    representable masters the packing is exact, so a weight rounding error — one fixed perturbation applied at every
    pixel, which per-channel sums over pixels (BatchNorm bias / scale gradients) do not average away — is not part of
    the comparison."""
    with torch.no_grad():
        for v in list(model.sd.values()) + list(model.backbone.parameters()):
            if v.dim() == 4:
                v.copy_(v.to(torch.bfloat16).float())
    return model


# ----------------------------------------------------------------------------------------------
# L0: ConvNormAct (external pytorch_modules.nn.ConvNormAct; constraints in SURVEY.md §8c)
# ----------------------------------------------------------------------------------------------
def conv_norm_act(sd, prefix, x, ksize, dilation=1, training=False, relu=True):
    """conv(bias=False, same padding) -> BN -> ReLU, parameters ``<prefix>.0.*`` / ``<prefix>.1.*``.

    Used at models/aspp.py:12,27,29,30; models/deeplabv3plus.py:20; models/unet.py:19-21.
    In training mode updates running_mean/var (momentum 0.1, unbiased var) and num_batches_tracked
    in ``sd`` exactly like nn.BatchNorm2d.
    """
    pad = ((ksize - 1) // 2) * dilation
    y = F.conv2d(x, sd[prefix + '.0.weight'], None, 1, pad, dilation)
    y = _subst(prefix + '.0.weight.raw', y)
    if training:
        sd[prefix + '.1.num_batches_tracked'] += 1
    y = F.batch_norm(y, sd[prefix + '.1.running_mean'], sd[prefix + '.1.running_var'],
                     sd[prefix + '.1.weight'], sd[prefix + '.1.bias'], training, BN_MOMENTUM, BN_EPS)
    if relu and _SUBST is not None:
        return _relu_adopt(y, prefix + '.0.weight.out', _SUBST, _RECORD)      # stored value AND mask
    return _subst(prefix + '.0.weight.out', F.relu(y) if relu else y)


# ----------------------------------------------------------------------------------------------
# L1: ASPP (models/aspp.py)
# ----------------------------------------------------------------------------------------------
def aspp_pooling(sd, prefix, x, training=False):
    """models/aspp.py:14-19 — GAP -> 1x1 ConvNormAct -> bilinear(align_corners=False) back to HxW.
    A 1x1 source interpolated with align_corners=False is a constant broadcast."""
    size = x.shape[-2:]
    g = F.adaptive_avg_pool2d(x, 1)
    g = conv_norm_act(sd, prefix + '.gap.1', g, 1, training=training)
    return F.interpolate(g, size=size, mode='bilinear', align_corners=False)


def aspp(sd, prefix, x, rates, training=False, return_branches=False):
    """models/aspp.py:32-37 — [pool, 1x1, 3x3 dil r...] -> cat(dim=1) -> 1x1 project."""
    outs = [aspp_pooling(sd, prefix + '.blocks.0', x, training),
            conv_norm_act(sd, prefix + '.blocks.1', x, 1, training=training)]
    for i, r in enumerate(rates):
        outs.append(conv_norm_act(sd, prefix + '.blocks.%d' % (i + 2), x, 3, dilation=r, training=training))
    cat = torch.cat(outs, dim=1)
    y = conv_norm_act(sd, prefix + '.project', cat, 1, training=training)
    return (y, outs) if return_branches else y


# ----------------------------------------------------------------------------------------------
# L1: DeepLabV3+ and UNet heads
# ----------------------------------------------------------------------------------------------
DEEPLAB_RATES = (6, 12, 18)   # models/deeplabv3plus.py:21


def deeplab_head(sd, low, high, training=False):
    """models/deeplabv3plus.py:30-43 given features[1] (``low``) and features[-1] (``high``)."""
    low = conv_norm_act(sd, 'project', low, 1, training=training)                    # :31
    x = aspp(sd, 'aspp', high, DEEPLAB_RATES, training)                             # :33
    x = F.interpolate(x, scale_factor=4, mode='bilinear', align_corners=True)       # :34-37
    x = torch.cat([x, low], 1)                                                      # :38
    x = F.conv2d(x, sd['cls_conv.weight'], sd['cls_conv.bias'], 1, 1)               # :39
    return F.interpolate(x, scale_factor=4, mode='bilinear', align_corners=True)    # :40-43


def unet_head(sd, feats, training=False):
    """models/unet.py:28-55 given the 5 backbone maps (x1 unused)."""
    _x1, x2, x3, x4, x = feats

    def up2(t):
        return F.interpolate(t, scale_factor=2, mode='bilinear', align_corners=True)

    x = up2(conv_norm_act(sd, 'up_convs.0', x, 3, training=training))               # :29-33
    x = torch.cat([x, x4], 1)                                                       # :34
    x = up2(conv_norm_act(sd, 'up_convs.1', x, 3, training=training))               # :35-39
    x = torch.cat([x, x3], 1)                                                       # :40
    x = up2(conv_norm_act(sd, 'up_convs.2', x, 3, training=training))               # :41-45
    x = torch.cat([x, x2], 1)                                                       # :46
    x = up2(x)                                                                      # :47-50
    x = F.conv2d(x, sd['cls_conv.weight'], sd['cls_conv.bias'], 1, 1)               # :51
    return up2(x)                                                                   # :52-55


class OracleModel:
    """Reference-keyed state dict + torchvision backbone (through the shim) + functional head.

    ``kind`` is 'deeplabv3plus' (models/deeplabv3plus.py:15-26) or 'unet' (models/unet.py:14-25).
    Head tensors live in ``self.sd`` (leaf tensors, requires_grad on parameters) so tests can read
    ``.grad`` by reference key.
    """

    def __init__(self, kind, num_classes, seed=0):
        bb = _shim_backbones()
        import pytorch_modules.nn as pmnn
        from pytorch_modules.utils import initialize_weights
        g = torch.random.get_rng_state()
        torch.manual_seed(seed)
        self.kind, self.num_classes = kind, num_classes
        heads = {}
        if kind == 'deeplabv3plus':
            self.backbone = bb.resnet50(pretrained=True, replace_stride_with_dilation=[False, False, True])
            heads['project'] = pmnn.ConvNormAct(256, 128, 1)
            blocks = [('aspp.blocks.0.gap.1', pmnn.ConvNormAct(2048, 256, 1)),
                      ('aspp.blocks.1', pmnn.ConvNormAct(2048, 256, 1))]
            for i, r in enumerate(DEEPLAB_RATES):
                blocks.append(('aspp.blocks.%d' % (i + 2), pmnn.ConvNormAct(2048, 256, dilation=r)))
            heads.update(blocks)
            heads['aspp.project'] = pmnn.ConvNormAct(256 * 5, 256, 1)
            heads['cls_conv'] = torch.nn.Conv2d(384, num_classes, 3, padding=1)
        elif kind == 'unet':
            self.backbone = bb.mobilenet_v2(pretrained=True)
            for i, (ci, co) in enumerate(((1280, 256), (352, 128), (160, 64))):
                heads['up_convs.%d' % i] = pmnn.ConvNormAct(ci, co)
            heads['cls_conv'] = torch.nn.Conv2d(88, num_classes, 3, padding=1)
        else:
            raise ValueError(kind)
        self.sd = {}
        for name, mod in heads.items():
            initialize_weights(mod)
            for k, v in mod.state_dict().items():
                v = v.detach().clone()
                # perturb BN affine / conv bias / running stats so scale, shift and eval-mode
                # statistics are all exercised (init alone gives gamma=1, beta=0, mean=0, var=1)
                if k.endswith('1.weight') or k.endswith('running_var'):
                    v = 1.0 + 0.2 * (torch.rand_like(v) - 0.5)
                elif k.endswith('.bias') or k.endswith('running_mean'):
                    v = 0.1 * torch.randn_like(v)
                self.sd[name + '.' + k] = v
        torch.random.set_rng_state(g)
        self.training = False
        self.requires_grad_(True)

    # -- nn.Module-like surface -------------------------------------------------------------
    def requires_grad_(self, flag=True):
        for k, v in self.sd.items():
            if v.is_floating_point() and 'running_' not in k:
                v.requires_grad_(flag)
        return self

    def train(self, mode=True):
        self.training = mode
        self.backbone.train(mode)
        return self

    def eval(self):
        return self.train(False)

    def state_dict(self):
        out = {'backbone.' + k: v.detach().clone() for k, v in self.backbone.state_dict().items()}
        out.update({k: v.detach().clone() for k, v in self.sd.items()})
        return out

    def load_state_dict(self, state):
        self.backbone.load_state_dict({k[len('backbone.'):]: v for k, v in state.items()
                                       if k.startswith('backbone.')})
        with torch.no_grad():
            for k in self.sd:
                self.sd[k].copy_(state[k])

    def to(self, device):
        """Moves the oracle (tests use the GPU copy under autocast to measure the bf16 noise floor of a
        library implementation; the parity reference itself stays fp32 on the CPU)."""
        self.backbone.to(device)
        self.sd = {k: v.detach().to(device).requires_grad_(v.requires_grad) for k, v in self.sd.items()}
        return self

    def zero_grad(self):
        self.backbone.zero_grad()
        for v in self.sd.values():
            v.grad = None

    def features(self, x):
        return self.backbone(x)

    def head(self, feats):
        if self.kind == 'deeplabv3plus':
            return deeplab_head(self.sd, feats[1], feats[-1], self.training)
        return unet_head(self.sd, feats, self.training)

    def __call__(self, x):
        return self.head(self.features(x))


# ----------------------------------------------------------------------------------------------
# L2: loss and metrics
# ----------------------------------------------------------------------------------------------
def compute_loss(outputs, targets, model=None):
    """utils/utils.py:17-24 — bilinear(align_corners=True) resize to the target size, then
    nn.CrossEntropyLoss() (mean over all N*H*W pixels, ignore_index -100, no class weights)."""
    outputs = F.interpolate(outputs, (targets.size(1), targets.size(2)), mode='bilinear', align_corners=True)
    return F.cross_entropy(outputs, targets)


def compute_metrics(tp, fn, fp):
    """utils/utils.py:51-65 — float32 vectors in, (T, P, R, miou, F1) out; zero denominators -> 1."""
    def guarded(d):
        d = d.clone()
        d[d <= 0] = 1
        return d
    miou = tp / guarded(tp + fp + fn)
    T = tp + fn
    P = tp / guarded(tp + fp)
    R = tp / guarded(tp + fn)
    F1 = 2 * tp / guarded(2 * tp + fp + fn)
    return T, P, R, miou, F1


def tp_fn_fp_loop(predicted, targets, num_classes):
    """test.py:34-46 — the per-class eq/mask/sum loop (int64 counts; the reference adds them into
    float32 accumulators, test.py:22-24, which is exact only below 2**24)."""
    predicted = predicted.reshape(-1)
    targets = targets.reshape(-1)
    eq = predicted.eq(targets)
    tp = torch.zeros(num_classes, dtype=torch.int64)
    fn = torch.zeros(num_classes, dtype=torch.int64)
    fp = torch.zeros(num_classes, dtype=torch.int64)
    for c in range(num_classes):
        idx = targets.eq(c)
        positive = int(idx.sum())
        tpi = int(eq[idx].sum())
        tp[c] += tpi
        fn[c] += positive - tpi
        fp[c] += int(predicted.eq(c).sum()) - tpi
    return tp, fn, fp


def confusion_matrix(predicted, targets, num_classes):
    """CM[t, p] — superset of test.py:34-46 (tp = diag, fn = rowsum - diag, fp = colsum - diag)."""
    idx = targets.reshape(-1).to(torch.int64) * num_classes + predicted.reshape(-1).to(torch.int64)
    return torch.bincount(idx, minlength=num_classes * num_classes).reshape(num_classes, num_classes)


def counts_from_confusion(cm):
    tp = cm.diag()
    return tp, cm.sum(1) - tp, cm.sum(0) - tp


def evaluate(model, batches, num_classes):
    """test.py:14-73 restated without tqdm / show_batch / printing: returns (mean IoU float,
    mean loss, (tp, fn, fp) int64).  ``batches`` yields (inputs, targets)."""
    model.eval()
    tp = torch.zeros(num_classes, dtype=torch.int64)
    fn = torch.zeros_like(tp)
    fp = torch.zeros_like(tp)
    loss_sum, nb = 0.0, 0
    with torch.no_grad():
        for inputs, targets in batches:
            outputs = model(inputs)
            loss_sum += float(compute_loss(outputs, targets, model))
            nb += 1
            predicted = outputs.max(1)[1]
            a, b, c = tp_fn_fp_loop(predicted, targets, num_classes)
            tp += a
            fn += b
            fp += c
    miou = compute_metrics(tp.float(), fn.float(), fp.float())[3]
    return miou.mean().item(), loss_sum / max(nb, 1), (tp, fn, fp)
