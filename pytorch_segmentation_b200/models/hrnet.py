"""HRNet — same constructor, module tree (state-dict keys) and forward as models/hrnet.py:255-406, on libb200seg.

The reference builds HRNet from its own ``BasicBlock`` / ``Bottleneck`` (plain ``nn.Conv2d`` + ``nn.BatchNorm2d`` pairs,
models/hrnet.py:27-103), ``ConvNormAct`` transitions and fuse layers (stride 1 and 2, models/hrnet.py:187-226,
304-326) and ``HRModule``s that keep 2-4 resolutions alive and exchange them after every module
(models/hrnet.py:228-251).  Here every conv + BN (+ residual) (+ ReLU) is one ``ops.conv_bn`` node (tcgen05 implicit
GEMM with the BN statistics in its epilogue, one apply pass that also adds the residual), the bilinear x2^k of the fuse
layers is the gather kernel pair, the cross-resolution sums + ReLU are one ``add_n_relu`` pass, and the 1x1 classifier
with its x4 bilinear (models/hrnet.py:400-404) is the fp32 class-map path shared with DeepLabV3+/UNet.

Submodule names follow the reference exactly (``stem.{0,1}.{0,1}``, ``stem.2.<i>.conv1`` ..., ``transition<k>.<i>``,
``stage<k>.0.branches.<b>.<i>.conv1`` ..., ``stage<k>.0.fuse_layers.<i>.<j>...``, ``final_layer``) so reference
checkpoints load unchanged.
"""
import torch
import torch.nn as nn

from .. import engine as E
from .. import ops
from ..modules import ConvNormAct, initialize_weights

BN_MOMENTUM = 0.1        # models/hrnet.py:14


class _Residual(nn.Module):
    """Shared tail of both block types: optional 1x1 conv + BN on the identity branch (models/hrnet.py:46-52,93-100)."""

    def _identity(self, x):
        if self.downsample is None:
            return x
        if self._ds_state is None:
            self._ds_state = ops.ConvBNState(self.downsample[0])
        return ops.conv_bn(x, self.downsample[0], self.downsample[1], False, self._ds_state)

    def _unit(self, i, x, relu, residual=None):
        conv, bn = getattr(self, 'conv%d' % i), getattr(self, 'bn%d' % i)
        if self._states[i] is None:
            self._states[i] = ops.ConvBNState(conv)
        return ops.conv_bn(x, conv, bn, relu, self._states[i], residual=residual)


class BasicBlock(_Residual):
    """models/hrnet.py:27-56 — 3x3 conv/BN/ReLU, 3x3 conv/BN, + identity, ReLU."""
    expansion = 1

    def __init__(self, inplanes, planes, stride=1, downsample=None):
        super().__init__()
        self.conv1 = nn.Conv2d(inplanes, planes, 3, stride, 1, bias=False)
        self.bn1 = nn.BatchNorm2d(planes, momentum=BN_MOMENTUM)
        self.relu = nn.ReLU(inplace=True)
        self.conv2 = nn.Conv2d(planes, planes, 3, 1, 1, bias=False)
        self.bn2 = nn.BatchNorm2d(planes, momentum=BN_MOMENTUM)
        self.downsample = downsample
        self.stride = stride
        self._states, self._ds_state = [None] * 4, None

    def forward(self, x):
        out = self._unit(1, x, True)
        return self._unit(2, out, True, residual=self._identity(x))


class Bottleneck(_Residual):
    """models/hrnet.py:59-103 — 1x1, 3x3 (stride), 1x1 (x4 channels), + identity, ReLU."""
    expansion = 4

    def __init__(self, inplanes, planes, stride=1, downsample=None):
        super().__init__()
        self.conv1 = nn.Conv2d(inplanes, planes, 1, bias=False)
        self.bn1 = nn.BatchNorm2d(planes, momentum=BN_MOMENTUM)
        self.conv2 = nn.Conv2d(planes, planes, 3, stride, 1, bias=False)
        self.bn2 = nn.BatchNorm2d(planes, momentum=BN_MOMENTUM)
        self.conv3 = nn.Conv2d(planes, planes * self.expansion, 1, bias=False)
        self.bn3 = nn.BatchNorm2d(planes * self.expansion, momentum=BN_MOMENTUM)
        self.relu = nn.ReLU(inplace=True)
        self.downsample = downsample
        self.stride = stride
        self._states, self._ds_state = [None] * 4, None

    def forward(self, x):
        out = self._unit(1, x, True)
        out = self._unit(2, out, True)
        return self._unit(3, out, True, residual=self._identity(x))


class Upsample(nn.Upsample):
    """nn.Upsample(scale_factor=2**k, mode='bilinear', align_corners=False) of the fuse layers
    (models/hrnet.py:205-207) on the bilinear gather kernels."""

    def forward(self, x):
        return ops.upsample(x, int(self.scale_factor), bool(self.align_corners))


def _shortcut(inplanes, outplanes, stride):
    return nn.Sequential(nn.Conv2d(inplanes, outplanes, 1, stride, bias=False),
                         nn.BatchNorm2d(outplanes, momentum=BN_MOMENTUM))


class HRModule(nn.Module):
    """models/hrnet.py:106-251 — parallel branches of residual blocks, then every output resolution i sums all
    branches j brought to its size: identity (j == i), 1x1 ConvNormAct + bilinear up (j > i), a chain of stride-2 3x3
    ConvNormActs (j < i; the last without activation), followed by one ReLU."""

    def __init__(self, num_branches, blocks, num_blocks, num_inchannels, num_channels, multi_scale_output=True):
        super().__init__()
        for what, seq in (('NUM_BLOCKS', num_blocks), ('NUM_CHANNELS', num_channels), ('NUM_INCHANNELS', num_inchannels)):
            if num_branches != len(seq):
                raise ValueError('NUM_BRANCHES({}) <> {}({})'.format(num_branches, what, len(seq)))
        self.num_inchannels = num_inchannels
        self.num_branches = num_branches
        self.multi_scale_output = multi_scale_output
        self.branches = nn.ModuleList(self._branch(b, blocks, num_blocks[b], num_channels[b])
                                      for b in range(num_branches))
        self.fuse_layers = self._fuse() if num_branches > 1 else None
        self.relu = nn.ReLU(True)
        initialize_weights(self)

    def _branch(self, b, block, depth, planes):
        width = planes * block.expansion
        shortcut = _shortcut(self.num_inchannels[b], width, 1) if self.num_inchannels[b] != width else None
        layers = [block(self.num_inchannels[b], planes, 1, shortcut)]
        self.num_inchannels[b] = width
        layers += [block(width, planes) for _ in range(1, depth)]
        return nn.Sequential(*layers)

    def _fuse(self):
        ch = self.num_inchannels
        rows = []
        for i in range(self.num_branches if self.multi_scale_output else 1):
            row = []
            for j in range(self.num_branches):
                if j > i:
                    row.append(nn.Sequential(ConvNormAct(ch[j], ch[i], 1),
                                             Upsample(scale_factor=2 ** (j - i), mode='bilinear', align_corners=False)))
                elif j == i:
                    row.append(None)
                else:
                    steps = [nn.Sequential(ConvNormAct(ch[j], ch[j], 3, 2)) for _ in range(i - j - 1)]
                    steps.append(ConvNormAct(ch[j], ch[i], 3, 2, activate=None))
                    row.append(nn.Sequential(*steps))
            rows.append(nn.ModuleList(row))
        return nn.ModuleList(rows)

    def get_num_inchannels(self):
        return self.num_inchannels

    def forward(self, x):
        if self.num_branches == 1:
            return [self.branches[0](x[0])]
        x = [self.branches[b](x[b]) for b in range(self.num_branches)]
        fused = []
        for i, row in enumerate(self.fuse_layers):
            terms = [x[j] if j == i else row[j](x[j]) for j in range(self.num_branches)]
            fused.append(ops.sum_relu(terms))
        return fused


class HRNet(nn.Module):
    def __init__(self, num_classes=2, num_branches_list=[2, 3, 4]):
        super().__init__()
        self.inplanes = 64
        block = BasicBlock
        self.stem = nn.Sequential(ConvNormAct(3, 64, 3, 2, activate=None), ConvNormAct(64, 64, 3, 2),
                                  self._make_layer(Bottleneck, 64, 4))
        pre = [256]
        for stage, nb in zip((2, 3, 4), num_branches_list):
            channels = [32 * (2 ** i) for i in range(nb)]
            inch = [c * block.expansion for c in channels]
            setattr(self, 'transition%d' % (stage - 1), self._make_transition_layer(pre, channels))
            mod, pre = self._make_stage(nb, [4] * nb, channels, block, inch, multi_scale_output=stage != 4)
            setattr(self, 'stage%d' % stage, mod)
        self.final_layer = nn.Conv2d(pre[0], num_classes, 1)
        self.num_classes = num_classes
        self._final_packed = E.PackedWeight()

    def _make_transition_layer(self, pre, cur):
        layers = []
        for i, c in enumerate(cur):
            if i < len(pre):
                layers.append(ConvNormAct(pre[i], c, 3) if c != pre[i] else None)
            else:
                steps = i + 1 - len(pre)
                layers.append(nn.Sequential(*[ConvNormAct(pre[-1], c if s == steps - 1 else pre[-1], 3, 2)
                                              for s in range(steps)]))
        return nn.ModuleList(layers)

    def _make_layer(self, block, planes, blocks, stride=1):
        width = planes * block.expansion
        shortcut = _shortcut(self.inplanes, width, stride) if (stride != 1 or self.inplanes != width) else None
        layers = [block(self.inplanes, planes, stride, shortcut)]
        self.inplanes = width
        layers += [block(width, planes) for _ in range(1, blocks)]
        return nn.Sequential(*layers)

    def _make_stage(self, num_branches, num_blocks, num_channels, block, num_inchannels, multi_scale_output=True,
                    num_modules=1):
        mods = []
        for i in range(num_modules):
            last = i == num_modules - 1
            mods.append(HRModule(num_branches, block, num_blocks, num_inchannels, num_channels,
                                 multi_scale_output or not last))
            num_inchannels = mods[-1].get_num_inchannels()
        return nn.Sequential(*mods), num_inchannels

    def forward(self, x):
        x = self.stem(x)
        ys = [x]
        for stage in (2, 3, 4):
            trans = getattr(self, 'transition%d' % (stage - 1))
            xs = []
            for i, t in enumerate(trans):
                if t is None:
                    xs.append(ys[i])
                else:
                    # new branches grow from the lowest resolution so far; a changed width re-projects its own branch
                    xs.append(t(ys[-1] if (stage > 2 or i >= len(ys)) else ys[i]))
            ys = getattr(self, 'stage%d' % stage)(xs)
        return ops.classify(ys[0], self.final_layer, self._final_packed, 4, False)
