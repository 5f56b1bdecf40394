"""Feature extractors behind the heads — stand-ins for ``pytorch_modules.backbones`` (external to the
reference; contract in SURVEY.md §8c): each returns the list of 5 feature maps the reference indexes
(models/deeplabv3plus.py:29-32, models/unet.py:28).

SCOPE NOTE: the backbone is *not* on this round's hot path (SURVEY.md §8f-1 marks it "next").  It runs
on torchvision modules in channels_last under bf16 autocast, so its feature maps are already the NHWC
bf16 buffers the tcgen05 heads consume (zero-copy hand-off) and the heads' input gradients flow back
into it through autograd.  ``pretrained`` is ignored: there is no network, and the benchmark
configuration is random init.
"""
import torch
import torch.nn as nn
import torchvision


class _Features(nn.Module):
    def _run(self, x):
        raise NotImplementedError

    def forward(self, x):
        if x.is_cuda:
            x = x.contiguous(memory_format=torch.channels_last)
            with torch.autocast('cuda', dtype=torch.bfloat16):
                return self._run(x)
        return self._run(x)


class _ResNetFeatures(_Features):
    """[relu(bn1(conv1 x)), layer1, layer2, layer3, layer4] — strides 2, 4, 8, 16, 16 (layer4 dilated)."""

    def __init__(self, net):
        super().__init__()
        self.conv1, self.bn1, self.relu, self.maxpool = net.conv1, net.bn1, net.relu, net.maxpool
        self.layer1, self.layer2, self.layer3, self.layer4 = net.layer1, net.layer2, net.layer3, net.layer4

    def _run(self, x):
        feats = []
        x = self.relu(self.bn1(self.conv1(x)))
        feats.append(x)
        x = self.layer1(self.maxpool(x))
        feats.append(x)
        x = self.layer2(x)
        feats.append(x)
        x = self.layer3(x)
        feats.append(x)
        x = self.layer4(x)
        feats.append(x)
        return feats


def resnet50(pretrained=False, replace_stride_with_dilation=None, native=True, **kw):
    """native=True (default): the tcgen05 path (resnet.py).  native=False: the torchvision/cuDNN stand-in kept for
    A/B measurements (scripts/)."""
    if native:
        from .resnet import ResNet50Features
        return ResNet50Features(replace_stride_with_dilation)
    net = torchvision.models.resnet50(weights=None, replace_stride_with_dilation=replace_stride_with_dilation)
    return _ResNetFeatures(net).to(memory_format=torch.channels_last)


class _MobileNetV2Features(_Features):
    """torchvision mobilenet_v2.features split so channels are 16, 24, 32, 96, 1280 at strides 2..32."""
    SPLITS = ((0, 2), (2, 4), (4, 7), (7, 14), (14, 19))

    def __init__(self, net):
        super().__init__()
        f = net.features
        self.stages = nn.ModuleList([nn.Sequential(*[f[i] for i in range(a, b)]) for a, b in self.SPLITS])

    def _run(self, x):
        feats = []
        for s in self.stages:
            x = s(x)
            feats.append(x)
        return feats


def mobilenet_v2(pretrained=False, **kw):
    return _MobileNetV2Features(torchvision.models.mobilenet_v2(weights=None)).to(memory_format=torch.channels_last)
