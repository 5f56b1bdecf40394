"""``pytorch_modules.backbones`` stand-in: torchvision nets returning 5 feature maps."""
import torch.nn as nn
import torchvision

from . import mobilenet  # noqa: F401  (models/unet.py:8 imports backbones.mobilenet.InvertedResidual)


class _ResNetFeatures(nn.Module):
    """[relu(bn1(conv1 x)), layer1, layer2, layer3, layer4] — strides 2,4,8,16,(32|16)."""

    def __init__(self, net):
        super().__init__()
        self.conv1, self.bn1, self.relu, self.maxpool = net.conv1, net.bn1, net.relu, net.maxpool
        self.layer1, self.layer2, self.layer3, self.layer4 = net.layer1, net.layer2, net.layer3, net.layer4

    def forward(self, x):
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


def resnet50(pretrained=False, replace_stride_with_dilation=None, **kw):
    # pretrained is ignored: no network, and BASELINE.json asks for random init.
    return _ResNetFeatures(torchvision.models.resnet50(
        weights=None, replace_stride_with_dilation=replace_stride_with_dilation))


def resnet34(pretrained=False, **kw):
    return _ResNetFeatures(torchvision.models.resnet34(weights=None))


class _MobileNetV2Features(nn.Module):
    """torchvision mobilenet_v2.features split so channels are 16,24,32,96,1280 at strides 2..32."""

    SPLITS = ((0, 2), (2, 4), (4, 7), (7, 14), (14, 19))

    def __init__(self, net):
        super().__init__()
        f = net.features
        self.stages = nn.ModuleList([nn.Sequential(*[f[i] for i in range(a, b)]) for a, b in self.SPLITS])

    def forward(self, x):
        feats = []
        for s in self.stages:
            x = s(x)
            feats.append(x)
        return feats


def mobilenet_v2(pretrained=False, **kw):
    return _MobileNetV2Features(torchvision.models.mobilenet_v2(weights=None))
