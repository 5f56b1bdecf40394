"""``pytorch_modules.nn`` stand-in: ConvNormAct and the names the reference imports beside it."""
import torch
import torch.nn as nn


def _same_pad(ksize, dilation):
    return ((ksize - 1) // 2) * dilation


class ConvNormAct(nn.Sequential):
    """conv(bias=False, pad=((k-1)/2)*dil) -> BatchNorm2d -> activation.

    Index 0 = conv, 1 = bn, 2 = act, so state_dict keys are ``<name>.0.weight`` and
    ``<name>.1.{weight,bias,running_mean,running_var,num_batches_tracked}``.
    ``activate=None`` drops the activation (models/hrnet.py:217,260).
    """

    _DEFAULT = object()

    def __init__(self, inplanes, planes, ksize=3, stride=1, groups=1, dilation=1,
                 activate=_DEFAULT):
        layers = [
            nn.Conv2d(inplanes, planes, ksize, stride=stride, padding=_same_pad(ksize, dilation),
                      dilation=dilation, groups=groups, bias=False),
            nn.BatchNorm2d(planes),
        ]
        if activate is ConvNormAct._DEFAULT:
            activate = nn.ReLU(inplace=True)
        if activate is not None:
            layers.append(activate)
        super().__init__(*layers)


class SeparableConvNormAct(nn.Sequential):
    """Imported (never instantiated) by models/aspp.py:5."""

    def __init__(self, inplanes, planes, ksize=3, stride=1, dilation=1, activate=ConvNormAct._DEFAULT):
        super().__init__(
            ConvNormAct(inplanes, inplanes, ksize, stride, groups=inplanes, dilation=dilation,
                        activate=None),
            ConvNormAct(inplanes, planes, 1, activate=activate))


class FocalBCELoss(nn.Module):
    """Instantiated at import time by utils/utils.py:14, never called on the hot path."""

    def __init__(self, alpha=0.25, gamma=2.0):
        super().__init__()
        self.alpha, self.gamma = alpha, gamma

    def forward(self, logits, targets):
        bce = nn.functional.binary_cross_entropy_with_logits(logits, targets, reduction='none')
        p = torch.exp(-bce)
        return (self.alpha * (1 - p) ** self.gamma * bce).mean()
