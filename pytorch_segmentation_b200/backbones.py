"""Feature extractors behind the heads — stand-ins for ``pytorch_modules.backbones`` (external to the reference;
contract in SURVEY.md §8c): each returns the list of 5 feature maps the reference indexes
(models/deeplabv3plus.py:29-32, models/unet.py:28).

* ``resnet50`` (DeepLabV3+) is ``resnet.ResNet50Features``: every conv, BatchNorm, ReLU, residual join and
  the max pool run on libb200seg — no cuDNN, no autocast.
* ``mobilenet_v2`` (UNet) is NOT on the native path: its depthwise 3x3 convs and ReLU6 are outside SURVEY §8 (row 3 of §2:
  "backbone (depthwise) OUT — not named in north_star"), so it stays torchvision modules in channels_last under bf16
  autocast (cuDNN); its feature maps are the NHWC bf16 buffers the native UNet head consumes without a copy, and the
  head's input gradients flow back into it through autograd.  CUDA only, like everything else here.
* ``pretrained=True``: see ``resnet.load_pretrained`` (local checkpoint, else random init with a warning).
"""
import torch
import torch.nn as nn
import torchvision

from .resnet import find_pretrained, load_pretrained, resnet50  # noqa: F401  (re-exported: models/deeplabv3plus.py imports it from here)
from .utils.utils import model_input


class _MobileNetV2Features(nn.Module):
    """torchvision mobilenet_v2.features split so channels are 16, 24, 32, 96, 1280 at strides 2..32."""
    SPLITS = ((0, 2), (2, 4), (4, 7), (7, 14), (14, 19))

    def __init__(self, net):
        super().__init__()
        f = net.features
        self.stages = nn.ModuleList([nn.Sequential(*[f[i] for i in range(a, b)]) for a, b in self.SPLITS])

    def forward(self, x):
        if not x.is_cuda:
            raise RuntimeError('mobilenet_v2 features: input on %s — this package runs on CUDA (sm_100a) only' % x.device)
        x = model_input(x).contiguous(memory_format=torch.channels_last)
        feats = []
        with torch.autocast('cuda', dtype=torch.bfloat16):
            for s in self.stages:
                x = s(x)
                feats.append(x)
        return feats


def mobilenet_v2(pretrained=False, **kw):
    full = torchvision.models.mobilenet_v2(weights=None)
    if pretrained and find_pretrained('mobilenet_v2', 'mobilenet_v2-b0353104.pth'):
        # the torchvision checkpoint is keyed features.<i>...: load it into the unsplit net, the stages share its modules
        load_pretrained(full, 'mobilenet_v2', 'mobilenet_v2-b0353104.pth')
    return _MobileNetV2Features(full).to(memory_format=torch.channels_last)
