"""pytorch_segmentation_b200 — the data-parallel hot path of WoodsGao/pytorch_segmentation, rebuilt for
B200 (sm_100a): tcgen05 implicit-GEMM convolutions and 128-bit bandwidth kernels behind the reference's
model / loss / metric API.  See DESIGN.md and INTEGRATION.md at the repository root.
"""
from . import _lib
from .evaluate import test
from .models import ASPP, ASPPPooling, DeepLabV3Plus, HRNet, UNet
from .modules import ConvNormAct, initialize_weights
from . import optim
from .trainer import TrainStep
from .utils.utils import compute_loss, compute_metrics

__all__ = ['ASPP', 'ASPPPooling', 'DeepLabV3Plus', 'HRNet', 'UNet', 'ConvNormAct', 'initialize_weights', 'compute_loss',
           'compute_metrics', 'test', 'TrainStep', 'optim']
