"""Drop-in for the reference's ``models`` package (models/__init__.py:1-3) on the sm_100a path."""
from .aspp import ASPP, ASPPPooling
from .deeplabv3plus import DeepLabV3Plus
from .hrnet import HRNet
from .unet import UNet

__all__ = ['ASPP', 'ASPPPooling', 'DeepLabV3Plus', 'HRNet', 'UNet']
