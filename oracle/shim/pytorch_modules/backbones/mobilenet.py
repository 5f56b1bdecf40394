from torchvision.models.mobilenetv2 import InvertedResidual  # noqa: F401
