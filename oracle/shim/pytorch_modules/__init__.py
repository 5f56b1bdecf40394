"""Stand-in for the third-party PyPI package ``pytorch-modules`` (woodsgao/pytorch_modules).

TEST INFRASTRUCTURE ONLY.  The reference (``/root/reference/requirements.txt:4`` pins
``pytorch-modules>=0.3.0``) imports its conv building block, backbones and trainer from that
package; it is not vendored, not installed, and cannot be fetched (no network).  This shim
restates only what the reference's call sites constrain:

* ``nn.ConvNormAct(in, out, ksize=3, stride=1, dilation=1, activate=ReLU)`` as
  ``nn.Sequential(Conv2d(bias=False, "same" padding), BatchNorm2d, act)``
  (call sites: models/aspp.py:12,27,29,30; models/deeplabv3plus.py:20; models/unet.py:19-21;
  models/hrnet.py:200-224,260-261).
* ``backbones.resnet50`` / ``mobilenet_v2`` returning the list of 5 feature maps the reference
  indexes (models/deeplabv3plus.py:29-32, models/unet.py:28), torchvision-backed, random init.
* ``utils.initialize_weights`` (models/deeplabv3plus.py:24-26, models/unet.py:24-25).

Nothing under ``pytorch_segmentation_b200/`` imports this.
"""
