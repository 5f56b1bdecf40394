"""Import the UNMODIFIED reference from /root/reference (build container only).

TEST INFRASTRUCTURE.  /root/reference does not exist on the GPU box, so nothing in the ``-m gpu``
tests, ``smoke()`` or ``bench.py`` may call this at run time; it is used by ``make_golden.py`` and by
the CPU tests that validate ``oracle/restatement.py`` (skipped when the reference is absent).

Two stand-ins are needed: the ``pytorch_modules`` shim on sys.path, and
``imgaug`` stubbed in sys.modules because utils/datasets.py:7,11-13 imports it at module level.
"""
import importlib
import os
import sys
import types
from unittest import mock

REFERENCE_ROOT = os.environ.get('B200SEG_REFERENCE_ROOT', '/root/reference')
_SHIM = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'shim')


def available():
    return os.path.isdir(os.path.join(REFERENCE_ROOT, 'models'))


def load():
    """Returns a namespace with the reference's DeepLabV3Plus, UNet, ASPP, compute_loss,
    compute_metrics and test() — the real code objects from /root/reference."""
    if not available():
        raise RuntimeError('reference tree not present at %s' % REFERENCE_ROOT)
    for name in ('imgaug', 'imgaug.augmenters', 'imgaug.augmentables', 'imgaug.augmentables.segmaps',
                 'imgaug.augmentables.polys'):
        sys.modules.setdefault(name, mock.MagicMock())
    for p in (_SHIM, REFERENCE_ROOT):
        if p not in sys.path:
            sys.path.insert(0, p)
    ns = types.SimpleNamespace()
    models = importlib.import_module('models')
    aspp = importlib.import_module('models.aspp')
    utils = importlib.import_module('utils.utils')
    test_mod = importlib.import_module('test')
    ns.DeepLabV3Plus, ns.UNet, ns.HRNet = models.DeepLabV3Plus, models.UNet, models.HRNet
    ns.ASPP, ns.ASPPPooling = aspp.ASPP, aspp.ASPPPooling
    ns.compute_loss, ns.compute_metrics = utils.compute_loss, utils.compute_metrics
    ns.test = test_mod.test
    return ns
