"""Shared helpers for the golden fixtures (TEST INFRASTRUCTURE): case table, seeded inputs,
tensor summaries and their comparison.  Used by oracle/make_golden.py (writer) and tests/ (reader)."""
import os

import numpy as np
import torch

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), 'tests', 'golden')

CASES = {
    # BASELINE.json config 1: UNet 2-class 320x320 batch 4 fwd+bwd
    'unet_320': dict(kind='unet', nc=2, n=4, h=320, w=320, seed=2),
    # BASELINE.json configs 2/4 per-image shape (512x512, 21 classes); N=2 is the smallest batch
    # train-mode BN on the 1x1 pooled map accepts
    'deeplab_512': dict(kind='deeplabv3plus', nc=21, n=2, h=512, w=512, seed=3),
    'deeplab_small': dict(kind='deeplabv3plus', nc=21, n=2, h=128, w=160, seed=4),
    # BASELINE.json config 3: ASPP(2048, 256, [6,12,18]) on 32x32
    'aspp_full': dict(cin=2048, cout=256, rates=(6, 12, 18), n=2, h=32, w=32, seed=5),
    'aspp_small': dict(cin=128, cout=64, rates=(2, 4, 6), n=3, h=16, w=24, seed=6),
    # BASELINE.json config 5 architecture (HRNet 19-class, W = 2H) at a size the CPU reference finishes in seconds
    'hrnet_small': dict(kind='hrnet', nc=19, n=2, h=64, w=128, seed=8),
}

SAMPLE = 384


def make_inputs(c):
    """x = randn(N,3,H,W) seed s, targets = randint(0,nc,(N,H,W)) seed s+1."""
    gx = torch.Generator().manual_seed(c['seed'] * 100)
    gt = torch.Generator().manual_seed(c['seed'] * 100 + 1)
    x = torch.randn(c['n'], 3, c['h'], c['w'], generator=gx)
    t = torch.randint(0, c['nc'], (c['n'], c['h'], c['w']), generator=gt)
    return x, t


def small_aspp_state(c):
    """Deterministic ASPP weights under the reference's keys (models/aspp.py:25-30)."""
    g = torch.Generator().manual_seed(c['seed'] + 100)
    cin, cout, nb = c['cin'], c['cout'], len(c['rates']) + 2
    sd = {}

    def cna(prefix, ci, co, k):
        fan = co * k * k
        sd[prefix + '.0.weight'] = torch.randn(co, ci, k, k, generator=g) * (2.0 / fan) ** 0.5
        sd[prefix + '.1.weight'] = 1.0 + 0.2 * (torch.rand(co, generator=g) - 0.5)
        sd[prefix + '.1.bias'] = 0.1 * torch.randn(co, generator=g)
        sd[prefix + '.1.running_mean'] = 0.1 * torch.randn(co, generator=g)
        sd[prefix + '.1.running_var'] = 1.0 + 0.2 * (torch.rand(co, generator=g) - 0.5)
        sd[prefix + '.1.num_batches_tracked'] = torch.tensor(0)

    cna('blocks.0.gap.1', cin, cout, 1)
    cna('blocks.1', cin, cout, 1)
    for i in range(len(c['rates'])):
        cna('blocks.%d' % (i + 2), cin, cout, 3)
    cna('project', cout * nb, cout, 1)
    return sd


def _sample_index(numel):
    if numel <= SAMPLE:
        return np.arange(numel)
    return np.linspace(0, numel - 1, SAMPLE).astype(np.int64)


def summarize(name, t):
    """{name/norm, name/mean, name/sample, name/shape} for a tensor."""
    t = t.detach().float().contiguous()
    flat = t.flatten()
    idx = torch.from_numpy(_sample_index(flat.numel()))
    return {name + '/norm': np.float64(flat.double().norm()),
            name + '/mean': np.float64(flat.double().mean()),
            name + '/absmax': np.float32(flat.abs().max()),
            name + '/sample': flat[idx].numpy(),
            name + '/shape': np.array(t.shape, dtype=np.int64)}


def load(name):
    with np.load(os.path.join(GOLDEN_DIR, name + '.npz')) as z:
        return {k: z[k] for k in z.files}


def summary_error(gold, name, t):
    """Error of tensor ``t`` against the stored summary, as max|delta| / absmax(ref) over the sample
    (the per-tensor relative error of SURVEY §8c) and the relative norm difference."""
    t = t.detach().float().contiguous().cpu()
    assert list(t.shape) == list(gold[name + '/shape']), (name, t.shape, gold[name + '/shape'])
    flat = t.flatten()
    idx = torch.from_numpy(_sample_index(flat.numel()))
    ref = torch.from_numpy(gold[name + '/sample'])
    scale = max(float(gold[name + '/absmax']), 1e-30)
    samp = float((flat[idx] - ref).abs().max()) / scale
    norm = abs(float(flat.double().norm()) - float(gold[name + '/norm'])) / max(float(gold[name + '/norm']), 1e-30)
    return samp, norm
