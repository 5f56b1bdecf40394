"""Optimizer step of the hot path: ``SGD``, ``Adam`` and ``AdamW`` with the constructors, state-dict
layouts and update rules of their ``torch.optim`` namesakes — what the reference's external ``Trainer`` builds for
train.py:61-68 (``--adam`` off / on, train.py:96) — whose ``step()`` is one or two multi-tensor launches
(``b200seg_sgd_multi`` / ``b200seg_adam_multi``) over all parameters instead of a dozen foreach kernels.

State layout is PyTorch's (``state[p]['momentum_buffer']``; ``'step'``, ``'exp_avg'``, ``'exp_avg_sq'``), so checkpoints
written by ``optimizer.state_dict()`` load into the torch optimizers and back.  CUDA fp32 parameters only: anything else
raises (no fallback).

Hyper-parameters under a CUDA graph: the kernels read learning rate, momentum / betas, weight decay, gradient scale and
Adam's bias corrections from a small DEVICE vector that ``step()`` refreshes from ``param_groups`` through a pinned host
mirror.  Captured into a graph (trainer.TrainStep), that refresh is a memcpy node which re-reads the pinned mirror on
every replay — ``refresh_hyper()`` (called by TrainStep before each replay) rewrites the mirror, so an LR scheduler or a
manual ``param_groups[i]['lr'] = ...`` takes effect on the next replay, and Adam's step count keeps advancing.
"""
import ctypes
import math

import torch

from . import _lib
from . import engine as _engine


def _dense(p):
    # the update is elementwise over the storage: any dense layout works (contiguous, or the channels_last conv weights
    # of a torchvision backbone) as long as p, g and the state buffers share it
    return p.is_contiguous() or (p.dim() == 4 and p.is_contiguous(memory_format=torch.channels_last))


def _check_param(name, p, g):
    if not p.is_cuda or p.dtype != torch.float32 or not _dense(p):
        raise RuntimeError('pytorch_segmentation_b200.optim.%s: parameters must be dense fp32 CUDA tensors (got %s %s '
                           'strides %s); there is no CPU fallback' % (name, p.device, p.dtype, tuple(p.stride())))
    if g.is_sparse or g.dtype != torch.float32:
        raise RuntimeError('pytorch_segmentation_b200.optim.%s: gradients must be dense fp32' % name)
    return g if g.stride() == p.stride() else torch.empty_like(p).copy_(g)


class _HyperMixin:
    """Per param-group device vector of hyper-parameters + its pinned host mirror."""
    HYPER_LEN = 12

    def _hyper(self, gi, device):
        store = self.__dict__.setdefault('_hyper_store', {})
        h = store.get(gi)
        if h is None or h[1].device != device:
            host = torch.zeros(self.HYPER_LEN, dtype=torch.float32).pin_memory()
            h = store[gi] = (host, torch.zeros(self.HYPER_LEN, dtype=torch.float32, device=device))
        return h

    def _upload(self, gi, device, values):
        host, dev = self._hyper(gi, device)
        host[:len(values)] = torch.tensor(values, dtype=torch.float32)
        dev.copy_(host, non_blocking=True)       # in a captured graph: a memcpy node that re-reads the pinned mirror
        return dev

    def refresh_hyper(self):
        """Rewrite the pinned mirrors from ``param_groups`` (and advance Adam's step count) WITHOUT launching anything:
        what a CUDA-graph replay of a captured ``step()`` needs done on the host first."""
        for gi, group in enumerate(self.param_groups):
            store = self.__dict__.get('_hyper_store', {})
            if gi in store:
                vals = self._hyper_values(group, advance=True)
                store[gi][0][:len(vals)] = torch.tensor(vals, dtype=torch.float32)


class SGD(_HyperMixin, torch.optim.Optimizer):
    def __init__(self, params, lr=1e-3, momentum=0.0, dampening=0.0, weight_decay=0.0, nesterov=False, grad_scale=1.0):
        if lr < 0.0:
            raise ValueError('Invalid learning rate: %r' % (lr,))
        if momentum < 0.0:
            raise ValueError('Invalid momentum value: %r' % (momentum,))
        if weight_decay < 0.0:
            raise ValueError('Invalid weight_decay value: %r' % (weight_decay,))
        if nesterov and (momentum <= 0 or dampening != 0):
            raise ValueError('Nesterov momentum requires a momentum and zero dampening')
        defaults = dict(lr=lr, momentum=momentum, dampening=dampening, weight_decay=weight_decay, nesterov=nesterov,
                        grad_scale=grad_scale)
        super().__init__(params, defaults)

    def _hyper_values(self, group, advance=False):
        return [float(group['lr']), float(group['momentum']), float(group['dampening']), float(group['weight_decay']),
                float(group.get('grad_scale', 1.0))]

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        lib = _lib.load()
        stream = _engine._stream()
        for gi, group in enumerate(self.param_groups):
            momentum = float(group['momentum'])
            with_buf, without_buf = [], []
            for p in group['params']:
                g = p.grad
                if g is None:
                    continue
                g = _check_param('SGD', p, g)
                if momentum != 0.0:
                    st = self.state[p]
                    buf = st.get('momentum_buffer')
                    if buf is None:
                        st['momentum_buffer'] = buf = torch.empty_like(p)          # preserves p's layout
                        without_buf.append((p, g, buf))
                    else:
                        with_buf.append((p, g, buf))
                else:
                    with_buf.append((p, g, None))
            if not with_buf and not without_buf:
                continue
            hyper = self._upload(gi, (with_buf or without_buf)[0][0].device, self._hyper_values(group))
            # parameters seeing their first step initialise the buffer with the gradient (torch.optim.SGD semantics)
            for items, first in ((with_buf, 0), (without_buf, 1)):
                if not items:
                    continue
                arr = (_lib.SgdDesc * len(items))()
                for i, (p, g, buf) in enumerate(items):
                    arr[i] = _lib.SgdDesc(p.data_ptr(), g.data_ptr(), buf.data_ptr() if buf is not None else None, p.numel())
                _lib.check(lib.b200seg_sgd_multi(arr, len(items), float(group['lr']), momentum, float(group['dampening']),
                                                 float(group['weight_decay']), int(bool(group['nesterov'])), first,
                                                 float(group.get('grad_scale', 1.0)), hyper.data_ptr(), stream), 'sgd_multi')
                _lib.count_launch((len(items) + 95) // 96)
                # the kernel wrote the parameters behind autograd's back: move their version counters so the packed
                # bf16 weight copies (engine.PackedWeight) see the change
                torch.autograd.graph.increment_version([p for p, _, _ in items])
        return loss


class Adam(_HyperMixin, torch.optim.Optimizer):
    """torch.optim.Adam (L2 weight decay added to the gradient); ``AdamW`` below decouples it.  amsgrad / maximize /
    capturable variants are not offered (the reference's Trainer exposes a single ``adam`` switch, train.py:96)."""
    _DECOUPLED = False
    _NAME = 'Adam'

    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.0, grad_scale=1.0):
        if lr < 0.0:
            raise ValueError('Invalid learning rate: %r' % (lr,))
        if eps < 0.0:
            raise ValueError('Invalid epsilon value: %r' % (eps,))
        if not 0.0 <= betas[0] < 1.0 or not 0.0 <= betas[1] < 1.0:
            raise ValueError('Invalid beta parameters: %r' % (betas,))
        if weight_decay < 0.0:
            raise ValueError('Invalid weight_decay value: %r' % (weight_decay,))
        super().__init__(params, dict(lr=lr, betas=tuple(betas), eps=eps, weight_decay=weight_decay, grad_scale=grad_scale))

    def _hyper_values(self, group, advance=False):
        # one step count per group (torch keeps one per parameter, all equal when every parameter has a gradient)
        if advance:
            group['_b200seg_step'] = group.get('_b200seg_step', 0) + 1
            for p in group['params']:
                st = self.state.get(p)
                if st and 'step' in st:
                    st['step'] += 1
        t = max(int(group.get('_b200seg_step', 0)), 1)
        b1, b2 = group['betas']
        lr, wd = float(group['lr']), float(group['weight_decay'])
        # derived scalars in double precision, as torch.optim forms them in Python before they meet an fp32 tensor
        return [lr, float(b1), float(b2), float(group['eps']), wd, float(group.get('grad_scale', 1.0)),
                lr / (1.0 - b1 ** t), math.sqrt(1.0 - b2 ** t), 1.0 - b1, 1.0 - b2, 1.0 - lr * wd]

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        lib = _lib.load()
        stream = _engine._stream()
        for gi, group in enumerate(self.param_groups):
            old, new = [], []
            for p in group['params']:
                g = p.grad
                if g is None:
                    continue
                g = _check_param(self._NAME, p, g)
                st = self.state[p]
                if 'exp_avg' not in st:
                    st['step'] = torch.tensor(0.0)
                    st['exp_avg'] = torch.empty_like(p)
                    st['exp_avg_sq'] = torch.empty_like(p)
                    new.append((p, g, st))
                else:
                    old.append((p, g, st))
            if not old and not new:
                continue
            hyper = self._upload(gi, (old or new)[0][0].device, self._hyper_values(group, advance=True))
            for items, first in ((old, 0), (new, 1)):
                if not items:
                    continue
                arr = (_lib.AdamDesc * len(items))()
                for i, (p, g, st) in enumerate(items):
                    arr[i] = _lib.AdamDesc(p.data_ptr(), g.data_ptr(), st['exp_avg'].data_ptr(), st['exp_avg_sq'].data_ptr(),
                                           p.numel())
                _lib.check(lib.b200seg_adam_multi(arr, len(items), hyper.data_ptr(), int(self._DECOUPLED), first, stream),
                           'adam_multi')
                _lib.count_launch((len(items) + 63) // 64)
                torch.autograd.graph.increment_version([p for p, _, _ in items])
        return loss


class AdamW(Adam):
    """torch.optim.AdamW: ``p *= 1 - lr * weight_decay`` ahead of the Adam update (default weight_decay 1e-2)."""
    _DECOUPLED = True
    _NAME = 'AdamW'

    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=1e-2, grad_scale=1.0):
        super().__init__(params, lr=lr, betas=betas, eps=eps, weight_decay=weight_decay, grad_scale=grad_scale)
