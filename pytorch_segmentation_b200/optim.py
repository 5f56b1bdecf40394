"""Optimizer step of the hot path: ``SGD`` with ``torch.optim.SGD``'s constructor, state-dict layout
and update rule — the optimizer the reference's external ``Trainer`` builds for train.py:61-68 (``--adam`` off) — whose
``step()`` is one or two ``b200seg_sgd_multi`` launches over all parameters instead of ~11 foreach kernels.

State layout is PyTorch's (``state[p]['momentum_buffer']``), so checkpoints written by ``optimizer.state_dict()``
load into ``torch.optim.SGD`` and back.  CUDA fp32 parameters only: anything else raises (no fallback).
"""
import ctypes

import torch

from . import _lib


class SGD(torch.optim.Optimizer):
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

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        lib = _lib.load()
        stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
        for group in self.param_groups:
            momentum = float(group['momentum'])
            with_buf, without_buf = [], []
            for p in group['params']:
                g = p.grad
                if g is None:
                    continue
                # the update is elementwise over the storage: any dense layout works (contiguous, or the channels_last
                # conv weights of a torchvision backbone) as long as p, g and the momentum buffer share it
                dense = p.is_contiguous() or (p.dim() == 4 and p.is_contiguous(memory_format=torch.channels_last))
                if not p.is_cuda or p.dtype != torch.float32 or not dense:
                    raise RuntimeError('pytorch_segmentation_b200.optim.SGD: parameters must be dense fp32 CUDA '
                                       'tensors (got %s %s strides %s); there is no CPU fallback'
                                       % (p.device, p.dtype, tuple(p.stride())))
                if g.is_sparse or g.dtype != torch.float32:
                    raise RuntimeError('pytorch_segmentation_b200.optim.SGD: gradients must be dense fp32')
                if g.stride() != p.stride():
                    g = torch.empty_like(p).copy_(g)
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
            # parameters seeing their first step initialise the buffer with the gradient (torch.optim.SGD semantics)
            for items, first in ((with_buf, 0), (without_buf, 1)):
                if not items:
                    continue
                arr = (_lib.SgdDesc * len(items))()
                for i, (p, g, buf) in enumerate(items):
                    arr[i] = _lib.SgdDesc(p.data_ptr(), g.data_ptr(), buf.data_ptr() if buf is not None else None, p.numel())
                _lib.check(lib.b200seg_sgd_multi(arr, len(items), float(group['lr']), momentum, float(group['dampening']),
                                                 float(group['weight_decay']), int(bool(group['nesterov'])), first,
                                                 float(group.get('grad_scale', 1.0)), stream), 'sgd_multi')
                _lib.count_launch((len(items) + 95) // 96)
                # the kernel wrote the parameters behind autograd's back: move their version counters so the packed
                # bf16 weight copies (engine.PackedWeight) see the change
                torch.autograd.graph.increment_version([p for p, _, _ in items])
        return loss
