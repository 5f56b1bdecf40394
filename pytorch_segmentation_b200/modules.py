"""``ConvNormAct`` — the reference's arithmetic leaf — on the tcgen05 path, plus the explicit
forward/backward records the model heads compose.

The reference imports ``ConvNormAct`` from the third-party ``pytorch_modules.nn`` (models/aspp.py:5,
models/deeplabv3plus.py:7, models/unet.py:7); its contract is reconstructed from the call sites
(SURVEY.md §8c): ``ConvNormAct(inplanes, planes, ksize=3, stride=1, ..., dilation=1, activate=...)`` =
``Conv2d(bias=False, padding=((k-1)/2)*dilation)`` -> ``BatchNorm2d`` -> ``ReLU``; an ``nn.Sequential``
so state-dict keys are ``<name>.0.weight`` and ``<name>.1.{weight,bias,running_mean,running_var,
num_batches_tracked}``.

Parameters stay ordinary fp32 ``nn.Parameter``s in the reference layout (OIHW); the bf16 packings the
kernels read are refreshed lazily when a parameter's version changes (optimizer step, load_state_dict).
"""
import torch
import torch.nn as nn

from . import engine as E

_DEFAULT = object()


class ConvNormAct(nn.Sequential):
    def __init__(self, inplanes, planes, ksize=3, stride=1, groups=1, dilation=1, activate=_DEFAULT):
        if stride not in (1, 2) or groups != 1 or (stride == 2 and dilation != 1):
            raise NotImplementedError(
                'ConvNormAct(stride=%r, groups=%r, dilation=%r): the sm_100a implicit-GEMM path covers dense convs of '
                'stride 1 (any dilation) and stride 2; there is no cuDNN fallback' % (stride, groups, dilation))
        if ksize not in (1, 3):
            raise NotImplementedError('ConvNormAct(ksize=%r): only 1x1 and 3x3 filters' % (ksize,))
        layers = [nn.Conv2d(inplanes, planes, ksize, stride=stride, padding=((ksize - 1) // 2) * dilation,
                            dilation=dilation, bias=False),
                  nn.BatchNorm2d(planes)]
        if activate is _DEFAULT:
            activate = nn.ReLU(inplace=True)
        if activate is not None:
            if not isinstance(activate, nn.ReLU):
                raise NotImplementedError('ConvNormAct: only ReLU (or activate=None) is fused on this path')
            layers.append(activate)
        super().__init__(*layers)
        self.relu = activate is not None
        self.ksize, self.dilation, self.stride = ksize, dilation, stride
        # stride-1 convs over >= 8 channels are composed by the heads through run_forward / run_backward_params; stride 2
        # (HRNet transitions and fuse layers, models/hrnet.py:213-224,320-324) and 3-channel stems (models/hrnet.py:260)
        # go through the op-level node in ops.py
        self._chain = stride == 1 and inplanes % 8 == 0
        self._packed = E.PackedWeight() if self._chain else None
        self._op_state = None

    # ---- parameter plumbing ---------------------------------------------------------------
    @property
    def conv(self):
        return self[0]

    @property
    def bn(self):
        return self[1]

    def _packed_weights(self):
        return [(self._packed, self[0].weight)] if self._chain else []

    def tensors(self):
        """The differentiable leaves in the order the heads pass them through autograd."""
        return [self[0].weight, self[1].weight, self[1].bias]

    # ---- explicit forward / backward over Act slices ----------------------------------------
    def run_forward(self, x, out, training, save):
        """x: Act input slice, out: Act output slice (written).  Returns the saved record (or None)."""
        conv, bn = self[0], self[1]
        k = conv.out_channels
        if k % 8:
            raise ValueError('ConvNormAct with %d output channels: need a multiple of 8' % k)
        pw = self._packed.update(conv.weight, need_dgrad=save)
        dev = x.t.device
        if training:
            if bn.momentum is None:
                raise NotImplementedError('BatchNorm2d(momentum=None) (cumulative average) is not supported')
            if x.pixels <= 1:
                raise ValueError('Expected more than 1 value per channel when training, got input size %s'
                                 % ([x.n, k, x.h, x.w],))
            raw = E.Act.empty(x.n, x.h, x.w, k, dev)
            stats = torch.zeros((2 * k,), dtype=torch.float32, device=dev)
            E.conv_fprop([(x, pw, self.dilation)], raw, stats=stats)
            track = bn.track_running_stats and bn.running_mean is not None
            vec = E.bn_finalize_lazy(stats, k, x.pixels, bn.weight, bn.bias, bn.eps, bn.momentum,
                                bn.running_mean if track else None, bn.running_var if track else None)
            if track and bn.num_batches_tracked is not None:
                bn.num_batches_tracked.add_(1)
            E.bn_apply(raw, out, vec, self.relu)
            E.capture(conv.weight, 'raw', raw)
            E.capture(conv.weight, 'out', out)
            return (x, raw, vec, True) if save else None
        vec = E.bn_eval_affine(k, bn.weight, bn.bias, bn.running_mean, bn.running_var, bn.eps)
        if not save:
            # Conv + BN + ReLU in one kernel: scale/shift/ReLU in the tcgen05 epilogue
            E.conv_fprop([(x, pw, self.dilation)], out, scale=vec[0], bias=vec[1], relu=self.relu)
            return None
        raw = E.Act.empty(x.n, x.h, x.w, k, dev)
        E.conv_fprop([(x, pw, self.dilation)], raw)
        E.bn_apply(raw, out, vec, self.relu)
        return (x, raw, vec, False)

    def run_backward_params(self, rec, dout):
        """dout: Act grad of the output slice.  Returns (draw: Act grad of the raw conv output,
        [dW, dgamma, dbeta])."""
        x, raw, vec, training = rec
        draw = E.Act.empty(raw.n, raw.h, raw.w, raw.c, raw.t.device)
        dgamma, dbeta = E.bn_backward(dout, raw, draw, vec, self.relu, training)
        E.capture(self[0].weight, 'graw', draw)          # parity tests: the stored gradient of the raw conv output
        dw = E.conv_wgrad(draw, x, self[0].out_channels, self.ksize, self.ksize, self.dilation, lane=True)
        return draw, [dw, dgamma, dbeta]

    def dgrad_src(self, draw):
        return (draw, self._packed, self.dilation)

    # ---- standalone nn.Module use ---------------------------------------------------------------
    def forward(self, x):
        if self._chain:
            return _ConvNormActFn.apply(self, x, *self.tensors())
        from . import ops
        if self._op_state is None:
            self._op_state = ops.ConvBNState(self[0])
        return ops.conv_bn(x, self[0], self[1], self.relu, self._op_state)


class _ConvNormActFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, mod, x, *params):
        xa = E.as_act(x, 'ConvNormAct input')
        need = any(ctx.needs_input_grad)
        out = E.Act.empty(xa.n, xa.h, xa.w, mod[0].out_channels, x.device)
        ctx.rec = mod.run_forward(xa, out, mod.training, need)
        ctx.mod, ctx.x = mod, x
        return out.nchw()

    @staticmethod
    def backward(ctx, dout):
        mod, rec = ctx.mod, ctx.rec
        da = E.as_act(dout, 'ConvNormAct grad')
        draw, pgrads = mod.run_backward_params(rec, da)
        dx = None
        if ctx.needs_input_grad[1]:
            g = E.Act.empty(da.n, da.h, da.w, mod[0].in_channels, dout.device)
            E.conv_dgrad([mod.dgrad_src(draw)], g)
            dx = E.grad_like(ctx.x, g)
        E.join_wgrad_lane()
        return (None, dx) + tuple(pgrads)


def initialize_weights(module):
    """Stand-in for ``pytorch_modules.utils.initialize_weights`` (models/deeplabv3plus.py:24-26,
    models/unet.py:24-25): kaiming-normal(fan_out) convs, zero conv bias, BN (1, 0).  The exact law of
    the absent third-party function is unknown; parity tests copy weights, so it does not matter."""
    for m in module.modules():
        if isinstance(m, nn.Conv2d):
            nn.init.kaiming_normal_(m.weight, mode='fan_out', nonlinearity='relu')
            if m.bias is not None:
                nn.init.zeros_(m.bias)
        elif isinstance(m, nn.BatchNorm2d):
            nn.init.ones_(m.weight)
            nn.init.zeros_(m.bias)
