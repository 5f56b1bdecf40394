"""ASPP on the tcgen05 path — same constructor, forward and state-dict keys as models/aspp.py.

Differences in *how* (not what): the five branches write their BN+ReLU result straight into channel
slices of one NHWC buffer (no torch.cat, models/aspp.py:36); the image-pooling branch's bilinear
1x1 -> HxW (align_corners=False, models/aspp.py:16-19) is a broadcast; in backward the gradients of
the four conv branches w.r.t. the shared input are summed inside one implicit-GEMM launch (TMEM
accumulation over sources) and the pooled branch adds its 1/(H*W) broadcast on top.
"""
import torch
import torch.nn as nn

from .. import engine as E
from ..modules import ConvNormAct


class ASPPPooling(nn.Module):
    def __init__(self, inplanes, planes):
        super().__init__()
        # models/aspp.py:11-12 — index 0 keeps the reference's key layout (gap.1.0.weight, gap.1.1.*)
        self.gap = nn.Sequential(nn.AdaptiveAvgPool2d(1), ConvNormAct(inplanes, planes, 1))

    def forward(self, x):
        return _PoolFn.apply(self, x, *self.gap[1].tensors())

    # explicit records used by ASPP.  The whole branch after the pool is fp32 on CUDA cores
    # (csrc/pool_branch.cu): N = batch rows, reads the fp32 master weight directly.
    def run_forward(self, x, out, training, save):
        cna = self.gap[1]
        conv, bn = cna[0], cna[1]
        if training:
            if bn.momentum is None:
                raise NotImplementedError('BatchNorm2d(momentum=None) (cumulative average) is not supported')
            if x.n <= 1:
                raise ValueError('Expected more than 1 value per channel when training, got input size %s'
                                 % ([x.n, conv.out_channels, 1, 1],))
        g = E.gap_fwd(x)                                        # fp32 [n, cin]
        raw = E.linear_f32_fwd(g, conv.weight)                  # fp32 [n, planes]
        z, bn_save = E.bn_rows_fwd(raw, bn, training, cna.relu)
        if training and bn.track_running_stats and bn.num_batches_tracked is not None:
            bn.num_batches_tracked.add_(1)
        E.broadcast_hw(z, out)
        E.capture(conv.weight, 'raw', raw)
        E.capture(conv.weight, 'out', z)
        return (g, raw, bn_save, training, x.h * x.w) if save else None

    def run_backward(self, rec, dout, dx, accumulate):
        """dout: grad slice of the broadcast output; adds (or writes) the branch's input grad into dx."""
        g, raw, bn_save, training, hw = rec
        cna = self.gap[1]
        conv, bn = cna[0], cna[1]
        dz = E.sum_hw(dout)                                     # fp32 [n, planes]
        draw, dgamma, dbeta = E.bn_rows_bwd(dz, raw, bn, bn_save, training, cna.relu)
        dg, dw = E.linear_f32_bwd(draw, g, conv.weight, need_dx=dx is not None)
        if dx is not None:
            E.broadcast_hw(dg, dx, scale=1.0 / hw, accumulate=accumulate)
        return [dw, dgamma, dbeta]


class _PoolFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, mod, x, *params):
        xa = E.as_act(x, 'ASPPPooling input')
        need = any(ctx.needs_input_grad)
        out = E.Act.empty(xa.n, xa.h, xa.w, mod.gap[1][0].out_channels, x.device)
        ctx.rec = mod.run_forward(xa, out, mod.training, need)
        ctx.mod, ctx.x = mod, x
        return out.nchw()

    @staticmethod
    def backward(ctx, dout):
        da = E.as_act(dout, 'ASPPPooling grad')
        dx = None
        g = None
        if ctx.needs_input_grad[1]:
            g = E.Act.empty(da.n, da.h, da.w, ctx.mod.gap[1][0].in_channels, dout.device)
        pgrads = ctx.mod.run_backward(ctx.rec, da, g, accumulate=False)
        if g is not None:
            dx = E.grad_like(ctx.x, g)
        return (None, dx) + tuple(pgrads)


class ASPP(nn.Module):
    def __init__(self, inplanes, planes, atrous_rates=[12, 24, 36]):
        super().__init__()
        self.blocks = nn.ModuleList([ASPPPooling(inplanes, planes), ConvNormAct(inplanes, planes, 1)])
        for rate in atrous_rates:
            self.blocks.append(ConvNormAct(inplanes, planes, dilation=rate))
        self.project = ConvNormAct(planes * len(self.blocks), planes, 1)
        self.inplanes, self.planes = inplanes, planes

    def tensors(self):
        ts = list(self.blocks[0].gap[1].tensors())
        for b in list(self.blocks)[1:]:
            ts += b.tensors()
        return ts + self.project.tensors()

    def forward(self, x):
        return _ASPPFn.apply(self, x, *self.tensors())

    # ---- explicit records ---------------------------------------------------------------------
    def run_forward(self, x, out, training, save):
        nb, planes = len(self.blocks), self.planes
        cat = E.Act.empty(x.n, x.h, x.w, planes * nb, x.t.device)
        recs = [self.blocks[0].run_forward(x, cat.slice(0, planes), training, save)]
        for i in range(1, nb):
            recs.append(self.blocks[i].run_forward(x, cat.slice(i * planes, planes), training, save))
        prec = self.project.run_forward(cat, out, training, save)
        return (recs, prec) if save else None

    def run_backward(self, rec, dout, dx):
        """dout: grad of the ASPP output; dx: Act to receive the grad of the input (or None).
        Returns param grads in ``tensors()`` order."""
        recs, prec = rec
        nb, planes = len(self.blocks), self.planes
        dproj_raw, proj_grads = self.project.run_backward_params(prec, dout)
        dcat = E.Act.empty(dout.n, dout.h, dout.w, planes * nb, dout.t.device)
        E.conv_dgrad([self.project.dgrad_src(dproj_raw)], dcat)
        grads = []
        srcs = []
        for i in range(1, nb):
            draw, g = self.blocks[i].run_backward_params(recs[i], dcat.slice(i * planes, planes))
            grads += g
            srcs.append(self.blocks[i].dgrad_src(draw))
        if dx is not None:
            # sum of the conv branches' data gradients: TMEM accumulation over <= 4 sources per launch
            first = True
            for j in range(0, len(srcs), E._lib.MAX_CONV_SRC):
                chunk = srcs[j:j + E._lib.MAX_CONV_SRC]
                if first:
                    E.conv_dgrad(chunk, dx)
                    first = False
                else:
                    tmp = E.Act.empty(dx.n, dx.h, dx.w, dx.c, dx.t.device)
                    E.conv_dgrad(chunk, tmp)
                    dx.nchw().add_(tmp.nchw())   # only reached with > 4 conv branches (never in the reference)
        pool_grads = self.blocks[0].run_backward(recs[0], dcat.slice(0, planes), dx, accumulate=True)
        return pool_grads + grads + proj_grads


class _ASPPFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, mod, x, *params):
        xa = E.as_act(x, 'ASPP input')
        need = any(ctx.needs_input_grad)
        out = E.Act.empty(xa.n, xa.h, xa.w, mod.planes, x.device)
        ctx.rec = mod.run_forward(xa, out, mod.training, need)
        ctx.mod, ctx.x = mod, x
        return out.nchw()

    @staticmethod
    def backward(ctx, dout):
        da = E.as_act(dout, 'ASPP grad')
        g = E.Act.empty(da.n, da.h, da.w, ctx.mod.inplanes, dout.device)
        pgrads = ctx.mod.run_backward(ctx.rec, da, g)
        dx = E.grad_like(ctx.x, g) if ctx.needs_input_grad[1] else None
        E.join_wgrad_lane()
        return (None, dx) + tuple(pgrads)
