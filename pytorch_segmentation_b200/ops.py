"""Op-level autograd nodes over NHWC bf16 buffers, for networks whose topology is a general DAG (HRNet's
multi-resolution branches, models/hrnet.py) rather than the fixed chains of DeepLabV3+/UNet, whose heads are single
hand-sequenced nodes.  Every node's forward and backward is a short sequence of libb200seg kernels; tensors that travel
between nodes are NCHW-shaped views of contiguous NHWC bf16 buffers (zero-copy in and out of ``engine.Act``).

  conv_bn(x, conv, bn, relu, residual)   Conv2d(no bias) -> BatchNorm2d [-> + residual] [-> ReLU]
                                         (ConvNormAct; the conv/bn pairs and the `out += residual; relu` tail of
                                         BasicBlock / Bottleneck, models/hrnet.py:27-56,59-103; stride 1 or 2, 1x1 or 3x3,
                                         3-channel input through the patch gather)
  upsample(x, scale, align_corners)      nn.Upsample(bilinear) of the fuse layers (models/hrnet.py:205-207)
  sum_relu(xs)                           relu(x0 + x1 + ...) of HRModule.forward (models/hrnet.py:240-249)
  classify(x, conv, scale, ac)           nn.Conv2d(bias) + F.interpolate(bilinear) -> fp32 NCHW logits
                                         (models/hrnet.py:400-404)
"""
import torch
import torch.nn as nn

from . import engine as E


class ConvBNState:
    """Per (conv, bn) pair: the packed-weight cache.  Lives on the owning module (not an nn.Module itself)."""

    def __init__(self, conv):
        k = conv.kernel_size[0]
        if conv.kernel_size not in ((1, 1), (3, 3)) or conv.groups != 1 or conv.bias is not None:
            raise NotImplementedError('%r is outside the tcgen05 path (1x1 / 3x3, dense, no bias)' % (conv,))
        if conv.stride not in ((1, 1), (2, 2)) or conv.padding[0] != (k // 2) * conv.dilation[0]:
            raise NotImplementedError('%r: stride / padding outside the tcgen05 path' % (conv,))
        if conv.stride == (2, 2) and conv.dilation != (1, 1):
            raise NotImplementedError('%r: stride 2 with dilation' % (conv,))
        # few-channel stem: stride 2 takes the space-to-depth form, stride 1 the patch gather + 1x1 GEMM
        self.gather = conv.in_channels % 8 != 0
        self.s2d = self.gather and E.PackedS2DWeight.applies(conv)
        self.packed = (E.PackedS2DWeight() if self.s2d else E.PackedIm2colWeight()) if self.gather else E.PackedWeight()


class _ConvBNFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight, gamma, beta, residual, conv, bn, relu, state):
        k, stride, dil = conv.kernel_size[0], conv.stride[0], conv.dilation[0]
        cout = conv.out_channels
        if cout % 8:
            raise ValueError('conv with %d output channels: need a multiple of 8' % cout)
        training = bn.training
        need_dx = ctx.needs_input_grad[0]
        save = any(ctx.needs_input_grad)
        dev = x.device
        E._require_cuda(x, 'conv_bn input')
        if state.gather:
            xf = x.detach()
            norm = getattr(x, '_b200seg_norm', (None, None))
            u8 = xf.dtype == torch.uint8        # utils.post_fetch_fn: 1-byte pixels, normalised by the kernel that reads them
            if u8 and not state.s2d:
                xf = E.normalize_u8(xf, *norm)
            elif not u8 and (xf.dtype != torch.float32 or not xf.is_contiguous()):
                xf = xf.float().contiguous()
            pw = state.packed.update(weight)
            if state.s2d and u8:
                xa, oh, ow = E.space_to_depth_u8(xf, k, k, conv.padding[0], *norm)
            elif state.s2d:
                xa, oh, ow = E.space_to_depth(xf, k, k, conv.padding[0])
            else:
                xa = E.im2col(xf, k, k, stride, conv.padding[0])
                oh, ow = xa.h, xa.w
            src = (xa, pw, 1, 1)
        else:
            xa = E.as_act(x, 'conv_bn input')
            pw = state.packed.update(weight, need_dgrad=need_dx)
            src = (xa, pw, dil, stride)
        if not state.gather:
            oh, ow = xa.h // stride, xa.w // stride
        res = E.as_act(residual, 'conv_bn residual') if residual is not None else None
        out = E.Act.empty(xa.n, oh, ow, cout, dev)
        track = bn.track_running_stats and bn.running_mean is not None
        if training:
            if bn.momentum is None:
                raise NotImplementedError('BatchNorm2d(momentum=None) is not supported')
            raw = E.Act.empty(xa.n, oh, ow, cout, dev)
            if raw.pixels <= 1:
                raise ValueError('Expected more than 1 value per channel when training, got input size %s'
                                 % ([xa.n, cout, oh, ow],))
            stats = torch.zeros((2 * cout,), dtype=torch.float32, device=dev)
            E.conv_fprop([src], raw, stats=stats)
            vec = E.bn_finalize_lazy(stats, cout, raw.pixels, gamma, beta, bn.eps, bn.momentum,
                                bn.running_mean if track else None, bn.running_var if track else None)
            if track and bn.num_batches_tracked is not None:
                bn.num_batches_tracked.add_(1)
            E.bn_apply(raw, out, vec, relu, res=res)
            E.capture(conv.weight, 'raw', raw)
            E.capture(conv.weight, 'out', out)
        else:
            if not track:
                raise RuntimeError('eval-mode BatchNorm without running statistics')
            vec = E.bn_eval_affine(cout, gamma, beta, bn.running_mean, bn.running_var, bn.eps)
            if not save and res is None:
                E.conv_fprop([src], out, scale=vec[0], bias=vec[1], relu=relu)      # Conv+BN+ReLU in one kernel
                raw = None
            else:
                raw = E.Act.empty(xa.n, oh, ow, cout, dev)
                E.conv_fprop([src], raw)
                E.bn_apply(raw, out, vec, relu, res=res)
        if save:
            ctx.rec = (xa, raw, vec, out if (res is not None and relu) else None, pw)
            ctx.cfg = (conv, relu, training, state, k, stride, dil, res is not None, x, residual)
        return out.nchw()

    @staticmethod
    def backward(ctx, dout):
        xa, raw, vec, out_saved, pw = ctx.rec
        conv, relu, training, state, k, stride, dil, has_res, x, residual = ctx.cfg
        dev = dout.device
        da = E.as_act(dout, 'conv_bn grad')
        dres = None
        if has_res:
            # out = relu?(bn(raw) + res): the gradient reaching both branches is dout * (out > 0)
            if relu:
                dz = E.Act.empty(da.n, da.h, da.w, da.c, dev)
                E.relu_bwd(da, out_saved, dz)
            else:
                dz = da
            if ctx.needs_input_grad[4]:
                dres = E.grad_like(residual, dz)
            draw = E.Act.empty(raw.n, raw.h, raw.w, raw.c, dev)
            dgamma, dbeta = E.bn_backward(dz, raw, draw, vec, False, training)
        else:
            draw = E.Act.empty(raw.n, raw.h, raw.w, raw.c, dev)
            dgamma, dbeta = E.bn_backward(da, raw, draw, vec, relu, training)
        E.capture(conv.weight, 'graw', draw)             # parity tests: the stored gradient of the raw conv output
        if state.gather:
            dwcol = E.conv_wgrad(draw, xa, raw.c, pw.r, pw.s, 1, lane=True)
            dw = (E.unpack_wgrad_s2d if state.s2d else E.unpack_wgrad_im2col)(dwcol, pw, lane=True)
            dx = None                                    # the image gets no gradient
        else:
            dw = E.conv_wgrad(draw, xa, conv.out_channels, k, k, dil, stride=stride, lane=True)
            dx = None
            if ctx.needs_input_grad[0]:
                g = E.Act.empty(xa.n, xa.h, xa.w, xa.c, dev)
                E.conv_dgrad([(draw, pw, dil, stride)], g)
                dx = E.grad_like(x, g)
        E.join_wgrad_lane()
        ctx.rec = None
        return dx, dw, dgamma, dbeta, dres, None, None, None, None


def conv_bn(x, conv, bn, relu, state, residual=None):
    return _ConvBNFn.apply(x, conv.weight, bn.weight, bn.bias, residual, conv, bn, relu, state)


class _UpsampleFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, scale, align_corners):
        xa = E.as_act(x, 'upsample input')
        y = E.Act.empty(xa.n, xa.h * scale, xa.w * scale, xa.c, x.device)
        E.upsample_fwd(xa, y, align_corners=align_corners)
        ctx.cfg = (xa.h, xa.w, align_corners, x)
        return y.nchw()

    @staticmethod
    def backward(ctx, dout):
        h, w, ac, x = ctx.cfg
        da = E.as_act(dout, 'upsample grad')
        dx = E.Act.empty(da.n, h, w, da.c, dout.device)
        E.upsample_bwd(da, dx, align_corners=ac)
        return E.grad_like(x, dx), None, None


def upsample(x, scale, align_corners=False):
    return _UpsampleFn.apply(x, int(scale), bool(align_corners))


class _SumReLUFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, relu, *xs):
        acts = [E.as_act(t, 'sum_relu input') for t in xs]
        a0 = acts[0]
        y = E.Act.empty(a0.n, a0.h, a0.w, a0.c, xs[0].device)
        cur = acts
        while True:                                   # up to four addends per pass
            chunk, rest = cur[:4], cur[4:]
            last = not rest
            tgt = y if last else E.Act.empty(a0.n, a0.h, a0.w, a0.c, xs[0].device)
            E.add_n_relu(chunk, tgt, relu=relu and last)
            if last:
                break
            cur = [tgt] + rest
        ctx.y = y if relu else None
        ctx.xs = xs
        E.capture_seq('sum_relu', y)
        return y.nchw()

    @staticmethod
    def backward(ctx, dout):
        da = E.as_act(dout, 'sum_relu grad')
        if ctx.y is not None:
            dz = E.Act.empty(da.n, da.h, da.w, da.c, dout.device)
            E.relu_bwd(da, ctx.y, dz)
        else:
            dz = da
        return (None,) + tuple(E.grad_like(x, dz) if need else None
                               for x, need in zip(ctx.xs, ctx.needs_input_grad[1:]))


def sum_relu(xs, relu=True):
    return _SumReLUFn.apply(bool(relu), *xs)


class _ClassifyFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight, bias, conv, packed, scale, align_corners):
        from .models.deeplabv3plus import cls_conv_forward
        xa = E.as_act(x, 'classifier input')
        save = any(ctx.needs_input_grad)
        lr = cls_conv_forward(conv, packed, xa, save)
        out = E.logits_upsample_fwd(lr, conv.out_channels, xa.h * scale, xa.w * scale, align_corners=align_corners)
        if save:
            ctx.cfg = (xa, conv, packed, lr.shape[3], align_corners, x)
            ctx.set_materialize_grads(False)
        return out, lr

    @staticmethod
    def backward(ctx, dout, dlr_direct):
        from .models.deeplabv3plus import cls_conv_backward
        xa, conv, packed, ncp, ac, x = ctx.cfg
        dlr = E.class_map_backward(dout, dlr_direct, xa.h, xa.w, ncp, ac)
        dx = E.Act.empty(xa.n, xa.h, xa.w, xa.c, xa.t.device) if ctx.needs_input_grad[0] else None
        dw, db = cls_conv_backward(conv, packed, xa, dlr, dx, lane=True)
        E.join_wgrad_lane()
        return (E.grad_like(x, dx) if dx is not None else None), dw, db, None, None, None, None


def classify(x, conv, packed, scale, align_corners):
    """nn.Conv2d (with bias, 1x1 or 3x3 'same') + bilinear x``scale`` -> fp32 NCHW logits."""
    if conv.bias is None:
        raise NotImplementedError('classifier conv without bias')
    out, lr = _ClassifyFn.apply(x, conv.weight, conv.bias, conv, packed, int(scale), bool(align_corners))
    return E.attach_class_map(out, lr, conv.out_channels, bool(align_corners))
