"""DeepLabV3+ — same constructor, forward and state-dict keys as models/deeplabv3plus.py:15-44.

The head (everything after the backbone) is one autograd node whose forward and backward are explicit
sequences of libb200seg kernels over NHWC bf16 buffers: ``project`` and the upsampled ASPP output are
written straight into the two channel slices of the decoder concat buffer (no torch.cat,
models/deeplabv3plus.py:38), ``cls_conv`` produces an fp32 class map that the final bilinear x4
(models/deeplabv3plus.py:40-43) turns into the NCHW fp32 logits the reference returns.
"""
import torch
import torch.nn as nn

from .. import engine as E
from ..backbones import resnet50
from ..modules import ConvNormAct, initialize_weights
from .aspp import ASPP


class DeepLabV3Plus(nn.Module):
    def __init__(self, num_classes):
        super().__init__()
        self.backbone = resnet50(pretrained=True, replace_stride_with_dilation=[False, False, True])
        self.project = ConvNormAct(256, 128, 1)
        self.aspp = ASPP(2048, 256, [6, 12, 18])
        self.cls_conv = nn.Conv2d(384, num_classes, 3, padding=1)
        initialize_weights(self.aspp)
        initialize_weights(self.project)
        initialize_weights(self.cls_conv)
        self.num_classes = num_classes
        self._cls_packed = make_cls_packed(self.cls_conv)
        self._pack_entries = None

    def head_tensors(self):
        return self.project.tensors() + self.aspp.tensors() + [self.cls_conv.weight, self.cls_conv.bias]

    def _packed_weights(self):
        return [] if isinstance(self._cls_packed, E.PackedTapWeight) else [(self._cls_packed, self.cls_conv.weight)]

    def forward(self, x):
        # every conv weight that went stale (all of them after an optimizer step) is re-packed in one launch
        if self._pack_entries is None:
            self._pack_entries = E.packed_entries(self)
        E.pack_all(self._pack_entries, need_dgrad=torch.is_grad_enabled())
        features = self.backbone(x)
        return self.head(features[1], features[-1])

    def head(self, low, high):
        """features[1] (256 ch, stride 4) and features[-1] (2048 ch, stride 16) -> N x nc x 4H4 x 4W4 logits."""
        out, lr = _DeepLabHeadFn.apply(self, low, high, *self.head_tensors())
        return E.attach_class_map(out, lr, self.num_classes, True)


def make_cls_packed(conv):
    """Narrow-output 3x3 convs take the taps-as-channels form (engine.PackedTapWeight, csrc/taps.cu)."""
    return E.PackedTapWeight() if E.PackedTapWeight.applies(conv) else E.PackedWeight()


def cls_conv_forward(conv, packed, x, save):
    """Plain 3x3 conv with bias, no BN/act (models/deeplabv3plus.py:22,39; models/unet.py:23,51):
    NHWC bf16 slice -> NHWC fp32 class map padded to a multiple of 8 channels."""
    nc = conv.out_channels
    ncp = (nc + 7) // 8 * 8
    bias = torch.zeros((ncp,), dtype=torch.float32, device=x.t.device)
    if conv.bias is not None:
        bias[:nc].copy_(conv.bias.detach())
    if isinstance(packed, E.PackedTapWeight):
        # the input is read once: one 1x1 GEMM produces the nine per-tap partial class maps, a gather sums them
        pw = packed.update(conv.weight)
        z = E.Act.empty(x.n, x.h, x.w, pw.zc, x.t.device)
        E.conv_fprop([(x, pw.fprop, 1)], z)
        return E.tap_scatter_add(z, ncp, pw.r, pw.s, conv.dilation[0], bias=bias)
    pw = packed.update(conv.weight, need_dgrad=save)
    out = torch.empty((x.n, x.h, x.w, ncp), dtype=torch.float32, device=x.t.device)
    E.conv_fprop_f32([(x, pw, conv.dilation[0])], out, ncp, bias=bias)
    return out


def cls_conv_backward(conv, packed, x, dlr, dx, lane=False):
    """dlr: Act grad of the padded class map (bf16).  Returns [dW, dbias]; writes dx if given.  ``lane``: put the
    weight gradient on the weight-gradient lane (the caller joins it)."""
    nc = conv.out_channels
    db = E.channel_sum(dlr)[:nc]
    if isinstance(packed, E.PackedTapWeight):
        pw = packed.update(conv.weight)
        col = E.tap_gather(dlr, pw.r, pw.s, conv.dilation[0])            # [p][tap*ncp + k] = dY[p + off(tap)][k]
        dwcol = E.conv_wgrad(col, x, pw.zc, 1, 1, 1, lane=lane)
        dw = E.unpack_wgrad_taps(dwcol, pw, lane=lane)
        if dx is not None:
            E.conv_dgrad([(col, pw.dgrad, 1)], dx)
        return [dw, db]
    dw = E.conv_wgrad(dlr, x, nc, conv.kernel_size[0], conv.kernel_size[1], conv.dilation[0], lane=lane)
    if dx is not None:
        E.conv_dgrad([(dlr, packed, conv.dilation[0])], dx)
    return [dw, db]


class _DeepLabHeadFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, m, low, high, *params):
        la, ha = E.as_act(low, 'features[1]'), E.as_act(high, 'features[-1]')
        training, save = m.training, any(ctx.needs_input_grad)
        dev = low.device
        n, h4, w4 = la.n, la.h, la.w
        if (ha.h * 4, ha.w * 4) != (h4, w4):
            raise ValueError('DeepLabV3Plus: features[-1] (%dx%d) x4 does not match features[1] (%dx%d)'
                             % (ha.h, ha.w, h4, w4))
        dec = E.Act.empty(n, h4, w4, 384, dev)                       # [ up(aspp) 256 | project(low) 128 ]
        rec_proj = m.project.run_forward(la, dec.slice(256, 128), training, save)
        aspp_out = E.Act.empty(n, ha.h, ha.w, 256, dev)
        rec_aspp = m.aspp.run_forward(ha, aspp_out, training, save)
        E.upsample_fwd(aspp_out, dec.slice(0, 256), align_corners=True)
        lr = cls_conv_forward(m.cls_conv, m._cls_packed, dec, save)
        out = E.logits_upsample_fwd(lr, m.num_classes, h4 * 4, w4 * 4, align_corners=True)
        if save:
            ctx.m, ctx.low, ctx.high = m, low, high
            ctx.rec = (rec_proj, rec_aspp, dec, (ha.h, ha.w), lr.shape[3])
            ctx.set_materialize_grads(False)
        return out, lr

    @staticmethod
    def backward(ctx, dout, dlr_direct):
        m = ctx.m
        rec_proj, rec_aspp, dec, (h16, w16), ncp = ctx.rec
        n, h4, w4 = dec.n, dec.h, dec.w
        dev = dec.t.device
        dlr = E.class_map_backward(dout, dlr_direct, h4, w4, ncp, True)
        ddec = E.Act.empty(n, h4, w4, 384, dev)
        cls_grads = cls_conv_backward(m.cls_conv, m._cls_packed, dec, dlr, ddec, lane=True)
        # project branch
        draw, proj_grads = m.project.run_backward_params(rec_proj, ddec.slice(256, 128))
        dlow = None
        if ctx.needs_input_grad[1]:
            g = E.Act.empty(n, h4, w4, 256, dev)
            E.conv_dgrad([m.project.dgrad_src(draw)], g)
            dlow = E.grad_like(ctx.low, g)
        # ASPP branch
        daspp = E.Act.empty(n, h16, w16, 256, dev)
        E.upsample_bwd(ddec.slice(0, 256), daspp, align_corners=True)
        g = E.Act.empty(n, h16, w16, 2048, dev) if ctx.needs_input_grad[2] else None
        aspp_grads = m.aspp.run_backward(rec_aspp, daspp, g)
        dhigh = E.grad_like(ctx.high, g) if g is not None else None
        E.join_wgrad_lane()
        return (None, dlow, dhigh) + tuple(proj_grads + aspp_grads + cls_grads)
