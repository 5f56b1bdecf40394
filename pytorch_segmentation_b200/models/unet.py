"""UNet (MobileNetV2 encoder) — same constructor, forward and state-dict keys as models/unet.py:14-56.

The decoder is one autograd node of libb200seg kernels.  There is no torch.cat
(models/unet.py:34,40,46): the x2 upsample writes its result into the first channel slice of the
concat buffer and the backbone skip map is placed in the second slice by the same resampling kernel
at scale 1 (align_corners=True with equal sizes is an exact copy).
"""
import torch
import torch.nn as nn

from .. import engine as E
from ..backbones import mobilenet_v2
from ..modules import ConvNormAct, initialize_weights
from .deeplabv3plus import cls_conv_backward, cls_conv_forward, make_cls_packed


class UNet(nn.Module):
    def __init__(self, num_classes):
        super().__init__()
        self.backbone = mobilenet_v2(pretrained=True)
        self.up_convs = nn.ModuleList([ConvNormAct(1280, 256), ConvNormAct(352, 128), ConvNormAct(160, 64)])
        self.cls_conv = nn.Conv2d(88, num_classes, 3, padding=1)
        initialize_weights(self.up_convs)
        initialize_weights(self.cls_conv)
        self.num_classes = num_classes
        self._cls_packed = make_cls_packed(self.cls_conv)
        self._pack_entries = None

    def head_tensors(self):
        ts = []
        for c in self.up_convs:
            ts += c.tensors()
        return ts + [self.cls_conv.weight, self.cls_conv.bias]

    def _packed_weights(self):
        return [] if isinstance(self._cls_packed, E.PackedTapWeight) else [(self._cls_packed, self.cls_conv.weight)]

    def forward(self, x):
        # every conv weight that went stale (all of them after an optimizer step) is re-packed in one launch
        if self._pack_entries is None:
            self._pack_entries = E.packed_entries(self)
        E.pack_all(self._pack_entries, need_dgrad=torch.is_grad_enabled())
        x1, x2, x3, x4, x5 = self.backbone(x)
        return self.head(x2, x3, x4, x5)

    def head(self, x2, x3, x4, x5):
        out, lr = _UNetHeadFn.apply(self, x2, x3, x4, x5, *self.head_tensors())
        return E.attach_class_map(out, lr, self.num_classes, True)


def _up2(x, out):
    E.upsample_fwd(x, out, align_corners=True)


class _UNetHeadFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, m, x2, x3, x4, x5, *params):
        a2, a3, a4, a5 = (E.as_act(t, 'backbone feature') for t in (x2, x3, x4, x5))
        training, save = m.training, any(ctx.needs_input_grad)
        dev = x5.device
        n = a5.n
        skips = (a4, a3, a2)
        x = a5
        recs = []
        cats = []
        for i, cna in enumerate(m.up_convs):
            k = cna[0].out_channels
            y = E.Act.empty(n, x.h, x.w, k, dev)
            recs.append(cna.run_forward(x, y, training, save))
            sk = skips[i]
            if (sk.h, sk.w) != (x.h * 2, x.w * 2):
                raise ValueError('UNet: skip map %dx%d does not match the x2 upsample of %dx%d'
                                 % (sk.h, sk.w, x.h, x.w))
            # concat buffer [ up(y) | skip ] (models/unet.py:30-34): the upsample writes its slice, the skip
            # is placed beside it by the same kernel at scale 1 (exact copy)
            cat = E.Act.empty(n, sk.h, sk.w, k + sk.c, dev)
            _up2(y, cat.slice(0, k))
            E.upsample_fwd(sk, cat.slice(k, sk.c), align_corners=True)
            cats.append(cat)
            x = cat
        # models/unet.py:47-55
        u = E.Act.empty(n, x.h * 2, x.w * 2, x.c, dev)
        _up2(x, u)
        lr = cls_conv_forward(m.cls_conv, m._cls_packed, u, save)
        out = E.logits_upsample_fwd(lr, m.num_classes, u.h * 2, u.w * 2, align_corners=True)
        if save:
            ctx.m, ctx.inputs = m, (x2, x3, x4, x5)
            ctx.rec = (recs, cats, u, lr.shape[3])
            ctx.set_materialize_grads(False)
        return out, lr

    @staticmethod
    def backward(ctx, dout, dlr_direct):
        m = ctx.m
        recs, cats, u, ncp = ctx.rec
        x2, x3, x4, x5 = ctx.inputs
        dev = u.t.device
        n = u.n
        dlr = E.class_map_backward(dout, dlr_direct, u.h, u.w, ncp, True)
        du = E.Act.empty(n, u.h, u.w, u.c, dev)
        cls_grads = cls_conv_backward(m.cls_conv, m._cls_packed, u, dlr, du, lane=True)
        dcat = E.Act.empty(n, cats[2].h, cats[2].w, cats[2].c, dev)
        E.upsample_bwd(du, dcat, align_corners=True)
        skip_inputs = (x4, x3, x2)
        skip_need = (ctx.needs_input_grad[3], ctx.needs_input_grad[2], ctx.needs_input_grad[1])
        dskips = [None, None, None]
        pgrads = [None, None, None]
        for i in (2, 1, 0):
            cna = m.up_convs[i]
            k = cna[0].out_channels
            sk_c = dcat.c - k
            if skip_need[i]:
                # the skip slice of the concat gradient, compacted to its own buffer (identity resample)
                gs = E.Act.empty(n, dcat.h, dcat.w, sk_c, dev)
                E.upsample_fwd(dcat.slice(k, sk_c), gs, align_corners=True)
                dskips[i] = E.grad_like(skip_inputs[i], gs)
            xin = recs[i][0]
            dy = E.Act.empty(n, xin.h, xin.w, k, dev)
            E.upsample_bwd(dcat.slice(0, k), dy, align_corners=True)
            draw, pgrads[i] = cna.run_backward_params(recs[i], dy)
            need_dx = True if i > 0 else ctx.needs_input_grad[4]
            if need_dx:
                dx = E.Act.empty(n, xin.h, xin.w, xin.c, dev)
                E.conv_dgrad([cna.dgrad_src(draw)], dx)
                dcat = dx
        dx5 = E.grad_like(x5, dcat) if ctx.needs_input_grad[4] else None
        grads = pgrads[0] + pgrads[1] + pgrads[2] + cls_grads
        E.join_wgrad_lane()
        return (None, dskips[2], dskips[1], dskips[0], dx5) + tuple(grads)
