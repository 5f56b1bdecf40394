"""ResNet-50 (output stride 16) feature extractor on the tcgen05 path — SURVEY.md §8f-1.

The reference gets its backbone from the external ``pytorch_modules.backbones.resnet50(pretrained=True,
replace_stride_with_dilation=[False, False, True])`` (models/deeplabv3plus.py:17-19) and indexes the list of five
feature maps it returns (models/deeplabv3plus.py:29-32).  Parameters live in an ordinary torchvision ResNet-50
module tree (same state-dict keys as the oracle shim); the arithmetic below is libb200seg:

* every 1x1 / 3x3 convolution (stride 1, stride 2, dilation 2) is the tcgen05 implicit GEMM with the train-mode
  BatchNorm statistics taken in its epilogue; BN-apply + ReLU (+ residual add, + the downsample branch's BN) is one
  bandwidth kernel; backward is the two-pass BN backward, the tcgen05 wgrad and dgrad (the block-input gradient of a
  downsampling bottleneck — conv1's and the downsample conv's data gradients — is one accumulation in TMEM);
* the 3x3 max pool is a gather kernel pair;
* the 7x7 stride-2 stem convolution over 3 input channels cannot feed a 64-wide K block tap by tap, so the fp32
  NCHW image is gathered once into a bf16 patch matrix (147 -> 152 columns per output pixel; the same pass is the
  fp32 -> bf16 / NCHW -> NHWC conversion of the input) and the stem becomes a 1x1 tcgen05 GEMM over it, its weight
  gradient the tcgen05 wgrad over the same matrix.  No cuDNN anywhere on the path.

Forward and backward are explicit kernel sequences inside one autograd node (no per-op autograd graph).
"""
import torch
import torch.nn as nn
import torchvision

from . import engine as E


class _Unit:
    """conv (no bias) + BatchNorm2d bound to the modules that own the parameters."""

    def __init__(self, conv, bn):
        k = conv.kernel_size[0]
        if conv.kernel_size not in ((1, 1), (3, 3)) or conv.groups != 1 or conv.bias is not None:
            raise NotImplementedError('backbone conv %r is outside the tcgen05 path' % (conv,))
        if conv.stride not in ((1, 1), (2, 2)) or conv.padding[0] != (k // 2) * conv.dilation[0]:
            raise NotImplementedError('backbone conv %r: stride/padding outside the tcgen05 path' % (conv,))
        self.conv, self.bn = conv, bn
        self.k, self.stride, self.dil = k, conv.stride[0], conv.dilation[0]
        self.cout = conv.out_channels
        self.packed = E.PackedWeight()

    def params(self):
        return [self.conv.weight, self.bn.weight, self.bn.bias]

    def forward_raw(self, x, training, stats, need_dgrad=True):
        """-> (raw conv output Act, vec[4, c] = scale/shift/mean/invstd)"""
        bn = self.bn
        pw = self.packed.update(self.conv.weight, need_dgrad=need_dgrad)
        raw = E.Act.empty(x.n, x.h // self.stride, x.w // self.stride, self.cout, x.t.device)
        if training:
            E.conv_fprop([(x, pw, self.dil, self.stride)], raw, stats=stats)
            track = bn.track_running_stats and bn.running_mean is not None
            vec = E.bn_finalize_lazy(stats, self.cout, raw.pixels, bn.weight, bn.bias, bn.eps, bn.momentum,
                                bn.running_mean if track else None, bn.running_var if track else None)
        else:
            E.conv_fprop([(x, pw, self.dil, self.stride)], raw)
            vec = E.bn_eval_affine(self.cout, bn.weight, bn.bias, bn.running_mean, bn.running_var, bn.eps)
        E.capture(self.conv.weight, 'raw', raw)
        return raw, vec

    def wgrad(self, draw, x):
        return E.conv_wgrad(draw, x, self.cout, self.k, self.k, self.dil, stride=self.stride, lane=True)

    def dgrad_src(self, draw):
        return (draw, self.packed, self.dil, self.stride)


class _Block:
    """torchvision Bottleneck: conv1 1x1 -> conv2 3x3 (stride / dilation) -> conv3 1x1, + identity or downsample."""

    def __init__(self, blk):
        self.u1, self.u2, self.u3 = _Unit(blk.conv1, blk.bn1), _Unit(blk.conv2, blk.bn2), _Unit(blk.conv3, blk.bn3)
        self.ud = _Unit(blk.downsample[0], blk.downsample[1]) if blk.downsample is not None else None

    def units(self):
        return [self.u1, self.u2, self.u3] + ([self.ud] if self.ud is not None else [])

    def forward(self, x, training, stats_of):
        dev = x.t.device
        raw1, v1 = self.u1.forward_raw(x, training, stats_of(self.u1))
        a1 = E.Act.empty(raw1.n, raw1.h, raw1.w, raw1.c, dev)
        E.bn_apply(raw1, a1, v1, True)
        raw2, v2 = self.u2.forward_raw(a1, training, stats_of(self.u2))
        a2 = E.Act.empty(raw2.n, raw2.h, raw2.w, raw2.c, dev)
        E.bn_apply(raw2, a2, v2, True)
        raw3, v3 = self.u3.forward_raw(a2, training, stats_of(self.u3))
        out = E.Act.empty(raw3.n, raw3.h, raw3.w, raw3.c, dev)
        if self.ud is not None:
            rawd, vd = self.ud.forward_raw(x, training, stats_of(self.ud))
            E.bn_apply(raw3, out, v3, True, res=rawd, res_vec=vd, want_bits=training)
        else:
            rawd = vd = None
            E.bn_apply(raw3, out, v3, True, res=x, want_bits=training)
        if E.CAPTURE is not None:
            E.capture(self.u1.conv.weight, 'out', a1)
            E.capture(self.u2.conv.weight, 'out', a2)
            E.capture(self.u3.conv.weight, 'out', out)
        return out, (x, raw1, v1, a1, raw2, v2, a2, raw3, v3, rawd, vd, out)

    def join_target(self, rec, red_of):
        """What a later block's gradient join needs to do this block's first BatchNorm-backward pass for it
        (engine.grad_join_reduce): (raw3, v3, red3) and, with a downsample branch, (rawd, vd, redd)."""
        x, raw1, v1, a1, raw2, v2, a2, raw3, v3, rawd, vd, out = rec
        t3 = (raw3, v3, red_of(self.u3))
        td = (rawd, vd, red_of(self.ud)) if self.ud is not None else None
        return t3, td

    def backward(self, rec, dout, training, grads, need_dx=True, dout_masked=False, mask_in=True, red_of=None,
                 red_done=False, join_to=None):
        """dout: Act grad of the block output — already multiplied by (out > 0) when ``dout_masked`` (it then came from
        the next block's ``backward``), raw otherwise (a head's gradient).  Fills grads[param] and returns the Act grad
        of the block input, multiplied by (x > 0) when ``mask_in`` (x is a ReLU output: the previous block's), so that
        block's BatchNorm backward never re-reads its own output for the mask."""
        x, raw1, v1, a1, raw2, v2, a2, raw3, v3, rawd, vd, out = rec
        dev = x.t.device
        self.fused_join = False
        E.capture(self.u3.conv.weight, 'gout', dout)       # parity tests: the stored gradient that enters this block

        def unit_bwd(u, dy, raw, vec, xin, relu=True, mask=None, reduced=False):
            draw = E.Act.empty(raw.n, raw.h, raw.w, raw.c, dev)
            dgamma, dbeta = E.bn_backward(dy, raw, draw, vec, relu, training, mask=mask,
                                          red=red_of(u) if red_of is not None else None, reduced=reduced)
            E.capture(u.conv.weight, 'graw', draw)         # parity tests: the stored gradient of the raw conv output
            grads[u.conv.weight] = u.wgrad(draw, xin)
            grads[u.bn.weight], grads[u.bn.bias] = dgamma, dbeta
            return draw

        # bn3 (and the downsample BN) sit under the block's final ReLU, whose mask is the sign of `out`
        # ``red_done``: the join that produced dout already reduced it against raw3 (and rawd) into the arena
        assert not red_done or (dout_masked and training and red_of is not None)
        tail = dict(relu=False, reduced=red_done) if dout_masked else dict(relu=True, mask=out)
        draw3 = unit_bwd(self.u3, dout, raw3, v3, a2, **tail)
        da2 = E.Act.empty(a2.n, a2.h, a2.w, a2.c, dev)
        E.conv_dgrad([self.u3.dgrad_src(draw3)], da2)
        draw2 = unit_bwd(self.u2, da2, raw2, v2, a1)
        da1 = E.Act.empty(a1.n, a1.h, a1.w, a1.c, dev)
        E.conv_dgrad([self.u2.dgrad_src(draw2)], da1)
        draw1 = unit_bwd(self.u1, da1, raw1, v1, x)
        drawd = unit_bwd(self.ud, dout, rawd, vd, x, **tail) if self.ud is not None else None
        if not need_dx:
            return None
        dx = E.Act.empty(x.n, x.h, x.w, x.c, dev)
        xin = x if mask_in else None
        # ``join_to`` = (t3, td) of the block whose output x is: the join then also does that block's first BN-backward pass
        fuse = join_to is not None and xin is not None and xin.bits is not None and (self.ud is not None or dout_masked)
        if self.ud is not None:
            E.conv_dgrad([self.u1.dgrad_src(draw1), self.ud.dgrad_src(drawd)], dx)
            if fuse:
                E.grad_join_reduce(dx, None, xin, *join_to)
            elif xin is not None:
                E.grad_add_masked(dx, xin=xin)
        else:
            E.conv_dgrad([self.u1.dgrad_src(draw1)], dx)
            # identity branch: + dout * (out > 0), the mask being already in dout when it is pre-masked
            if fuse:
                E.grad_join_reduce(dx, dout, xin, *join_to)
            else:
                E.grad_add_masked(dx, dy=dout, y=None if dout_masked else out, xin=xin)
        self.fused_join = fuse
        return dx


class ResNet50Features(nn.Module):
    """[relu(bn1(conv1 x)), layer1, layer2, layer3, layer4] — strides 2, 4, 8, 16, 16 with layer4 dilated, the list
    the reference indexes at models/deeplabv3plus.py:29-32.  ``forward`` returns NCHW-shaped bf16 tensors that are
    views of NHWC buffers (what the heads consume without a copy)."""

    def __init__(self, replace_stride_with_dilation=None):
        super().__init__()
        net = torchvision.models.resnet50(weights=None, replace_stride_with_dilation=replace_stride_with_dilation)
        self.conv1, self.bn1, self.relu, self.maxpool = net.conv1, net.bn1, net.relu, net.maxpool
        self.layer1, self.layer2, self.layer3, self.layer4 = net.layer1, net.layer2, net.layer3, net.layer4
        self._plan = None
        # stride-2 few-channel stem: space-to-depth form (no gathered matrix); anything else: patch gather
        self._stem_packed = E.PackedS2DWeight() if E.PackedS2DWeight.applies(self.conv1) else E.PackedIm2colWeight()

    def _build_plan(self):
        layers = [[_Block(b) for b in layer] for layer in (self.layer1, self.layer2, self.layer3, self.layer4)]
        units = [u for layer in layers for blk in layer for u in blk.units()]
        offs, off = {}, 0
        for u in units:
            offs[id(u)] = off
            off += 2 * u.cout
        self._plan = (layers, units, offs, off)
        return self._plan

    def _packed_weights(self):
        units = (self._plan or self._build_plan())[1]
        stem = [] if isinstance(self._stem_packed, E.PackedS2DWeight) else [(self._stem_packed, self.conv1.weight)]
        return stem + [(u.packed, u.conv.weight) for u in units]

    def forward(self, x):
        params = list(self.parameters())
        feats = _ResNetFn.apply(self, x, *params)
        return list(feats)

    # ---- explicit forward / backward -------------------------------------------------------------------------
    def run_forward(self, x, training, save):
        E._require_cuda(x, 'backbone input')
        layers, units, offs, total = self._plan or self._build_plan()
        dev = x.device
        # stem: patch gather of the image + 1x1 tcgen05 GEMM (see module docstring), then BN/ReLU/max-pool
        conv1 = self.conv1
        if conv1.bias is not None or conv1.groups != 1 or conv1.dilation != (1, 1):
            raise NotImplementedError('backbone stem %r is outside the patch-gather path' % (conv1,))
        xf = x.detach()
        pw0 = self._stem_packed.update(conv1.weight)
        u8 = xf.dtype == torch.uint8          # utils.post_fetch_fn: 1-byte pixels, normalised by the first kernel that reads them
        if u8 and not isinstance(pw0, E.PackedS2DWeight):
            xf = E.normalize_u8(xf, *getattr(x, '_b200seg_norm', (None, None)))
        elif not u8 and (xf.dtype != torch.float32 or not xf.is_contiguous()):
            xf = xf.float().contiguous()
        if isinstance(pw0, E.PackedS2DWeight) and u8:
            col, oh0, ow0 = E.space_to_depth_u8(xf, conv1.kernel_size[0], conv1.kernel_size[1], conv1.padding[0],
                                                *getattr(x, '_b200seg_norm', (None, None)))
        elif isinstance(pw0, E.PackedS2DWeight):
            col, oh0, ow0 = E.space_to_depth(xf, conv1.kernel_size[0], conv1.kernel_size[1], conv1.padding[0])
        else:
            col = E.im2col(xf, conv1.kernel_size[0], conv1.kernel_size[1], conv1.stride[0], conv1.padding[0])
            oh0, ow0 = col.h, col.w
        raw0 = E.Act.empty(col.n, oh0, ow0, conv1.out_channels, dev)
        bn = self.bn1
        if training:
            stats0 = torch.zeros((2 * raw0.c,), dtype=torch.float32, device=dev)
            E.conv_fprop([(col, pw0, 1)], raw0, stats=stats0)
            track = bn.track_running_stats and bn.running_mean is not None
            v0 = E.bn_finalize_lazy(stats0, raw0.c, raw0.pixels, bn.weight, bn.bias, bn.eps, bn.momentum,
                               bn.running_mean if track else None, bn.running_var if track else None)
        else:
            E.conv_fprop([(col, pw0, 1)], raw0)
            v0 = E.bn_eval_affine(raw0.c, bn.weight, bn.bias, bn.running_mean, bn.running_var, bn.eps)
        a0 = E.Act.empty(raw0.n, raw0.h, raw0.w, raw0.c, dev)
        E.bn_apply(raw0, a0, v0, True)
        E.capture(conv1.weight, 'raw', raw0)
        E.capture(conv1.weight, 'out', a0)
        p0, idx = E.maxpool_fwd(a0)

        stats_all = torch.zeros((total,), dtype=torch.float32, device=dev) if training else None

        def stats_of(u):
            if stats_all is None:
                return None
            o = offs[id(u)]
            return stats_all[o:o + 2 * u.cout]

        feats, recs = [a0], []
        cur = p0
        for layer in layers:
            lrec = []
            for blk in layer:
                cur, rec = blk.forward(cur, training, stats_of)
                lrec.append(rec if save else None)
            recs.append(lrec)
            feats.append(cur)
        if training:
            nbt = [m.num_batches_tracked for m in self.modules()
                   if isinstance(m, nn.BatchNorm2d) and m.track_running_stats and m.num_batches_tracked is not None]
            if nbt:
                torch._foreach_add_(nbt, 1)
        saved = (col, pw0, raw0, v0, a0, idx, recs, training) if save else None
        return feats, saved

    def run_backward(self, saved, gfeats):
        """gfeats: gradients of the five feature maps (None where unused).  -> {parameter: gradient}"""
        col, pw0, raw0, v0, a0, idx, recs, training = saved
        layers = self._plan[0]
        grads = {}
        # one zeroed arena for the per-channel reductions of every BatchNorm backward (same offsets as the forward stats)
        units, offs, total = self._plan[1], self._plan[2], self._plan[3]
        red_all = torch.zeros((total,), dtype=torch.float32, device=raw0.t.device)

        def red_of(u):
            o = offs[id(u)]
            return red_all[o:o + 2 * u.cout]

        d = None
        masked = False       # whether d already carries the ReLU mask of the tensor it is the gradient of
        red_done = False     # whether d was already reduced against the next block's raw3 / rawd (fused join)
        fuse_ok = E.JOIN_REDUCE and training
        for li in (3, 2, 1, 0):
            g = gfeats[li + 1]
            if g is not None:
                ga = E.as_act(g, 'backbone feature gradient')
                if d is None:
                    d, masked = ga, False
                else:
                    # d (masked) + external gradient (raw): mask the sum with the layer output's sign.  This join is the
                    # last word on d, so it is the one that may carry the reduction for the layer's last block
                    out_last = recs[li][-1][-1]
                    if fuse_ok and out_last.bits is not None:
                        E.grad_join_reduce(d, ga, out_last, *layers[li][-1].join_target(recs[li][-1], red_of))
                        red_done = True
                    else:
                        E.grad_add_masked(d, dy=ga, xin=out_last)
            if d is None:
                continue
            for bi in range(len(layers[li]) - 1, -1, -1):
                # the block this one's input gradient goes to (None: the max-pool output, or a layer boundary where an
                # external feature gradient still has to be added — handled above)
                prev = None
                if fuse_ok:
                    if bi > 0:
                        prev = (li, bi - 1)
                    elif li > 0 and gfeats[li] is None:
                        prev = (li - 1, len(layers[li - 1]) - 1)
                join_to = layers[prev[0]][prev[1]].join_target(recs[prev[0]][prev[1]], red_of) if prev is not None else None
                # the input of layer1's first block is the max-pool output, not a ReLU output: no mask there
                blk = layers[li][bi]
                d = blk.backward(recs[li][bi], d, training, grads, dout_masked=masked,
                                 mask_in=not (li == 0 and bi == 0), red_of=red_of, red_done=red_done, join_to=join_to)
                masked = not (li == 0 and bi == 0)
                red_done = blk.fused_join
            # this stage's gradients are final: a data-parallel sink starts their all-reduce under the remaining stages
            E.deliver_grads(grads, [p for blk in layers[li] for u in blk.units() for p in u.params()])
        if d is not None:
            da0 = E.maxpool_bwd(d, idx, a0.h, a0.w)
            if gfeats[0] is not None:
                da0.nchw().add_(E.as_act(gfeats[0], 'backbone feature gradient').nchw())
        elif gfeats[0] is not None:
            da0 = E.as_act(gfeats[0], 'backbone feature gradient')
        else:
            return grads
        draw0 = E.Act.empty(raw0.n, raw0.h, raw0.w, raw0.c, raw0.t.device)
        dgamma, dbeta = E.bn_backward(da0, raw0, draw0, v0, True, training)
        E.capture(self.conv1.weight, 'graw', draw0)
        # dW of the stem = the 1x1 wgrad over the gathered matrix, scattered back to OIHW
        dwcol = E.conv_wgrad(draw0, col, raw0.c, pw0.r, pw0.s, 1, lane=True)
        unpack = E.unpack_wgrad_s2d if isinstance(pw0, E.PackedS2DWeight) else E.unpack_wgrad_im2col
        grads[self.conv1.weight] = unpack(dwcol, pw0, lane=True)
        grads[self.bn1.weight], grads[self.bn1.bias] = dgamma, dbeta
        return grads


class _ResNetFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, mod, x, *params):
        save = any(ctx.needs_input_grad[2:])
        # unused feature maps (the heads read two of the five) must reach backward as None, not as zero tensors
        ctx.set_materialize_grads(False)
        feats, saved = mod.run_forward(x, mod.training, save)
        ctx.mod, ctx.saved, ctx.params = mod, saved, params
        outs = tuple(f.nchw() for f in feats)
        return outs

    @staticmethod
    def backward(ctx, *gfeats):
        grads = ctx.mod.run_backward(ctx.saved, list(gfeats))
        E.join_wgrad_lane()
        ctx.saved = None
        out = []
        for p, need in zip(ctx.params, ctx.needs_input_grad[2:]):
            g = grads.get(p) if need else None
            if g is not None and g.shape != p.shape:
                g = g.reshape(p.shape)
            out.append(g)
        # the image itself gets no gradient (the reference never asks for one: inputs come from the data loader)
        return (None, None) + tuple(out)


def load_pretrained(module, arch, filename):
    """``pretrained=True`` (models/deeplabv3plus.py:17, models/unet.py:17): the reference's backbones start from ImageNet
    weights.  There is no network here, so the torchvision checkpoint is taken from a local file — ``$B200SEG_PRETRAINED_DIR``
    or the torch hub cache (``<hub>/checkpoints/<filename>``) — and when none exists the backbone stays randomly
    initialised WITH A WARNING (never silently)."""
    path = find_pretrained(arch, filename)
    if path is None:
        return False
    state = torch.load(path, map_location='cpu', weights_only=True)
    own = module.state_dict()
    missing = [k for k in own if k not in state]
    if missing:
        raise RuntimeError('%s: %s lacks %d tensors (e.g. %s)' % (arch, path, len(missing), missing[0]))
    module.load_state_dict({k: state[k] for k in own})
    return True


def find_pretrained(arch, filename):
    """Path of the local checkpoint, or None after warning that the backbone stays randomly initialised."""
    import os
    import warnings
    dirs = [os.environ.get('B200SEG_PRETRAINED_DIR'), os.path.join(torch.hub.get_dir(), 'checkpoints')]
    for d in dirs:
        path = os.path.join(d, filename) if d else None
        if path and os.path.exists(path):
            return path
    warnings.warn('%s(pretrained=True): no local ImageNet checkpoint (%s under $B200SEG_PRETRAINED_DIR or %s) and no '
                  'network — the backbone is RANDOMLY INITIALISED, unlike the reference\'s' % (arch, filename, dirs[1]),
                  RuntimeWarning, stacklevel=4)
    return None


def resnet50(pretrained=False, replace_stride_with_dilation=None, **kw):
    """Drop-in for ``pytorch_modules.backbones.resnet50`` (models/deeplabv3plus.py:17-19)."""
    net = ResNet50Features(replace_stride_with_dilation)
    if pretrained:
        load_pretrained(net, 'resnet50', 'resnet50-0676ba61.pth')
    return net
