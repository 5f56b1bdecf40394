"""Loss and metric functions of the hot path — drop-ins for utils/utils.py:17-24 and :51-65."""
import random

import torch

from .. import engine as E

IGNORE_INDEX = -100      # nn.CrossEntropyLoss() default (utils/utils.py:12)


class _CrossEntropyFn(torch.autograd.Function):
    """bilinear(align_corners=True) resize to the target size + mean softmax cross-entropy."""

    @staticmethod
    def forward(ctx, outputs, targets):
        E._require_cuda(outputs, 'compute_loss(outputs)')
        E._require_cuda(targets, 'compute_loss(targets)')
        if targets.dtype not in (torch.int64, torch.int32, torch.uint8):
            raise RuntimeError('compute_loss: expected int64 (or int32 / uint8) targets, got %s' % targets.dtype)
        x = outputs.detach()
        if x.dtype != torch.float32 or not x.is_contiguous():
            x = x.float().contiguous()
        th, tw = targets.shape[1], targets.shape[2]
        resized = (x.shape[2], x.shape[3]) != (th, tw)
        if resized:          # utils/utils.py:18-20; identity (skipped) when the sizes match
            x = E.resize_nchw_fwd(x, th, tw, align_corners=True)
        need = ctx.needs_input_grad[0]
        targets = targets.contiguous()
        # forward reads the logits once and writes nothing but three scalars; the gradient is recomputed in backward
        # with the upstream gradient and the pixel count read on the device (one pass, no scaling pass)
        acc, _ = E.softmax_ce(x, targets, IGNORE_INDEX, want_grad=False)
        count_i = acc[1:2].view(torch.int64)
        # mean over the valid pixels; an out-of-range target raises in the reference, here it poisons the loss with NaN
        # (no host sync)
        loss = E.ce_finalize(acc)
        if need:
            ctx.save_for_backward(x, targets, count_i)
            ctx.resized, ctx.in_hw, ctx.in_dtype = resized, (outputs.shape[2], outputs.shape[3]), outputs.dtype
        return loss

    @staticmethod
    def backward(ctx, g):
        x, targets, count_i = ctx.saved_tensors
        g = g.detach().to(torch.float32).contiguous()
        dl = E.softmax_ce_bwd(x, targets, g, count_i, IGNORE_INDEX)
        if ctx.resized:
            dl = E.resize_nchw_bwd(dl, ctx.in_hw[0], ctx.in_hw[1], align_corners=True)
        return dl.to(ctx.in_dtype), None


class _ClassMapCrossEntropyFn(torch.autograd.Function):
    """The same loss taken from the head's low-resolution fp32 class map: the logits the model returned are that map's
    bilinear upsample, so loss and gradient are computed through the interpolation without reading the logits or
    writing their gradient (csrc/loss_metrics.cu, lowres_ce_kernel)."""

    @staticmethod
    def forward(ctx, lr, targets, nc, align_corners):
        if ctx.needs_input_grad[0] and E.FUSED_CE:
            # training: loss and (unscaled) class-map gradient from ONE pass over the target tiles; backward only scales
            acc, dlr = E.lowres_ce_fused(lr.detach(), nc, targets, align_corners, IGNORE_INDEX)
            ctx.save_for_backward(dlr, acc[1:2].view(torch.int64))
            ctx.fused = True
            return E.ce_finalize(acc)
        acc = E.lowres_ce_fwd(lr.detach(), nc, targets, align_corners, IGNORE_INDEX)
        if ctx.needs_input_grad[0]:
            ctx.save_for_backward(lr.detach(), targets, acc[1:2].view(torch.int64))
            ctx.nc, ctx.ac, ctx.fused = nc, align_corners, False
        return E.ce_finalize(acc)

    @staticmethod
    def backward(ctx, g):
        g = g.detach().to(torch.float32).contiguous()
        if ctx.fused:
            dlr, count_i = ctx.saved_tensors
            if getattr(ctx, 'scaled', False):
                raise RuntimeError('compute_loss: backward through the class-map loss a second time')
            ctx.scaled = True                   # the saved gradient is scaled in place
            return E.lowres_ce_scale(dlr, g, count_i), None, None, None
        lr, targets, count_i = ctx.saved_tensors
        return E.lowres_ce_bwd(lr, ctx.nc, targets, g, count_i, ctx.ac, IGNORE_INDEX), None, None, None


def compute_loss(outputs, targets, model=None):
    """utils/utils.py:17-24: logits N x nc x h x w, int64 targets N x H x W -> 0-dim loss tensor."""
    cm = getattr(outputs, '_b200seg_class_map', None)
    if (cm is not None and cm[3] == outputs._version and targets.is_cuda
            and targets.dtype in (torch.int64, torch.int32, torch.uint8)
            and tuple(outputs.shape[2:]) == tuple(targets.shape[1:]) and E.LOWRES_CE
            and E.lowres_ce_supported(cm[0].shape, cm[1], targets.shape[1], targets.shape[2], cm[2])):
        return _ClassMapCrossEntropyFn.apply(cm[0], targets.contiguous(), cm[1], cm[2])
    E.materialize_logits(outputs)           # a head may have deferred writing them (engine.DEFER_LOGITS)
    return _CrossEntropyFn.apply(outputs, targets)


def post_fetch_fn(batch, mean=None, std=None, multi_scale=False, resize_masks=False):
    """Drop-in for ``BasicDataset.post_fetch_fn`` (utils/datasets.py:199-213) on device-resident batches:
    ``(uint8 N x 3 x H x W images, uint8 N x H x W masks)`` -> ``(images, masks)`` ready for ``model(images)`` /
    ``compute_loss(outputs, masks, model)``.

    The reference computes ``imgs.float(); imgs -= mean; imgs /= std`` and ``segs.long()``: 4-byte pixels and 8-byte
    labels.  Here both stay 1-byte: the images are returned as they are, tagged with the normalisation constants, and
    the first kernel that touches them — the stem's space-to-depth pass for DeepLabV3+ / HRNet
    (``engine.space_to_depth_u8``), ``engine.normalize_u8`` for any other backbone — applies the same fp32 arithmetic;
    loss and metric kernels read uint8 masks directly.  float images pass through unchanged.

    ``multi_scale`` (utils/datasets.py:206-212): one ``random.uniform(0.7, 1.5)`` draw per batch (the same call on the
    same ``random`` state as the reference), sizes rounded down to multiples of 32, nearest ``F.interpolate`` — done on
    the 1-byte images, which commutes with the normalisation bit for bit.  Like the reference it resizes the images only
    (its masks keep their size, which its own loss then rejects); ``resize_masks=True`` resizes the masks the same way."""
    imgs, segs = batch
    if multi_scale:
        h, w = imgs.shape[2], imgs.shape[3]
        scale = random.uniform(0.7, 1.5)
        h, w = int(h * scale / 32) * 32, int(w * scale / 32) * 32
        if imgs.dtype == torch.uint8 and imgs.is_cuda:
            imgs = E.resize_nearest_u8(imgs, h, w)
            if resize_masks and segs.dtype == torch.uint8:
                segs = E.resize_nearest_u8(segs, h, w)
            elif resize_masks:
                segs = torch.nn.functional.interpolate(segs[:, None].float(), (h, w))[:, 0].to(segs.dtype)
        else:
            imgs = torch.nn.functional.interpolate(imgs.float(), (h, w))
            if resize_masks:
                segs = torch.nn.functional.interpolate(segs[:, None].float(), (h, w))[:, 0].to(segs.dtype)
    if imgs.dtype == torch.uint8:
        imgs._b200seg_norm = (tuple(E.INPUT_MEAN if mean is None else mean), tuple(E.INPUT_STD if std is None else std))
    return imgs, segs


def model_input(x):
    """What a model's forward does with its input first: uint8 images (``post_fetch_fn`` above) become the normalised
    fp32 NCHW tensor of the reference; floats pass through.  (DeepLabV3+'s stem bypasses this and fuses the
    normalisation into its space-to-depth pass.)"""
    if x.dtype == torch.uint8:
        mean, std = getattr(x, '_b200seg_norm', (None, None))
        return E.normalize_u8(x, mean, std)
    return x


def compute_metrics(tp, fn, fp):
    """utils/utils.py:51-65, same float32 op order (so identical integer counts give bit-identical
    IoU / P / R / F1).  Host-side vectors of length num_classes; inputs are not modified."""
    union = tp + fp + fn
    union[union <= 0] = 1
    miou = tp / union
    T = tp + fn
    P = tp + fp
    P[P <= 0] = 1
    P = tp / P
    R = tp + fn
    R[R <= 0] = 1
    R = tp / R
    F1 = (2 * tp + fp + fn)
    F1[F1 <= 0] = 1
    F1 = 2 * tp / F1
    return T, P, R, miou, F1


def counts_from_confusion(cm):
    """tp / fn / fp of test.py:38-46 from the confusion matrix CM[target, predicted]."""
    tp = cm.diag()
    return tp, cm.sum(1) - tp, cm.sum(0) - tp
