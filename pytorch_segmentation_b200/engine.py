"""Host side of the hot path: NHWC bf16 activation slices and one thin wrapper per C-ABI entry point.

PyTorch is used for device memory (caching allocator), streams and autograd plumbing only; every
arithmetic step below is a call into libb200seg.so on ``torch.cuda.current_stream()``.  There is no
CPU / cuDNN fallback: tensors that are not on a CUDA device raise.
"""
import ctypes

import torch

from . import _lib
from ._lib import ConvOut, ConvSrc, WgradDesc

BF16 = torch.bfloat16


# torch.cuda.current_stream() costs ~4 us of Python per call (device-index resolution, a Stream object) and an eager
# training step asks ~500 times; the raw handle is one C call.
_raw_stream = torch._C._cuda_getCurrentRawStream
_cur_device = torch._C._cuda_getDevice
_stream_objs = {}


def _stream():
    return ctypes.c_void_p(_raw_stream(_cur_device()))


def current_stream():
    """torch.cuda.current_stream(), with the Stream objects cached by (device, raw handle)."""
    dev = _cur_device()
    key = (dev, _raw_stream(dev))
    st = _stream_objs.get(key)
    if st is None:
        st = _stream_objs[key] = torch.cuda.current_stream()
    return st


def _require_cuda(t, what):
    if not t.is_cuda:
        raise RuntimeError('%s: tensor is on %s — pytorch_segmentation_b200 runs on CUDA (sm_100a) only and has '
                           'no CPU fallback' % (what, t.device))


class Act:
    """A channel slice [c_off, c_off + c) of an NHWC bf16 buffer ``t`` of shape [n, h, w, c_total]."""
    __slots__ = ('t', 'c_off', 'c', 'bits')

    def __init__(self, t, c_off=0, c=None):
        assert t.dim() == 4 and t.dtype == BF16 and t.is_contiguous(), (t.shape, t.dtype, t.stride())
        self.t, self.c_off = t, c_off
        self.bits = None          # optional uint8 [pixels, c/8] sign mask written by bn_apply(want_bits=True)
        self.c = t.shape[3] - c_off if c is None else c
        assert self.c_off % 8 == 0 and self.c % 8 == 0 and t.shape[3] % 8 == 0, \
            'channel counts and offsets must be multiples of 8 (got c=%d off=%d total=%d)' % (self.c, c_off, t.shape[3])

    n = property(lambda s: s.t.shape[0])
    h = property(lambda s: s.t.shape[1])
    w = property(lambda s: s.t.shape[2])
    ct = property(lambda s: s.t.shape[3])
    pixels = property(lambda s: s.t.shape[0] * s.t.shape[1] * s.t.shape[2])
    ptr = property(lambda s: s.t.data_ptr())

    def slice(self, off, c):
        return Act(self.t, self.c_off + off, c)

    @staticmethod
    def empty(n, h, w, c, device):
        return Act(torch.empty((n, h, w, c), dtype=BF16, device=device))

    def nchw(self):
        """NCHW-shaped (channels_last-strided) view of the slice — what module boundaries hand out."""
        v = self.t.permute(0, 3, 1, 2)
        return v if (self.c_off == 0 and self.c == self.ct) else v[:, self.c_off:self.c_off + self.c]


class WindowAct(Act):
    """An anchored window source (include/b200seg.h, b200seg_conv_src.x_rows): ``c`` is the window length in elements
    and may exceed the buffer's channels per pixel — the window runs on into the following pixels.  Only conv_fprop /
    conv_wgrad take it."""
    __slots__ = ('algo_k',)

    def __init__(self, t, window, algo_k=None):
        assert t.dim() == 4 and t.dtype == BF16 and t.is_contiguous() and window % 8 == 0 and t.shape[3] % 8 == 0
        self.t, self.c_off, self.c = t, 0, window
        self.algo_k = algo_k        # real multiply-adds per output element and output channel (profiling only)


def as_act(x, what='input'):
    """NCHW-shaped tensor (any float dtype / memory format) -> contiguous NHWC bf16 Act.
    bf16 channels_last input is a zero-copy view; fp32 NCHW goes through the layout kernel."""
    _require_cuda(x, what)
    if x.dim() != 4:
        raise ValueError('%s: expected an N x C x H x W tensor, got shape %s' % (what, tuple(x.shape)))
    n, c, h, w = x.shape
    if c % 8:
        raise ValueError('%s: %d channels — the tcgen05 conv path needs a multiple of 8' % (what, c))
    if x.dtype == BF16 and x.permute(0, 2, 3, 1).is_contiguous():
        return Act(x.permute(0, 2, 3, 1))
    if x.dtype == torch.float32 and x.is_contiguous():
        out = Act.empty(n, h, w, c, x.device)
        _lib.check(_lib.load().b200seg_nchw_to_nhwc(x.data_ptr(), out.ptr, n, c, h, w, c, 0, _stream()),
                   'nchw_to_nhwc')
        _lib.count_launch()
        return out
    # other dtypes / strides: plumbing copy
    return Act(x.detach().to(BF16).permute(0, 2, 3, 1).contiguous())


def grad_like(x, g):
    """Gradient Act ``g`` (full buffer) -> tensor shaped/typed like the forward input ``x``."""
    if x.dtype == BF16:
        return g.nchw()
    if x.dtype == torch.float32:
        n, c, h, w = x.shape
        out = torch.empty((n, c, h, w), dtype=torch.float32, device=x.device)
        _lib.check(_lib.load().b200seg_nhwc_to_nchw(g.ptr, out.data_ptr(), n, c, h, w, g.ct, g.c_off, _stream()),
                   'nhwc_to_nchw')
        _lib.count_launch()
        return out
    return g.nchw().to(x.dtype)


# ------------------------------------------------------------------------------------------------
# forward-value capture (parity tests)
# ------------------------------------------------------------------------------------------------
# tests/test_gpu_gradparity.py sets CAPTURE to a dict: every Conv->BN boundary then leaves a copy of the tensor the
# kernels actually stored (the bf16-rounded raw conv output, and the BN/ReLU/residual output) under
# (id(conv weight parameter), 'raw' | 'out').  The fp32 CPU oracle consumes those values at the same boundaries, so both
# sides differentiate the same piecewise-linear function (same ReLU masks, same BN statistics) and gradients can be
# compared to the 2e-2 tolerance instead of to a bf16 mask-flip floor.  None in normal operation (one global lookup).
CAPTURE = None


def capture(param, kind, value):
    if CAPTURE is not None:
        t = value.nchw() if isinstance(value, Act) else value
        CAPTURE[(id(param), kind)] = t.detach().float().clone()


def capture_seq(family, value):
    """The same for tensors no parameter names (the fused sums of HRNet's branches): keyed by call order within
    ``family`` — CAPTURE[('seq', family)] is the list, in the order the forward produced them."""
    if CAPTURE is not None:
        t = value.nchw() if isinstance(value, Act) else value
        CAPTURE.setdefault(('seq', family), []).append(t.detach().float().clone())


# ------------------------------------------------------------------------------------------------
# weights
# ------------------------------------------------------------------------------------------------
def _round_up(a, b):
    return (a + b - 1) // b * b


# bumped by TrainStep after every CUDA-graph replay (parameters change without their version counters moving)
WEIGHT_EPOCH = 0


class PackedWeight:
    """bf16 packings of an fp32 OIHW conv weight for fprop ([taps][k_rows][c]) and dgrad
    ([taps mirrored][c_rows][k_cols]); refreshed when the parameter's version changes.  The buffers are allocated
    once per shape and rewritten in place (stream order makes that safe; a CUDA graph needs it)."""
    __slots__ = ('k', 'c', 'r', 's', 'k_rows', 'c_rows', 'k_cols', 'wf', 'wd', 'key', '_shape')
    im2col = False

    def __init__(self):
        self.key = None
        self.wf = self.wd = self._shape = None

    def _key(self, w):
        return (w.data_ptr(), w._version, tuple(w.shape), WEIGHT_EPOCH)

    def fresh(self, w, need_dgrad=True):
        return self.key == self._key(w) and (self.wd is not None or not need_dgrad)

    def prepare(self, w, need_dgrad=True):
        """Allocates the buffers for ``w`` if needed and returns (source tensor, filled PackDesc)."""
        _require_cuda(w, 'conv weight')
        src = w.detach()
        if src.dtype != torch.float32 or not src.is_contiguous():
            src = src.float().contiguous()
        k, c, r, s = src.shape
        if c % 8:
            raise ValueError('conv weight with %d input channels: need a multiple of 8' % c)
        if self._shape != (k, c, r, s, src.device):
            self.k, self.c, self.r, self.s = k, c, r, s
            self.k_rows, self.c_rows, self.k_cols = _round_up(k, 8), _round_up(c, 8), _round_up(k, 8)
            self.wf = torch.empty((r * s, self.k_rows, c), dtype=BF16, device=src.device)
            self.wd = None
            self._shape = (k, c, r, s, src.device)
        if need_dgrad and self.wd is None:
            self.wd = torch.empty((r * s, self.c_rows, self.k_cols), dtype=BF16, device=src.device)
        d = _lib.PackDesc(src.data_ptr(), k, c, r, s, self.wf.data_ptr(),
                          self.wd.data_ptr() if self.wd is not None else None, self.k_rows, self.c_rows, self.k_cols,
                          0, 0, 0)
        return src, d

    def update(self, w, need_dgrad=True):
        if self.fresh(w, need_dgrad):
            return self
        src, d = self.prepare(w, need_dgrad)
        _lib.check(_lib.load().b200seg_pack_weights(d.w, d.k, d.c, d.r, d.s, d.w_fprop, d.k_rows, d.w_dgrad, d.c_rows,
                                                    d.k_cols, _stream()), 'pack_weights')
        _lib.count_launch()
        self.key = self._key(w)
        return self


class PackedIm2colWeight:
    """bf16 [1][k_rows][c_pad] packing of an fp32 OIHW weight for the patch-gather path (few-channel stems): the
    conv is a 1x1 GEMM over the gathered matrix, so this quacks like a PackedWeight with r = s = 1 and c = c_pad."""
    __slots__ = ('k', 'c', 'r', 's', 'k_rows', 'wf', 'wd', 'key', 'cin', 'kr', 'ks', '_shape')
    im2col = True

    def __init__(self):
        self.key = None
        self.wf = self.wd = self._shape = None

    def _key(self, w):
        return (w.data_ptr(), w._version, tuple(w.shape), WEIGHT_EPOCH)

    def fresh(self, w, need_dgrad=False):
        return self.key == self._key(w)

    def prepare(self, w, need_dgrad=False):
        _require_cuda(w, 'conv weight')
        src = w.detach()
        if src.dtype != torch.float32 or not src.is_contiguous():
            src = src.float().contiguous()
        k, cin, kr, ks = src.shape
        if self._shape != (k, cin, kr, ks, src.device):
            self.k, self.cin, self.kr, self.ks = k, cin, kr, ks
            self.r = self.s = 1
            self.c = _round_up(cin * kr * ks, 8)
            self.k_rows = _round_up(k, 8)
            self.wf = torch.empty((1, self.k_rows, self.c), dtype=BF16, device=src.device)
            self._shape = (k, cin, kr, ks, src.device)
        d = _lib.PackDesc(src.data_ptr(), k, cin, kr, ks, self.wf.data_ptr(), None, self.k_rows, 0, 0, 1, self.c, 0)
        return src, d

    def update(self, w, need_dgrad=False):
        if self.fresh(w):
            return self
        src, d = self.prepare(w)
        _lib.check(_lib.load().b200seg_pack_weights_im2col(d.w, d.k, d.c, d.r, d.s, d.w_fprop, d.k_rows, d.c_pad,
                                                           _stream()), 'pack_weights_im2col')
        _lib.count_launch()
        self.key = self._key(w)
        return self


class PackedS2DWeight:
    """bf16 [ceil(r/2)][k_rows][ceil(s/2)*cs] packing of a stride-2 few-channel OIHW conv weight for the
    space-to-depth form (csrc/im2col.cu): quacks like a PackedWeight with ceil(r/2) x 1 taps over windows of
    ceil(s/2)*cs elements."""
    __slots__ = ('k', 'c', 'r', 's', 'k_rows', 'wf', 'wd', 'key', 'cin', 'kr', 'ks', 'cs', '_shape')

    def __init__(self):
        self.key = None
        self.wf = self.wd = self._shape = None

    @staticmethod
    def applies(conv):
        return (conv.stride == (2, 2) and conv.in_channels <= 4 and conv.groups == 1 and conv.dilation == (1, 1)
                and conv.bias is None and (conv.kernel_size[0] + 1) // 2 <= 9)

    def update(self, w):
        _require_cuda(w, 'conv weight')
        key = (w.data_ptr(), w._version, tuple(w.shape), WEIGHT_EPOCH)
        if key == self.key:
            return self
        src = w.detach()
        if src.dtype != torch.float32 or not src.is_contiguous():
            src = src.float().contiguous()
        k, cin, kr, ks = src.shape
        if self._shape != (k, cin, kr, ks, src.device):
            self.k, self.cin, self.kr, self.ks = k, cin, kr, ks
            self.cs = _round_up(4 * cin, 8)
            self.r, self.s = (kr + 1) // 2, 1
            self.c = (ks + 1) // 2 * self.cs
            self.k_rows = _round_up(k, 8)
            self.wf = torch.empty((self.r, self.k_rows, self.c), dtype=BF16, device=src.device)
            self._shape = (k, cin, kr, ks, src.device)
        _lib.check(_lib.load().b200seg_pack_weights_s2d(src.data_ptr(), k, cin, kr, ks, self.wf.data_ptr(), self.k_rows,
                                                        self.cs, _stream()), 'pack_weights_s2d')
        _lib.count_launch()
        self.key = key
        return self


class _GemmView:
    """What conv_fprop / conv_dgrad read from a packed weight, for a 1x1 GEMM over a re-arranged operand."""
    __slots__ = ('wf', 'wd', 'k_rows', 'c_rows', 'k_cols', 'r', 's', 'k', 'c')

    def __init__(self, wf=None, wd=None, k_rows=0, c_rows=0, k_cols=0, k=0, c=0):
        self.wf, self.wd, self.k_rows, self.c_rows, self.k_cols, self.k, self.c = wf, wd, k_rows, c_rows, k_cols, k, c
        self.r = self.s = 1


class PackedTapWeight:
    """Packings of a narrow-output 3x3 conv weight (cls_conv) for the taps-as-channels form (csrc/taps.cu):
    ``fprop`` = Wz [taps*ncp][c] (the 1x1 GEMM that produces all per-tap partial class maps at once), ``dgrad`` =
    Wcol [c][taps*ncp] (the 1x1 GEMM over the tap-gathered gradient)."""
    __slots__ = ('k', 'c', 'r', 's', 'ncp', 'zc', 'wz', 'wcol', 'key', '_shape', 'fprop', 'dgrad')
    MAX_COLS = 512

    def __init__(self):
        self.key = None
        self.wz = self.wcol = self._shape = None

    @staticmethod
    def applies(conv):
        k, c, r, s = conv.weight.shape
        ncp = _round_up(k, 8)
        return (r, s) == (3, 3) and conv.stride == (1, 1) and c % 8 == 0 and ncp <= 32 and c >= 4 * ncp

    def update(self, w):
        _require_cuda(w, 'conv weight')
        key = (w.data_ptr(), w._version, tuple(w.shape), WEIGHT_EPOCH)
        if key == self.key:
            return self
        src = w.detach()
        if src.dtype != torch.float32 or not src.is_contiguous():
            src = src.float().contiguous()
        k, c, r, s = src.shape
        if self._shape != (k, c, r, s, src.device):
            self.k, self.c, self.r, self.s = k, c, r, s
            self.ncp = _round_up(k, 8)
            self.zc = r * s * self.ncp
            self.wz = torch.empty((1, self.zc, c), dtype=BF16, device=src.device)
            self.wcol = torch.empty((1, c, self.zc), dtype=BF16, device=src.device)
            self.fprop = _GemmView(wf=self.wz, k_rows=self.zc, k=self.zc, c=c)
            self.dgrad = _GemmView(wd=self.wcol, c_rows=c, k_cols=self.zc, k=self.zc, c=c)
            self._shape = (k, c, r, s, src.device)
        _lib.check(_lib.load().b200seg_pack_weights_taps(src.data_ptr(), k, c, r, s, self.ncp, self.wz.data_ptr(), self.zc,
                                                         self.wcol.data_ptr(), c, _stream()), 'pack_weights_taps')
        _lib.count_launch()
        self.key = key
        return self


def tap_gather(x, r, s, dil):
    """col[p][tap*c + ch] = x[p + off(tap)][ch] -> Act with r*s*c channels."""
    col = Act.empty(x.n, x.h, x.w, r * s * x.c, x.t.device)
    _lib.check(_lib.load().b200seg_tap_gather(x.ptr, x.ct, x.c_off, x.c, x.n, x.h, x.w, r, s, dil, col.ptr, col.ct,
                                              _stream()), 'tap_gather')
    _lib.count_launch()
    return col


def tap_scatter_add(z, ncp, r, s, dil, bias=None):
    """y[p][k] = bias[k] + sum_tap z[p + off(tap)][tap*ncp + k] -> fp32 NHWC tensor [n, h, w, ncp]."""
    assert z.c_off == 0 and z.c == z.ct
    y = torch.empty((z.n, z.h, z.w, ncp), dtype=torch.float32, device=z.t.device)
    _lib.check(_lib.load().b200seg_tap_scatter_add(z.ptr, z.ct, ncp, z.n, z.h, z.w, r, s, dil,
                                                   bias.data_ptr() if bias is not None else None, y.data_ptr(), _stream()),
               'tap_scatter_add')
    _lib.count_launch()
    return y


def unpack_wgrad_taps(dwcol, pw, lane=False):
    return on_wgrad_lane(_unpack_wgrad_taps, dwcol, pw, keep=(dwcol,), chain=True) if lane else _unpack_wgrad_taps(dwcol, pw)


def _unpack_wgrad_taps(dwcol, pw):
    dw = torch.empty((pw.k, pw.c, pw.r, pw.s), dtype=torch.float32, device=dwcol.device)
    _lib.check(_lib.load().b200seg_unpack_wgrad_taps(dwcol.data_ptr(), pw.k, pw.c, pw.r, pw.s, pw.ncp, dw.data_ptr(), 0,
                                                     _stream()), 'unpack_wgrad_taps')
    _lib.count_launch()
    return dw


def pack_all(entries, need_dgrad=True):
    """entries: iterable of (PackedWeight | PackedIm2colWeight, weight parameter).  Re-packs every stale one in a
    single multi-tensor launch (after an optimizer step that is all of them) — the per-layer ``update`` calls that
    follow find fresh keys and do nothing."""
    stale = [(pw, w) for pw, w in entries if not pw.fresh(w, need_dgrad)]
    if not stale:
        return 0
    arr = (_lib.PackDesc * len(stale))()
    keep = []
    for i, (pw, w) in enumerate(stale):
        src, arr[i] = pw.prepare(w, need_dgrad)
        keep.append(src)
    _lib.check(_lib.load().b200seg_pack_weights_multi(arr, len(stale), _stream()), 'pack_weights_multi')
    _lib.count_launch((len(stale) + 47) // 48)
    for pw, w in stale:
        pw.key = pw._key(w)
    return len(stale)


def packed_entries(model):
    """(packed, weight) pairs of every module under ``model`` that owns packed conv weights."""
    out = []
    for m in model.modules():
        fn = getattr(m, '_packed_weights', None)
        if fn is not None:
            out.extend(fn())
    return out


def space_to_depth(x, r, s, pad):
    """fp32 NCHW image (c <= 4) -> WindowAct over the bf16 space-to-depth tensor of the padded image: the operand of a
    stride-2 r x s conv as ceil(r/2) row taps (csrc/im2col.cu)."""
    _require_cuda(x, 'space_to_depth input')
    if x.dim() != 4 or x.dtype != torch.float32 or not x.is_contiguous():
        raise ValueError('space_to_depth: expected a contiguous fp32 N x C x H x W tensor, got %s %s' % (tuple(x.shape), x.dtype))
    n, c, h, w = x.shape
    oh, ow = (h + 2 * pad - r) // 2 + 1, (w + 2 * pad - s) // 2 + 1
    cs = _round_up(4 * c, 8)
    z = torch.empty((n, oh + (r + 1) // 2 - 1, ow + (s + 1) // 2 - 1, cs), dtype=BF16, device=x.device)
    _lib.check(_lib.load().b200seg_space_to_depth_nchw(x.data_ptr(), n, c, h, w, r, s, pad, z.data_ptr(), cs, _stream()),
               'space_to_depth')
    _lib.count_launch()
    return WindowAct(z, (s + 1) // 2 * cs, algo_k=r * s * c), oh, ow


# utils/datasets.py:201-205 — the reference's normalisation constants (ImageNet mean / std on the 0..255 scale)
INPUT_MEAN = (123.675, 116.28, 103.53)
INPUT_STD = (58.395, 57.12, 57.375)


def _norm_args(c, mean, std):
    mean = tuple(INPUT_MEAN if mean is None else mean)
    std = tuple(INPUT_STD if std is None else std)
    if len(mean) != c or len(std) != c:
        raise ValueError('normalisation constants for %d channels, image has %d' % (len(mean), c))
    return (ctypes.c_float * c)(*mean), (ctypes.c_float * c)(*std)


def _u8_image(x, what):
    _require_cuda(x, what)
    if x.dim() != 4 or x.dtype != torch.uint8:
        raise ValueError('%s: expected an N x C x H x W uint8 tensor, got %s %s' % (what, tuple(x.shape), x.dtype))
    if x.is_contiguous():
        return x, 0
    if x.permute(0, 2, 3, 1).is_contiguous():
        return x, 1          # channels_last uint8 (what cv2 decodes to): read in place
    return x.contiguous(), 0


def normalize_u8(x, mean=None, std=None):
    """uint8 N x C x H x W (NCHW or channels_last) -> fp32 NCHW (x - mean) / std: post_fetch_fn of utils/datasets.py:199-206
    on the device, bit-identical to the reference's three fp32 tensor ops."""
    x, layout = _u8_image(x, 'normalize_u8')
    n, c, h, w = x.shape
    m, s = _norm_args(c, mean, std)
    y = torch.empty((n, c, h, w), dtype=torch.float32, device=x.device)
    _lib.check(_lib.load().b200seg_input_u8_normalize(x.data_ptr(), layout, n, c, h, w, m, s, y.data_ptr(), _stream()),
               'input_u8_normalize')
    _lib.count_launch()
    return y


def resize_nearest_u8(x, oh, ow):
    """uint8 [..., H, W] (contiguous) -> [..., oh, ow], F.interpolate(mode='nearest') indexing (csrc/input.cu)."""
    _require_cuda(x, 'resize_nearest_u8 input')
    if x.dtype != torch.uint8 or x.dim() < 2:
        raise ValueError('resize_nearest_u8: expected a uint8 tensor [..., H, W], got %s %s' % (x.dtype, tuple(x.shape)))
    x = x.contiguous()
    h, w = x.shape[-2:]
    y = torch.empty(tuple(x.shape[:-2]) + (int(oh), int(ow)), dtype=torch.uint8, device=x.device)
    planes = x.numel() // (h * w)
    _lib.check(_lib.load().b200seg_resize_nearest_u8(x.data_ptr(), planes, h, w, y.data_ptr(), int(oh), int(ow), _stream()),
               'resize_nearest_u8')
    _lib.count_launch()
    return y


def space_to_depth_u8(x, r, s, pad, mean=None, std=None):
    """uint8 image batch -> the stem's WindowAct in one pass: normalise + bf16 + NHWC + 2x2 space-to-depth (csrc/input.cu)."""
    x, layout = _u8_image(x, 'space_to_depth_u8')
    n, c, h, w = x.shape
    m, sd = _norm_args(c, mean, std)
    oh, ow = (h + 2 * pad - r) // 2 + 1, (w + 2 * pad - s) // 2 + 1
    cs = _round_up(4 * c, 8)
    z = torch.empty((n, oh + (r + 1) // 2 - 1, ow + (s + 1) // 2 - 1, cs), dtype=BF16, device=x.device)
    _lib.check(_lib.load().b200seg_input_u8_s2d(x.data_ptr(), layout, n, c, h, w, r, s, pad, m, sd, z.data_ptr(), cs,
                                                _stream()), 'input_u8_s2d')
    _lib.count_launch()
    return WindowAct(z, (s + 1) // 2 * cs, algo_k=r * s * c), oh, ow


def unpack_wgrad_s2d(dwin, pw, lane=False):
    return on_wgrad_lane(_unpack_wgrad_s2d, dwin, pw, keep=(dwin,), chain=True) if lane else _unpack_wgrad_s2d(dwin, pw)


def _unpack_wgrad_s2d(dwin, pw):
    dw = torch.empty((pw.k, pw.cin, pw.kr, pw.ks), dtype=torch.float32, device=dwin.device)
    _lib.check(_lib.load().b200seg_unpack_wgrad_s2d(dwin.data_ptr(), pw.k, pw.cin, pw.kr, pw.ks, pw.cs, dw.data_ptr(), 0,
                                                    _stream()), 'unpack_wgrad_s2d')
    _lib.count_launch()
    return dw


def im2col(x, r, s, stride, pad):
    """fp32 NCHW image -> Act [n, oh, ow, round_up(r*s*c, 8)] of gathered patches (bf16)."""
    _require_cuda(x, 'im2col input')
    if x.dim() != 4 or x.dtype != torch.float32 or not x.is_contiguous():
        raise ValueError('im2col: expected a contiguous fp32 N x C x H x W tensor, got %s %s' % (tuple(x.shape), x.dtype))
    n, c, h, w = x.shape
    oh, ow = (h + 2 * pad - r) // stride + 1, (w + 2 * pad - s) // stride + 1
    col = Act.empty(n, oh, ow, _round_up(r * s * c, 8), x.device)
    _lib.check(_lib.load().b200seg_im2col_nchw(x.data_ptr(), n, c, h, w, r, s, stride, pad, col.ptr, col.c, _stream()),
               'im2col')
    _lib.count_launch()
    return col


def unpack_wgrad_im2col(dwcol, pw, lane=False):
    """fp32 [k, c_pad, 1, 1] gradient of the gathered-matrix GEMM -> fp32 OIHW gradient of the original filter
    (on the weight-gradient lane, behind the wgrad that produced ``dwcol``)."""
    return on_wgrad_lane(_unpack_wgrad_im2col, dwcol, pw, keep=(dwcol,), chain=True) if lane else _unpack_wgrad_im2col(dwcol, pw)


def _unpack_wgrad_im2col(dwcol, pw):
    dw = torch.empty((pw.k, pw.cin, pw.kr, pw.ks), dtype=torch.float32, device=dwcol.device)
    _lib.check(_lib.load().b200seg_unpack_wgrad_im2col(dwcol.data_ptr(), pw.k, pw.cin, pw.kr, pw.ks, pw.c, dw.data_ptr(), 0,
                                                       _stream()), 'unpack_wgrad_im2col')
    _lib.count_launch()
    return dw


# ------------------------------------------------------------------------------------------------
# convolution
# ------------------------------------------------------------------------------------------------
# bench.py sets this to a list to time every tcgen05 conv launch with CUDA events on the launching stream:
# entries are (kind, nominal dense flops, start event, end event)
PROFILE = None


def _conv(fn, what, n, h, w, srcs, out_ptr, y_ct, y_off, k, out_fp32, scale, bias, relu, stats):
    if PROFILE is not None:
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        # nominal dense flops: fprop counts output pixels; a stride-2 dgrad source has h*w/4 dY pixels
        flops = 2.0 * n * h * w * k * sum((getattr(x, 'algo_k', None) or x.c * r * s) / (st * st if what == 'conv_dgrad' else 1)
                                          for x, _, _, r, s, _, st in srcs)
        tag = '%s %dx%dx%d k%d %s' % (what, n, h, w, k, '+'.join('c%d r%d d%d%s' % (x.c, r, d, ' s2' if st == 2 else '')
                                                              for x, _, _, r, _, d, st in srcs))
        ev0.record()
        _conv_launch(fn, what, n, h, w, srcs, out_ptr, y_ct, y_off, k, out_fp32, scale, bias, relu, stats)
        ev1.record()
        run, dense = ctypes.c_int64(0), ctypes.c_int64(0)
        _lib.load().b200seg_conv_last_plan(ctypes.byref(run), ctypes.byref(dense))
        # algorithmic bytes: every input slice once + the output slice + every packed weight once
        nbytes = n * h * w * k * (4 if out_fp32 else 2)
        for x, _, w_rows, r, s, _, st in srcs:
            px = x.pixels if not isinstance(x, WindowAct) else x.t.numel() // max(x.ct, 1)
            nbytes += 2 * px * min(x.c, x.ct) + 2 * r * s * w_rows * x.c
        PROFILE.append((tag, flops, ev0, ev1, {'bytes': float(nbytes),
                                               'run_share': run.value / dense.value if dense.value else 1.0}))
        return
    _conv_launch(fn, what, n, h, w, srcs, out_ptr, y_ct, y_off, k, out_fp32, scale, bias, relu, stats)


def _conv_launch(fn, what, n, h, w, srcs, out_ptr, y_ct, y_off, k, out_fp32, scale, bias, relu, stats):
    arr = (ConvSrc * len(srcs))()
    for i, (x, wptr, w_rows, r, s, dil, stride) in enumerate(srcs):
        arr[i] = ConvSrc(x.ptr, x.ct, x.c_off, x.c, wptr, w_rows, r, s, dil, stride,
                         *((x.h, x.w) if isinstance(x, WindowAct) else (0, 0)))
    out = ConvOut(out_ptr, y_ct, y_off, k, int(out_fp32),
                  scale.data_ptr() if scale is not None else None,
                  bias.data_ptr() if bias is not None else None, int(relu),
                  stats.data_ptr() if stats is not None else None)
    _lib.check(fn(n, h, w, len(srcs), arr, ctypes.byref(out), _stream()), what)
    # a dgrad with a stride-2 source is one launch per parity sub-grid of dX
    _lib.count_launch(4 if what == 'conv_dgrad' and any(t[6] == 2 for t in srcs) else 1)


def _src(t):
    """(act, pw, dil) or (act, pw, dil, stride) -> (act, pw, dil, stride)"""
    return t if len(t) == 4 else (t[0], t[1], t[2], 1)


def conv_fprop(srcs, out, scale=None, bias=None, relu=False, stats=None):
    """srcs: list of (x: Act, pw: PackedWeight, dil[, stride]).  out: Act (bf16) on the output grid —
    y = act(scale * sum conv + bias); a stride-2 source's x is twice the size of out."""
    srcs = [_src(t) for t in srcs]
    for x, pw, dil, stride in srcs:
        assert isinstance(x, WindowAct) or (x.n, x.h, x.w) == (out.n, out.h * stride, out.w * stride), \
            'conv_fprop: input/output size mismatch'
    _conv(_lib.load().b200seg_conv_fprop, 'conv_fprop', out.n, out.h, out.w,
          [(x, pw.wf.data_ptr(), pw.k_rows, pw.r, pw.s, dil, stride) for x, pw, dil, stride in srcs],
          out.ptr, out.ct, out.c_off, out.c, False, scale, bias, relu, stats)


def conv_fprop_f32(srcs, out_t, k, bias=None):
    """Same, fp32 NHWC output tensor [n, h, w, k] (the class map of cls_conv)."""
    srcs = [_src(t) for t in srcs]
    _conv(_lib.load().b200seg_conv_fprop, 'conv_fprop', out_t.shape[0], out_t.shape[1], out_t.shape[2],
          [(x, pw.wf.data_ptr(), pw.k_rows, pw.r, pw.s, dil, stride) for x, pw, dil, stride in srcs],
          out_t.data_ptr(), out_t.shape[3], 0, k, True, None, bias, False, None)


def conv_dgrad(srcs, dx):
    """srcs: list of (dy: Act, pw: PackedWeight, dil[, stride]); dx: Act receiving the sum of the data gradients
    (a stride-2 source's dy is half the size of dx)."""
    srcs = [_src(t) for t in srcs]
    for dy, pw, dil, stride in srcs:
        assert (dy.n, dy.h * stride, dy.w * stride) == (dx.n, dx.h, dx.w), 'conv_dgrad: gradient size mismatch'
    _conv(_lib.load().b200seg_conv_dgrad, 'conv_dgrad', dx.n, dx.h, dx.w,
          [(dy, pw.wd.data_ptr(), pw.c_rows, pw.r, pw.s, dil, stride) for dy, pw, dil, stride in srcs],
          dx.ptr, dx.ct, dx.c_off, dx.c, False, None, None, False, None)


# ------------------------------------------------------------------------------------------------
# the weight-gradient lane
# ------------------------------------------------------------------------------------------------
# In backward the critical path is  BN-backward -> dgrad -> BN-backward -> ...  (HBM-bound passes alternating with
# tensor-core GEMMs); a layer's weight gradient hangs off that chain and nothing downstream needs it before the
# optimizer.  conv_wgrad therefore runs on a second stream forked after the BN backward that produced its operand:
# the tcgen05 wgrad of layer i overlaps the bandwidth-bound BN backward of layer i-1 (complementary resources), and
# small GEMMs that leave SMs idle overlap each other.  Every autograd Function joins the lane before it returns.
# Operands stay referenced until the join so the caching allocator cannot recycle them under the side stream.
# B200SEG_WGRAD_LANE=0 turns it off (wgrad then runs in line).
import os as _os

_LANE_ON = _os.environ.get('B200SEG_WGRAD_LANE', '1') != '0'
_lanes = {}
# Several lanes, taken round-robin: weight-gradient kernels hold one CTA per SM, and a launch rarely fills the GPU evenly
# (128 CTAs of an 8-cluster launch leave 20 SMs idle; the corner-tap CTAs of a dilated 3x3 finish in a quarter of the
# centre tap's time) — with the next wgrad on another stream its CTAs take those SMs instead of waiting for the tail.
_N_LANES = max(1, int(_os.environ.get('B200SEG_WGRAD_LANES', '1')))      # measured: 1 / 2 / 3 / 4 lanes = 11.66 / 11.64 / 11.64 / 11.64 ms per step: one is enough


class _Lane:
    def __init__(self, device):
        self.stream = torch.cuda.Stream(device=device)
        self.keep = []
        self.dirty = False


class _LaneSet:
    def __init__(self, device):
        self.lanes = [_Lane(device) for _ in range(_N_LANES)]
        self.next = 0
        self.last = self.lanes[0]

    def take(self, chain=False):
        """The next lane in turn; ``chain``: the lane the previous call used (an op that consumes its result)."""
        if not chain:
            self.last = self.lanes[self.next]
            self.next = (self.next + 1) % len(self.lanes)
        return self.last


def _laneset(device):
    if not _LANE_ON:
        return None
    key = device.index if device.index is not None else torch.cuda.current_device()
    ls = _lanes.get(key)
    if ls is None:
        ls = _lanes[key] = _LaneSet(device)
    return ls


def on_wgrad_lane(fn, *args, keep=(), chain=False):
    """Runs ``fn(*args)`` on a weight-gradient lane, ordered after everything enqueued so far on the current stream.
    ``chain=True``: on the same lane as the previous call (its input is that call's output)."""
    dev = next((t.device for t in keep if isinstance(t, torch.Tensor)), None)
    ls = _laneset(dev) if dev is not None else None
    if ls is None:
        return fn(*args)
    ln = ls.take(chain)
    cur = current_stream()
    ln.stream.wait_stream(cur)
    torch.cuda.set_stream(ln.stream)          # (the torch.cuda.stream() context manager costs ~10 us per use)
    try:
        out = fn(*args)
    finally:
        torch.cuda.set_stream(cur)
    ln.keep.extend(keep)
    ln.dirty = True
    return out


def join_wgrad_lane():
    """The current stream waits for the weight-gradient lanes; called before an autograd Function hands its gradients back."""
    for ls in _lanes.values():
        for ln in ls.lanes:
            if ln.dirty:
                current_stream().wait_stream(ln.stream)
                ln.keep.clear()
                ln.dirty = False


def conv_wgrad(dy, x, k_out, r, s, dil, dw=None, stride=1, lane=False):
    """dW fp32 OIHW [k_out, x.c, r, s] = sum_pixels dy (x) shifted x.  ``lane=True`` (the models' backward passes)
    enqueues it on the weight-gradient lane: the result may then be touched again only through ``on_wgrad_lane`` or
    after ``join_wgrad_lane()``."""
    if not lane:
        return _conv_wgrad(dy, x, k_out, r, s, dil, dw, stride)
    return on_wgrad_lane(_conv_wgrad, dy, x, k_out, r, s, dil, dw, stride, keep=(dy.t, x.t) + ((dw,) if dw is not None else ()))


def _conv_wgrad(dy, x, k_out, r, s, dil, dw=None, stride=1):
    desc = WgradDesc()
    if isinstance(x, WindowAct):
        desc.x_rows, desc.x_cols = x.h, x.w
    else:
        assert (x.n, x.h, x.w) == (dy.n, dy.h * stride, dy.w * stride), 'conv_wgrad: size mismatch'
    desc.n, desc.h, desc.w = dy.n, dy.h, dy.w
    desc.dy, desc.dy_c_total, desc.dy_c_offset, desc.k = dy.ptr, dy.ct, dy.c_off, dy.c
    desc.x, desc.x_c_total, desc.x_c_offset, desc.c = x.ptr, x.ct, x.c_off, x.c
    desc.r, desc.s, desc.dil, desc.stride = r, s, dil, stride
    accumulate = dw is not None
    if dw is None:
        dw = torch.empty((k_out, x.c, r, s), dtype=torch.float32, device=x.t.device)
    desc.dw, desc.k_out, desc.accumulate = dw.data_ptr(), k_out, int(accumulate)
    desc.workspace, desc.workspace_bytes = None, 0
    lib = _lib.load()
    # the query validates everything except the workspace itself
    nbytes = lib.b200seg_conv_wgrad_workspace(ctypes.byref(desc))
    if nbytes == 0:
        raise RuntimeError('conv_wgrad: %s' % _lib.last_error())
    ws = torch.empty((nbytes,), dtype=torch.uint8, device=x.t.device)
    desc.workspace, desc.workspace_bytes = ws.data_ptr(), nbytes
    if PROFILE is not None:
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ev0.record()
    _lib.check(lib.b200seg_conv_wgrad(ctypes.byref(desc), _stream()), 'conv_wgrad')
    if PROFILE is not None:
        ev1.record()
        run, dense = ctypes.c_int64(0), ctypes.c_int64(0)
        lib.b200seg_conv_wgrad_plan_info(ctypes.byref(desc), ctypes.byref(run), ctypes.byref(dense))
        px = x.pixels if not isinstance(x, WindowAct) else x.t.numel() // max(x.ct, 1)
        PROFILE.append(('conv_wgrad %dx%dx%d k%d c%d r%d d%d%s' % (dy.n, dy.h, dy.w, dy.c, x.c, r, dil,
                                                                   ' s2' if stride == 2 else ''),
                        2.0 * dy.pixels * dy.c * (getattr(x, 'algo_k', None) or x.c * r * s), ev0, ev1,
                        {'bytes': 2.0 * dy.pixels * dy.c + 2.0 * px * min(x.c, x.ct) + 4.0 * k_out * x.c * r * s,
                         'run_share': run.value / dense.value if dense.value else 1.0}))
    _lib.count_launch(2)
    return dw


# ------------------------------------------------------------------------------------------------
# batch norm
# ------------------------------------------------------------------------------------------------
# ------------------------------------------------------------------------------------------------
# early gradient delivery (data parallelism)
# ------------------------------------------------------------------------------------------------
# A hand-sequenced autograd node (the whole ResNet-50 is one) returns its parameter gradients when its backward ends,
# so a bucketed all-reduce driven by per-parameter hooks would see the whole backbone at once, at the very end of
# backward, and run exposed.  A node calls ``deliver_grads`` whenever a stage's gradients are complete; with a sink
# installed (parallel.GradBucketer.attach) they move into their all-reduce buckets right then and the collectives run
# under the rest of backward.  Without a sink this is a no-op.
GRAD_SINK = None
SINK_ON_LANE = _os.environ.get('B200SEG_SINK_ON_LANE', '1') != '0'
MASK_BITS = _os.environ.get('B200SEG_MASK_BITS', '1') != '0'      # ReLU masks of block outputs as bits (bn_apply / grad_add_masked)


def deliver_grads(grads, params):
    """grads: {parameter: gradient tensor}; params: the parameters whose gradients are final.  The sink may replace
    entries of ``grads`` (by views of its flat buffers)."""
    if GRAD_SINK is None:
        return
    ready = [p for p in params if grads.get(p) is not None]
    if not ready:
        return
    ls = _laneset(ready[0].device) if (SINK_ON_LANE and ready[0].is_cuda) else None
    if ls is None:
        join_wgrad_lane()            # weight gradients live on the side streams until joined
        GRAD_SINK(grads, ready)
        return
    # The weight gradients of this stage are still being computed on the weight-gradient lanes.  Joining them here
    # would stall the compute stream at every stage boundary; instead the sink — the bucket copy and the all-reduce
    # launch — runs ON a lane, behind those wgrads, and only that lane waits (for the other lanes and for the BN
    # parameter gradients the compute stream produced).  The node joins the lanes before it returns.
    ln = ls.lanes[0]
    ln.stream.wait_stream(current_stream())
    for other in ls.lanes[1:]:
        if other.dirty:
            ln.stream.wait_stream(other.stream)
    with torch.cuda.stream(ln.stream):
        GRAD_SINK(grads, ready)
    ln.dirty = True


def bn_stats(x):
    stats = torch.empty((2 * x.c,), dtype=torch.float32, device=x.t.device)
    _lib.check(_lib.load().b200seg_bn_stats(x.ptr, x.ct, x.c_off, x.c, x.pixels, stats.data_ptr(), _stream()),
               'bn_stats')
    _lib.count_launch()
    return stats


def bn_finalize(stats, c, count, gamma, beta, eps, momentum, running_mean, running_var):
    """-> vec[4, c] = (scale, shift, mean, invstd); updates the running statistics in place."""
    vec = torch.empty((4, c), dtype=torch.float32, device=stats.device)
    p = vec.data_ptr()
    _lib.check(_lib.load().b200seg_bn_finalize(
        stats.data_ptr(), c, count, gamma.data_ptr(), beta.data_ptr(), eps, momentum,
        running_mean.data_ptr() if running_mean is not None else None,
        running_var.data_ptr() if running_var is not None else None,
        p, p + 4 * c, p + 8 * c, p + 12 * c, _stream()), 'bn_finalize')
    _lib.count_launch()
    return vec


def bn_finalize_lazy(stats, c, count, gamma, beta, eps, momentum, running_mean, running_var):
    """Like ``bn_finalize`` but defers the work to the ``bn_apply`` that follows (one launch less per layer): the
    returned vec is unfilled and carries what is needed to fill it.  ``bn_resolve`` forces it."""
    ok = (gamma.dtype == torch.float32 and beta.dtype == torch.float32 and gamma.is_contiguous() and beta.is_contiguous()
          and gamma.data_ptr() % 16 == 0 and beta.data_ptr() % 16 == 0 and stats.data_ptr() % 16 == 0 and c % 8 == 0)
    if not ok:
        return bn_finalize(stats, c, count, gamma, beta, eps, momentum, running_mean, running_var)
    vec = torch.empty((4, c), dtype=torch.float32, device=stats.device)
    vec._b200seg_pending = (stats, c, count, gamma, beta, eps, momentum, running_mean, running_var)
    return vec


def bn_resolve(vec):
    """Fill a vec from ``bn_finalize_lazy`` that no ``bn_apply`` consumed yet (the downsample BN of a bottleneck is applied
    inside another layer's pass)."""
    pend = getattr(vec, '_b200seg_pending', None)
    if pend is not None:
        stats, c, count, gamma, beta, eps, momentum, rm, rv = pend
        p = vec.data_ptr()
        _lib.check(_lib.load().b200seg_bn_finalize(
            stats.data_ptr(), c, count, gamma.data_ptr(), beta.data_ptr(), eps, momentum,
            rm.data_ptr() if rm is not None else None, rv.data_ptr() if rv is not None else None,
            p, p + 4 * c, p + 8 * c, p + 12 * c, _stream()), 'bn_finalize')
        _lib.count_launch()
        vec._b200seg_pending = None
    return vec


def bn_eval_affine(c, gamma, beta, running_mean, running_var, eps):
    vec = torch.empty((4, c), dtype=torch.float32, device=gamma.device)
    p = vec.data_ptr()
    _lib.check(_lib.load().b200seg_bn_eval_affine(c, gamma.data_ptr(), beta.data_ptr(), running_mean.data_ptr(),
                                                  running_var.data_ptr(), eps, p, p + 4 * c, p + 8 * c, p + 12 * c,
                                                  _stream()), 'bn_eval_affine')
    _lib.count_launch()
    return vec


def bn_apply(x, y, vec, relu, res=None, res_vec=None, want_bits=False):
    """y = relu?(x*scale + shift [+ res | + res*res_scale + res_shift]).  ``want_bits``: also leave the sign mask of y as
    bits in ``y.bits`` (train-mode fused path, full-buffer outputs only) for ``grad_add_masked``."""
    c = x.c
    p = vec.data_ptr()
    if res_vec is not None:
        bn_resolve(res_vec)
    pend = getattr(vec, '_b200seg_pending', None)
    if pend is not None:
        stats, pc, count, gamma, beta, eps, momentum, rm, rv = pend
        assert pc == c and count == x.pixels, 'bn_apply: statistics of another tensor'
        rp = res_vec.data_ptr() if res_vec is not None else None
        bits = None
        if want_bits and MASK_BITS and y.c_off == 0 and y.c == y.ct:
            bits = torch.empty((y.pixels, c // 8), dtype=torch.uint8, device=y.t.device)
        _lib.check(_lib.load().b200seg_bn_apply_train(
            x.ptr, x.ct, x.c_off, y.ptr, y.ct, y.c_off, c, x.pixels, stats.data_ptr(), gamma.data_ptr(), beta.data_ptr(),
            eps, momentum, rm.data_ptr() if rm is not None else None, rv.data_ptr() if rv is not None else None, p,
            int(relu), res.ptr if res is not None else None, res.ct if res is not None else 0,
            res.c_off if res is not None else 0, rp, rp + 4 * c if rp is not None else None,
            bits.data_ptr() if bits is not None else None, _stream()), 'bn_apply_train')
        y.bits = bits
        vec._b200seg_pending = None
        _lib.count_launch()
        return
    if res is None:
        _lib.check(_lib.load().b200seg_bn_apply(x.ptr, x.ct, x.c_off, y.ptr, y.ct, y.c_off, c, x.pixels, p, p + 4 * c,
                                                int(relu), _stream()), 'bn_apply')
    else:
        rp = res_vec.data_ptr() if res_vec is not None else None
        _lib.check(_lib.load().b200seg_bn_apply_add(
            x.ptr, x.ct, x.c_off, y.ptr, y.ct, y.c_off, c, x.pixels, p, p + 4 * c, int(relu), res.ptr, res.ct, res.c_off,
            rp, rp + 4 * c if rp is not None else None, _stream()), 'bn_apply_add')
    _lib.count_launch()


def bn_backward(dy, x, dx, vec, relu, training, mask=None, red=None, reduced=False):
    """dy: grad of the BN(+ReLU) output, x: saved raw conv output, dx: receives the grad of x.  ``mask``: stored
    block output whose sign is the ReLU mask (residual blocks); None recomputes it from x.  ``red``: a ZEROED fp32
    [2c] slice of a per-backward arena (saves the per-layer memset).  Returns (dgamma, dbeta)."""
    c = x.c
    p = vec.data_ptr()
    lib = _lib.load()
    reduce_fn = lib.b200seg_bn_bwd_reduce_acc if red is not None else lib.b200seg_bn_bwd_reduce_masked
    if red is None:
        red = torch.empty((2 * c,), dtype=torch.float32, device=x.t.device)
    m = (mask.ptr, mask.ct, mask.c_off) if mask is not None else (None, 0, 0)
    if not reduced:        # ``reduced``: ``red`` already holds the sums (grad_join_reduce did the first pass)
        _lib.check(reduce_fn(dy.ptr, dy.ct, dy.c_off, x.ptr, x.ct, x.c_off, c, x.pixels, p, p + 4 * c,
                             p + 8 * c, p + 12 * c, int(relu), red.data_ptr(), m[0], m[1], m[2], _stream()), 'bn_bwd_reduce')
    _lib.check(lib.b200seg_bn_bwd_apply_masked(dy.ptr, dy.ct, dy.c_off, x.ptr, x.ct, x.c_off, dx.ptr, dx.ct, dx.c_off, c,
                                               x.pixels, p, p + 4 * c, p + 8 * c, p + 12 * c, int(relu), int(training),
                                               red.data_ptr(), None, None, 0, m[0], m[1], m[2], _stream()),
               'bn_bwd_apply')
    _lib.count_launch(1 if reduced else 2)      # reduce + apply
    return red[c:], red[:c]


def relu_bwd_add(dx, dy, y):
    """dx += dy * (y > 0)"""
    _lib.check(_lib.load().b200seg_relu_bwd_add(dx.ptr, dx.ct, dx.c_off, dy.ptr, dy.ct, dy.c_off, y.ptr, y.ct, y.c_off,
                                                dx.c, dx.pixels, _stream()), 'relu_bwd_add')
    _lib.count_launch()


def grad_add_masked(dx, dy=None, y=None, xin=None):
    """dx = (dx + dy * [y > 0]) * [xin > 0]; dy, y, xin optional Acts."""
    def trip(a):
        return (a.ptr, a.ct, a.c_off) if a is not None else (None, 0, 0)
    if xin is not None and xin.bits is not None and xin.c_off == 0 and xin.c == dx.c:
        _lib.check(_lib.load().b200seg_grad_add_masked_bits(dx.ptr, dx.ct, dx.c_off, *trip(dy), *trip(y),
                                                            xin.bits.data_ptr(), dx.c, dx.pixels, _stream()),
                   'grad_add_masked_bits')
        _lib.count_launch()
        return
    _lib.check(_lib.load().b200seg_grad_add_masked(dx.ptr, dx.ct, dx.c_off, *trip(dy), *trip(y), *trip(xin), dx.c, dx.pixels,
                                                   _stream()), 'grad_add_masked')
    _lib.count_launch()


JOIN_REDUCE = _os.environ.get('B200SEG_JOIN_REDUCE', '1') != '0'      # resnet.py: fuse the join with the next BN reduce


def grad_join_reduce(dx, dy, xin, t3, td=None):
    """dx = (dx + dy) * [xin > 0] (dy optional, xin's mask as bits) AND the first pass of the BatchNorm backward(s) that
    consume dx as their output gradient: t3 / td = (raw conv output Act, vec[4, c], zeroed red[2c]) of the previous
    block's bn3 / downsample BN.  Afterwards ``bn_backward(..., red=red, reduced=True)`` runs the apply pass only."""
    c = dx.c
    assert xin.bits is not None and xin.c_off == 0 and xin.c == c

    def trip(a):
        return (a.ptr, a.ct, a.c_off) if a is not None else (None, 0, 0)

    def target(t):
        if t is None:
            return (None, 0, 0, None, None, None)
        x, vec, red = t
        bn_resolve(vec)
        assert x.c == c and x.pixels == dx.pixels
        p = vec.data_ptr()
        return (x.ptr, x.ct, x.c_off, p + 8 * c, p + 12 * c, red.data_ptr())

    _lib.check(_lib.load().b200seg_grad_join_reduce(dx.ptr, dx.ct, dx.c_off, *trip(dy), xin.bits.data_ptr(), c, dx.pixels,
                                                    *target(t3), *target(td), _stream()), 'grad_join_reduce')
    _lib.count_launch()


def add_n_relu(xs, y, relu=True):
    """y = relu?(sum of 1..4 Acts)"""
    n = len(xs)
    ptrs = (ctypes.c_void_p * n)(*[a.ptr for a in xs])
    cts = (ctypes.c_int * n)(*[a.ct for a in xs])
    offs = (ctypes.c_int * n)(*[a.c_off for a in xs])
    _lib.check(_lib.load().b200seg_add_n_relu(n, ptrs, cts, offs, y.ptr, y.ct, y.c_off, y.c, y.pixels, int(relu), _stream()),
               'add_n_relu')
    _lib.count_launch()


def relu_bwd(dy, y, dx):
    """dx = dy * [y > 0]"""
    _lib.check(_lib.load().b200seg_relu_bwd(dy.ptr, dy.ct, dy.c_off, y.ptr, y.ct, y.c_off, dx.ptr, dx.ct, dx.c_off, dx.c,
                                            dx.pixels, _stream()), 'relu_bwd')
    _lib.count_launch()


def maxpool_fwd(x):
    """3x3 stride-2 pad-1 max pool -> (y: Act, idx: uint8 tensor of window positions)"""
    oh, ow = (x.h + 1) // 2, (x.w + 1) // 2
    y = Act.empty(x.n, oh, ow, x.c, x.t.device)
    idx = torch.empty((x.n, oh, ow, x.c), dtype=torch.uint8, device=x.t.device)
    _lib.check(_lib.load().b200seg_maxpool3x3s2_fwd(x.ptr, x.ct, x.c_off, x.n, x.h, x.w, y.ptr, y.ct, y.c_off,
                                                    idx.data_ptr(), x.c, _stream()), 'maxpool_fwd')
    _lib.count_launch()
    return y, idx


def maxpool_bwd(dy, idx, h, w):
    dx = Act.empty(dy.n, h, w, dy.c, dy.t.device)
    _lib.check(_lib.load().b200seg_maxpool3x3s2_bwd(dy.ptr, dy.ct, dy.c_off, idx.data_ptr(), dy.n, h, w, dx.ptr, dx.ct,
                                                    dx.c_off, dy.c, _stream()), 'maxpool_bwd')
    _lib.count_launch()
    return dx


# ------------------------------------------------------------------------------------------------
# resampling / pooling
# ------------------------------------------------------------------------------------------------
def upsample_fwd(x, y, align_corners=True):
    _lib.check(_lib.load().b200seg_upsample_fwd(x.ptr, x.ct, x.c_off, x.n, x.h, x.w, y.ptr, y.ct, y.c_off, y.h, y.w,
                                                x.c, int(align_corners), _stream()), 'upsample_fwd')
    _lib.count_launch()


def upsample_bwd(dy, dx, align_corners=True):
    _lib.check(_lib.load().b200seg_upsample_bwd(dy.ptr, dy.ct, dy.c_off, dy.n, dy.h, dy.w, dx.ptr, dx.ct, dx.c_off,
                                                dx.h, dx.w, dy.c, int(align_corners), _stream()), 'upsample_bwd')
    _lib.count_launch()


def _reduce_hw(fn, what, x):
    y = torch.empty((x.n, x.c), dtype=torch.float32, device=x.t.device)
    _lib.check(fn(x.ptr, x.ct, x.c_off, x.n, x.h * x.w, x.c, y.data_ptr(), _stream()), what)
    _lib.count_launch(2)
    return y


def gap_fwd(x):
    """mean over H*W -> fp32 [n, c]"""
    return _reduce_hw(_lib.load().b200seg_gap_fwd, 'gap_fwd', x)


def sum_hw(x):
    """sum over H*W -> fp32 [n, c]"""
    return _reduce_hw(_lib.load().b200seg_sum_hw, 'sum_hw', x)


def broadcast_hw(v, y, scale=1.0, accumulate=False):
    """v: fp32 [n, c] broadcast over y's H*W into the bf16 slice y (optionally y += scale * v)."""
    assert v.dtype == torch.float32 and v.is_contiguous() and tuple(v.shape) == (y.n, y.c), (v.shape, y.n, y.c)
    _lib.check(_lib.load().b200seg_broadcast_hw(v.data_ptr(), y.n, y.c, scale, y.ptr, y.ct, y.c_off, y.h * y.w,
                                                int(accumulate), _stream()), 'broadcast_hw')
    _lib.count_launch()


def linear_f32_fwd(x, w):
    """x fp32 [n, c], w fp32 [k, c, 1, 1] (the master weight) -> fp32 [n, k]"""
    n, c = x.shape
    k = w.shape[0]
    w = w.detach()
    assert w.dtype == torch.float32 and w.is_contiguous() and w.numel() == k * c
    y = torch.empty((n, k), dtype=torch.float32, device=x.device)
    _lib.check(_lib.load().b200seg_linear_f32_fwd(x.data_ptr(), w.data_ptr(), y.data_ptr(), n, c, k, _stream()),
               'linear_f32_fwd')
    _lib.count_launch()
    return y


def linear_f32_bwd(dy, x, w, need_dx=True):
    """-> (dx fp32 [n, c] or None, dw fp32 shaped like w)"""
    n, k = dy.shape
    c = x.shape[1]
    w = w.detach()
    lib = _lib.load()
    dw = torch.empty_like(w)
    _lib.check(lib.b200seg_linear_f32_bwd_weight(dy.data_ptr(), x.data_ptr(), dw.data_ptr(), n, c, k, 0, _stream()),
               'linear_f32_bwd_weight')
    dx = None
    if need_dx:
        dx = torch.empty((n, c), dtype=torch.float32, device=dy.device)
        _lib.check(lib.b200seg_linear_f32_bwd_data(dy.data_ptr(), w.data_ptr(), dx.data_ptr(), n, c, k, _stream()),
                   'linear_f32_bwd_data')
    _lib.count_launch(2 if need_dx else 1)
    return dx, dw


def bn_rows_fwd(x, bn, training, relu):
    """BatchNorm(+ReLU) over the rows of x fp32 [n, k] with an nn.BatchNorm2d's parameters.
    -> (y fp32 [n, k], save [2, k] = mean / invstd)"""
    n, k = x.shape
    y = torch.empty_like(x)
    save = torch.empty((2, k), dtype=torch.float32, device=x.device)
    track = bn.track_running_stats and bn.running_mean is not None
    if not training and not track:
        raise RuntimeError('eval-mode BatchNorm without running statistics')
    _lib.check(_lib.load().b200seg_bn_rows_fwd(
        x.data_ptr(), n, k, bn.weight.data_ptr(), bn.bias.data_ptr(), bn.eps, bn.momentum if bn.momentum else 0.0,
        bn.running_mean.data_ptr() if track else None, bn.running_var.data_ptr() if track else None, int(training),
        int(relu), y.data_ptr(), save.data_ptr(), save.data_ptr() + 4 * k, _stream()), 'bn_rows_fwd')
    _lib.count_launch()
    return y, save


def bn_rows_bwd(dy, x, bn, save, training, relu):
    """-> (dx, dgamma, dbeta)"""
    n, k = x.shape
    dx = torch.empty_like(x)
    dgb = torch.empty((2, k), dtype=torch.float32, device=x.device)
    _lib.check(_lib.load().b200seg_bn_rows_bwd(
        dy.data_ptr(), x.data_ptr(), n, k, bn.weight.data_ptr(), bn.bias.data_ptr(), save.data_ptr(),
        save.data_ptr() + 4 * k, int(training), int(relu), dx.data_ptr(), dgb.data_ptr(), dgb.data_ptr() + 4 * k,
        _stream()), 'bn_rows_bwd')
    _lib.count_launch()
    return dx, dgb[0], dgb[1]


def channel_sum(x):
    out = torch.empty((x.c,), dtype=torch.float32, device=x.t.device)
    _lib.check(_lib.load().b200seg_channel_sum(x.ptr, x.ct, x.c_off, x.c, x.pixels, out.data_ptr(), 0, _stream()),
               'channel_sum')
    _lib.count_launch()
    return out


# ------------------------------------------------------------------------------------------------
# class map, loss, metrics
# ------------------------------------------------------------------------------------------------
# trainer.TrainStep sets this around the model's forward: the step's loss is taken from the class map (compute_loss's
# shortcut), so the 352 MB of full-resolution logits the head would write are never read.  The head then returns the
# logits tensor ALLOCATED BUT UNWRITTEN, marked deferred in its class-map tag; compute_loss fills it first
# (``materialize_logits``) if for any reason it has to take the logits path after all.
DEFER_LOGITS = False
_deferred = set()


def logits_upsample_fwd(x_nhwc_f32, nc, oh, ow, align_corners=True, out=None):
    n, h, w, ct = x_nhwc_f32.shape
    y = out if out is not None else torch.empty((n, nc, oh, ow), dtype=torch.float32, device=x_nhwc_f32.device)
    if (out is None and DEFER_LOGITS and LOWRES_CE
            and lowres_ce_supported(x_nhwc_f32.shape, nc, oh, ow, align_corners)):
        _deferred.add(y.data_ptr())
        return y
    _lib.check(_lib.load().b200seg_logits_upsample_fwd(x_nhwc_f32.data_ptr(), ct, n, h, w, nc, y.data_ptr(), oh, ow,
                                                       int(align_corners), _stream()), 'logits_upsample_fwd')
    _lib.count_launch()
    return y


def materialize_logits(out):
    """Write the logits a head deferred (see DEFER_LOGITS); no-op for ordinary outputs."""
    tag = getattr(out, '_b200seg_class_map', None)
    if tag is not None and len(tag) > 4 and tag[4]:
        lr, nc, ac = tag[0], tag[1], tag[2]
        logits_upsample_fwd(lr, nc, out.shape[2], out.shape[3], ac, out=out.detach())
        out._b200seg_class_map = tag[:4] + (False,)
    return out


def logits_upsample_bwd(dy_nchw_f32, h, w, ct, align_corners=True):
    n, nc, oh, ow = dy_nchw_f32.shape
    dx = Act.empty(n, h, w, ct, dy_nchw_f32.device)
    _lib.check(_lib.load().b200seg_logits_upsample_bwd(dy_nchw_f32.data_ptr(), n, nc, oh, ow, dx.ptr, ct, h, w,
                                                       int(align_corners), _stream()), 'logits_upsample_bwd')
    _lib.count_launch()
    return dx


def resize_nchw_fwd(x, oh, ow, align_corners=True):
    n, c, h, w = x.shape
    y = torch.empty((n, c, oh, ow), dtype=torch.float32, device=x.device)
    _lib.check(_lib.load().b200seg_resize_nchw_fwd(x.data_ptr(), n * c, h, w, y.data_ptr(), oh, ow,
                                                   int(align_corners), _stream()), 'resize_nchw_fwd')
    _lib.count_launch()
    return y


def resize_nchw_bwd(dy, h, w, align_corners=True):
    n, c, oh, ow = dy.shape
    dx = torch.empty((n, c, h, w), dtype=torch.float32, device=dy.device)
    _lib.check(_lib.load().b200seg_resize_nchw_bwd(dy.data_ptr(), n * c, oh, ow, dx.data_ptr(), h, w,
                                                   int(align_corners), _stream()), 'resize_nchw_bwd')
    _lib.count_launch()
    return dx


_TARGET_KINDS = {torch.int64: 0, torch.int32: 1, torch.uint8: 2}      # B200SEG_TARGET_* (include/b200seg.h)


def target_kind(targets):
    """Masks are read as they are: int64 (post_fetch_fn's ``segs.long()``, utils/datasets.py:213) or the dataset's own
    uint8 (utils/datasets.py:161-194) / int32 — no widening copy."""
    k = _TARGET_KINDS.get(targets.dtype)
    if k is None:
        raise RuntimeError('targets must be int64, int32 or uint8 (got %s)' % targets.dtype)
    _require_cuda(targets, 'targets')
    if not targets.is_contiguous():
        raise RuntimeError('targets must be contiguous')
    return k


def softmax_ce(logits, targets, ignore_index=-100, want_grad=True):
    """logits NCHW fp32, targets int64 [n, h, w].  Returns (acc, dlogits): ``acc`` is a 3-element fp64
    device tensor (sum of -log p[target], number of valid pixels as int64 bits, status flag as int32
    bits); dlogits = softmax - onehot (unscaled; 0 on ignored pixels) or None."""
    n, nc, h, w = logits.shape
    acc = torch.zeros((3,), dtype=torch.float64, device=logits.device)
    dl = torch.empty_like(logits) if want_grad else None
    p = acc.data_ptr()
    _lib.check(_lib.load().b200seg_softmax_ce(logits.data_ptr(), targets.data_ptr(), target_kind(targets), n, nc, h * w, ignore_index, 1.0,
                                              p, p + 8, dl.data_ptr() if want_grad else None, p + 16, _stream()),
               'softmax_ce')
    _lib.count_launch()
    return acc, dl


LOWRES_CE = _os.environ.get('B200SEG_LOWRES_CE', '1') != '0'   # loss through the class map (utils.compute_loss)
# one fused loss + gradient launch when the class map needs a gradient (A/B: B200SEG_FUSED_CE=0 recomputes in backward)
FUSED_CE = _os.environ.get('B200SEG_FUSED_CE', '1') != '0'


def _lowres_geometry(lr, targets):
    n, h, w, ct = lr.shape
    assert lr.dtype == torch.float32 and lr.is_contiguous() and targets.dtype in _TARGET_KINDS and targets.is_contiguous()
    assert targets.shape[0] == n
    return n, h, w, ct, targets.shape[1], targets.shape[2]


def lowres_ce_supported(lr_shape, nc, th, tw, align_corners=True):
    """Mirror of launch_lowres_ce's checks (csrc/loss_metrics.cu), slightly conservative: an upsample of x1.5 to ~x5
    (a 4-pixel run touches <= 4 cells, <= 4 patch windows cover a cell) and the footprint of a 32 x 128 target tile plus
    the exchange slots in 110 KB of shared memory.  Anything else takes the logits path."""
    n, h, w, ct = lr_shape
    if nc > 32 or th <= h or tw <= w or h < 2 or w < 2:
        return False
    sh, sw = ((h - 1) / (th - 1), (w - 1) / (tw - 1)) if align_corners else (h / th, w / tw)
    ncell = int(3 * max(sh, sw) * (1 + 1e-6)) + 3
    if ncell > 4 or int((ncell - 1) / (4 * min(sh, sw) * (1 - 1e-6))) + 2 > (4 if ncell == 3 else 3):
        return False
    fh, fw = min(int(32 * sh) + 3, h), min(int(128 * sw) + 3, w)
    cells = fh * fw
    return fh <= 64 and fw <= 160 and 4 * ((cells * (ct | 1) + 3) // 4 * 4 + cells * (20 if ncell == 3 else 9)) <= 110 * 1024


def lowres_ce_fwd(lr, nc, targets, align_corners=True, ignore_index=-100):
    """Loss accumulators (as ``softmax_ce``) of the bilinear upsample of the fp32 NHWC class map ``lr`` to the target
    size, without the upsampled logits."""
    n, h, w, ct, th, tw = _lowres_geometry(lr, targets)
    acc = torch.zeros((3,), dtype=torch.float64, device=lr.device)
    p = acc.data_ptr()
    _lib.check(_lib.load().b200seg_lowres_ce_fwd(lr.data_ptr(), n, h, w, ct, nc, targets.data_ptr(), target_kind(targets), th, tw,
                                                 int(align_corners), ignore_index, p, p + 8, p + 16, _stream()),
               'lowres_ce_fwd')
    _lib.count_launch()
    return acc


def lowres_ce_bwd(lr, nc, targets, grad_out, count, align_corners=True, ignore_index=-100):
    """fp32 gradient of the class map: upsample^T((softmax - onehot) * grad_out / count)."""
    n, h, w, ct, th, tw = _lowres_geometry(lr, targets)
    assert grad_out.dtype == torch.float32 and count.dtype == torch.int64
    dlr = torch.empty_like(lr)
    _lib.check(_lib.load().b200seg_lowres_ce_bwd(lr.data_ptr(), n, h, w, ct, nc, targets.data_ptr(), target_kind(targets), th, tw,
                                                 int(align_corners), ignore_index, grad_out.data_ptr(), count.data_ptr(),
                                                 dlr.data_ptr(), _stream()), 'lowres_ce_bwd')
    _lib.count_launch()
    return dlr


def lowres_ce_fused(lr, nc, targets, align_corners=True, ignore_index=-100):
    """Training: loss accumulators AND the unscaled class-map gradient upsample^T(softmax - onehot) in one launch
    (``lowres_ce_scale`` applies grad_out / count once both are known)."""
    n, h, w, ct, th, tw = _lowres_geometry(lr, targets)
    acc = torch.zeros((3,), dtype=torch.float64, device=lr.device)
    dlr = torch.empty_like(lr)
    p = acc.data_ptr()
    _lib.check(_lib.load().b200seg_lowres_ce_fused(lr.data_ptr(), n, h, w, ct, nc, targets.data_ptr(), target_kind(targets), th, tw,
                                                   int(align_corners), ignore_index, p, p + 8, p + 16, dlr.data_ptr(),
                                                   _stream()), 'lowres_ce_fused')
    _lib.count_launch()
    return acc, dlr


def lowres_ce_scale(dlr, grad_out, count):
    """dlr *= grad_out / count, in place; both scalars read on the device."""
    assert grad_out.dtype == torch.float32 and count.dtype == torch.int64 and dlr.dtype == torch.float32
    _lib.check(_lib.load().b200seg_scale_by_ratio(dlr.data_ptr(), dlr.numel(), grad_out.data_ptr(), count.data_ptr(), _stream()),
               'scale_by_ratio')
    _lib.count_launch()
    return dlr


def class_map_grad(dlr_f32, extra=None):
    """fp32 NHWC class-map gradient (+ an optional bf16 Act addend) -> the bf16 Act the classifier backward takes."""
    n, h, w, ct = dlr_f32.shape
    out = Act.empty(n, h, w, ct, dlr_f32.device)
    _lib.check(_lib.load().b200seg_cast_add_bf16(dlr_f32.data_ptr(), extra.ptr if extra is not None else None, out.ptr,
                                                 dlr_f32.numel(), _stream()), 'cast_add_bf16')
    _lib.count_launch()
    return out


def class_map_backward(dout, dlr_direct, h, w, ncp, align_corners):
    """Gradient of a head's class map as a bf16 Act from either / both of its consumers: the full-resolution logits
    (``dout``, NCHW, through the upsample adjoint) and the class-map loss (``dlr_direct``, fp32 NHWC)."""
    extra = None
    if dout is not None:
        extra = logits_upsample_bwd(dout.contiguous().float(), h, w, ncp, align_corners=align_corners)
    if dlr_direct is None:
        return extra
    return class_map_grad(dlr_direct.contiguous(), extra)


def attach_class_map(out, lr, nc, align_corners):
    """Called by the model heads (read by utils.compute_loss): remember which class map ``out`` is the upsample of (and ``out``'s version, so an
    in-place edit of the logits turns the shortcut off)."""
    deferred = out.data_ptr() in _deferred
    if deferred:
        _deferred.discard(out.data_ptr())
    out._b200seg_class_map = (lr, nc, align_corners, out._version, deferred)
    return out


def ce_finalize(acc):
    """acc from ``softmax_ce`` -> 0-dim fp32 loss = sum / count (NaN if a target was out of range), one launch."""
    loss = torch.empty((), dtype=torch.float32, device=acc.device)
    p = acc.data_ptr()
    _lib.check(_lib.load().b200seg_ce_finalize(p, p + 8, p + 16, loss.data_ptr(), _stream()), 'ce_finalize')
    _lib.count_launch()
    return loss


def softmax_ce_bwd(logits, targets, grad_out, count, ignore_index=-100):
    """dlogits = (softmax - onehot) * grad_out / count, recomputed from the saved logits; ``grad_out`` (0-dim fp32)
    and ``count`` (1-element int64) are read on the device."""
    n, nc, h, w = logits.shape
    assert grad_out.dtype == torch.float32 and count.dtype == torch.int64
    dl = torch.empty_like(logits)
    _lib.check(_lib.load().b200seg_softmax_ce_bwd(logits.data_ptr(), targets.data_ptr(), target_kind(targets), n, nc, h * w, ignore_index,
                                                  grad_out.data_ptr(), count.data_ptr(), dl.data_ptr(), _stream()),
               'softmax_ce_bwd')
    _lib.count_launch()
    return dl


def argmax(logits):
    n, nc, h, w = logits.shape
    pred = torch.empty((n, h, w), dtype=torch.int64, device=logits.device)
    _lib.check(_lib.load().b200seg_argmax(logits.data_ptr(), n, nc, h * w, pred.data_ptr(), _stream()), 'argmax')
    _lib.count_launch()
    return pred


def _check_cm(cm, nc, device):
    """The kernels index the matrix with nc: a matrix of another size would be written out of bounds / mis-binned."""
    if cm is None:
        return torch.zeros((nc, nc), dtype=torch.int64, device=device)
    if tuple(cm.shape) != (nc, nc) or cm.dtype != torch.int64 or not cm.is_contiguous() or not cm.is_cuda:
        raise ValueError('confusion matrix must be a contiguous CUDA int64 %d x %d tensor (got %s %s)'
                         % (nc, nc, tuple(cm.shape), cm.dtype))
    return cm


def confusion(pred, targets, nc, cm=None):
    cm = _check_cm(cm, nc, pred.device)
    if pred.numel() != targets.numel():
        raise ValueError('confusion: %d predictions vs %d targets' % (pred.numel(), targets.numel()))
    _lib.check(_lib.load().b200seg_confusion(pred.data_ptr(), targets.data_ptr(), target_kind(targets), pred.numel(), nc,
                                             cm.data_ptr(), _stream()), 'confusion')
    _lib.count_launch()
    return cm


def argmax_confusion(logits, targets, cm=None):
    n, nc, h, w = logits.shape
    cm = _check_cm(cm, nc, logits.device)
    if targets.numel() != n * h * w:
        raise ValueError('argmax_confusion: logits %s vs targets %s' % (tuple(logits.shape), tuple(targets.shape)))
    _lib.check(_lib.load().b200seg_argmax_confusion(logits.data_ptr(), targets.data_ptr(), target_kind(targets), n, nc, h * w,
                                                    cm.data_ptr(), _stream()), 'argmax_confusion')
    _lib.count_launch()
    return cm


def class_map_confusion(lr, nc, targets, align_corners=True, cm=None, want_pred=False):
    """Evaluation from a head's low-resolution fp32 NHWC class map (the tag ``attach_class_map`` leaves on the logits):
    argmax of its bilinear upsample to the target size + confusion histogram, without reading the full-resolution logits.
    Bit-identical to ``argmax_confusion`` over the logits the head wrote (same interpolation instructions; ties -> lowest
    class index).  Returns (cm, pred or None)."""
    n, h, w, ct = lr.shape
    assert lr.dtype == torch.float32 and lr.is_contiguous() and targets.shape[0] == n
    th, tw = targets.shape[1], targets.shape[2]
    cm = _check_cm(cm, nc, lr.device)
    pred = torch.empty((n, th, tw), dtype=torch.int64, device=lr.device) if want_pred else None
    _lib.check(_lib.load().b200seg_class_map_confusion(lr.data_ptr(), n, h, w, ct, nc, targets.data_ptr(), target_kind(targets),
                                                       th, tw, int(align_corners), cm.data_ptr(),
                                                       pred.data_ptr() if pred is not None else None, _stream()),
               'class_map_confusion')
    _lib.count_launch()
    return cm, pred


# ------------------------------------------------------------------------------------------------
# per-op CUDA-event timing (bench.py --detail): every wrapper above except the convs, which time themselves with
# their nominal FLOPs.  OP_PROFILE is None in normal operation: the wrappers then add one global lookup per call.
# ------------------------------------------------------------------------------------------------
OP_PROFILE = None


def _act_bytes(a):
    return 2.0 * a.pixels * a.c if isinstance(a, Act) else 0.0


# algorithmic HBM bytes of the BatchNorm-family passes ("BN apply: 2 x tensor bytes"; backward = the two
# passes' reads of dy and x plus the dx write; the residual join reads what it is given and rewrites dx)
_OP_BYTES = {
    'bn_apply': lambda a, kw: _act_bytes(a[0]) * (2 + (kw.get('res') is not None or (len(a) > 4 and a[4] is not None))),
    'bn_backward': lambda a, kw: _act_bytes(a[1]) * (3 if kw.get('reduced') else 5),
    'grad_join_reduce': lambda a, kw: _act_bytes(a[0]) * (2 + (a[1] is not None) + 1.0 / 16 + 1 + ((a[4] if len(a) > 4 else kw.get('td')) is not None)),
    'grad_add_masked': lambda a, kw: _act_bytes(a[0]) * (2 + sum(1 for v in list(a[1:3]) + [kw.get('dy'), kw.get('y')]
                                                                 if v is not None)
                                                         + (1.0 / 16 if (kw.get('xin') if 'xin' in kw else (a[3] if len(a) > 3 else None)) is not None else 0)),
}


def _timed(name, fn):
    def wrapper(*a, **kw):
        if OP_PROFILE is None:
            return fn(*a, **kw)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        out = fn(*a, **kw)
        e1.record()
        act = next((v for v in a if isinstance(v, Act)), None)
        tag = name if act is None else '%s %dx%dx%dx%d%s' % (name, act.n, act.h, act.w, act.c,
                                                            ' masked' if kw.get('mask') is not None else '')
        nbytes = _OP_BYTES[name](a, kw) if name in _OP_BYTES else 0.0
        OP_PROFILE.append((tag, e0, e1, nbytes))
        return out
    wrapper.__name__, wrapper.__doc__ = fn.__name__, fn.__doc__
    return wrapper


for _name in ('as_act', 'grad_like', 'im2col', 'unpack_wgrad_im2col', 'tap_gather', 'tap_scatter_add', 'unpack_wgrad_taps', 'bn_stats', 'bn_finalize', 'bn_eval_affine',
              'bn_apply', 'bn_backward', 'relu_bwd_add', 'grad_add_masked', 'grad_join_reduce', 'maxpool_fwd', 'maxpool_bwd', 'upsample_fwd', 'upsample_bwd',
              'gap_fwd', 'sum_hw', 'broadcast_hw', 'linear_f32_fwd', 'linear_f32_bwd', 'bn_rows_fwd', 'bn_rows_bwd',
              'channel_sum', 'logits_upsample_fwd', 'logits_upsample_bwd', 'resize_nchw_fwd', 'resize_nchw_bwd',
              'softmax_ce', 'softmax_ce_bwd', 'argmax', 'confusion', 'argmax_confusion'):
    globals()[_name] = _timed(_name, globals()[_name])
del _name
