"""ctypes binding of libb200seg.so (the C ABI declared in include/b200seg.h).

The library is built in-tree by ``__graft_entry__.build()`` / ``make -C pytorch_segmentation_b200/csrc``.
There is no fallback: if the shared object is missing or a call fails, a RuntimeError is raised.
"""
import ctypes
import os
from ctypes import POINTER, Structure, c_char_p, c_double, c_float, c_int, c_int32, c_int64, c_size_t, c_void_p

_HERE = os.path.dirname(os.path.abspath(__file__))
# B200SEG_LIB_PATH: another build of the library (same-box A/B of a kernel change: scripts/gpu_ab_lib.sh)
LIB_PATH = os.environ.get('B200SEG_LIB_PATH') or os.path.join(_HERE, 'libb200seg.so')

MAX_CONV_SRC = 4
NOALLOC_WGRAD_MAX_BYTES = 16 << 20      # B200SEG_NOALLOC_WGRAD_MAX_BYTES


class ConvSrc(Structure):
    """struct b200seg_conv_src"""
    _fields_ = [('x', c_void_p), ('x_c_total', c_int32), ('x_c_offset', c_int32), ('c', c_int32),
                ('w', c_void_p), ('w_rows', c_int32), ('r', c_int32), ('s', c_int32), ('dil', c_int32),
                ('stride', c_int32), ('x_rows', c_int32), ('x_cols', c_int32)]


class ConvOut(Structure):
    """struct b200seg_conv_out"""
    _fields_ = [('y', c_void_p), ('y_c_total', c_int32), ('y_c_offset', c_int32), ('k', c_int32),
                ('out_fp32', c_int32), ('scale', c_void_p), ('bias', c_void_p), ('relu', c_int32),
                ('stats', c_void_p)]


class WgradDesc(Structure):
    """struct b200seg_wgrad_desc"""
    _fields_ = [('n', c_int32), ('h', c_int32), ('w', c_int32),
                ('dy', c_void_p), ('dy_c_total', c_int32), ('dy_c_offset', c_int32), ('k', c_int32),
                ('x', c_void_p), ('x_c_total', c_int32), ('x_c_offset', c_int32), ('c', c_int32),
                ('r', c_int32), ('s', c_int32), ('dil', c_int32), ('stride', c_int32),
                ('dw', c_void_p), ('k_out', c_int32), ('accumulate', c_int32),
                ('workspace', c_void_p), ('workspace_bytes', c_size_t), ('x_rows', c_int32), ('x_cols', c_int32)]


class PackDesc(Structure):
    """struct b200seg_pack_desc"""
    _fields_ = [('w', c_void_p), ('k', c_int32), ('c', c_int32), ('r', c_int32), ('s', c_int32),
                ('w_fprop', c_void_p), ('w_dgrad', c_void_p), ('k_rows', c_int32), ('c_rows', c_int32),
                ('k_cols', c_int32), ('im2col', c_int32), ('c_pad', c_int32), ('reserved', c_int32)]


class SgdDesc(Structure):
    """struct b200seg_sgd_desc"""
    _fields_ = [('p', c_void_p), ('g', c_void_p), ('m', c_void_p), ('numel', c_int64)]


class AdamDesc(Structure):
    """struct b200seg_adam_desc"""
    _fields_ = [('p', c_void_p), ('g', c_void_p), ('m', c_void_p), ('v', c_void_p), ('numel', c_int64)]


_P = c_void_p
_I = c_int
_L = c_int64
_F = c_float

# name -> (restype, argtypes).  Must list every function of include/b200seg.h (tests/test_abi.py checks).
PROTOTYPES = {
    'b200seg_last_error': (c_char_p, []),
    'b200seg_version': (c_int, []),
    'b200seg_device_info': (c_int, [POINTER(c_int), POINTER(c_int), POINTER(c_int)]),
    'b200seg_debug_timing': (c_int, [_P, _I]),
    'b200seg_nchw_to_nhwc': (c_int, [_P, _P, _I, _I, _I, _I, _I, _I, _P]),
    'b200seg_nhwc_to_nchw': (c_int, [_P, _P, _I, _I, _I, _I, _I, _I, _P]),
    'b200seg_pack_weights': (c_int, [_P, _I, _I, _I, _I, _P, _I, _P, _I, _I, _P]),
    'b200seg_im2col_nchw': (c_int, [_P, _I, _I, _I, _I, _I, _I, _I, _I, _P, _I, _P]),
    'b200seg_pack_weights_im2col': (c_int, [_P, _I, _I, _I, _I, _P, _I, _I, _P]),
    'b200seg_unpack_wgrad_im2col': (c_int, [_P, _I, _I, _I, _I, _I, _P, _I, _P]),
    'b200seg_space_to_depth_nchw': (c_int, [_P, _I, _I, _I, _I, _I, _I, _I, _P, _I, _P]),
    'b200seg_pack_weights_s2d': (c_int, [_P, _I, _I, _I, _I, _P, _I, _I, _P]),
    'b200seg_unpack_wgrad_s2d': (c_int, [_P, _I, _I, _I, _I, _I, _P, _I, _P]),
    'b200seg_tap_gather': (c_int, [_P, _I, _I, _I, _I, _I, _I, _I, _I, _I, _P, _I, _P]),
    'b200seg_tap_scatter_add': (c_int, [_P, _I, _I, _I, _I, _I, _I, _I, _I, _P, _P, _P]),
    'b200seg_pack_weights_taps': (c_int, [_P, _I, _I, _I, _I, _I, _P, _I, _P, _I, _P]),
    'b200seg_unpack_wgrad_taps': (c_int, [_P, _I, _I, _I, _I, _I, _P, _I, _P]),
    'b200seg_pack_weights_multi': (c_int, [POINTER(PackDesc), _I, _P]),
    'b200seg_sgd_multi': (c_int, [POINTER(SgdDesc), _I, _F, _F, _F, _F, _I, _I, _F, _P, _P]),
    'b200seg_adam_multi': (c_int, [POINTER(AdamDesc), _I, _P, _I, _I, _P]),
    'b200seg_copy_multi': (c_int, [POINTER(SgdDesc), _I, _P]),
    'b200seg_conv_fprop': (c_int, [_I, _I, _I, _I, POINTER(ConvSrc), POINTER(ConvOut), _P]),
    'b200seg_conv_dgrad': (c_int, [_I, _I, _I, _I, POINTER(ConvSrc), POINTER(ConvOut), _P]),
    'b200seg_conv_wgrad_workspace': (c_size_t, [POINTER(WgradDesc)]),
    'b200seg_conv_last_plan': (c_int, [POINTER(c_int64), POINTER(c_int64)]),
    'b200seg_conv_wgrad_plan_info': (c_int, [POINTER(WgradDesc), POINTER(c_int64), POINTER(c_int64)]),
    'b200seg_conv_wgrad': (c_int, [POINTER(WgradDesc), _P]),
    'b200seg_conv_wgrad_noalloc': (c_int, [POINTER(WgradDesc), _P]),
    'b200seg_bn_stats': (c_int, [_P, _I, _I, _I, _L, _P, _P]),
    'b200seg_bn_finalize': (c_int, [_P, _I, _L, _P, _P, _F, _F, _P, _P, _P, _P, _P, _P, _P]),
    'b200seg_bn_eval_affine': (c_int, [_I, _P, _P, _P, _P, _F, _P, _P, _P, _P, _P]),
    'b200seg_bn_apply': (c_int, [_P, _I, _I, _P, _I, _I, _I, _L, _P, _P, _I, _P]),
    'b200seg_bn_apply_train': (c_int, [_P, _I, _I, _P, _I, _I, _I, _L, _P, _P, _P, _F, _F, _P, _P, _P, _I, _P, _I, _I, _P, _P, _P, _P]),
    'b200seg_grad_add_masked_bits': (c_int, [_P, _I, _I, _P, _I, _I, _P, _I, _I, _P, _I, _L, _P]),
    'b200seg_grad_join_reduce': (c_int, [_P, _I, _I, _P, _I, _I, _P, _I, _L, _P, _I, _I, _P, _P, _P, _P, _I, _I, _P, _P, _P, _P]),
    'b200seg_bn_apply_add': (c_int, [_P, _I, _I, _P, _I, _I, _I, _L, _P, _P, _I, _P, _I, _I, _P, _P, _P]),
    'b200seg_relu_bwd_add': (c_int, [_P, _I, _I, _P, _I, _I, _P, _I, _I, _I, _L, _P]),
    'b200seg_add_n_relu': (c_int, [_I, POINTER(c_void_p), POINTER(c_int), POINTER(c_int), _P, _I, _I, _I, _L, _I, _P]),
    'b200seg_relu_bwd': (c_int, [_P, _I, _I, _P, _I, _I, _P, _I, _I, _I, _L, _P]),
    'b200seg_grad_add_masked': (c_int, [_P, _I, _I, _P, _I, _I, _P, _I, _I, _P, _I, _I, _I, _L, _P]),
    'b200seg_bn_bwd_reduce_masked': (c_int, [_P, _I, _I, _P, _I, _I, _I, _L, _P, _P, _P, _P, _I, _P, _P, _I, _I, _P]),
    'b200seg_bn_bwd_reduce_acc': (c_int, [_P, _I, _I, _P, _I, _I, _I, _L, _P, _P, _P, _P, _I, _P, _P, _I, _I, _P]),
    'b200seg_bn_bwd_apply_masked': (c_int, [_P, _I, _I, _P, _I, _I, _P, _I, _I, _I, _L, _P, _P, _P, _P, _I, _I, _P, _P,
                                            _P, _I, _P, _I, _I, _P]),
    'b200seg_bn_bwd_reduce': (c_int, [_P, _I, _I, _P, _I, _I, _I, _L, _P, _P, _P, _P, _I, _P, _P]),
    'b200seg_bn_bwd_apply': (c_int, [_P, _I, _I, _P, _I, _I, _P, _I, _I, _I, _L, _P, _P, _P, _P, _I, _I, _P, _P,
                                     _P, _I, _P]),
    'b200seg_upsample_fwd': (c_int, [_P, _I, _I, _I, _I, _I, _P, _I, _I, _I, _I, _I, _I, _P]),
    'b200seg_upsample_bwd': (c_int, [_P, _I, _I, _I, _I, _I, _P, _I, _I, _I, _I, _I, _I, _P]),
    'b200seg_maxpool3x3s2_fwd': (c_int, [_P, _I, _I, _I, _I, _I, _P, _I, _I, _P, _I, _P]),
    'b200seg_maxpool3x3s2_bwd': (c_int, [_P, _I, _I, _P, _I, _I, _I, _P, _I, _I, _I, _P]),
    'b200seg_gap_fwd': (c_int, [_P, _I, _I, _I, _I, _I, _P, _P]),
    'b200seg_sum_hw': (c_int, [_P, _I, _I, _I, _I, _I, _P, _P]),
    'b200seg_linear_f32_fwd': (c_int, [_P, _P, _P, _I, _I, _I, _P]),
    'b200seg_linear_f32_bwd_data': (c_int, [_P, _P, _P, _I, _I, _I, _P]),
    'b200seg_linear_f32_bwd_weight': (c_int, [_P, _P, _P, _I, _I, _I, _I, _P]),
    'b200seg_bn_rows_fwd': (c_int, [_P, _I, _I, _P, _P, _F, _F, _P, _P, _I, _I, _P, _P, _P, _P]),
    'b200seg_bn_rows_bwd': (c_int, [_P, _P, _I, _I, _P, _P, _P, _P, _I, _I, _P, _P, _P, _P]),
    'b200seg_broadcast_hw': (c_int, [_P, _I, _I, _F, _P, _I, _I, _I, _I, _P]),
    'b200seg_channel_sum': (c_int, [_P, _I, _I, _I, _L, _P, _I, _P]),
    'b200seg_logits_upsample_fwd': (c_int, [_P, _I, _I, _I, _I, _I, _P, _I, _I, _I, _P]),
    'b200seg_logits_upsample_bwd': (c_int, [_P, _I, _I, _I, _I, _P, _I, _I, _I, _I, _P]),
    'b200seg_resize_nchw_fwd': (c_int, [_P, _L, _I, _I, _P, _I, _I, _I, _P]),
    'b200seg_resize_nchw_bwd': (c_int, [_P, _L, _I, _I, _P, _I, _I, _I, _P]),
    'b200seg_softmax_ce': (c_int, [_P, _P, _I, _I, _I, _L, _L, _F, _P, _P, _P, _P, _P]),
    'b200seg_lowres_ce_fwd': (c_int, [_P, _I, _I, _I, _I, _I, _P, _I, _I, _I, _I, _L, _P, _P, _P, _P]),
    'b200seg_lowres_ce_bwd': (c_int, [_P, _I, _I, _I, _I, _I, _P, _I, _I, _I, _I, _L, _P, _P, _P, _P]),
    'b200seg_lowres_ce_fused': (c_int, [_P, _I, _I, _I, _I, _I, _P, _I, _I, _I, _I, _L, _P, _P, _P, _P, _P]),
    'b200seg_scale_by_ratio': (c_int, [_P, _L, _P, _P, _P]),
    'b200seg_cast_add_bf16': (c_int, [_P, _P, _P, _L, _P]),
    'b200seg_ce_finalize': (c_int, [_P, _P, _P, _P, _P]),
    'b200seg_softmax_ce_bwd': (c_int, [_P, _P, _I, _I, _I, _L, _L, _P, _P, _P, _P]),
    'b200seg_argmax': (c_int, [_P, _I, _I, _L, _P, _P]),
    'b200seg_confusion': (c_int, [_P, _P, _I, _L, _I, _P, _P]),
    'b200seg_argmax_confusion': (c_int, [_P, _P, _I, _I, _I, _L, _P, _P]),
    'b200seg_class_map_confusion': (c_int, [_P, _I, _I, _I, _I, _I, _P, _I, _I, _I, _I, _P, _P, _P]),
    'b200seg_input_u8_normalize': (c_int, [_P, _I, _I, _I, _I, _I, POINTER(c_float), POINTER(c_float), _P, _P]),
    'b200seg_input_u8_s2d': (c_int, [_P, _I, _I, _I, _I, _I, _I, _I, _I, POINTER(c_float), POINTER(c_float), _P, _I, _P]),
    'b200seg_resize_nearest_u8': (c_int, [_P, _L, _I, _I, _P, _I, _I, _P]),
}

_lib = None


def load():
    """Loads the shared object (once) and installs the prototypes."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            'libb200seg.so is missing (%s). Build it with `python -c "import __graft_entry__ as g; g.build()"` '
            'or `make -C pytorch_segmentation_b200/csrc`. There is no CPU / cuDNN fallback for this path.' % LIB_PATH)
    lib = ctypes.CDLL(LIB_PATH)
    for name, (res, args) in PROTOTYPES.items():
        fn = getattr(lib, name)      # AttributeError here = header/library mismatch: fail loudly
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def last_error():
    msg = load().b200seg_last_error()
    return msg.decode('utf-8', 'replace') if msg else ''


def check(rc, what):
    if rc != 0:
        raise RuntimeError('%s failed: %s' % (what, last_error()))


_calls = 0


def count_launch(n=1):
    """Launch counter for bench.py's ``gpu_launches`` (incremented by the op wrappers)."""
    global _calls
    _calls += n


def launches():
    return _calls
