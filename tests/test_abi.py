"""The C-ABI boundary: libb200seg.so loads and exports exactly what include/b200seg.h declares, and the
ctypes prototypes cover every declaration (no compute calls — runs without a GPU)."""
import ctypes
import os
import re
import subprocess

from pytorch_segmentation_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, 'include', 'b200seg.h')).read()
    src = re.sub(r'/\*.*?\*/', '', src, flags=re.S)
    return sorted(set(re.findall(r'\b(b200seg_[a-z0-9_]+)\s*\(', src)))


def test_header_declares_the_hot_path():
    names = _declared()
    for required in ('b200seg_conv_fprop', 'b200seg_conv_dgrad', 'b200seg_conv_wgrad', 'b200seg_bn_stats',
                     'b200seg_upsample_fwd', 'b200seg_gap_fwd', 'b200seg_softmax_ce', 'b200seg_argmax_confusion'):
        assert required in names


def test_library_exports_every_declared_symbol(lib):
    for name in _declared():
        assert hasattr(lib, name), 'libb200seg.so does not export %s' % name


def test_prototypes_cover_header():
    assert sorted(_lib.PROTOTYPES) == _declared()


def test_struct_layouts_match_header(tmp_path):
    # compile the header with the host C compiler and compare struct sizes / key offsets with ctypes
    src = tmp_path / 'sz.c'
    src.write_text('#include <stdio.h>\n#include <stddef.h>\n#include "b200seg.h"\n'
                   'int main(void){printf("%zu %zu %zu %zu %zu %zu %zu %zu %zu %zu\\n", sizeof(b200seg_conv_src), '
                   'sizeof(b200seg_conv_out), sizeof(b200seg_wgrad_desc), offsetof(b200seg_conv_src, w), '
                   'offsetof(b200seg_conv_out, stats), offsetof(b200seg_wgrad_desc, workspace), '
                   'sizeof(b200seg_pack_desc), offsetof(b200seg_pack_desc, k_rows), sizeof(b200seg_sgd_desc), '
                   'offsetof(b200seg_sgd_desc, numel));return 0;}\n')
    exe = tmp_path / 'sz'
    subprocess.check_call(['gcc', '-I', os.path.join(ROOT, 'include'), str(src), '-o', str(exe)])
    got = [int(v) for v in subprocess.check_output([str(exe)]).split()]
    want = [ctypes.sizeof(_lib.ConvSrc), ctypes.sizeof(_lib.ConvOut), ctypes.sizeof(_lib.WgradDesc),
            _lib.ConvSrc.w.offset, _lib.ConvOut.stats.offset, _lib.WgradDesc.workspace.offset,
            ctypes.sizeof(_lib.PackDesc), _lib.PackDesc.k_rows.offset, ctypes.sizeof(_lib.SgdDesc),
            _lib.SgdDesc.numel.offset]
    assert got == want


def test_version_and_error_string(lib):
    assert lib.b200seg_version() >= 100
    assert isinstance(_lib.last_error(), str)


def test_argument_validation_without_gpu(lib):
    # pure host-side checks return an error code and a message before touching the device
    rc = lib.b200seg_conv_fprop(1, 8, 8, 0, None, None, None)
    assert rc != 0 and 'n_src' in _lib.last_error()
    rc = lib.b200seg_bn_stats(None, 8, 0, 8, 10, None, None)
    assert rc != 0 and 'null' in _lib.last_error()
    rc = lib.b200seg_upsample_fwd(ctypes.c_void_p(16), 12, 0, 1, 4, 4, ctypes.c_void_p(16), 8, 0, 8, 8, 8, 1, None)
    assert rc != 0 and 'multiples of 8' in _lib.last_error()
