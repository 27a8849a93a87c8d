"""The C-ABI library loads and exports every symbol include/pifpaf_b200.h
declares; struct layouts agree between the header (gcc) and the ctypes mirror;
host-only entry points and argument validation work without a GPU."""
import ctypes as C
import os
import re
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, 'include', 'pifpaf_b200.h')


@pytest.fixture(scope='module')
def lib():
    from openpifpaf_b200 import _lib
    return _lib.lib()


def declared_symbols():
    text = open(HEADER).read()
    text = re.sub(r'/\*.*?\*/', '', text, flags=re.S)
    return sorted(set(re.findall(r'\b(pp_[a-z0-9_]+)\s*\(', text)))


def test_every_declared_symbol_is_exported(lib):
    from openpifpaf_b200 import _lib
    names = declared_symbols()
    assert len(names) >= 20
    for name in names:
        assert hasattr(lib, name), 'libpifpaf_b200.so does not export %s' % name
    assert set(_lib.EXPORTS) == set(names), set(_lib.EXPORTS) ^ set(names)


def test_library_is_plain_c_abi_without_torch():
    from openpifpaf_b200 import _lib
    out = subprocess.run(['ldd', _lib.LIB_PATH], stdout=subprocess.PIPE, text=True).stdout
    assert 'torch' not in out and 'python' not in out.lower(), out


def test_struct_layouts_match_header(tmp_path):
    from openpifpaf_b200 import _lib
    src = tmp_path / 'sizes.c'
    src.write_text('#include <stdio.h>\n#include <stddef.h>\n#include "pifpaf_b200.h"\n'
                   'int main(void){printf("%zu %zu %zu %zu %zu %zu %zu %zu\\n", sizeof(pp_shape),'
                   'sizeof(pp_caps), sizeof(pp_params), sizeof(pp_skeleton), sizeof(pp_outputs),'
                   'offsetof(pp_params, profile), offsetof(pp_params, seed_mask),'
                   'offsetof(pp_skeleton, confidence_scales));return 0;}\n')
    exe = tmp_path / 'sizes'
    subprocess.run(['gcc', '-std=c99', '-Wall', '-Werror', '-I', os.path.join(ROOT, 'include'),
                    str(src), '-o', str(exe)], check=True)
    got = [int(v) for v in subprocess.run([str(exe)], stdout=subprocess.PIPE, text=True,
                                          check=True).stdout.split()]
    want = [C.sizeof(_lib.Shape), C.sizeof(_lib.Caps), C.sizeof(_lib.Params),
            C.sizeof(_lib.Skeleton), C.sizeof(_lib.Outputs), _lib.Params.profile.offset,
            _lib.Params.seed_mask.offset, _lib.Skeleton.confidence_scales.offset]
    assert got == want


def test_host_only_entries(lib):
    from openpifpaf_b200 import _lib
    p = _lib.Params()
    lib.pp_default_params(C.byref(p))
    assert abs(p.cif_threshold - 0.1) < 1e-7 and abs(p.seed_threshold - 0.2) < 1e-7
    assert p.nms == 1 and p.exp_mode == 1 and p.connection_method == 0 and not p.profile
    s = _lib.Shape(64, 17, 19, 81, 81, 8)
    c = _lib.Caps()
    assert lib.pp_default_caps(C.byref(s), C.byref(c)) == 0
    assert (c.seed_cap, c.caf_cap, c.src_cap) == (17 * 81 * 81, 81 * 81, 81 * 81)
    ws = lib.pp_workspace_bytes(C.byref(s), C.byref(c))
    assert ws > 64 * 17 * 641 * 641 * 4          # includes the hr maps
    sk = _lib.Skeleton()
    assert lib.pp_coco_skeleton(C.byref(sk)) == 0 and sk.L == 19
    from openpifpaf_b200.datasets.constants import COCO_PERSON_SKELETON
    assert [(sk.joints[i][0], sk.joints[i][1]) for i in range(19)] == \
        [tuple(l) for l in COCO_PERSON_SKELETON]
    assert lib.pp_error_string(0) == b'ok'
    assert b'capacity' in lib.pp_error_string(-2)
    assert lib.pp_decode_launch_count(None) == _lib.PP_N_STAGES + 3   # 8 kernels + 2 memsets
    assert [lib.pp_stage_name(i) for i in range(7)][1] == b'cifhr_splat'


def test_argument_validation_returns_codes(lib):
    from openpifpaf_b200 import _lib
    bad = _lib.Shape(0, 17, 19, 81, 81, 8)
    c = _lib.Caps()
    assert lib.pp_default_caps(C.byref(bad), C.byref(c)) == _lib.PP_ERR_ARG
    assert lib.pp_workspace_bytes(C.byref(bad), C.byref(c)) == 0
    s = _lib.Shape(1, 17, 19, 11, 11, 8)
    lib.pp_default_caps(C.byref(s), C.byref(c))
    p = _lib.Params()
    lib.pp_default_params(C.byref(p))
    sk = _lib.Skeleton()
    lib.pp_coco_skeleton(C.byref(sk))
    o = _lib.Outputs()
    # NULL device pointers are rejected before any CUDA call
    assert lib.pp_decode_batch(None, None, C.byref(sk), C.byref(o), None, 0, C.byref(s),
                               C.byref(c), C.byref(p), None) == _lib.PP_ERR_ARG
    assert lib.pp_cifhr_accumulate(None, None, None, 0, C.byref(s), C.byref(c), C.byref(p),
                                   None, None) == _lib.PP_ERR_ARG
    too_many = _lib.Shape(1, 33, 19, 11, 11, 8)      # > PP_MAX_F keypoint types: pose entries refuse,
    assert lib.pp_default_caps(C.byref(too_many), C.byref(c)) == _lib.PP_OK    # detection may have 256
    assert lib.pp_workspace_bytes(C.byref(too_many), C.byref(c)) == 0
    assert lib.pp_cifdet_workspace_bytes(C.byref(too_many), C.byref(c)) > 0
    assert lib.pp_default_caps(C.byref(_lib.Shape(1, 257, 0, 11, 11, 8)), C.byref(c)) == _lib.PP_ERR_ARG


def test_no_cpu_fallback_in_product():
    """The product package never imports the oracle and has no CPU decode path."""
    pkg = os.path.join(ROOT, 'openpifpaf_b200')
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith(('.py', '.cu', '.cuh', '.h')):
                text = open(os.path.join(dirpath, f)).read()
                assert not re.search(r'^\s*(from|import)\s+oracle\b', text, flags=re.M), f
                assert 'cifcaf_oracle' not in text, f


def test_missing_library_fails_loudly(monkeypatch):
    from openpifpaf_b200 import _lib
    monkeypatch.setattr(_lib, '_lib', None)
    monkeypatch.setattr(_lib, 'LIB_PATH', '/nonexistent/libpifpaf_b200.so')
    with pytest.raises(ImportError):
        _lib.lib()
