"""CPU: libdfu3d.so builds for sm_100a, loads without a GPU and exports every entry point include/dfu3d.h
declares; the product path refuses to run without the library / a device (no CPU fallback)."""
import ctypes
import os
import re
import subprocess

import pytest

from dfu3d_b200 import _lib, build

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared():
    text = open(os.path.join(ROOT, 'include', 'dfu3d.h')).read()
    text = re.sub(r'/\*.*?\*/', '', text, flags=re.S)
    return sorted(set(re.findall(r'\b(dfu_[a-z0-9_]+)\s*\(', text)))


def test_header_and_binding_agree():
    assert declared() == _lib.EXPORTED


def test_library_exports_every_declared_symbol(lib_built):
    lib = ctypes.CDLL(lib_built)
    for name in declared():
        assert hasattr(lib, name), name
    _lib.load()


def test_built_for_sm100a_only(lib_built):
    out = subprocess.run(['/usr/local/cuda/bin/cuobjdump', '-lelf', lib_built], capture_output=True, text=True).stdout
    archs = set(re.findall(r'sm_(\d+a?)', out))
    assert archs == {'100a'}, archs


def test_struct_sizes_match_header():
    assert _lib.VIEW_DTYPE.itemsize == 248 and _lib.FUSION_DTYPE.itemsize == 104 and _lib.INST_DTYPE.itemsize == 104


def test_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip('GPU present')
    with pytest.raises(RuntimeError):
        _lib.Context(0)


def test_product_never_imports_oracle():
    for dp, _, files in os.walk(os.path.join(ROOT, 'dfu3d_b200')):
        for f in files:
            if f.endswith('.py'):
                src = open(os.path.join(dp, f)).read()
                assert not re.search(r'^\s*(from|import)\s+oracle', src, flags=re.M), f


def test_host_label_text_equals_python_str_of_floats():
    """dfu_repr_f64 / dfu_format_label_lines (host-only routines of the library): the reference writes label fields with
    Python's str() (my_loader.py:680); str(np.float64(v)) == repr(float(v)) == the library's text, on 1e5 random values
    over 28 decades plus the special cases, and whole lines equal pipeline.format_label_line."""
    import numpy as np
    from dfu3d_b200 import _lib, pipeline
    lib = _lib.load()
    r = np.random.default_rng(1)
    x = np.concatenate([r.normal(size=100000) * 10.0 ** r.integers(-9, 19, 100000),
                        [0.0, -0.0, 1e16, 1e-5, 1e-4, 123456.0, 1e15, 9999999999999998.0, np.inf, -np.inf, np.nan, 5e-324,
                         1.7976931348623157e308, 1e22, 100.0, 0.5]])
    out = np.empty(x.size * 48, np.uint8)
    nb = lib.dfu_repr_f64(_lib.hptr(x), x.size, _lib.hptr(out), out.size)
    got = out[:nb].tobytes().decode().split('\n')[:-1]
    assert got == [repr(float(v)) for v in x]
    assert all(str(np.float64(v)) == repr(float(v)) for v in x[:20000])
    K = 500
    boxes = r.normal(size=(K, 24)) * 10.0 ** r.integers(-6, 6, (K, 1))
    boxes[:, 0] = np.sort(r.integers(0, 4, K)); boxes[:, 1] = r.integers(0, 50, K); boxes[:, 3] = r.integers(0, 10, K)
    boxes[:, 20] = (r.random(K) < 0.1)
    b2 = r.uniform(0, 1600, (50, 4)).astype(np.float32)

    class VB(object):
        pass
    got = pipeline.format_label_lines(pipeline.NUSC_NAMES, boxes, VB(), b2, 4)
    want = [[] for _ in range(4)]
    for rec in boxes:
        if rec[20] == 0:
            want[int(rec[0])].append(pipeline.format_label_line(pipeline.NUSC_NAMES[int(rec[3])], rec, b2[int(rec[1])]))
    assert got == want
    # float32 tokens of the 2-D boxes: str(np.float32(v)) on 4e4 values over 24 decades plus special cases
    f = np.concatenate([(r.normal(size=40000) * 10.0 ** r.integers(-8, 16, 40000)).astype(np.float32),
                        np.array([0.0, -0.0, 1e16, 1e-5, 1e-4, 123456.0, 1e15, 16777216.0, 0.1, 1600.0, 899.5, np.inf, -np.inf, np.nan,
                                  1e-45, 3.4028235e38, 1e22, 100.0, 0.5, 1e8], np.float32)])
    f = f[:f.size // 4 * 4].reshape(-1, 4)
    out = np.empty(f.shape[0] * 100, np.uint8)
    off = np.zeros(f.shape[0], np.int32)
    nb = lib.dfu_box2d_tokens(_lib.hptr(f), f.shape[0], _lib.hptr(out), out.size, _lib.hptr(off))
    assert nb > 0
    toks = out[:nb].tobytes().decode().split('\0')[:-1]
    assert len(toks) == f.shape[0] and off[1] == len(toks[0]) + 1
    assert toks == [' '.join(str(np.float32(v)) for v in row) for row in f]
