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
    assert _lib.VIEW_DTYPE.itemsize == 248 and _lib.FUSION_DTYPE.itemsize == 104 and _lib.INST_DTYPE.itemsize == 72


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
