"""Builds tuning variants of libdfu3d.so (one source recompiled with -D flags, the other objects reused) into
gpurun_variants/ and prints the bench commands; select one at run time with DFU3D_LIB=<path>.
    python tools/variants.py dfu_densify.cu name1:-DS2_TILES=8 name2:-DS2_MINB=6,-DS2_TILES=4"""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from dfu3d_b200 import build as b   # noqa: E402

b.build()
src = os.path.join(b.CSRC, sys.argv[1])
out = os.path.join(ROOT, 'gpurun_variants')
os.makedirs(out, exist_ok=True)
others = [s[:-3] + '.o' for s in b.sources() if s != src]
for spec in sys.argv[2:]:
    name, flags = spec.split(':', 1)
    obj = os.path.join(out, name + '.o')
    subprocess.check_call([b.NVCC] + b.FLAGS + flags.split(',') + ['-c', src, '-o', obj])
    lib = os.path.join(out, 'libdfu3d_%s.so' % name)
    subprocess.check_call([b.NVCC, '-gencode', 'arch=compute_100a,code=sm_100a', '-shared', '-cudart', 'static', '-o', lib, obj] + others)
    os.remove(obj)
    print(lib)
