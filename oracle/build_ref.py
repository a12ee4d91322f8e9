"""TEST INFRASTRUCTURE ONLY — compiles the reference's OWN C++ source of `points_in_boxes_cpu`
(/root/reference/pcdet/ops/roiaware_pool3d/src/roiaware_pool3d.cpp:112-168) where it lies, into oracle/_ref/ (git-ignored,
travels to the GPU box).  No reference source is copied; the CUDA launchers the file declares stay unresolved (they are
never called by the CPU entry point).  Usable only where /root/reference exists; the built module is imported by
oracle/ref_points_in_boxes.py."""
import os
import subprocess
import sys
import sysconfig

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(os.environ.get('DFU3D_REFERENCE', '/root/reference'), 'pcdet', 'ops', 'roiaware_pool3d', 'src',
                   'roiaware_pool3d.cpp')
OUT = os.path.join(HERE, '_ref', 'roiaware_pool3d_cuda' + (sysconfig.get_config_var('EXT_SUFFIX') or '.so'))


def build(force=False):
    if not os.path.isfile(SRC):
        return None
    if os.path.exists(OUT) and not force and os.path.getmtime(OUT) >= os.path.getmtime(SRC):
        return OUT
    import torch
    from torch.utils import cpp_extension as ce
    os.makedirs(os.path.dirname(OUT), exist_ok=True)
    inc = ['-I' + p for p in ce.include_paths()] + ['-I' + sysconfig.get_paths()['include']]
    cmd = ['g++', '-O2', '-std=c++17', '-shared', '-fPIC', '-w', '-DTORCH_EXTENSION_NAME=roiaware_pool3d_cuda',
           '-DTORCH_API_INCLUDE_EXTENSION_H', '-D_GLIBCXX_USE_CXX11_ABI=%d' % int(torch._C._GLIBCXX_USE_CXX11_ABI), SRC, '-o', OUT,
           '-Wl,--unresolved-symbols=ignore-all'] + inc
    libdir = os.path.join(os.path.dirname(torch.__file__), 'lib')
    cmd += ['-L' + libdir, '-Wl,-rpath,' + libdir, '-lc10', '-ltorch_cpu', '-ltorch', '-ltorch_python']
    subprocess.check_call(cmd)
    return OUT


if __name__ == '__main__':
    print(build(force='--force' in sys.argv))
