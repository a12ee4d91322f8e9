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


# second module: the reference's CPU rotated-BEV IoU (pcdet/ops/iou3d_nms/src/iou3d_cpu.cpp:232-273) behind its own pybind
# table (iou3d_nms_api.cpp); the CUDA entry points the table takes the address of are aborting stand-ins (oracle/iou3d_gpu_stubs.cpp), never called
IOU_DIR = os.path.join(os.environ.get('DFU3D_REFERENCE', '/root/reference'), 'pcdet', 'ops', 'iou3d_nms', 'src')
IOU_SRC = [os.path.join(IOU_DIR, f) for f in ('iou3d_cpu.cpp', 'iou3d_nms_api.cpp')]
IOU_OUT = os.path.join(HERE, '_ref', 'iou3d_nms_cuda' + (sysconfig.get_config_var('EXT_SUFFIX') or '.so'))


def build_iou(force=False):
    if not all(os.path.isfile(f) for f in IOU_SRC):
        return None
    if os.path.exists(IOU_OUT) and not force and os.path.getmtime(IOU_OUT) >= max(os.path.getmtime(f) for f in IOU_SRC):
        return IOU_OUT
    import torch
    from torch.utils import cpp_extension as ce
    os.makedirs(os.path.dirname(IOU_OUT), exist_ok=True)
    inc = ['-I' + p for p in ce.include_paths()] + ['-I' + sysconfig.get_paths()['include'], '-I/usr/local/cuda/include']
    cmd = ['g++', '-O2', '-std=c++17', '-shared', '-fPIC', '-w', '-DTORCH_EXTENSION_NAME=iou3d_nms_cuda',
           '-DTORCH_API_INCLUDE_EXTENSION_H', '-D_GLIBCXX_USE_CXX11_ABI=%d' % int(torch._C._GLIBCXX_USE_CXX11_ABI)] + IOU_SRC + \
          [os.path.join(HERE, 'iou3d_gpu_stubs.cpp'), '-o', IOU_OUT, '-Wl,--unresolved-symbols=ignore-all'] + inc
    libdir = os.path.join(os.path.dirname(torch.__file__), 'lib')
    cmd += ['-L' + libdir, '-Wl,-rpath,' + libdir, '-lc10', '-ltorch_cpu', '-ltorch', '-ltorch_python']
    subprocess.check_call(cmd)
    return IOU_OUT


if __name__ == '__main__':
    print(build(force='--force' in sys.argv))
    print(build_iou(force='--force' in sys.argv))
