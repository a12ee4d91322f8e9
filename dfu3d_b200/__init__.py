"""dfu3d_b200 — B200-native (sm_100a) drop-in for DFU3D's fusion / noise-suppression /
pseudo-box path (reference: tools/PENet/vis_utils.py:136-166,
tools/PENet/dataloaders/my_loader.py:502-684, tools/PENet/rectangle_fitting/rectangle_fitting.py).

The compute path is the C-ABI library ``dfu3d_b200/csrc/libdfu3d.so`` (include/dfu3d.h); there
is no CPU fallback: any compute entry point raises if the library or a CUDA device is missing.
"""
__version__ = '0.1.0'
