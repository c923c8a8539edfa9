"""vision_b200 - B200-native (sm_100a) implementation of Vision's megakernel path-tracing hot path.

The product is libvision_b200.so (hand-written CUDA behind the C ABI of include/vision_b200.h) and
the C++ host layer; this package is the Python harness used by tests and bench.py.
"""
from .capi import Context, VbError, load_library, LIB_PATH  # noqa: F401
