"""terra_b200 — B200-native (sm_100a) implementation of Terra's path-tracing hot path behind
the reference's C API. The product is terra_b200/libterra_b200.so (CUDA, built from csrc/);
this package is the ctypes plumbing the tests and bench.py use to call it."""
from .api import TerraLib, Scene, make_camera, PRODUCT_LIB  # noqa: F401
