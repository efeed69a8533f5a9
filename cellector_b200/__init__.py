"""cellector_b200 — B200-native (sm_100a) implementation of cellector's per-cell genotype-scoring EM loop.

The product is `libcellector_cuda.so` (C ABI: include/cellector_cuda.h) and the `cellector` CLI built on it;
this package holds the CUDA/C++ sources (csrc/), the build recipe and a ctypes binding used by tests/bench.
"""
from .api import Cellector, IterStats, LABELS, load_library  # noqa: F401
