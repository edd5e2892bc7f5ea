"""B200-native batched llama2 forward pass behind the thaBLAS / thaDNN call surface.

Layout:
  csrc/        CUDA kernels for sm_100a + the extern "C" boundary (include/tha_b200.h)
  host/        C++ host side mirroring the reference headers (thaBLAS.hpp, thaDNN.hpp, ...)
  cabi.py      ctypes binding of the C-ABI library (used by tests and bench.py)
  checkpoint.py llama2.c v0 / v2 checkpoint formats
  build.py     nvcc / g++ build recipes (in-tree, sm_100a only)
"""
