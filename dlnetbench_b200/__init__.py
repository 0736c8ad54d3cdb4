"""dlnetbench_b200 — B200-native data-parallel hot path of DLNetBench.

The product is libdlnb.so (csrc/, C ABI in include/dlnb.h) and the C++ host side in
cpp/. `capi` is the ctypes binding used by tests and bench.py. No CPU fallback.
"""
