"""optimo_b200 -- B200-native objective(+gradient) engine for OptiMo's optimisation hot path.

The product is two in-tree shared libraries built by `optimo_b200/Makefile`:

  csrc/libomo.so          hand-written sm_100a kernels behind the C ABI of include/omo.h
  host/liboptimo_core.so  C++ headless `Core` mirroring the reference's operator API (include/optimo_core.h)

This Python package is only a ctypes view of those two for tests and the benchmark harness. It contains no
arithmetic and no fallback: importing `optimo_b200.core` without the built libraries raises.
"""
from .core import Core, Engine, EngineError, ForcedStop, lib_paths  # noqa: F401
