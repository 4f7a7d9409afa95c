"""elemental_b200 — B200-native (sm_100a) implementation of Hydrogen/Elemental's distributed
`El::Gemm` hot path: tcgen05 3xTF32 / DMMA local GEMMs, 128-bit BLAS-1 kernels, one-launch
lattice pack/unpack and NCCL redistributions behind the reference's Grid / DistMatrix / Gemm API.

  elemental_b200.el      Python mirror of the El:: interface (handles into libelb200.so)
  elemental_b200._lib    ctypes binding of include/elb200.h
  include/El.hpp         C++ facade with the reference's names for C++ callers
"""
from . import _lib  # noqa: F401  (fails loudly if libelb200.so is missing)
from . import el  # noqa: F401

__all__ = ["el", "_lib"]
