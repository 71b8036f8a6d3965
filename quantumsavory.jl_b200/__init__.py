"""qsb200 -- B200-native batched register-state backend for QuantumSavory's backend API.

Layout
  csrc/        CUDA kernels (sm_100a) + the C-ABI shim  -> libqsb200.so   (include/qsb200.h)
  _lib.py      ctypes binding of the C-ABI (the same calls the Julia `ccall` glue makes)
  backend.py   CudaBatchRepr / CudaBatchState: the backend generics (newstate, subsystemcompose,
               apply!, uptotime!, observable, project_traceout!, traceout!) on batched states
  registers.py Register / RegRef / StateRef bookkeeping, mirroring src/baseops/*.jl of the reference
  circuits.py  CircuitZoo circuits used by the configs (EntanglementSwap, Purify2to1, ...)
  symbols.py   kets, gates, backgrounds, Clifford lowering tables
"""
from . import symbols  # noqa: F401

__all__ = ["symbols"]
