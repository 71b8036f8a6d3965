"""ctypes wrapper of oracle/qsb_bitsliced.cpp: the optimised (bit-sliced, OpenMP) CPU replay of a recorded tableau program.
TEST / BENCH INFRASTRUCTURE ONLY (bench.py's `cpu_baseline_optimised`, tests/test_gpu_bitsliced_cpu.py)."""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None


def lib():
    global _LIB
    if _LIB is None:
        so = os.path.join(HERE, "libqsb_bitsliced.so")
        src = [os.path.join(HERE, "qsb_bitsliced.cpp")] + [os.path.join(HERE, "..", "quantumsavory.jl_b200", "csrc", f) for f in ("qsb_tab_core.cuh", "qsb_exec_kernel.cuh")]
        if not os.path.exists(so) or any(os.path.getmtime(s) > os.path.getmtime(so) for s in src):
            subprocess.run(["make", "-C", HERE, "libqsb_bitsliced.so"], check=True, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
        L = C.CDLL(so)
        L.bs_run.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64, C.c_uint64, C.c_uint64, C.c_int64,
                             C.POINTER(C.c_double), C.POINTER(C.c_int)]
        _LIB = L
    return _LIB


def run(export: dict, seed: int, traj0: int, ntraj: int, draw_base: int = 0):
    """export: Program.export() of a tableau program.  Returns (stats rows [nstats, 4], threads used)."""
    nstats = int(export["meta"][6])
    out = np.zeros((max(1, nstats), 4))
    thr = C.c_int(0)
    keep = [np.frombuffer(export[k], np.uint8) if len(export[k]) else np.zeros(1, np.uint8) for k in ("ops", "specs", "obs", "aux")]
    stats = np.ascontiguousarray(export["stats"]) if len(export["stats"]) else np.zeros(1, np.uint64)
    meta = np.ascontiguousarray(export["meta"])
    rc = lib().bs_run(*[k.ctypes.data_as(C.c_void_p) for k in keep], stats.ctypes.data_as(C.c_void_p), meta.ctypes.data_as(C.c_void_p),
                      int(seed), int(traj0), int(draw_base), int(ntraj), out.ctypes.data_as(C.POINTER(C.c_double)), C.byref(thr))
    if rc:
        raise RuntimeError(f"bs_run failed: {rc}")
    return out[:nstats], thr.value
