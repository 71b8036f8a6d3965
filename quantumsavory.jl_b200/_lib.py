"""ctypes binding of libqsb200.so -- the same symbols the Julia `ccall` glue binds.

There is no CPU fallback: if the shared library is missing or no CUDA device is present, the first
call raises.  The oracle under oracle/ is never imported from here.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(_HERE, "csrc")
SO_PATH = os.path.join(CSRC, "libqsb200.so")
HEADER = os.path.join(os.path.dirname(_HERE), "include", "qsb200.h")

QSB_OK, QSB_ERR_CUDA, QSB_ERR_BAD_INDEX, QSB_ERR_UNSUPPORTED, QSB_ERR_NOT_NORMALIZED, QSB_ERR_BAD_ARG, QSB_ERR_OOM, QSB_ERR_TRACE_ALL = range(8)
KET, DM, MCKET, TAB = 0, 1, 2, 3
BASIS_X, BASIS_Y, BASIS_Z, BASIS_KETS = 1, 2, 3, 4
OBS_PAULI, OBS_OP, OBS_PROJ_KET, OBS_PROJ_STAB = 0, 1, 2, 3
MASK_AND, MASK_OR, MASK_XOR, MASK_ANDNOT, MASK_NOT, MASK_XNOR = range(6)
RUN_KEEP_STATS = 1


class QsbError(RuntimeError):
    """Maps qsb_status codes onto the exception the reference raises at the same place."""

    def __init__(self, code, msg):
        super().__init__(msg)
        self.code = code


class MethodError(QsbError, TypeError):
    pass


class BoundsError(QsbError, IndexError):
    pass


class ArgumentError(QsbError, ValueError):
    pass


def build(force: bool = False, verbose: bool = False) -> str:
    """Compile csrc/*.cu for sm_100a into csrc/libqsb200.so (nvcc cross-compiles without a GPU)."""
    srcs = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cu", ".cuh"))] + [HEADER]
    stale = not os.path.exists(SO_PATH) or any(os.path.getmtime(s) > os.path.getmtime(SO_PATH) for s in srcs)
    if force or stale:
        cmd = ["make", "-C", CSRC, "-j8"] + (["-B"] if force else [])
        out = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
        if verbose or out.returncode:
            print(out.stdout)
        if out.returncode:
            raise RuntimeError("building libqsb200.so failed")
    return SO_PATH


_LIB = None


def _declare(L):
    vp, vpp = C.c_void_p, C.POINTER(C.c_void_p)
    ip, dp, u8p, u32p, i64p, u64p = (C.POINTER(t) for t in (C.c_int, C.c_double, C.c_uint8, C.c_uint32, C.c_int64, C.c_uint64))
    sig = {
        "qsb_last_error": (C.c_char_p, []),
        "qsb_version": (C.c_int, []),
        "qsb_ctx_create": (C.c_int, [C.c_int, C.c_int64, C.c_int64, C.c_uint64, vp, vpp]),
        "qsb_ctx_destroy": (C.c_int, [vp]),
        "qsb_ctx_set_stream": (C.c_int, [vp, vp]),
        "qsb_ctx_sync": (C.c_int, [vp]),
        "qsb_ctx_ntraj": (C.c_int64, [vp]),
        "qsb_ctx_set_traj_offset": (C.c_int, [vp, C.c_int64]),
        "qsb_ctx_counters": (C.c_int, [vp, u64p]),
        "qsb_ctx_reset_counters": (C.c_int, [vp]),
        "qsb_ctx_profile": (C.c_int, [vp, C.c_int]),
        "qsb_ctx_profile_report": (C.c_int, [vp, C.c_char_p, C.c_size_t]),
        "qsb_state_new": (C.c_int, [vp, C.c_int, C.c_int, vpp]),
        "qsb_state_from_host": (C.c_int, [vp, C.c_int, C.c_int, dp, vpp]),
        "qsb_tab_from_host": (C.c_int, [vp, C.c_int, C.c_int, u8p, u8p, u8p, vpp]),
        "qsb_state_select": (C.c_int, [vp, vpp, dp, C.c_int, C.c_uint64, vpp]),
        "qsb_state_copy": (C.c_int, [vp, vpp]),
        "qsb_state_free": (C.c_int, [vp]),
        "qsb_state_kind": (C.c_int, [vp]),
        "qsb_state_nsubsystems": (C.c_int, [vp]),
        "qsb_state_promote_dm": (C.c_int, [vp]),
        "qsb_state_to_host": (C.c_int, [vp, C.c_int64, dp]),
        "qsb_tab_to_host": (C.c_int, [vp, C.c_int64, u8p, u8p, u8p, u8p]),
        "qsb_tensor": (C.c_int, [vp, vpp, C.c_int, vpp]),
        "qsb_apply": (C.c_int, [vp, ip, C.c_int, dp, vp]),
        "qsb_apply_clifford": (C.c_int, [vp, ip, C.c_int, u8p, vp]),
        "qsb_channel": (C.c_int, [vp, C.c_int, C.c_int, dp, C.c_double, C.c_uint64]),
        "qsb_evolve": (C.c_int, [vp, ip, C.c_int, dp, C.c_int, ip, dp, C.c_double, C.c_uint64]),
        "qsb_measure": (C.c_int, [vp, C.c_int, C.c_int, dp, C.c_uint64, vpp, vpp, i64p]),
        "qsb_ptrace": (C.c_int, [vp, C.c_int, C.c_uint64]),
        "qsb_expect": (C.c_int, [vp, ip, C.c_int, C.c_int, vp, C.c_int, vp, vp, dp]),
        "qsb_expect_pauli_static": (C.c_int, [vp, ip, C.c_int, u8p, C.c_int, vp, vp]),
        "qsb_mask_create": (C.c_int, [vp, C.c_int, vpp]),
        "qsb_mask_free": (C.c_int, [vp]),
        "qsb_mask_logic": (C.c_int, [vp, vp, vp, C.c_int]),
        "qsb_mask_count": (C.c_int, [vp, i64p]),
        "qsb_mask_to_host": (C.c_int, [vp, u32p]),
        "qsb_mask_from_host": (C.c_int, [vp, u32p]),
        "qsb_mask_random": (C.c_int, [vp, C.c_double, C.c_uint64, vpp]),
        "qsb_record_begin": (C.c_int, [vp]),
        "qsb_record_end": (C.c_int, [vp, vpp]),
        "qsb_program_run": (C.c_int, [vp, C.c_uint64, C.c_uint64]),
        "qsb_program_run_lanes": (C.c_int, [vp, vp, C.c_int64, C.c_uint64, C.c_uint64, C.c_int]),
        "qsb_wave_run": (C.c_int, [vp, vpp, vpp, i64p, C.c_int, C.c_uint64, C.c_uint64, C.c_int]),
        "qsb_ctx_retire": (C.c_int, [vp, vp, C.c_int32, vp]),
        "qsb_cohort_count": (C.c_int, [vp, vp, vp, C.c_int, C.c_int, C.c_int32, vp]),
        "qsb_cohort_scatter": (C.c_int, [vp, vp, vp, C.c_int, C.c_int, C.c_int32, vp, vp, vp]),
        "qsb_program_info": (C.c_int, [vp, i64p]),
        "qsb_program_upload": (C.c_int, [vp, i64p]),
        "qsb_program_table_bytes": (C.c_int, [vp, C.POINTER(C.c_uint64)]),
        "qsb_program_export": (C.c_int, [vp, C.c_int, vp, C.c_size_t, C.POINTER(C.c_size_t)]),
        "qsb_program_free": (C.c_int, [vp]),
        "qsb_state_alive": (C.c_int, [vp, vpp]),
        "qsb_state_kill": (C.c_int, [vp, vp]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(L, name)  # AttributeError if the symbol is not exported
        fn.restype = res
        fn.argtypes = args
    return sig


EXPORTS = None


def lib():
    """Load libqsb200.so (building it if stale).  Raises if it cannot be loaded -- no fallback."""
    global _LIB, EXPORTS
    if _LIB is None:
        path = os.environ.get("QSB200_LIB") or SO_PATH   # QSB200_LIB: a differently tuned build of the same library
        if path == SO_PATH and not os.path.exists(SO_PATH):
            build()
        L = C.CDLL(path)
        EXPORTS = _declare(L)
        _LIB = L
    return _LIB


def check(rc: int):
    if rc == QSB_OK:
        return
    msg = lib().qsb_last_error().decode()
    if rc == QSB_ERR_UNSUPPORTED:
        raise MethodError(rc, msg)
    if rc == QSB_ERR_BAD_INDEX:
        raise BoundsError(rc, msg)
    if rc in (QSB_ERR_BAD_ARG, QSB_ERR_TRACE_ALL):
        raise ArgumentError(rc, msg)
    raise QsbError(rc, msg)
