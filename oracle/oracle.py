"""ctypes front-end of the CPU oracle (oracle/qsb_oracle.c).

TEST INFRASTRUCTURE ONLY.  Importable from tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs; the product package never imports this module.
Indices are 0-based here (the reference's stateindex minus one).
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None


def build(force: bool = False) -> str:
    so = os.path.join(_HERE, "libqsb_oracle.so")
    src = os.path.join(_HERE, "qsb_oracle.c")
    if force or not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-s"] + (["-B"] if force else []))
    return so


def lib():
    global _LIB
    if _LIB is None:
        L = C.CDLL(build())
        vp, ip, dp, u8p = C.c_void_p, C.POINTER(C.c_int), C.POINTER(C.c_double), C.POINTER(C.c_uint8)
        L.qso_draw.restype = C.c_double
        L.qso_draw.argtypes = [C.c_uint64, C.c_uint64, C.c_uint64]
        L.qso_philox.argtypes = [C.POINTER(C.c_uint32)] * 3
        for name in ("qo_alloc", "qo_copy", "qo_newstate", "qo_from_data", "qo_dm", "qo_tensor",
                     "qc_alloc", "qc_copy", "qc_newstate", "qc_from_generators", "qc_tensor"):
            getattr(L, name).restype = vp
        L.qo_alloc.argtypes = [C.c_int, C.c_int]
        L.qo_copy.argtypes = [vp]
        L.qo_free.argtypes = [vp]
        L.qo_from_data.argtypes = [C.c_int, C.c_int, dp]
        L.qo_get.argtypes = [vp, dp]
        L.qo_dm.argtypes = [vp]
        L.qo_tensor.argtypes = [vp, vp]
        L.qo_kind.argtypes = [vp]
        L.qo_nsubsystems.argtypes = [vp]
        L.qo_apply.argtypes = [vp, ip, C.c_int, dp]
        L.qo_uptotime_dm.argtypes = [vp, C.c_int, C.c_int, dp, C.c_double]
        L.qo_uptotime_mc.argtypes = [vp, C.c_int, C.c_int, dp, C.c_double, C.c_double, ip]
        L.qo_project_traceout.argtypes = [vp, C.c_int, dp, C.c_double, ip]
        L.qo_ptrace.argtypes = [vp, C.c_int]
        L.qo_expect_op.argtypes = [vp, ip, C.c_int, dp, dp]
        L.qo_evolve_master.argtypes = [vp, ip, C.c_int, dp, C.c_int, ip, dp, C.c_double]
        L.qo_evolve_mcwf.argtypes = [vp, ip, C.c_int, dp, C.c_int, ip, dp, C.c_double, C.c_uint64, C.c_uint64, C.c_uint64, ip]
        L.qo_expect_proj.argtypes = [vp, ip, C.c_int, dp, dp]
        L.qc_copy.argtypes = [vp]
        L.qc_free.argtypes = [vp]
        L.qc_nqubits.argtypes = [vp]
        L.qc_rank.argtypes = [vp]
        L.qc_from_generators.argtypes = [C.c_int, C.c_int, u8p, u8p, u8p]
        L.qc_get_generators.argtypes = [vp, u8p, u8p, u8p]
        L.qc_tensor.argtypes = [vp, vp]
        L.qc_apply.argtypes = [vp, ip, C.c_int, u8p]
        L.qc_uptotime.argtypes = [vp, C.c_int, C.c_int, dp, C.c_double, C.c_double]
        L.qc_canonicalize.argtypes = [vp]
        L.qc_ptrace.argtypes = [vp, C.c_int]
        L.qc_project_traceout.argtypes = [vp, C.c_int, C.c_int, C.c_double, ip]
        L.qc_expect_pauli.argtypes = [vp, ip, C.c_int, u8p, C.c_int]
        L.qc_expect_proj.restype = C.c_double
        L.qc_expect_proj.argtypes = [vp, ip, C.c_int, u8p, u8p, u8p]
        _LIB = L
    return _LIB


def draw(seed: int, traj: int, draw_id: int) -> float:
    return lib().qso_draw(seed, traj, draw_id)


def philox(ctr, key):
    c = (C.c_uint32 * 4)(*ctr)
    k = (C.c_uint32 * 2)(*key)
    o = (C.c_uint32 * 4)()
    lib().qso_philox(c, k, o)
    return list(o)


def _ri(a) -> np.ndarray:
    a = np.ascontiguousarray(np.asarray(a, dtype=np.complex128))
    return a.view(np.float64).reshape(-1)


def _dp(a: np.ndarray):
    return a.ctypes.data_as(C.POINTER(C.c_double))


def _ip(idx):
    return (C.c_int * len(idx))(*[int(i) for i in idx])


def _u8(a):
    a = np.ascontiguousarray(np.asarray(a, dtype=np.uint8))
    return a, a.ctypes.data_as(C.POINTER(C.c_uint8))


class OracleError(RuntimeError):
    def __init__(self, code):
        super().__init__({3: "unsupported", 4: "State not normalized. Could be due to passing wrong state to `initialize!`",
                          7: "Partial trace can't be used to trace out all subsystems - use tr() instead."}.get(code, f"oracle error {code}"))
        self.code = code


class Dense:
    """One trajectory's Ket / Operator (kind 0 / 1)."""

    def __init__(self, ptr):
        self.p = ptr

    @classmethod
    def newstate(cls):
        return cls(lib().qo_newstate())

    @classmethod
    def from_array(cls, a):
        a = np.asarray(a, dtype=np.complex128)
        if a.ndim == 1:
            n = int(round(np.log2(a.size)))
            return cls(lib().qo_from_data(0, n, _dp(_ri(a))))
        n = int(round(np.log2(a.shape[0])))
        return cls(lib().qo_from_data(1, n, _dp(_ri(np.asfortranarray(a).T.copy()))))

    def __del__(self):
        if getattr(self, "p", None):
            lib().qo_free(self.p)
            self.p = None

    def copy(self):
        return Dense(lib().qo_copy(self.p))

    @property
    def kind(self):
        return lib().qo_kind(self.p)

    @property
    def n(self):
        return lib().qo_nsubsystems(self.p)

    def array(self) -> np.ndarray:
        L = (1 << self.n) if self.kind == 0 else (1 << (2 * self.n))
        out = np.empty(2 * L)
        lib().qo_get(self.p, _dp(out))
        v = out.view(np.complex128)
        if self.kind == 0:
            return v.copy()
        D = 1 << self.n
        return v.reshape(D, D).T.copy()  # stored column-major

    def dm(self):
        return Dense(lib().qo_dm(self.p))

    def tensor(self, other):
        return Dense(lib().qo_tensor(self.p, other.p))

    def apply(self, idx, U):
        U = np.asarray(U, dtype=np.complex128)
        lib().qo_apply(self.p, _ip(idx), len(idx), _dp(_ri(U.T.copy())))
        return self

    def uptotime(self, idx, bg, dt):
        prm = np.array(bg.params, dtype=np.float64)
        rc = lib().qo_uptotime_dm(self.p, int(idx), bg.kind, _dp(prm), float(dt))
        if rc:
            raise OracleError(rc)
        return self

    def uptotime_mc(self, idx, bg, dt, u):
        prm = np.array(bg.params, dtype=np.float64)
        br = C.c_int(0)
        rc = lib().qo_uptotime_mc(self.p, int(idx), bg.kind, _dp(prm), float(dt), float(u), C.byref(br))
        if rc:
            raise OracleError(rc)
        return br.value

    @staticmethod
    def _evo_args(idx, H, lind):
        """lind: [(position in idx, 2x2 collapse operator)]"""
        Hri = None if H is None else _ri(np.asarray(H, dtype=np.complex128).T.copy())
        pos = np.array([p for p, _ in lind], dtype=np.int32)
        Lri = _ri(np.stack([np.asarray(L, dtype=np.complex128).T.copy() for _, L in lind])) if lind else None
        return (_ip(idx), len(idx), None if Hri is None else _dp(Hri), len(lind),
                pos.ctypes.data_as(C.POINTER(C.c_int)) if lind else None, None if Lri is None else _dp(Lri)), (Hri, pos, Lri)

    def evolve_master(self, idx, H, lind, duration):
        """timeevolution.master on the embedded operators (adaptive DP5, reltol 1e-6 / abstol 1e-8)."""
        args, keep = self._evo_args(idx, H, lind)
        rc = lib().qo_evolve_master(self.p, *args, float(duration))
        del keep
        if rc:
            raise OracleError(rc)
        return self

    def evolve_mcwf(self, idx, H, lind, duration, seed, traj, draw_id):
        """timeevolution.schroedinger / mcwf on a trajectory ket; returns the number of jumps."""
        args, keep = self._evo_args(idx, H, lind)
        nj = C.c_int(0)
        rc = lib().qo_evolve_mcwf(self.p, *args, float(duration), int(seed), int(traj), int(draw_id), C.byref(nj))
        del keep
        if rc:
            raise OracleError(rc)
        return nj.value

    def project_traceout(self, idx, psis, u):
        ps = np.stack([np.asarray(p, dtype=np.complex128) for p in psis])
        j = C.c_int(0)
        rc = lib().qo_project_traceout(self.p, int(idx), _dp(_ri(ps)), float(u), C.byref(j))
        if rc:
            raise OracleError(rc)
        return j.value

    def ptrace(self, idx):
        rc = lib().qo_ptrace(self.p, int(idx))
        if rc:
            raise OracleError(rc)
        return self

    def expect_op(self, idx, O):
        O = np.asarray(O, dtype=np.complex128)
        out = np.empty(2)
        lib().qo_expect_op(self.p, _ip(idx), len(idx), _dp(_ri(O.T.copy())), _dp(out))
        return complex(out[0], out[1])

    def expect_proj(self, idx, psi):
        out = np.empty(2)
        lib().qo_expect_proj(self.p, _ip(idx), len(idx), _dp(_ri(psi)), _dp(out))
        return complex(out[0], out[1])


class Stab:
    """One trajectory's (possibly mixed) stabilizer state."""

    def __init__(self, ptr):
        self.p = ptr

    @classmethod
    def newstate(cls):
        return cls(lib().qc_newstate())

    @classmethod
    def from_generators(cls, n, xs, zs, signs):
        xs_, xp = _u8(xs)
        zs_, zp = _u8(zs)
        sg_, sp = _u8(signs)
        return cls(lib().qc_from_generators(int(n), int(len(sg_)), xp, zp, sp))

    @classmethod
    def from_spec(cls, spec, n=None):
        from importlib import import_module
        sym = import_module("qsb200.symbols")
        xs, zs, signs = sym.stabilizer_generators(spec)
        return cls.from_generators(xs.shape[1] if n is None else n, xs, zs, signs)

    @classmethod
    def maximally_mixed(cls, n):
        return cls.from_generators(n, np.zeros((0, n), np.uint8), np.zeros((0, n), np.uint8), np.zeros(0, np.uint8))

    def __del__(self):
        if getattr(self, "p", None):
            lib().qc_free(self.p)
            self.p = None

    def copy(self):
        return Stab(lib().qc_copy(self.p))

    @property
    def n(self):
        return lib().qc_nqubits(self.p)

    @property
    def rank(self):
        return lib().qc_rank(self.p)

    def generators(self, canonical=True):
        s = self.copy() if canonical else self
        if canonical:
            lib().qc_canonicalize(s.p)
        n, r = s.n, s.rank
        xs = np.zeros((max(r, 1), max(n, 1)), np.uint8)
        zs = np.zeros_like(xs)
        sg = np.zeros(max(r, 1), np.uint8)
        lib().qc_get_generators(s.p, xs.ctypes.data_as(C.POINTER(C.c_uint8)), zs.ctypes.data_as(C.POINTER(C.c_uint8)),
                                sg.ctypes.data_as(C.POINTER(C.c_uint8)))
        # C side writes with row stride n: re-read with the right stride
        flatx = xs.reshape(-1)[: r * n].reshape(r, n) if n else np.zeros((r, 0), np.uint8)
        flatz = zs.reshape(-1)[: r * n].reshape(r, n) if n else np.zeros((r, 0), np.uint8)
        return flatx.copy(), flatz.copy(), sg[:r].copy()

    def tensor(self, other):
        return Stab(lib().qc_tensor(self.p, other.p))

    def apply(self, idx, table):
        t, tp = _u8(table)
        lib().qc_apply(self.p, _ip(idx), len(idx), tp)
        return self

    def uptotime(self, idx, bg, dt, u):
        prm = np.array(bg.params, dtype=np.float64)
        rc = lib().qc_uptotime(self.p, int(idx), bg.kind, _dp(prm), float(dt), float(u))
        if rc < 0:
            raise OracleError(-rc)
        return rc

    def project_traceout(self, idx, basis, u):
        """basis in 'X','Y','Z'; returns (outcome 1|2, was_random)."""
        wr = C.c_int(0)
        j = lib().qc_project_traceout(self.p, int(idx), {"X": 1, "Y": 2, "Z": 3}[basis], float(u), C.byref(wr))
        return j, bool(wr.value)

    def ptrace(self, idx):
        lib().qc_ptrace(self.p, int(idx))
        return self

    def expect_pauli(self, idx, paulis, sign=0):
        codes = np.array([{"I": 0, "X": 1, "Y": 2, "Z": 3}[p] for p in paulis], np.uint8)
        c, cp = _u8(codes)
        return lib().qc_expect_pauli(self.p, _ip(idx), len(idx), cp, int(sign))

    def expect_proj(self, idx, spec):
        from importlib import import_module
        sym = import_module("qsb200.symbols")
        xs, zs, signs = sym.stabilizer_generators(spec)
        assert xs.shape[0] == xs.shape[1] == len(idx)
        x_, xp = _u8(xs)
        z_, zp = _u8(zs)
        s_, sp = _u8(signs)
        return lib().qc_expect_proj(self.p, _ip(idx), len(idx), xp, zp, sp)
