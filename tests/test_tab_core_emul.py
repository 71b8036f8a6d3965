"""CPU-only check of the bit-sliced tableau core (csrc/qsb_tab_core.cuh) against the oracle.

The CUDA kernels in csrc/qsb_tab.cu are one-line wrappers around the inline functions compiled here
with g++ (tests/cpu_emul/tab_emul.cpp).  This does NOT replace the GPU parity tests
(tests/test_gpu_*.py call the real kernels through the C-ABI); it lets the lane-masked pivot logic be
exercised on every commit in the CPU-only suite: 64 lanes, each lane a different trajectory because
noise and measurement draws differ per lane, each lane compared bit-exactly with its own oracle run.
"""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

from oracle import oracle as orc
from qsb200 import symbols as S

HERE = os.path.dirname(os.path.abspath(__file__))
EMU_DIR = os.path.join(HERE, "cpu_emul")


def _build():
    so = os.path.join(EMU_DIR, "libtab_emul.so")
    srcs = [os.path.join(EMU_DIR, "tab_emul.cpp"), os.path.join(HERE, "..", "quantumsavory.jl_b200", "csrc", "qsb_tab_core.cuh")]
    if not os.path.exists(so) or any(os.path.getmtime(s) > os.path.getmtime(so) for s in srcs):
        subprocess.check_call(["g++", "-O1", "-fPIC", "-shared", "-std=c++17", "-o", so, srcs[0]])
    L = C.CDLL(so)
    vp = C.c_void_p
    L.et_new.restype = vp; L.et_new.argtypes = [C.c_int, C.c_size_t, C.c_int]
    L.et_free.argtypes = [vp]
    L.et_n.argtypes = [vp]
    L.et_copy.restype = vp; L.et_copy.argtypes = [vp]
    L.et_force.argtypes = [vp, C.c_uint32, C.c_uint32, C.c_int]
    L.et_tensor.restype = vp; L.et_tensor.argtypes = [vp, vp]
    L.et_gate.argtypes = [vp, C.POINTER(C.c_int), C.c_int, C.POINTER(C.c_uint8), C.POINTER(C.c_uint32)]
    L.et_noise.argtypes = [vp, C.c_int, C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_int), C.c_uint64, C.c_uint64, C.c_uint64]
    L.et_measure.argtypes = [vp, C.c_int, C.c_int, C.c_int, C.c_uint64, C.c_uint64, C.c_uint64, C.POINTER(C.c_uint32), C.POINTER(C.c_uint32)]
    L.et_ptrace.argtypes = [vp, C.c_int]
    L.et_expect.argtypes = [vp, C.c_int, C.POINTER(C.c_uint32), C.POINTER(C.c_uint32), C.POINTER(C.c_uint8), C.POINTER(C.c_int)]
    L.et_lane.argtypes = [vp, C.c_size_t] + [C.POINTER(C.c_uint8)] * 4
    L.et_draw.restype = C.c_double; L.et_draw.argtypes = [C.c_uint64] * 3
    return L


EMU = _build()
W = 2
LANES = 32 * W
SEED = 0x5153


def T(name):
    return S.clifford_table(S.NAMED_GATES[name])


class Emu:
    def __init__(self, ptr):
        self.p = ptr

    @classmethod
    def from_spec(cls, spec, n=None):
        xs, zs, sg = S.stabilizer_generators(spec)
        n = xs.shape[1] if n is None else n
        t = cls(EMU.et_new(n, W, 0))
        for g in range(xs.shape[0]):
            px = sum(int(xs[g, q]) << q for q in range(n))
            pz = sum(int(zs[g, q]) << q for q in range(n))
            EMU.et_force(t.p, px, pz, int(sg[g]))
        return t

    @classmethod
    def zero(cls, n):
        return cls(EMU.et_new(n, W, 1))

    @classmethod
    def mixed(cls, n):
        return cls(EMU.et_new(n, W, 0))

    @property
    def n(self):
        return EMU.et_n(self.p)

    def tensor(self, o):
        return Emu(EMU.et_tensor(self.p, o.p))

    def gate(self, idx, name, mask=None):
        t = T(name)
        idx_c = (C.c_int * len(idx))(*idx)
        m = None if mask is None else (C.c_uint32 * W)(*mask)
        rc = EMU.et_gate(self.p, idx_c, len(idx), t.ctypes.data_as(C.POINTER(C.c_uint8)), m)
        assert rc == 0

    def noise(self, q, bg, dt, draw_id):
        import math
        if bg.kind == S.BG_T2:
            p = 1 - math.exp(-dt / bg.t2); probs, ops = [p / 2], [3]
        else:
            p = 1 - math.exp(-dt / bg.tau); probs, ops = [p / 4] * 3, [1, 2, 3]
        cum, c = [], 1.0
        for pr in probs:
            c -= pr; cum.append(c)
        EMU.et_noise(self.p, q, len(ops), (C.c_double * 3)(*(cum + [0] * (3 - len(cum)))), (C.c_int * 3)(*(ops + [0] * (3 - len(ops)))), SEED, 0, draw_id)

    def measure(self, q, basis, draw_id):
        tx, tz = {"X": (1, 0), "Y": (1, 1), "Z": (0, 1)}[basis]
        o = (C.c_uint32 * W)(); r = (C.c_uint32 * W)()
        EMU.et_measure(self.p, q, tx, tz, SEED, 0, draw_id, o, r)
        bits = lambda a: [(a[l // 32] >> (l % 32)) & 1 for l in range(LANES)]
        return bits(o), bits(r)

    def ptrace(self, q):
        EMU.et_ptrace(self.p, q)

    def expect_pauli(self, idx, paulis, sign=0):
        px = sum(1 << q for q, c in zip(idx, paulis) if c in "XY")
        pz = sum(1 << q for q, c in zip(idx, paulis) if c in "ZY")
        vals = (C.c_int * LANES)()
        EMU.et_expect(self.p, 1, (C.c_uint32 * 1)(px), (C.c_uint32 * 1)(pz), (C.c_uint8 * 1)(sign), vals)
        return list(vals)

    def lane(self, l):
        n = self.n
        xs = np.zeros((2 * n, n), np.uint8); zs = np.zeros_like(xs); sg = np.zeros(2 * n, np.uint8); ac = np.zeros(n, np.uint8)
        u8 = lambda a: a.ctypes.data_as(C.POINTER(C.c_uint8))
        EMU.et_lane(self.p, l, u8(xs), u8(zs), u8(sg), u8(ac))
        return xs, zs, sg, ac


def check_lane_against_oracle(emu, l, ora):
    n = emu.n
    assert n == ora.n
    xs, zs, sg, ac = emu.lane(l)
    # symplectic structure of the pair tableau
    def comm(a, b):
        return int(np.sum(xs[a] & zs[b]) + np.sum(zs[a] & xs[b])) & 1
    for a in range(2 * n):
        for b in range(2 * n):
            want = 1 if (a < n and b == a + n) or (b < n and a == b + n) else 0
            assert comm(a, b) == want, (l, a, b)
    act = [i for i in range(n) if ac[i]]
    assert len(act) == ora.rank
    st = orc.Stab.from_generators(n, xs[n:][act], zs[n:][act], sg[n:][act]) if act else orc.Stab.maximally_mixed(n)
    gx, gz, gs = st.generators(canonical=True)
    ox, oz, os_ = ora.generators(canonical=True)
    assert np.array_equal(gx, ox) and np.array_equal(gz, oz) and np.array_equal(gs, os_), l


def test_draws_agree_with_oracle():
    for t, d in ((0, 0), (5, 9), (2**33 + 7, 2**40 + 3)):
        assert EMU.et_draw(SEED, t, d) == orc.draw(SEED, t, d)


def test_build_from_generators_and_expect():
    for spec in ("XX ZZ", "-XX ZZ", "XXX ZZI IZZ", "X_Z _XZ ZZX".replace("_", "I"), "ZZ", "YY"):
        e = Emu.from_spec(spec)
        o = orc.Stab.from_spec(spec)
        for l in (0, 17, 63):
            check_lane_against_oracle(e, l, o)
    e = Emu.from_spec("XXX ZZI IZZ")
    assert set(e.expect_pauli([0, 1, 2], "XXX")) == {1} and set(e.expect_pauli([0, 1], "ZZ")) == {1}
    assert set(e.expect_pauli([0], "X")) == {0} and set(e.expect_pauli([0, 1, 2], "YYX")) == {-1}


GATES1 = ["X", "Y", "Z", "H", "S"]
GATES2 = ["CNOT", "CPHASE", "XCZ", "ZCY", "XCY", "XCX", "YCZ"]


ORA_LANES = (0, 9, 21, 31, 32, 47, 63)  # lanes followed by an oracle twin (both words, both edges)


@pytest.mark.parametrize("seed", range(8))
def test_random_circuits_lanewise(seed):
    rng = np.random.default_rng(1000 + seed)
    n = int(rng.integers(2, 7))
    emu = Emu.zero(n)
    oras = {}
    for l in ORA_LANES:
        o = orc.Stab.newstate()
        for _ in range(n - 1):
            o = o.tensor(orc.Stab.newstate())
        oras[l] = o
    if seed % 2:
        emu = Emu.mixed(1).tensor(emu)
        oras = {l: orc.Stab.maximally_mixed(1).tensor(o) for l, o in oras.items()}
    draw_id = 0
    bgs = [S.Depolarization(0.4), S.T2Dephasing(0.3)]
    for step in range(50):
        cur = emu.n
        r = rng.random()
        if r < 0.25:
            g = GATES1[rng.integers(len(GATES1))]; q = [int(rng.integers(cur))]
            emu.gate(q, g)
            for o in oras.values(): o.apply(q, T(g))
        elif r < 0.55 and cur >= 2:
            g = GATES2[rng.integers(len(GATES2))]; q = [int(x) for x in rng.choice(cur, 2, replace=False)]
            emu.gate(q, g)
            for o in oras.values(): o.apply(q, T(g))
        elif r < 0.75:
            q = int(rng.integers(cur)); bg = bgs[rng.integers(2)]; dt = float(rng.random())
            emu.noise(q, bg, dt, draw_id)
            for l, o in oras.items(): o.uptotime(q, bg, dt, orc.draw(SEED, l, draw_id))
            draw_id += 1
        elif r < 0.88 and cur >= 2:
            q = int(rng.integers(cur)); basis = "XYZ"[rng.integers(3)]
            out, rnd = emu.measure(q, basis, draw_id)
            for l, o in oras.items():
                j, wr = o.project_traceout(q, basis, orc.draw(SEED, l, draw_id))
                assert (j == 2) == bool(out[l]) and wr == bool(rnd[l]), (step, l)
            draw_id += 1
        elif r < 0.94 and cur >= 2:
            q = int(rng.integers(cur))
            emu.ptrace(q)
            for o in oras.values(): o.ptrace(q)
        elif cur <= 5:
            spec = ["XX ZZ", "Z", "ZZ"][rng.integers(3)]
            add = Emu.from_spec(spec)
            emu = emu.tensor(add)
            oras = {l: o.tensor(orc.Stab.from_spec(spec)) for l, o in oras.items()}
        cur = emu.n
        k = int(rng.integers(1, cur + 1))
        idx = [int(x) for x in rng.choice(cur, k, replace=False)]
        ps = "".join("XYZ"[rng.integers(3)] for _ in range(k))
        vals = emu.expect_pauli(idx, ps)
        for l in ORA_LANES:
            assert vals[l] == oras[l].expect_pauli(idx, ps), (step, l)
    for l in ORA_LANES:
        check_lane_against_oracle(emu, l, oras[l])


def test_masked_gate_only_touches_selected_lanes():
    e = Emu.from_spec("XX ZZ")
    e.gate([0], "Z", mask=[0x0000FFFF, 0])
    v = e.expect_pauli([0, 1], "XX")
    assert v[:16] == [-1] * 16 and v[16:] == [1] * 48
