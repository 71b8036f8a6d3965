"""Pins the CPU oracle (oracle/qsb_oracle.c) before anything is compared against it.

Sources of the known answers (all under /root/reference, cited per test):
  * Random123 known-answer vectors for Philox4x32-10 (published with the generator)
  * closed forms / properties asserted by test/general/noninstant_and_backgrounds_qubit_tests.jl
  * doctest values in src/baseops/uptotime.jl:11-17 and src/states_registers.jl:58-61
  * an independent numpy kron/tensordot model (tests/np_dense.py) for index conventions
  * dense <-> stabilizer agreement on random Clifford circuits (physics pins the Clifford part)
"""
import itertools
import math

import numpy as np
import pytest

from oracle import oracle as orc
from qsb200 import symbols as S
from tests import np_dense as npd

RNG = np.random.default_rng(20261104)


def rand_ket(n):
    v = RNG.normal(size=2**n) + 1j * RNG.normal(size=2**n)
    return v / np.linalg.norm(v)


def rand_dm(n):
    a = RNG.normal(size=(2**n, 2**n)) + 1j * RNG.normal(size=(2**n, 2**n))
    rho = a @ a.conj().T
    return rho / np.trace(rho)


def rand_unitary(k):
    a = RNG.normal(size=(2**k, 2**k)) + 1j * RNG.normal(size=(2**k, 2**k))
    q, r = np.linalg.qr(a)
    return q * (np.diag(r) / np.abs(np.diag(r)))


# ---------------------------------------------------------------- RNG
def test_philox_known_answers():
    assert orc.philox([0, 0, 0, 0], [0, 0]) == [0x6627E8D5, 0xE169C58D, 0xBC57AC4C, 0x9B00DBD8]
    assert orc.philox([0xFFFFFFFF] * 4, [0xFFFFFFFF] * 2) == [0x408F276D, 0x41C83B0E, 0xA20BC7C6, 0x6D5451FD]
    assert orc.philox([0x243F6A88, 0x85A308D3, 0x13198A2E, 0x03707344], [0xA4093822, 0x299F31D0]) == \
        [0xD16CFE09, 0x94FDCCEB, 0x5001E420, 0x24126EA1]


def test_draw_maps_two_uniforms_onto_one_philox_block():
    seed, traj = 0x123456789ABCDEF0, 0x1_0000_0007
    for d in (0, 1, 6, 7, (1 << 33) + 5):
        b = d >> 1
        o = orc.philox([traj & 0xFFFFFFFF, traj >> 32, b & 0xFFFFFFFF, b >> 32], [seed & 0xFFFFFFFF, seed >> 32])
        lo, hi = (o[2], o[3]) if d & 1 else (o[0], o[1])
        assert orc.draw(seed, traj, d) == float(((hi << 32) | lo) >> 11) / 2.0 ** 53


def test_draw_is_uniform_53bit():
    us = np.array([orc.draw(0x5153, t, 7) for t in range(4000)])
    assert (us >= 0).all() and (us < 1).all()
    assert abs(us.mean() - 0.5) < 0.03 and abs(us.var() - 1 / 12) < 0.01
    assert orc.draw(1, 2, 3) != orc.draw(1, 2, 4) != orc.draw(1, 3, 3)


# ---------------------------------------------------------------- dense conventions
def test_tensor_is_kron_b_a():
    a, b = rand_ket(1), rand_ket(2)
    t = orc.Dense.from_array(a).tensor(orc.Dense.from_array(b)).array()
    assert np.allclose(t, np.kron(b, a), atol=1e-16)
    ra, rb = rand_dm(1), rand_dm(2)
    t = orc.Dense.from_array(ra).tensor(orc.Dense.from_array(rb)).array()
    assert np.allclose(t, np.kron(rb, ra), atol=1e-16)
    # Ket (x) Operator promotes the ket with dm(): quantumoptics.jl:5-6
    t = orc.Dense.from_array(a).tensor(orc.Dense.from_array(rb))
    assert t.kind == 1 and np.allclose(t.array(), np.kron(rb, np.outer(a, a.conj())), atol=1e-16)


@pytest.mark.parametrize("n,idx", [(1, [0]), (2, [1]), (3, [2, 0]), (3, [0, 1]), (4, [3, 1]), (4, [1, 2])])
def test_apply_matches_numpy_embed(n, idx):
    U = rand_unitary(len(idx))
    E = npd.embed(n, idx, U)
    psi = rand_ket(n)
    assert np.allclose(orc.Dense.from_array(psi).apply(idx, U).array(), E @ psi, atol=1e-14)
    rho = rand_dm(n)
    assert np.allclose(orc.Dense.from_array(rho).apply(idx, U).array(), E @ rho @ E.conj().T, atol=1e-14)


def test_cnot_first_index_is_control():
    # CNOT = |0><0| (x) I + |1><1| (x) X, control = first index (uptotime.jl:83)
    s = orc.Dense.from_array(S.tensor(S.Z2, S.Z1))  # qubit0=|1>, qubit1=|0>
    s.apply([0, 1], S.CNOT)
    assert np.allclose(s.array(), S.tensor(S.Z2, S.Z2))
    s = orc.Dense.from_array(S.tensor(S.Z2, S.Z1))
    s.apply([1, 0], S.CNOT)  # control is qubit1=|0>: nothing happens
    assert np.allclose(s.array(), S.tensor(S.Z2, S.Z1))


def test_doctest_values():
    # src/states_registers.jl:58-61 : observable(reg[1], X) on X1 -> 0.9999999999999998
    v = orc.Dense.from_array(S.X1).expect_op([0], S.X)
    assert abs(v - 1) < 3e-16 and v.imag == 0
    # src/baseops/uptotime.jl:11-17 : T1Decay(1.0), X1, uptotime 10 -> <Z> = 0.9999546000702374
    s = orc.Dense.from_array(S.X1)
    assert abs(s.expect_op([0], S.Z)) < 1e-16
    s.uptotime(0, S.T1Decay(1.0), 10.0)
    assert s.kind == 1
    assert abs(s.expect_op([0], S.Z).real - 0.9999546000702374) < 2e-16


@pytest.mark.parametrize("bg", [S.T1Decay(1.3), S.T2Dephasing(0.7), S.Depolarization(2.1), S.T1T2Noise(1.5, 0.9),
                                S.T1T2Noise(1.0, 5.0)])
def test_kraus_sum_matches_closed_form(bg):
    """Closed forms of SURVEY 8(a) a5 on the 2x2 block of the target qubit."""
    n, q, dt = 3, 1, 0.37
    rho = rand_dm(n)
    out = orc.Dense.from_array(rho).uptotime(q, bg, dt).array()
    t = rho.reshape(2, 2, 2, 2, 2, 2).copy()  # r2 r1 r0 c2 c1 c0
    blk = lambda a, b: (slice(None), a, slice(None), slice(None), b, slice(None))
    if bg.kind == S.BG_T1:
        p = math.exp(-dt / bg.t1)
        t00 = t[blk(0, 0)] + (1 - p) * t[blk(1, 1)]
        t[blk(1, 1)] *= p; t[blk(0, 1)] *= math.sqrt(p); t[blk(1, 0)] *= math.sqrt(p); t[blk(0, 0)] = t00
    elif bg.kind == S.BG_T2:
        f = math.exp(-dt / bg.t2)
        t[blk(0, 1)] *= f; t[blk(1, 0)] *= f
    elif bg.kind == S.BG_DEPOL:
        p = 1 - math.exp(-dt / bg.tau)
        a, b = t[blk(0, 0)].copy(), t[blk(1, 1)].copy()
        t[blk(0, 0)] = (1 - p / 2) * a + p / 2 * b
        t[blk(1, 1)] = (1 - p / 2) * b + p / 2 * a
        t[blk(0, 1)] *= 1 - p; t[blk(1, 0)] *= 1 - p
    else:
        p = math.exp(-dt / bg.t1)
        tphi_inv = 1 / bg.t2 - 1 / (2 * bg.t1)
        f = math.sqrt(p) * (math.exp(-dt * tphi_inv) if tphi_inv > 0 else 1.0)
        t00 = t[blk(0, 0)] + (1 - p) * t[blk(1, 1)]
        t[blk(1, 1)] *= p; t[blk(0, 1)] *= f; t[blk(1, 0)] *= f; t[blk(0, 0)] = t00
    assert np.allclose(out, t.reshape(8, 8), atol=1e-15)
    assert abs(np.trace(out) - 1) < 1e-14


def test_kraus_semigroup_and_t1t2_limit():
    # noninstant_and_backgrounds_qubit_tests.jl:72-96 (composition) and :149-170 (T2 > 2 T1 -> T1 only)
    rho = rand_dm(2)
    for bg in (S.T1Decay(2.0), S.T2Dephasing(3.0), S.Depolarization(1.0), S.T1T2Noise(2.0, 1.0)):
        a = orc.Dense.from_array(rho).uptotime(0, bg, 0.2).uptotime(0, bg, 0.5).array()
        b = orc.Dense.from_array(rho).uptotime(0, bg, 0.7).array()
        assert np.allclose(a, b, atol=1e-15)
    a = orc.Dense.from_array(rho).uptotime(1, S.T1T2Noise(1.0, 2.5), 0.3).array()
    b = orc.Dense.from_array(rho).uptotime(1, S.T1Decay(1.0), 0.3).array()
    assert np.allclose(a, b, atol=1e-16)


def test_measure_and_ptrace_match_numpy():
    n = 3
    rho = rand_dm(n)
    for q in range(n):
        assert np.allclose(orc.Dense.from_array(rho).ptrace(q).array(), npd.ptrace_dm(rho, n, q), atol=1e-15)
        for basis in "XYZ":
            psis = S.EIGVECS[basis]
            P = [np.trace(npd.project_dm(rho, n, q, p)).real for p in psis]
            assert abs(sum(P) - 1) < 1e-12
            for u, want in ((P[0] * 0.999, 1), (min(P[0] * 1.001, 0.999999), 2)):
                s = orc.Dense.from_array(rho)
                j = s.project_traceout(q, psis, u)
                assert j == want
                ref = npd.project_dm(rho, n, q, psis[j - 1])
                assert np.allclose(s.array(), ref / np.trace(ref), atol=1e-14)
    psi = rand_ket(n)
    s = orc.Dense.from_array(psi)
    j = s.project_traceout(1, S.EIGVECS["Y"], 0.0)
    assert j == 1 and s.kind == 0 and abs(np.linalg.norm(s.array()) - 1) < 1e-15
    full = orc.Dense.from_array(np.outer(psi, psi.conj()))
    full.project_traceout(1, S.EIGVECS["Y"], 0.0)
    assert np.allclose(np.outer(s.array(), s.array().conj()), full.array(), atol=1e-14)
    # ptrace of a Ket returns a density operator
    k = orc.Dense.from_array(psi).ptrace(0)
    assert k.kind == 1 and np.allclose(k.array(), npd.ptrace_dm(np.outer(psi, psi.conj()), n, 0), atol=1e-15)


def test_single_subsystem_measurement_and_error_string():
    # test/general/project_traceout_tests.jl:37-38,51-53
    s = orc.Dense.from_array(S.X1)
    assert s.project_traceout(0, S.EIGVECS["X"], 0.73) == 1
    bad = orc.Dense.from_array(np.array([[0.3, 0], [0, 0.3]], dtype=complex))
    with pytest.raises(orc.OracleError, match="State not normalized"):
        bad.project_traceout(0, S.EIGVECS["Z"], 0.5)


def test_observables():
    # test/general/observable_tests.jl:50-58 (Bell), :61-82 (projector probabilities)
    bell = orc.Dense.from_array(S.BELL)
    assert abs(bell.expect_proj([0, 1], S.BELL) - 1) < 1e-15
    assert abs(bell.expect_op([0, 1], S.tensor(S.X, S.X)) - 1) < 1e-15
    bell.apply([0], S.Y)
    assert abs(bell.expect_proj([0, 1], S.BELL)) < 1e-15
    assert abs(bell.expect_op([0, 1], S.tensor(S.X, S.X)) + 1) < 1e-15
    assert abs(orc.Dense.from_array(S.X1).expect_proj([0], S.Z1) - 0.5) < 1e-15
    ghz = orc.Dense.from_array(S.ghz_ket(3))
    assert abs(ghz.expect_proj([0, 2], S.BELL) - 0.5) < 1e-15
    # order flipping: projector qubit m sits on subsystem idx[m]
    ab = orc.Dense.from_array(S.tensor(S.X1, S.Z1))
    assert abs(ab.expect_proj([0, 1], S.tensor(S.X1, S.Z1)) - 1) < 1e-15
    assert abs(ab.expect_proj([1, 0], S.tensor(S.Z1, S.X1)) - 1) < 1e-15
    rho = rand_dm(3)
    O = rand_unitary(2)
    assert abs(orc.Dense.from_array(rho).expect_op([2, 0], O) - np.trace(npd.embed(3, [2, 0], O) @ rho)) < 1e-14


def test_mc_kraus_sampling():
    # uptotime.jl:20-42: one branch per draw; ensemble mean reproduces the Kraus map
    psi = rand_ket(2)
    bg = S.T1Decay(1.0)
    exact = orc.Dense.from_array(psi).uptotime(0, bg, 0.4).array()
    acc = np.zeros((4, 4), complex)
    N = 4000
    for t in range(N):
        s = orc.Dense.from_array(psi)
        s.uptotime_mc(0, bg, 0.4, orc.draw(1, t, 0))
        v = s.array()
        assert abs(np.linalg.norm(v) - 1) < 1e-14
        acc += np.outer(v, v.conj())
    assert np.abs(acc / N - exact).max() < 0.03


# ---------------------------------------------------------------- stabilizer part
def T(name):
    return S.clifford_table(S.NAMED_GATES[name])


def test_clifford_tables():
    assert list(T("X")) == [0, 1, 2 | 4, 3 | 4]  # I, X -> X, Z -> -Z, Y -> -Y
    assert list(T("H")) == [0, 2, 1, 3 | 4]      # X <-> Z, Y -> -Y
    t = T("CNOT")  # X(x)I -> X(x)X ; I(x)Z -> Z(x)Z
    assert t[0b0001] == 0b0101 and t[0b1000] == 0b1010 and t[0b0010] == 0b0010 and t[0b0100] == 0b0100
    with pytest.raises(ValueError):
        S.clifford_table(np.diag([1, np.exp(0.3j)]))
    for name in ("CNOT", "CPHASE", "XCZ", "ZCY", "XCY", "H", "S", "Y"):
        S.clifford_linear_form(T(name))


def test_stab_kats():
    # test/general/apply_tests.jl:6-15 : X(x)Z on Z1,X1 -> <Z>=<X>=-1
    s = orc.Stab.newstate().tensor(orc.Stab.from_spec("X"))
    s.apply([0], T("X")); s.apply([1], T("Z"))
    assert s.expect_pauli([0], "Z") == -1 and s.expect_pauli([1], "X") == -1
    # representations_dispatch_tests.jl:25-39 : CNOT on X1,Z1 -> xx=1, zz=1, x1=0, z2=0
    s = orc.Stab.from_spec("X").tensor(orc.Stab.newstate())
    s.apply([0, 1], T("CNOT"))
    assert (s.expect_pauli([0, 1], "XX"), s.expect_pauli([0, 1], "ZZ"), s.expect_pauli([0], "X"), s.expect_pauli([1], "Z")) == (1, 1, 0, 0)
    # traceout_tests.jl:29-70
    s = orc.Stab.from_spec("XII IYI IIZ")
    s.ptrace(1)
    assert s.n == 2 and s.expect_pauli([0], "X") == 1 and s.expect_pauli([1], "Z") == 1
    s = orc.Stab.from_spec("XXX ZZI IZZ")
    s.ptrace(1)
    assert s.n == 2 and s.rank == 1
    assert s.expect_pauli([0, 1], "ZZ") == 1 and s.expect_pauli([0], "X") == 0 and s.expect_pauli([1], "X") == 0
    # observable_tests.jl:61-82 projector probabilities
    assert orc.Stab.from_spec("X").expect_proj([0], "Z") == 0.5
    assert orc.Stab.from_spec("XXX ZZI IZZ").expect_proj([0, 2], "XX ZZ") == 0.5
    assert orc.Stab.from_generators(2, *S.stabilizer_generators("ZZ")).expect_proj([0, 1], "XX ZZ") == 0.5
    bell = orc.Stab.from_spec("XX ZZ")
    assert bell.expect_proj([0, 1], "XX ZZ") == 1.0
    bell.apply([0], T("Y"))
    assert bell.expect_proj([0, 1], "XX ZZ") == 0.0 and bell.expect_pauli([0, 1], "XX") == -1
    # project_traceout_tests.jl:11-38 : Bell measured in Y on both halves -> anticorrelated
    for u1, u2 in itertools.product((0.1, 0.9), repeat=2):
        b = orc.Stab.from_spec("XX ZZ")
        m1, r1 = b.project_traceout(0, "Y", u1)
        m2, r2 = b.project_traceout(0, "Y", u2)
        assert r1 and not r2 and m1 != m2 and m1 == (2 if u1 >= 0.5 else 1)
    assert orc.Stab.from_spec("X").project_traceout(0, "X", 0.99) == (1, False)
    # maximally mixed qubit: outcome random, rank grows (then the qubit is removed)
    mm = orc.Stab.maximally_mixed(2)
    assert mm.project_traceout(1, "Z", 0.7) == (2, True) and mm.n == 1 and mm.rank == 0


def test_stab_pauli_noise_thresholds():
    # clifford/uptotime.jl:1-12,26-39
    dt, tau = 0.5, 1.0
    p = 1 - math.exp(-dt / tau)
    s = orc.Stab.from_spec("Z")
    assert s.uptotime(0, S.Depolarization(tau), dt, 1 - p / 8) == 1      # X
    assert s.uptotime(0, S.Depolarization(tau), dt, 1 - 3 * p / 8) == 2  # Y
    assert s.uptotime(0, S.Depolarization(tau), dt, 1 - 5 * p / 8) == 3  # Z
    assert s.uptotime(0, S.Depolarization(tau), dt, 1 - 7 * p / 8) == 0
    assert s.uptotime(0, S.T2Dephasing(tau), dt, 1 - p / 4) == 3
    assert s.uptotime(0, S.T2Dephasing(tau), dt, 1 - 3 * p / 4) == 0
    with pytest.raises(orc.OracleError):
        s.uptotime(0, S.T1Decay(1.0), dt, 0.5)  # no paulinoise method in the reference


GATES1 = ["X", "Y", "Z", "H", "S"]
GATES2 = ["CNOT", "CPHASE", "XCZ", "ZCY", "XCY", "XCX", "YCZ"]


@pytest.mark.parametrize("seed", range(12))
def test_stabilizer_vs_dense_random_circuits(seed):
    """The physics pin for the Clifford oracle: on random Clifford circuits with mid-circuit
    measurements and partial traces, outcomes flagged 'determined' must have probability 1 in the
    dense model, 'random' ones probability 1/2, and all Pauli expectations must agree exactly."""
    rng = np.random.default_rng(seed)
    n = int(rng.integers(2, 6))
    st = orc.Stab.newstate()
    dn = orc.Dense.newstate()
    for _ in range(n - 1):
        st = st.tensor(orc.Stab.newstate())
        dn = dn.tensor(orc.Dense.newstate())
    if seed % 3 == 0:  # start from a mixed state: first qubit maximally mixed
        st = orc.Stab.maximally_mixed(1).tensor(orc.Stab.from_generators(n - 1, *S.stabilizer_generators(" ".join(
            "I" * i + "Z" + "I" * (n - 2 - i) for i in range(n - 1)))))
        dn = orc.Dense.from_array(np.eye(2) / 2)
        for _ in range(n - 1):
            dn = dn.tensor(orc.Dense.newstate())
    for step in range(40):
        cur = st.n
        r = rng.random()
        if r < 0.35 or cur == 1:
            g = GATES1[rng.integers(len(GATES1))]
            q = [int(rng.integers(cur))]
        elif r < 0.8:
            g = GATES2[rng.integers(len(GATES2))]
            q = [int(x) for x in rng.choice(cur, 2, replace=False)]
        elif r < 0.92 and cur > 1:
            q = int(rng.integers(cur)); basis = "XYZ"[rng.integers(3)]
            u = float(rng.random())
            j, was_random = st.project_traceout(q, basis, u)
            psis = S.EIGVECS[basis]
            probe = dn.copy()
            # probability of outcome j in the dense model
            full = probe.dm().array() if probe.kind == 0 else probe.array()
            pj = np.trace(npd.project_dm(full, cur, q, psis[j - 1])).real
            assert abs(pj - (0.5 if was_random else 1.0)) < 1e-9
            jd = dn.project_traceout(q, psis, 0.0 if j == 1 else 1.0 - 1e-12)
            assert jd == j
            continue
        else:
            if cur > 1:
                q = int(rng.integers(cur))
                st.ptrace(q); dn.ptrace(q)
            continue
        st.apply(q, T(g)); dn.apply(q, S.NAMED_GATES[g])
        # compare a few random Pauli strings
        cur = st.n
        for _ in range(3):
            k = int(rng.integers(1, cur + 1))
            idx = [int(x) for x in rng.choice(cur, k, replace=False)]
            ps = "".join("XYZ"[rng.integers(3)] for _ in range(k))
            ev = dn.expect_op(idx, S.tensor(*[S.PAULI[c] for c in ps]))
            assert abs(ev - st.expect_pauli(idx, ps)) < 1e-9
    # rank <-> purity: tr(rho^2) = 2^(r-n)
    full = dn.dm().array() if dn.kind == 0 else dn.array()
    assert abs(np.trace(full @ full).real - 2.0 ** (st.rank - st.n)) < 1e-9
