"""The reference's own hot-path tests, re-expressed once and run against two backends:
the CPU oracle (tests/test_kats_oracle.py, every commit) and the CUDA batched backend through the
C-ABI (tests/test_gpu_kats.py, on the B200).  Each function cites the reference test it restates
(paths relative to /root/reference/test/general/).

`env` supplies: env.repr(formalism) -> representation object; env.values(x) -> per-trajectory numpy
values of an observable result; env.outcomes(j) -> per-trajectory outcomes; env.truth(b) -> bools.
Outcomes differ between trajectories; every assertion below must hold on EVERY trajectory.
"""
import numpy as np
import pytest

from qsb200 import ops
from qsb200 import symbols as S
from qsb200.circuits import EntanglementSwap, LocalEntanglementSwap, Purify2to1, Purify2to1Node, Purify3to1
from qsb200.ops import CNOT, CPHASE, X, Y, Z, tensor
from qsb200.registers import (Register, apply, initialize, isassigned, observable, project_traceout,
                              stateof, traceout, uptotime)

bell = ops.StabilizerState("XX ZZ")
ALL_FORMALISMS = ("qo", "mc", "clifford")


def approx(env, x, want, atol=1e-12):
    v = env.values(x)
    assert np.allclose(v, want, atol=atol), (v[:8], want)


# register_interface_tests.jl:1-82 -------------------------------------------------------------------
def register_interface(env):
    for f in ALL_FORMALISMS:
        rep = env.repr(f)
        for bg in (None, S.T2Dephasing(1.0)):
            if f == "mc" and bg is not None:
                pass
            reg = Register(3, rep, bg)
            initialize(reg[1])
            initialize(reg[2], S.X1 if f != "clifford" else ops.StabilizerState("X"))
            initialize(reg[3])
            apply((reg[1], reg[2]), CNOT)
            assert stateof(reg[1]) is stateof(reg[2]) and stateof(reg[1]).state.nsubsystems == 2
            assert stateof(reg[3]).state.nsubsystems == 1
            apply((reg[3], reg[1]), CNOT, time=0.5)
            assert stateof(reg[3]).state.nsubsystems == 3
            approx(env, observable(reg[1], Z), 1.0)


# apply_tests.jl:6-15 ------------------------------------------------------------------------------
def apply_tensor_gate(env):
    for f in ("qo", "clifford"):
        reg = Register(2, env.repr(f))
        initialize(reg[1], ops.StabilizerState("Z"))
        initialize(reg[2], ops.StabilizerState("X"))
        apply(reg[1:2], tensor(X, Z))
        approx(env, observable(reg[1], Z), -1)
        approx(env, observable(reg[2], X), -1)


# representations_dispatch_tests.jl:25-39 ------------------------------------------------------------
def cnot_bell_contracts(env):
    for f in ("qo", "clifford"):
        regs = [Register(1, env.repr(f)), Register(1, env.repr(f))]
        initialize(regs[0][1], ops.StabilizerState("X"))
        initialize(regs[1][1], ops.StabilizerState("Z"))
        apply((regs[0][1], regs[1][1]), CNOT)
        refs = (regs[0][1], regs[1][1])
        approx(env, observable(refs, tensor(X, X)), 1)
        approx(env, observable(refs, tensor(Z, Z)), 1)
        approx(env, observable(refs[0], X), 0)
        approx(env, observable(refs[1], Z), 0)


# project_traceout_tests.jl:10-53 ----------------------------------------------------------------------
def project_traceout_kats(env):
    for f in ALL_FORMALISMS:
        rep = env.repr(f)
        a = Register(2, rep)
        initialize(a[1:2], bell)
        m1 = project_traceout(a[1], Y)
        assert set(np.unique(env.outcomes(m1))) <= {1, 2}
        assert not isassigned(a, 1) and isassigned(a, 2)
        m2 = project_traceout(a[2], Y)
        assert (env.outcomes(m1) != env.outcomes(m2)).all()
        assert not isassigned(a, 2)

        a = Register(4, rep)
        with pytest.raises(ValueError, match="correct number of subsystems"):
            initialize(a[1:2], ops.tensor(bell, bell))
        initialize(a[1:4], ops.tensor(bell, bell))
        m = [env.outcomes(project_traceout(a[i], Y)) for i in (1, 2, 3, 4)]
        assert (m[0] != m[1]).all() and (m[2] != m[3]).all()

        a = Register(2, rep)
        initialize(a[1], ops.StabilizerState("X"))
        assert (env.outcomes(project_traceout(a[1], ops.σˣ)) == 1).all()
    empty = Register(1, env.repr("qo"))
    t0 = empty.accesstimes[0]
    with pytest.raises(ValueError, match="Cannot project and trace out an unassigned register slot."):
        project_traceout(empty[1], Z, time=1)
    assert empty.accesstimes[0] == t0 and not isassigned(empty, 1)


def unnormalized_state_error(env):
    r = Register(1, env.repr("qo"))
    initialize(r[1], np.array([[0.3, 0], [0, 0.3]], dtype=complex))  # the test initializes with the operator Z
    with pytest.raises(Exception, match="State not normalized"):
        j = project_traceout(r[1], (S.L0, S.L1))
        env.sync()


# observable_tests.jl:44-108 ----------------------------------------------------------------------------
def observable_kats(env):
    for f in ("qo", "clifford"):
        rep = env.repr(f)
        a = Register(2, rep)
        initialize(a[1:2], bell)
        approx(env, observable(a[1:2], ops.SProjector(bell)), 1.0)
        approx(env, observable(a[1:2], tensor(X, X)), 1.0)
        apply(a[1], Y)
        approx(env, observable(a[1:2], ops.SProjector(bell)), 0.0, atol=1e-5)
        approx(env, observable(a[1:2], tensor(X, X)), -1.0)

        reg = Register(1, rep)
        initialize(reg[1], ops.StabilizerState("X"))
        approx(env, observable(reg[1], ops.SProjector(ops.StabilizerState("Z"))), 0.5)
        reg = Register(3, rep)
        initialize(reg[1:3], ops.StabilizerState("XXX ZZI IZZ"))
        approx(env, observable((reg[1], reg[3]), ops.SProjector(bell)), 0.5)

        A, B = ops.StabilizerState("X"), ops.StabilizerState("Z")
        AB, BA = ops.StabilizerState("XI IZ"), ops.StabilizerState("ZI IX")
        r1, r2, r12, r21 = (Register(2, rep) for _ in range(4))
        initialize(r1[1], A); initialize(r2[2], B); initialize(r12[1:2], AB); initialize((r21[2], r21[1]), BA)
        approx(env, observable(r12[1:2], ops.SProjector(AB)), 1.0)
        approx(env, observable(r12[1], ops.SProjector(A)), 1.0)
        approx(env, observable(r21[1:2], ops.SProjector(AB)), 1.0)
        approx(env, observable((r1[1], r2[2]), ops.SProjector(AB)), 1.0)
        approx(env, observable(r2[2], ops.SProjector(B)), 1.0)
        assert observable(r1[2], Z, something=0.25) == 0.25  # empty slot -> `something`
    reg = Register(2, env.repr("clifford"))
    initialize(reg[1:2], ops.MixedStabilizer("ZZ"))
    approx(env, observable(reg[1:2], ops.SProjector(bell)), 0.5)


# observable_tests.jl:110-194 : graph state through measurement + conditional corrections ---------------
def permuted_graph_state(env):
    pm = 3
    ref_manual = ops.StabilizerState("XIZ IXZ ZZX")
    for f in ("qo", "clifford"):
        for cz_order in ([1, 2], [2, 1]):
            rep = env.repr(f)
            regA, regB = Register(3, rep), Register(3, rep)
            for i in (1, 2, 3):
                initialize((regA[i], regB[i]), bell)
            for i in cz_order:
                apply((regA[pm], regA[i]), ops.ZCZ)
            for i in (1, 2, 3):
                zmeas = project_traceout(regA[i], X)
                apply(regB[i], Z, mask=(zmeas == 2))
            approx(env, observable(regB[1:3], ops.SProjector(ref_manual)), 1.0)


# traceout_tests.jl:29-70 ----------------------------------------------------------------------------------
def clifford_traceout(env):
    reg = Register(3, env.repr("clifford"))
    stateref = initialize(reg[1:3], ops.StabilizerState("XII IYI IIZ"))
    approx(env, observable(reg[1], X), 1); approx(env, observable(reg[2], Y), 1); approx(env, observable(reg[3], Z), 1)
    traceout(reg[2])
    assert [isassigned(reg, i) for i in (1, 2, 3)] == [True, False, True]
    assert reg.stateindices == [1, 0, 2] and stateref.registerindices == [1, 3]
    assert stateref.state.nsubsystems == 2
    approx(env, observable(reg[1], X), 1); approx(env, observable(reg[3], Z), 1)
    traceout(reg[1])
    assert reg.stateindices == [0, 0, 1] and stateref.state.nsubsystems == 1
    approx(env, observable(reg[3], Z), 1)

    reg = Register(3, env.repr("clifford"))
    stateref = initialize(reg[1:3], ops.StabilizerState("XXX ZZI IZZ"))
    traceout(reg[2])
    assert stateref.state.nsubsystems == 2
    approx(env, observable((reg[1], reg[3]), tensor(Z, Z)), 1)
    approx(env, observable(reg[1], X), 0); approx(env, observable(reg[3], X), 0)


def dense_traceout(env):
    reg = Register(3, env.repr("qo"))
    initialize(reg[1:3], ops.StabilizerState("XXX ZZI IZZ"))
    traceout(reg[2])
    approx(env, observable((reg[1], reg[3]), tensor(Z, Z)), 1)
    approx(env, observable(reg[1], X), 0)
    # whole-group delete skips the backend (traceout_tests.jl:159-180)
    regs = [Register(1, env.repr("qo")) for _ in range(2)]
    initialize((regs[0][1], regs[1][1]), bell)
    traceout(regs[0][1], regs[1][1])
    assert not isassigned(regs[0], 1) and not isassigned(regs[1], 1)


# circuitzoo_ent_swap_tests.jl:11-59 ---------------------------------------------------------------------
def entanglement_swap_chains(env, sizes=(3, 4, 6)):
    for f in ("qo", "clifford"):
        rep = env.repr(f)
        net = [Register(1, rep), Register(2, rep), Register(1, rep)]
        initialize((net[0][1], net[1][1]), bell)
        initialize((net[2][1], net[1][2]), bell)
        EntanglementSwap()(net[1][1], net[0][1], net[1][2], net[2][1])
        assert not isassigned(net[1], 1) and not isassigned(net[1], 2)
        approx(env, observable((net[0][1], net[2][1]), tensor(Z, Z)), 1)
        approx(env, observable((net[0][1], net[2][1]), tensor(X, X)), 1)
        for n in sizes:
            net = [Register(2, rep) for _ in range(n)]
            for i in range(n - 1):
                initialize((net[i][1], net[i + 1][2]), bell)
            for i in range(1, n - 1):
                EntanglementSwap()(net[i][2], net[0][1], net[i][1], net[i + 1][2])
            assert all(not isassigned(net[i], 1) and not isassigned(net[i], 2) for i in range(1, n - 1))
            approx(env, observable((net[0][1], net[n - 1][2]), tensor(Z, Z)), 1)
            approx(env, observable((net[0][1], net[n - 1][2]), tensor(X, X)), 1)
            net = [Register(2, rep) for _ in range(n)]
            for i in range(n - 1):
                initialize((net[i][1], net[i + 1][2]), bell)
            for i in range(1, n - 1):
                mx, mz = LocalEntanglementSwap()(net[i][2], net[i][1])
                apply(net[0][1], Z, mask=(mx == 2))
                apply(net[i + 1][2], X, mask=(mz == 2))
            approx(env, observable((net[0][1], net[n - 1][2]), tensor(Z, Z)), 1)
            approx(env, observable((net[0][1], net[n - 1][2]), tensor(X, X)), 1)


# circuitzoo_purification_tests.jl:43-140 ------------------------------------------------------------------
PAULI = {"X": X, "Y": Y, "Z": Z}


def purification_2to1(env):
    for f in ("qo", "clifford"):
        rep = env.repr(f)
        for leaveout in "XYZ":
            r = Register(4, rep)
            initialize(r[1:4], ops.tensor(bell, bell))
            assert env.truth(Purify2to1(leaveout)(*r[1:4])).all()
            approx(env, observable(r[1:2], ops.SProjector(bell)), 1.0)
            for error in "XYZ":
                for target in (1, 2, 3, 4):
                    r = Register(4, rep)
                    initialize(r[1:4], ops.tensor(bell, bell))
                    apply(r[target], PAULI[error])
                    ok = env.truth(Purify2to1(leaveout)(*r[1:4]))
                    if error == leaveout:
                        assert ok.all()
                        approx(env, observable(r[1:2], ops.SProjector(bell)), 0.0, atol=1e-5)
                    else:
                        assert not ok.any()
            # node variant
            r = Register(4, rep)
            initialize(r[1:4], ops.tensor(bell, bell))
            ma = Purify2to1Node(leaveout)(r[1], r[3]); mb = Purify2to1Node(leaveout)(r[2], r[4])
            assert (env.outcomes(ma) == env.outcomes(mb)).all()
            approx(env, observable(r[1:2], ops.SProjector(bell)), 1.0)


def purification_3to1(env, combos=(("X", "Z"), ("Y", "X"), ("Z", "Y"))):
    for f in ("qo", "clifford"):
        rep = env.repr(f)
        for l1, l2 in combos:
            r = Register(6, rep)
            initialize(r[1:6], ops.tensor(bell, bell, bell))
            assert env.truth(Purify3to1(l1, l2)(r[1], r[2], r[3], r[5], r[4], r[6])).all()
            approx(env, observable(r[1:2], ops.SProjector(bell)), 1.0)
            for error in "XYZ":
                for target in (3, 4, 5, 6):
                    r = Register(6, rep)
                    for i in (1, 2, 3):
                        initialize(r[(2 * i - 1):(2 * i)], bell)
                    apply(r[target], PAULI[error])
                    assert not env.truth(Purify3to1(l1, l2)(r[1], r[2], r[3], r[5], r[4], r[6])).any()
                for target in (1, 2):
                    r = Register(6, rep)
                    for i in (1, 2, 3):
                        initialize(r[(2 * i - 1):(2 * i)], bell)
                    apply(r[target], PAULI[error])
                    ok = env.truth(Purify3to1(l1, l2)(r[1], r[2], r[3], r[5], r[4], r[6]))
                    if error == l1:
                        assert ok.all()
                        approx(env, observable(r[1:2], ops.SProjector(bell)), 0.0, atol=1e-5)
                    else:
                        assert not ok.any()


# test_stateszoo_depolarized.jl:9-40 ---------------------------------------------------------------------
def depolarized_bell_pair(env):
    rep = env.repr("qo")
    for p, zz in ((0.0, 0.0), (1.0, 1.0), (0.6, 0.6)):
        r = Register(2, rep)
        initialize(r[1:2], ops.DepolarizedBellPair(p))
        approx(env, observable(r[1:2], tensor(Z, Z)), zz)
        approx(env, observable(r[1:2], ops.SProjector(bell)), (3 * p + 1) / 4)


# doctests: states_registers.jl:58-61, baseops/uptotime.jl:5-17 -------------------------------------------
def doctest_values(env):
    r = Register(2, env.repr("qo"))
    initialize(r[1], S.X1)
    approx(env, observable(r[1], X), 0.9999999999999998, atol=3e-16)
    reg = Register(2, env.repr("qo"), S.T1Decay(1.0))
    initialize(reg[1], S.X1)
    approx(env, observable(reg[1], Z), 0.0, atol=1e-16)
    uptotime(reg[1], 10)
    approx(env, observable(reg[1], Z), 0.9999546000702374, atol=3e-16)
    with pytest.raises(RuntimeError, match="simulation time is higher"):
        apply(reg[1], X, time=5)  # src/baseops/apply.jl:12-14


# noninstant_and_backgrounds_qubit_tests.jl:8-20 (time bookkeeping) + src/noninstant.jl:9-13 (apply! then uptotime!) ---
def noninstant_gate(env):
    from qsb200.ops import H
    from qsb200.registers import NonInstantGate
    for f in ("qo", "clifford"):
        reg = Register(2, env.repr(f))
        initialize(reg[1]); initialize(reg[2])
        uptotime(reg[1], 0.2)
        uptotime(reg[1], 0.2)
        with pytest.raises(RuntimeError, match="simulation time is higher"):
            uptotime(reg[1], 0.1)
        apply(reg[1], H, time=0.3)
        _, t1 = apply(reg[1], NonInstantGate(H, 0.1), time=0.4)
        assert abs(t1 - 0.5) < 1e-15 and reg.accesstimes[0] == t1
        _, t2 = apply((reg[1], reg[2]), NonInstantGate(CNOT, 0.1))
        assert abs(t2 - 0.6) < 1e-15 and reg.accesstimes == [t2, t2]
        with pytest.raises(RuntimeError, match="simulation time is higher"):
            uptotime(reg[1], 0.55)
        with pytest.raises(RuntimeError, match="simulation time is higher"):
            apply(reg[1], H, time=0.55)
        approx(env, observable(reg[1:2], tensor(Z, Z)), 1.0)      # H.H = 1, CNOT|00> = |00>, no background
    # the gate followed by waiting under a background == the two calls made by hand (density matrices: deterministic)
    for bg in (S.T1Decay(1.0), S.T2Dephasing(0.7), S.T1T2Noise(1.0, 0.8), S.Depolarization(2.0)):
        a, b = Register(2, env.repr("qo"), bg), Register(2, env.repr("qo"), bg)
        for r in (a, b):
            initialize(r[1], S.X1); initialize(r[2])
        apply(a[1], NonInstantGate(H, 0.25), time=0.1)
        apply((a[1], a[2]), NonInstantGate(CNOT, 0.5))
        _, t = apply(b[1], H, time=0.1)
        uptotime(b[1], t + 0.25)
        _, t = apply((b[1], b[2]), CNOT)
        uptotime((b[1], b[2]), t + 0.5)
        assert a.accesstimes == b.accesstimes and abs(a.accesstimes[0] - 0.85) < 1e-15
        for obs in (tensor(Z, Z), tensor(X, X), tensor(Z, X), tensor(Y, Y)):
            va, vb = env.values(observable(a[1:2], obs)), env.values(observable(b[1:2], obs))
            assert np.allclose(va, vb, atol=1e-15), (bg, va, vb)
        # and the closed form of the first step: H X1 = Z1 stays |0> under T1/T2 (no decay from the ground state)
    reg = Register(1, env.repr("qo"), S.T2Dephasing(1.0))
    initialize(reg[1])
    apply(reg[1], NonInstantGate(H, 0.3))                          # |+> dephases for 0.3: <X> = exp(-0.3)
    approx(env, observable(reg[1], X), np.exp(-0.3), atol=1e-15)


def _second_moment(env, x):
    """E[x^2] over the trajectories (env.values gives mean -/+ sigma for an ensemble, the value itself for one trajectory)."""
    v = env.values(x)
    return float(v[0]) ** 2 if len(v) == 1 else float(v[1]) ** 2 + ((float(v[2]) - float(v[0])) / 2) ** 2


# noninstant_and_backgrounds_qubit_tests.jl:24-67: Kraus evolution == Lindblad evolution, composed == not composed -----
def kraus_lindblad(env):
    from qsb200.registers import ConstantHamiltonianEvolution
    tol = getattr(env, "ode_atol", 1e-12)   # the oracle integrates with the reference's adaptive solver (reltol 1e-6)
    singles = [(i, P) for i in (1, 2) for P in (X, Y, Z)]
    pairs = [tensor(P, Q) for P in (X, Y, Z) for Q in (X, Y, Z)]

    def same_single(ra, ia, rb, ib, atol):
        for P in (X, Y, Z):
            va, vb = env.values(observable(ra[ia], P)), env.values(observable(rb[ib], P))
            assert np.allclose(va, vb, atol=atol), (P, va, vb)

    def product_matches(reg, regb, atol):   # reg[1] (x) reg[2] vs the composed register
        for i, P in singles:
            assert np.allclose(env.values(observable(reg[i], P)), env.values(observable(regb[i], P)), atol=atol)
        for k, PQ in enumerate(pairs):
            P, Q = (X, Y, Z)[k // 3], (X, Y, Z)[k % 3]
            prod = env.values(observable(reg[1], P))[len(env.values(observable(reg[1], P))) // 2] * \
                env.values(observable(reg[2], Q))[len(env.values(observable(reg[2], Q))) // 2]
            assert np.allclose(env.values(observable(regb[1:2], PQ)), prod, atol=atol), (k, prod)

    for bg in (S.T1Decay(1.0), S.T2Dephasing(1.0), S.T1T2Noise(1.0, 1.0), S.T1T2Noise(1.0, 3.0)):
        for init in (S.Z1, S.Z2, S.X1, S.X2):
            reg = Register(2, env.repr("qo"), bg)
            initialize(reg[1], init); initialize(reg[2], init)
            uptotime(reg[1], 0.2)
            uptotime(reg[2], 0.1); uptotime(reg[2], 0.2)
            same_single(reg, 1, reg, 2, 1e-14)                       # Kraus weights compose
            regb = Register(2, env.repr("qo"), bg)
            initialize(regb[1], init); initialize(regb[2], init)
            from qsb200.registers import subsystemcompose as compose
            compose((regb[1], regb[2]))
            uptotime(regb[1], 0.2)
            uptotime(regb[2], 0.1); uptotime(regb[2], 0.2)
            product_matches(reg, regb, 1e-14)
            _, t_end = apply(reg[1], ConstantHamiltonianEvolution(ops.I, 0.1), time=0.3)   # Lindblad evolution
            assert abs(t_end - 0.4) < 1e-15 and reg.accesstimes[0] == t_end                # overwritetime!, noninstant.jl:33
            uptotime(reg[2], 0.4)                                                          # Kraus evolution
            same_single(reg, 1, reg, 2, tol)
            apply(regb[1], ConstantHamiltonianEvolution(ops.I, 0.1), time=0.3)
            uptotime(regb[2], 0.4)
            product_matches(reg, regb, tol)
            # uptotime!'s own Lindblad route (krausops === nothing, uptotime.jl:51-54) against the Kraus closed form
            regc = Register(1, env.repr("qo"), S.LindbladOnly(bg))
            initialize(regc[1], init)
            uptotime(regc[1], 0.2); uptotime(regc[1], 0.4)
            same_single(regc, 1, reg, 2, tol)
    with pytest.raises(RuntimeError, match="lindblad operators implemented for Depolarization"):
        r = Register(1, env.repr("qo"), S.Depolarization(1.0))
        initialize(r[1], S.X1)
        apply(r[1], ConstantHamiltonianEvolution(ops.I, 0.1))


# quantummc_repr_tests.jl:57-83: ConstantHamiltonianEvolution on trajectory kets (schroedinger / mcwf) --------------------
def constant_hamiltonian_mc(env):
    from qsb200.registers import ConstantHamiltonianEvolution
    evolution = ConstantHamiltonianEvolution(np.pi / 2 * S.Z, 1.0)
    reg = Register(1, env.repr("mc"))
    initialize(reg[1], S.X1)
    apply(reg[1], evolution)
    approx(env, observable(reg[1], X), -1.0, atol=1e-6)                 # exp(-i pi/2 Z)|+> = -i|->
    noisy = Register(1, env.repr("mc"), S.T2Dephasing(1.0))
    initialize(noisy[1], S.X1)
    apply(noisy[1], evolution)
    assert abs(_second_moment(env, observable(noisy[1], X)) - 1.0) < 1e-6   # every trajectory stays pure: |<X>| = 1
    comp = Register(2, env.repr("mc"), [S.T2Dephasing(1.0), None])
    initialize(comp[1:2], ops.StabilizerState("XX ZZ"))
    _, t = apply(comp[1], evolution)
    assert t == 1.0 and stateof(comp[1]).state.nsubsystems == 2
    assert abs(_second_moment(env, observable(comp[1:2], tensor(X, X))) - 1.0) < 1e-6   # Z-rotation by pi and Z jumps: XX -> +/- XX
    # the same evolution on a density matrix: <X> = -exp(-1) (dephasing for one time unit at T2 = 1)
    d = Register(1, env.repr("qo"), S.T2Dephasing(1.0))
    initialize(d[1], S.X1)
    apply(d[1], evolution)
    approx(env, observable(d[1], X), -np.exp(-1.0), atol=max(getattr(env, "ode_atol", 1e-12), 1e-12))


ALL = [register_interface, apply_tensor_gate, cnot_bell_contracts, project_traceout_kats, unnormalized_state_error,
       observable_kats, permuted_graph_state, clifford_traceout, dense_traceout, entanglement_swap_chains,
       purification_2to1, purification_3to1, depolarized_bell_pair, doctest_values, noninstant_gate, kraus_lindblad,
       constant_hamiltonian_mc]
