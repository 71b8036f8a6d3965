"""GPU: qsb_evolve (ConstantHamiltonianEvolution / Lindblad fallback) through the C-ABI against the exact
propagator (1e-12), the oracle's restatement of the reference's adaptive solver (2e-6: the solver's tolerance) and, for
trajectory kets, the master equation as the ensemble mean."""
import numpy as np
import pytest

from tests.test_oracle_evolve import exact_master, rand_dm, rand_herm

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx():
    from qsb200.backend import BatchContext
    c = BatchContext(70, seed=0xE501)
    yield c
    c.close()


@pytest.mark.parametrize("k", [1, 2])
def test_dm_evolution_matches_exact_propagator_and_oracle(ctx, k):
    from oracle import oracle as orc
    from qsb200 import symbols as S
    from qsb200._lib import DM
    from qsb200.backend import CudaBatchState
    rng = np.random.default_rng(40 + k)
    n = 3
    for trial in range(4):
        idx = [int(q) for q in rng.permutation(n)[:k]]
        H = rand_herm(rng, 2 ** k)
        bgs = [S.T1T2Noise(1.3, 0.9), S.T1Decay(0.7)][:k]
        lind = [(p, L) for p, bg in enumerate(bgs) for L in S.lindbladop(bg)]
        rho = rand_dm(rng, n)
        t = float(rng.uniform(0.05, 1.5))
        st = CudaBatchState.from_array(ctx, rho, DM)
        st._evolve([q + 1 for q in idx], H, lind, t)
        want = exact_master(rho, n, idx, H, lind, t)
        for traj in (0, 33, 69):
            got = st.trajectory(traj).reshape(2 ** n, 2 ** n, order="F")
            assert np.abs(got - want).max() < 1e-12
        twin = orc.Dense.from_array(rho).evolve_master(idx, H, lind, t).array()
        assert np.abs(st.trajectory(5).reshape(2 ** n, 2 ** n, order="F") - twin).max() < 2e-6
        st.free()


def test_ket_promotes_and_mcket_without_collapse_stays_a_ket(ctx):
    from scipy.linalg import expm
    from qsb200 import symbols as S
    from qsb200._lib import DM, KET, MCKET
    from qsb200.backend import CudaBatchState
    rng = np.random.default_rng(2)
    psi = rng.normal(size=4) + 1j * rng.normal(size=4)
    psi /= np.linalg.norm(psi)
    H = rand_herm(rng, 4)
    U = expm(-1j * H * 0.6)
    k = CudaBatchState.from_array(ctx, psi, KET)
    k._evolve([2, 1], H, [], 0.6)                         # factor 0 of H acts on subsystem 2
    assert k.kind == DM
    P = np.eye(4)[[0, 2, 1, 3]]                           # swaps the two tensor factors
    Uf = P @ U @ P.T
    assert np.abs(k.trajectory(9).reshape(4, 4, order="F") - Uf @ np.outer(psi, psi.conj()) @ Uf.conj().T).max() < 1e-12
    m = CudaBatchState.from_array(ctx, psi, MCKET)
    d0 = ctx._draw
    m._evolve([2, 1], H, [], 0.6)
    assert m.kind == MCKET and ctx._draw == d0            # schroedinger: no draw consumed
    assert np.abs(m.trajectory(9) - Uf @ psi).max() < 1e-12
    k.free(); m.free()


def test_mcket_ensemble_mean_is_the_master_equation():
    from qsb200 import ops
    from qsb200 import symbols as S
    from qsb200.backend import BatchContext, CudaBatchRepr
    from qsb200.registers import ConstantHamiltonianEvolution, Register, apply, initialize, observable
    N = 20000
    c = BatchContext(N, seed=0xE502)
    H = 0.7 * S.X + 0.2 * S.Z
    bg = S.T1T2Noise(1.0, 0.8)
    reg = Register(2, CudaBatchRepr(c, "mc"), [bg, None])
    initialize(reg[1], S.X1); initialize(reg[2], S.Z1)
    apply((reg[1], reg[2]), ops.CNOT)                     # entangle first: the jump acts on half of a pair
    apply(reg[1], ConstantHamiltonianEvolution(H, 0.9))
    rho0 = S.dm(S.CNOT @ S.tensor(S.X1, S.Z1))
    want = exact_master(rho0, 2, [0], H, [(0, L) for L in S.lindbladop(bg)], 0.9)
    for P, M in ((ops.tensor(ops.X, ops.X), S.tensor(S.X, S.X)), (ops.tensor(ops.Z, ops.Z), S.tensor(S.Z, S.Z)),
                 (ops.tensor(ops.Y, ops.X), S.tensor(S.Y, S.X)), (ops.tensor(ops.Z, ops.I), S.tensor(S.Z, S.Id))):
        v = observable(reg[1:2], P)
        assert abs(v.mean - np.trace(want @ M).real) < 4 / np.sqrt(N), (P, v.mean, np.trace(want @ M).real)
    # trajectories stay normalised pure states: purity 1 <=> sum over the Pauli basis of <P>^2 = 4 per trajectory
    tot = 0.0
    for a in "IXYZ":
        for b in "IXYZ":
            v = observable(reg[1:2], ops.tensor(getattr(ops, a), getattr(ops, b)))
            tot += v.mean ** 2 + v.var
    assert abs(tot - 4.0) < 1e-9
    c.close()


def test_refusals(ctx):
    from qsb200 import ops
    from qsb200 import symbols as S
    from qsb200._lib import MethodError
    from qsb200.backend import CudaBatchRepr
    from qsb200.registers import ConstantHamiltonianEvolution, Register, apply, initialize
    reg = Register(1, CudaBatchRepr(ctx, "clifford"))
    initialize(reg[1])
    with pytest.raises(MethodError):                      # the reference has no apply_noninstant! method for tableaux either
        apply(reg[1], ConstantHamiltonianEvolution(ops.Z, 0.1))
    with ctx.record():
        r = Register(1, CudaBatchRepr(ctx, "qo"), S.T1Decay(1.0))
        initialize(r[1], S.X1)
        with pytest.raises(MethodError):
            apply(r[1], ConstantHamiltonianEvolution(ops.Z, 0.1))


def test_two_qubit_mcket_evolution_with_collapse_operators():
    """mcwf of a 2-qubit ConstantHamiltonianEvolution: sampled branch of the canonical Kraus form of the 16x16 propagator
    (k_rdm2_ket -> k_mc2_finalize -> k_apply_kraus2_sel); the ensemble mean is the master-equation solution, kets stay normalised."""
    from qsb200 import ops
    from qsb200 import symbols as S
    from qsb200.backend import BatchContext, CudaBatchRepr
    from qsb200.registers import ConstantHamiltonianEvolution, Register, apply, initialize, observable
    N = 20000
    c = BatchContext(N, seed=0xE503)
    H = 0.6 * S.tensor(S.Z, S.Z) + 0.4 * S.tensor(S.X, S.Id) + 0.25 * S.tensor(S.Id, S.Y)
    bgs = [S.T1T2Noise(1.0, 0.8), S.T2Dephasing(1.5), None]
    reg = Register(3, CudaBatchRepr(c, "mc"), bgs)
    initialize(reg[1], S.X1); initialize(reg[2], S.Y1); initialize(reg[3], S.Z1)
    apply((reg[2], reg[3]), ops.CNOT)                      # a spectator entangled with the second evolving qubit
    d0 = c._draw
    _, t_end = apply((reg[1], reg[2]), ConstantHamiltonianEvolution(H, 0.7))
    assert t_end == 0.7 and c._draw == d0 + 1              # one draw per call
    psi0 = S.tensor(S.X1, S.CNOT @ S.tensor(S.Y1, S.Z1))
    lind = [(0, L) for L in S.lindbladop(bgs[0])] + [(1, L) for L in S.lindbladop(bgs[1])]
    want = exact_master(S.dm(psi0), 3, [0, 1], H, lind, 0.7)
    for names in ("XII", "ZZI", "IYX", "ZIZ", "YXZ", "IIZ"):
        P = ops.tensor(*[getattr(ops, ch) for ch in names])
        M = S.tensor(*[S.PAULI[ch] for ch in names])
        v = observable(reg[1:3], P)
        assert abs(v.mean - np.trace(want @ M).real) < 4 / np.sqrt(N), (names, v.mean, np.trace(want @ M).real)
    tot = 0.0                                               # purity 1 on every trajectory: sum over the Pauli basis of <P>^2 = 2^n
    for a in "IXYZ":
        for b in "IXYZ":
            for d in "IXYZ":
                v = observable(reg[1:3], ops.tensor(getattr(ops, a), getattr(ops, b), getattr(ops, d)))
                tot += v.mean ** 2 + v.var
    assert abs(tot - 8.0) < 1e-9
    c.close()
