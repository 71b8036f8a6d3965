"""GPU parity proper: the CUDA path (through the C-ABI) against the CPU oracle on the same seeded inputs.

A random trace of backend calls is issued to a batched CUDA state (N trajectories) and, op for op, to
oracle twins of a few trajectories fed the same (seed, trajectory, draw_id) draws.
  * dense kinds: states / probabilities / observables within 1e-12 relative (north_star tolerance),
    outcomes identical;
  * Clifford: bit-exact outcomes, was-random flags, expectation values, rank and canonical generators.
"""
import numpy as np
import pytest

from oracle import oracle as orc
from qsb200 import ops
from qsb200 import symbols as S
from tests.oracle_backend import OracleContext, OracleRepr

pytestmark = pytest.mark.gpu
SEED = 0xC0FFEE
RTOL = 1e-12


def rand_unitary(rng, k):
    a = rng.normal(size=(2**k, 2**k)) + 1j * rng.normal(size=(2**k, 2**k))
    q, r = np.linalg.qr(a)
    return q * (np.diag(r) / np.abs(np.diag(r)))


def rand_ket(rng, n):
    v = rng.normal(size=2**n) + 1j * rng.normal(size=2**n)
    return v / np.linalg.norm(v)


def rand_dm(rng, n):
    a = rng.normal(size=(2**n, 2**n)) + 1j * rng.normal(size=(2**n, 2**n))
    rho = a @ a.conj().T
    return rho / np.trace(rho)


class Twins:
    """One CUDA batch + oracle twins of selected trajectories, driven with identical calls."""

    def __init__(self, ntraj, formalism, follow):
        from qsb200.backend import BatchContext, CudaBatchRepr
        self.ctx = BatchContext(ntraj, seed=SEED)
        self.rep = CudaBatchRepr(self.ctx, formalism)
        self.follow = list(follow)
        self.octx = {t: OracleContext(SEED, t) for t in self.follow}
        self.orep = {t: OracleRepr(self.octx[t], formalism) for t in self.follow}
        self.formalism = formalism

    def express(self, spec):
        return self.rep.express(spec), {t: self.orep[t].express(spec) for t in self.follow}

    def newstate(self):
        return self.rep.newstate(), {t: self.orep[t].newstate() for t in self.follow}

    def close(self):
        self.ctx.close()


def compare_dense(g, o, follow):
    for t in follow:
        a, b = g.trajectory(t), o[t].s.array()
        assert a.shape == b.shape, (a.shape, b.shape)
        scale = max(np.abs(b).max(), 1e-300)
        assert np.abs(a - b).max() <= RTOL * scale, (t, np.abs(a - b).max())


NAMED1 = [ops.X, ops.Y, ops.Z, ops.H]
NAMED2 = [ops.CNOT, ops.CPHASE, ops.XCZ, ops.ZCY, ops.XCY]
BGS = [S.T1Decay(1.3), S.T2Dephasing(0.7), S.Depolarization(2.1), S.T1T2Noise(1.5, 0.9), S.T1T2Noise(1.0, 5.0)]


@pytest.mark.parametrize("formalism,seed", [("qo", 0), ("qo", 1), ("qo", 2), ("mc", 3), ("mc", 4)])
def test_dense_random_trace(formalism, seed):
    rng = np.random.default_rng(seed)
    ntraj = 70
    follow = (0, 31, 32, 69)
    tw = Twins(ntraj, formalism, follow)
    n0 = 3
    if seed % 2 == 0:
        g, o = tw.express(rand_ket(rng, n0))
    else:
        g, o = tw.express(rand_dm(rng, n0) if formalism == "qo" else rand_ket(rng, n0))
    for step in range(45):
        n = g.nsubsystems
        r = rng.random()
        if r < 0.2:
            q = [int(rng.integers(n)) + 1]
            op = NAMED1[rng.integers(4)] if rng.random() < 0.5 else rand_unitary(rng, 1)
            g.apply(q, op)
            for t in follow: o[t].apply(q, op)
        elif r < 0.4 and n >= 2:
            q = [int(x) + 1 for x in rng.choice(n, 2, replace=False)]
            op = NAMED2[rng.integers(5)] if rng.random() < 0.5 else rand_unitary(rng, 2)
            g.apply(q, op)
            for t in follow: o[t].apply(q, op)
        elif r < 0.58:
            q = int(rng.integers(n)) + 1
            bg = BGS[rng.integers(len(BGS))]; dt = float(rng.random())
            g.uptotime(q, bg, dt)
            for t in follow: o[t].uptotime(q, bg, dt)
        elif r < 0.72 and n >= 2:
            q = int(rng.integers(n)) + 1
            basis = [ops.X, ops.Y, ops.Z, None][rng.integers(4)]
            if basis is None:
                u = rand_unitary(rng, 1)
                basis = (u[:, 0].copy(), u[:, 1].copy())
            out, g2 = g.project_traceout(q, basis)
            got = out.numpy()
            for t in follow:
                j, _ = o[t].project_traceout(q, basis)
                assert got[t] == j, (step, t)
            # outcome-conditioned correction on another qubit (the mask path)
            q2 = [int(rng.integers(g.nsubsystems)) + 1]
            g.apply(q2, ops.Z, mask=(out == 2))
            for t in follow:
                if got[t] == 2: o[t].apply(q2, ops.Z)
        elif r < 0.8 and n >= 2:
            q = int(rng.integers(n)) + 1
            g.traceout(q)
            for t in follow: o[t].traceout(q)
        elif n <= 4:
            add = [S.BELL, S.X1, rand_ket(rng, 1), S.depolarized_bell_dm(0.8)][rng.integers(4)]
            if formalism == "mc" and np.asarray(add).ndim == 2:
                add = S.BELL
            ga, oa = tw.express(add)
            if rng.random() < 0.5:
                g = g.subsystemcompose(ga); o = {t: o[t].subsystemcompose(oa[t]) for t in follow}
            else:
                g = ga.subsystemcompose(g); o = {t: oa[t].subsystemcompose(o[t]) for t in follow}
        assert g.nsubsystems == o[follow[0]].nsubsystems
        compare_dense(g, o, follow)
        # observables: Pauli string, dense operator, projector (full and partial)
        n = g.nsubsystems
        k = int(rng.integers(1, min(n, 3) + 1))
        idx = [int(x) + 1 for x in rng.choice(n, k, replace=False)]
        ps = "".join("XYZ"[rng.integers(3)] for _ in range(k))
        obs_list = [ops.tensor(*[{"X": ops.X, "Y": ops.Y, "Z": ops.Z}[c] for c in ps]), rand_unitary(rng, k),
                    ops.SProjector(rand_ket(rng, k))]
        if n <= 4:
            full = [int(x) + 1 for x in rng.permutation(n)]
            obs_list.append((full, ops.SProjector(rand_ket(rng, n))))
        for ob in obs_list:
            ii, ob = (ob if isinstance(ob, tuple) else (idx, ob))
            val = g.observable(ii, ob, per_trajectory=True).values()
            for t in follow:
                want = complex(o[t].observable(ii, ob))
                assert abs(val[t] - want) <= RTOL * max(1.0, abs(want)), (step, t, val[t], want)
    tw.ctx.sync()
    tw.close()


def canonical_from_tableau(xs, zs, sg, ac):
    n = len(ac)
    act = [i for i in range(n) if ac[i]]
    st = orc.Stab.from_generators(n, xs[n:][act], zs[n:][act], sg[n:][act]) if act else orc.Stab.maximally_mixed(n)
    return st.generators(canonical=True), len(act)


def compare_tab(g, o, follow):
    for t in follow:
        (gx, gz, gs), rank = canonical_from_tableau(*g.tableau(t))
        ox, oz, os_ = o[t].s.generators(canonical=True)
        assert rank == o[t].s.rank, (t, rank, o[t].s.rank)
        assert np.array_equal(gx, ox) and np.array_equal(gz, oz) and np.array_equal(gs, os_), t


CL1 = [ops.X, ops.Y, ops.Z, ops.H]
CL2 = [ops.CNOT, ops.CPHASE, ops.XCZ, ops.ZCY, ops.XCY, ops.XCX, ops.YCZ]


@pytest.mark.parametrize("seed", range(6))
def test_clifford_random_trace_bit_exact(seed):
    rng = np.random.default_rng(100 + seed)
    ntraj = 100
    follow = (0, 31, 32, 63, 64, 99)
    tw = Twins(ntraj, "clifford", follow)
    g, o = tw.express(ops.StabilizerState("XX ZZ") if seed % 2 else ops.Mixture([(0.7, ops.StabilizerState("ZZ XX")), (0.3, ops.MixedState(2))]))
    bgs = [S.Depolarization(0.4), S.T2Dephasing(0.3)]
    for step in range(60):
        n = g.nsubsystems
        r = rng.random()
        if r < 0.2:
            q = [int(rng.integers(n)) + 1]; op = CL1[rng.integers(4)]
            g.apply(q, op)
            for t in follow: o[t].apply(q, op)
        elif r < 0.45 and n >= 2:
            q = [int(x) + 1 for x in rng.choice(n, 2, replace=False)]; op = CL2[rng.integers(len(CL2))]
            g.apply(q, op)
            for t in follow: o[t].apply(q, op)
        elif r < 0.65:
            q = int(rng.integers(n)) + 1; bg = bgs[rng.integers(2)]; dt = float(rng.random())
            g.uptotime(q, bg, dt)
            for t in follow: o[t].uptotime(q, bg, dt)
        elif r < 0.8 and n >= 2:
            q = int(rng.integers(n)) + 1; basis = [ops.X, ops.Y, ops.Z][rng.integers(3)]
            out, _ = g.project_traceout(q, basis)
            got, rnd = out.numpy(), out.was_random.numpy()
            for t in follow:
                j, _ = o[t].project_traceout(q, basis)
                assert got[t] == j and bool(rnd[t]) == o[t].last_was_random, (step, t)
            q2 = [int(rng.integers(g.nsubsystems)) + 1]
            g.apply(q2, ops.X, mask=(out == 2))
            for t in follow:
                if got[t] == 2: o[t].apply(q2, ops.X)
        elif r < 0.88 and n >= 2:
            q = int(rng.integers(n)) + 1
            g.traceout(q)
            for t in follow: o[t].traceout(q)
        elif n <= 6:
            spec = [ops.StabilizerState("XX ZZ"), ops.StabilizerState("Z"), ops.MixedStabilizer("ZZ"),
                    ops.Mixture([(0.6, ops.StabilizerState("ZZ XX")), (0.4, ops.MixedState(2))])][rng.integers(4)]
            ga, oa = tw.express(spec)
            g = g.subsystemcompose(ga); o = {t: o[t].subsystemcompose(oa[t]) for t in follow}
        n = g.nsubsystems
        k = int(rng.integers(1, n + 1))
        idx = [int(x) + 1 for x in rng.choice(n, k, replace=False)]
        ps = "".join("XYZ"[rng.integers(3)] for _ in range(k))
        ob = ops.tensor(*[{"X": ops.X, "Y": ops.Y, "Z": ops.Z}[c] for c in ps])
        val = g.observable(idx, ob, per_trajectory=True).values()
        for t in follow:
            assert val[t].real == o[t].observable(idx, ob) and val[t].imag == 0, (step, t)
        if n >= 2:
            idx2 = [int(x) + 1 for x in rng.choice(n, 2, replace=False)]
            pr = ops.SProjector(ops.StabilizerState(["XX ZZ", "-XX ZZ", "XI IZ", "YY ZZ"][rng.integers(4)]))
            val = g.observable(idx2, pr, per_trajectory=True).values()
            for t in follow:
                assert val[t].real == o[t].observable(idx2, pr), (step, t)
        compare_tab(g, o, follow)
    tw.close()


def test_clifford_ensemble_vs_density_matrix():
    """test/general/noninstant_and_backgrounds_clifford_tests.jl:10-49 as ONE batched run per case:
    mean over N Clifford trajectories vs one dm run, |delta| < 4/sqrt(N)."""
    from qsb200.backend import BatchContext, CudaBatchRepr
    from qsb200.registers import Register, apply, initialize, observable, uptotime
    N = 20000
    ctx = BatchContext(N, seed=7)
    one = BatchContext(1, seed=7)

    def run(reg, init, gate, obs):
        if isinstance(init, tuple):
            initialize(reg[1], init[0]); initialize(reg[2], init[1])
        else:
            initialize((reg[1], reg[2]), init)
        uptotime(reg[2], 0.1); uptotime(reg[2], 0.2)
        apply((reg[1], reg[2]), gate)
        uptotime(reg[1], 0.4); uptotime(reg[2], 0.3); uptotime(reg[2], 0.4)
        return observable((reg[1], reg[2]), obs)

    names = ["X", "-X", "Z", "-Z", "Y", "-Y"]
    rng = np.random.default_rng(0)
    cases = [(l, r, o, g) for l in names for r in names for o in ("XX", "ZZ", "XY") for g in (ops.CNOT, ops.CPHASE)]
    for ci in rng.choice(len(cases), 24, replace=False):
        l, r, o, gate = cases[ci]
        obs = ops.tensor(*[{"X": ops.X, "Y": ops.Y, "Z": ops.Z}[c] for c in o])
        for bg in (S.T2Dephasing(1.0), S.Depolarization(1.0)):
            for composed in (False, True):
                sl, sr = ops.StabilizerState(l), ops.StabilizerState(r)
                init = ops.tensor(sl, sr) if composed else (sl, sr)
                v_qo = run(Register(2, CudaBatchRepr(one, "qo"), bg), init, gate, obs).mean
                v_qc = run(Register(2, CudaBatchRepr(ctx, "clifford"), bg), init, gate, obs).mean
                assert abs(v_qo - v_qc) < 4 / np.sqrt(N), (l, r, o, gate.name, bg, v_qo, v_qc)
    ctx.close(); one.close()


def test_mc_ensemble_matches_density_matrix():
    """quantummc_repr_tests.jl:105-132 / traceout_tests.jl:95-110: trajectory mean ~ dm value (atol 0.05)."""
    from qsb200.backend import BatchContext, CudaBatchRepr
    from qsb200.registers import Register, apply, initialize, observable, traceout, uptotime
    N = 20000
    ctx = BatchContext(N, seed=0x490)
    for f in ("mc", "qo"):
        reg = Register(2, CudaBatchRepr(ctx, f), S.T2Dephasing(1.0))
        initialize(reg[1], S.X1); initialize(reg[2], S.X1)
        apply((reg[1], reg[2]), ops.CPHASE)
        uptotime(reg[1], 0.3)
        v = observable((reg[1], reg[2]), ops.tensor(ops.X, ops.Z))
        if f == "mc":
            mc = v.mean
        else:
            assert abs(v.mean - mc) < 0.02 and abs(v.mean - np.exp(-0.3)) < 1e-12
    reg = Register(2, CudaBatchRepr(ctx, "mc"))
    initialize(reg[1:2], ops.StabilizerState("XX ZZ"))
    traceout(reg[1])
    assert abs(observable(reg[2], ops.Z).mean) < 0.03  # maximally mixed on average
    ctx.close()


def test_full_size_properties():
    """Size-independent properties at a large register (n=10 density matrix, 16 MiB per trajectory):
    U then U^dagger is the identity, channels preserve the trace, measurement renormalises, and the
    two halves of every measured Bell pair are perfectly correlated."""
    from qsb200.backend import BatchContext, CudaBatchRepr
    from qsb200.registers import Register, apply, initialize, observable, project_traceout, uptotime
    ctx = BatchContext(32, seed=3)
    rep = CudaBatchRepr(ctx, "qo")
    n = 10
    reg = Register(n, rep, S.T1T2Noise(100.0, 10.0))
    for i in range(1, n, 2):
        initialize((reg[i], reg[i + 1]), ops.DepolarizedBellPair(F=0.95))
    rng = np.random.default_rng(5)
    apply((reg[1], reg[3]), ops.CNOT)
    for i in range(3, n - 1, 2):
        apply((reg[i], reg[i + 2]), ops.CNOT)   # all pairs now share one 10-qubit dm
    ident = ops.tensor(*[ops.I] * n)
    assert abs(observable(reg[1:n], ident).mean - 1) < 1e-12
    U = rand_unitary(rng, 2)
    before = observable((reg[2], reg[7]), ops.tensor(ops.X, ops.Z)).mean
    apply((reg[2], reg[7]), U); apply((reg[2], reg[7]), U.conj().T)
    after = observable((reg[2], reg[7]), ops.tensor(ops.X, ops.Z)).mean
    assert abs(before - after) < 1e-12
    uptotime(reg[1:n], 0.5)
    assert abs(observable(reg[1:n], ident).mean - 1) < 1e-12
    m = project_traceout(reg[n], ops.Z)
    assert abs(observable(reg[1:n - 1], ops.tensor(*[ops.I] * (n - 1))).mean - 1) < 1e-12
    c = m.counts()
    assert c[1] + c[2] == 32
    ctx.close()


def test_tma_staged_gate_equals_register_path():
    """k_apply_dm2_tma (cp.async.bulk + mbarrier staged tiles, QSB_APPLY_TMA) must reproduce k_apply_dm<2> bit for bit:
    same arithmetic per element, only the data movement differs.  Ragged trajectory count: the last tile is partial."""
    import os
    from qsb200._lib import DM
    from qsb200.backend import BatchContext, CudaBatchState
    rng = np.random.default_rng(11)
    ctx = BatchContext(300, seed=9)
    a = rng.normal(size=(32, 32)) + 1j * rng.normal(size=(32, 32))
    rho = a @ a.conj().T
    rho /= np.trace(rho)
    q, r = np.linalg.qr(rng.normal(size=(4, 4)) + 1j * rng.normal(size=(4, 4)))
    got = {}
    try:
        for v in ("0", "2", "3"):
            os.environ["QSB_APPLY_TMA"] = v
            s = CudaBatchState.from_array(ctx, rho, DM)
            for pair in ([5, 2], [1, 3], [4, 5]):
                s.apply(pair, q)
            got[v] = np.stack([s.trajectory(t) for t in (0, 127, 128, 255, 256, 299)])
            s.free()
    finally:
        os.environ.pop("QSB_APPLY_TMA", None)
    assert np.array_equal(got["0"], got["2"]) and np.array_equal(got["0"], got["3"])
    assert abs(np.trace(got["0"][0].reshape(32, 32, order="F")) - 1) < 1e-12
    ctx.close()


def test_dense_observable_on_clifford_states_equals_the_dense_path():
    """observable(::MixedDestabilizer, idx, ::Operator) (src/backends/clifford/express.jl:22-46): here a Pauli-basis expansion of
    the operator, so it must agree with the QuantumOptics path on the same (also mixed) states, trajectory by trajectory."""
    import warnings
    from qsb200 import ops
    from qsb200.backend import BatchContext, CudaBatchRepr
    from qsb200.registers import Register, apply, initialize, observable, project_traceout
    rng = np.random.default_rng(5)
    a = rng.normal(size=(4, 4)) + 1j * rng.normal(size=(4, 4))
    herm, general = (a + a.conj().T) / 2, a
    one = rng.normal(size=(2, 2)) + 1j * rng.normal(size=(2, 2))
    ctx = BatchContext(70, seed=77)
    res = {}
    for f in ("clifford", "qo"):
        ctx._draw = 0
        reg = Register(4, CudaBatchRepr(ctx, f))
        initialize(reg[1:3], ops.StabilizerState("XXX ZZI IZZ"))
        initialize(reg[4], ops.MixedState(1))                       # rank-deficient: the reference would throw, the expansion does not care
        apply((reg[3], reg[4]), ops.CNOT)
        apply(reg[2], ops.H)
        with warnings.catch_warnings():
            warnings.simplefilter("ignore")
            vals = [observable((reg[1], reg[3]), herm, per_trajectory=True).values(), observable((reg[4], reg[2]), general, per_trajectory=True).values(),
                    observable(reg[2], one, per_trajectory=True).values()]
        m = project_traceout(reg[1], ops.X)                          # outcomes differ between trajectories from here on
        with warnings.catch_warnings():
            warnings.simplefilter("ignore")
            v = observable((reg[2], reg[3]), herm, per_trajectory=True)
        vals.append(v.values())
        res[f] = (vals, m.numpy(), v.mean, v.var)
    for x, y in zip(res["clifford"][0], res["qo"][0]):
        assert np.abs(x - y).max() < 1e-12
    assert (res["clifford"][1] == res["qo"][1]).all()
    assert abs(res["clifford"][2] - res["qo"][2]) < 1e-12 and abs(res["clifford"][3] - res["qo"][3]) < 1e-12
    ctx.close()
