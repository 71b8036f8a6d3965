"""GPU: error behaviour and edge sizes of the C-ABI (the reference's BoundsError / MethodError / ArgumentError cases and the
shapes a lockstep ensemble adds: one trajectory, a ragged last word, an empty lane subset, the largest dense state)."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx():
    from qsb200.backend import BatchContext
    c = BatchContext(33, seed=0xED6E)          # one full word + one lane
    yield c
    c.close()


def test_index_errors_are_bounds_errors(ctx):
    from qsb200 import ops
    from qsb200 import symbols as S
    from qsb200._lib import DM, KET, BoundsError
    from qsb200.backend import CudaBatchState
    st = CudaBatchState.zeros(ctx, KET, 3)
    with pytest.raises(BoundsError):
        st.apply([4], S.X)                     # subsystem 4 of 3
    with pytest.raises(BoundsError):
        st.apply([2, 2], S.CNOT)               # repeated index
    with pytest.raises(BoundsError):
        st.uptotime(0, S.T1Decay(1.0), 0.1)    # 1-based indices: 0 is out of range
    with pytest.raises(BoundsError):
        st.project_traceout(5, ops.Z)
    # the failed calls left the state untouched
    assert np.allclose(st.trajectory(32), np.eye(8)[0])
    st.free()
    d = CudaBatchState.zeros(ctx, DM, 2)
    with pytest.raises(BoundsError):
        d._evolve([1, 3], np.eye(4), [], 0.1)
    d.free()


def test_unsupported_operations_are_method_errors(ctx):
    from qsb200 import ops
    from qsb200 import symbols as S
    from qsb200._lib import KET, ArgumentError, MethodError
    from qsb200.backend import CudaBatchRepr, CudaBatchState, subsystemcompose
    from qsb200.registers import Register, apply, initialize, traceout, uptotime
    st = CudaBatchState.zeros(ctx, KET, 3)
    with pytest.raises(MethodError):
        st.apply([1, 2, 3], np.eye(8))         # dense operators on more than two subsystems
    with pytest.raises(ArgumentError):
        st.apply([1, 2], S.X)                  # operator size does not match the indices
    reg = Register(2, CudaBatchRepr(ctx, "clifford"), S.T1Decay(1.0))
    initialize(reg[1]); initialize(reg[2], ops.StabilizerState("X"))
    with pytest.raises(MethodError):
        uptotime(reg[1], 0.5)                  # no Clifford method for T1 (src/backends/clifford/uptotime.jl:26-43)
    with pytest.raises(MethodError):
        apply(reg[2], ops.Gate("T", np.diag([1, np.exp(0.25j * np.pi)])))   # not a Clifford gate
    with pytest.raises(MethodError):
        subsystemcompose(st, reg.staterefs[1].state)       # a tableau and a ket
    mc = Register(1, CudaBatchRepr(ctx, "mc"))
    initialize(mc[1], S.X1)
    with pytest.raises(ArgumentError, match="trace out all subsystems"):
        mc.staterefs[0].state.traceout(1)      # mcket.jl:47-49
    st.free()


def test_largest_dense_state_and_the_size_limit():
    from qsb200 import symbols as S
    from qsb200._lib import KET, MethodError
    from qsb200.backend import BatchContext, CudaBatchState, subsystemcompose
    c = BatchContext(2, seed=1)
    big = CudaBatchState.zeros(c, KET, 14)                 # 2^14 amplitudes per trajectory
    big.apply([14], S.H); big.apply([14, 1], S.CNOT)
    v = big.trajectory(1)
    assert abs(v[0] - 2 ** -0.5) < 1e-15 and abs(v[(1 << 13) | 1] - 2 ** -0.5) < 1e-15 and abs(np.vdot(v, v) - 1) < 1e-14
    one = CudaBatchState.zeros(c, KET, 1)
    with pytest.raises(MethodError, match="exceed"):
        subsystemcompose(big, one)                         # 15 qubits: past QSB_DENSE_MAXN
    big.free(); one.free(); c.close()


def test_one_trajectory_program_and_empty_lane_subset():
    """ntraj = 1 (the reference's scalar case) through a recorded program; a lane subset of zero lanes is a no-op."""
    import torch
    from qsb200 import ops
    from qsb200 import symbols as S
    from qsb200.backend import BatchContext, CudaBatchRepr
    from qsb200.registers import Register, apply, dispose, initialize, observable, project_traceout
    from qsb200.sharding import StatsBuffer
    for ntraj in (1, 31):
        c = BatchContext(ntraj, seed=7)
        stats = StatsBuffer(2, "cuda:0")

        def run():
            reg = Register(2, CudaBatchRepr(c, "clifford"), S.Depolarization(0.5))
            initialize(reg[1:2], ops.StabilizerState("XX ZZ"))
            observable(reg[1:2], ops.tensor(ops.Z, ops.Z), time=0.4, stats_dev=stats.slot_ptr(0), sync=False)
            m = project_traceout(reg[1], ops.X)
            if ntraj > 1:                                   # one trajectory: `m == 2` is a plain Bool the host branches on, which a
                apply(reg[2], ops.Z, mask=(m == 2))         # plain recording refuses (host-visible); cohort trees cover that case
            observable(reg[2], ops.X, something=0.0, time=0.4, stats_dev=stats.slot_ptr(1), sync=False)
            dispose([reg])
        run(); c.sync()
        want = stats.t.cpu().numpy().copy()
        c._draw = 0
        with c.record() as rec:
            run()
        stats.t.zero_(); torch.cuda.synchronize()
        rec.program.run(); c.sync()
        assert np.array_equal(stats.t.cpu().numpy()[:, :3], want[:, :3]) and want[0, 0] == ntraj
        lanes = torch.zeros(4, dtype=torch.int32, device="cuda:0")
        rec.program.run_lanes(lanes.data_ptr(), 0, keep_stats=True); c.sync()       # zero lanes: nothing changes
        assert np.array_equal(stats.t.cpu().numpy()[:, :3], want[:, :3])
        rec.program.free(); c.close()


def test_dense_operator_observable_on_tableau_states_is_a_pauli_expansion():
    """observable(::MixedDestabilizer, indices, ::Operator), src/backends/clifford/express.jl:22-46: the k-qubit operator is expanded
    in Pauli strings, every term one tableau expectation accumulated per trajectory (the accumulation is done on the host from
    blocking copies -- the device row is rewritten by the next term on another stream)."""
    import warnings
    from qsb200 import ops
    from qsb200 import symbols as S
    from qsb200.backend import BatchContext, CudaBatchRepr
    from qsb200.registers import Register, apply, initialize, observable, uptotime
    ctx = BatchContext(1000, seed=0xDE25)
    try:
        reg = Register(2, CudaBatchRepr(ctx, "clifford"), S.Depolarization(2.0))
        initialize(reg[1:2], ops.StabilizerState("XX ZZ"))
        bell = np.array([1, 0, 0, 1]) / np.sqrt(2)
        proj = np.outer(bell, bell.conj())                        # |Phi+><Phi+| = (II + XX - YY + ZZ) / 4: four terms share one device row
        with warnings.catch_warnings():
            warnings.simplefilter("ignore")
            assert abs(complex(observable(reg[1:2], proj)) - 1.0) < 1e-12
            uptotime(reg[1], 0.5); uptotime(reg[2], 0.5)          # Pauli noise: lanes now differ
            st = reg.staterefs[0].state
            got = st.observable([1, 2], proj, per_trajectory=True)
            terms = [st.observable([1, 2], ops.Gate(p, S.tensor(*[{"X": S.X, "Y": S.Y, "Z": S.Z}[c] for c in p]), p), per_trajectory=True).values()
                     for p in ("XX", "YY", "ZZ")]
        want = (1 + terms[0] - terms[1] + terms[2]) / 4
        assert np.max(np.abs(got.values() - want)) < 1e-12
        assert 0.05 < 1 - got.mean < 0.6                           # some lanes were hit, not all
        assert set(np.round(got.values().real, 12)) <= {0.0, 1.0}  # a Bell projector on a Bell-diagonal stabilizer state
    finally:
        ctx.close()
