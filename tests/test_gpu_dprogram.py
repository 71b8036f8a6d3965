"""GPU: a recorded DENSE program (one fused launch, every live density matrix of a tile of trajectories resident in
shared memory) must reproduce the per-op path: identical measurement outcomes (same Philox draws), per-trajectory
observables and states within 1e-12 relative (north_star tolerance; the fused kernels sum in a different order),
the same ensemble statistics, and `traj_shift` must equal shifting the context's trajectory ids."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu
RTOL = 1e-12
NOBS = 7


def schedule(rep, bg, sink):
    """swap + purification + noisy waiting on a 3-node chain with kets, dms, masked gates, partial traces and every
    observable mode; sink(i, refs, obs, time) records observable i."""
    from qsb200 import ops
    from qsb200.circuits import EntanglementSwap, Purify2to1
    from qsb200.registers import Register, apply, initialize, project_traceout, traceout, uptotime
    regs = [Register(3, rep, bg) for _ in range(3)]
    pair = ops.DepolarizedBellPair(F=0.93)
    bell = ops.StabilizerState("XX ZZ")
    initialize((regs[0][1], regs[1][1]), pair, time=0.0)
    initialize((regs[2][1], regs[1][2]), pair, time=0.1)
    initialize((regs[0][2], regs[2][2]), pair, time=0.3)
    initialize((regs[0][3], regs[2][3]), bell, time=0.3)          # a pure ket pair
    uptotime((regs[1][1], regs[0][1], regs[1][2], regs[2][1]), 0.5)
    EntanglementSwap()(regs[1][1], regs[0][1], regs[1][2], regs[2][1])
    sink(0, (regs[0][1], regs[2][1]), ops.tensor(ops.Z, ops.Z), 0.8)
    Purify2to1("X")(regs[0][1], regs[2][1], regs[0][2], regs[2][2])
    sink(1, (regs[0][1], regs[2][1]), ops.tensor(ops.X, ops.X), 1.0)
    sink(2, (regs[0][1], regs[2][1]), ops.SProjector(bell), 1.0)
    sink(3, (regs[2][1], regs[0][1]), np.kron(ops.H.mat, ops.Y.mat) + 0.3 * np.eye(4), 1.0)   # dense operator, swapped order
    apply((regs[0][3], regs[0][1]), ops.XCZ)                       # ket (x) dm -> 4-qubit dm
    m = project_traceout(regs[0][3], ops.Y)
    apply(regs[2][3], ops.H, mask=(m == 2))
    sink(4, (regs[2][3],), ops.Y, 1.5)
    traceout(regs[2][3])
    sink(5, (regs[0][1], regs[2][1]), ops.tensor(ops.Y, ops.Y), 2.0)
    initialize(regs[1][3], time=2.0)                               # fresh |0> ket
    apply(regs[1][3], ops.H)
    m2 = project_traceout(regs[1][3], [np.array([1, 1j]) / np.sqrt(2), np.array([1, -1j]) / np.sqrt(2)])
    apply(regs[0][1], ops.Z, mask=(m2 == 2))
    sink(6, (regs[0][1],), ops.X, 2.5)
    return regs


@pytest.mark.parametrize("ntraj,bgname", [(1000, "t1t2"), (40000, "depol"), (333, "t2")])
def test_dense_program_matches_per_op_path(ntraj, bgname):
    import torch
    from qsb200 import symbols as S
    from qsb200.backend import BatchContext, CudaBatchRepr
    from qsb200.registers import dispose, observable
    from qsb200.sharding import StatsBuffer
    bg = {"t1t2": S.T1T2Noise(3.0, 2.0), "depol": S.Depolarization(3.0), "t2": S.T2Dephasing(1.5)}[bgname]
    npad = (ntraj + 31) // 32 * 32

    def run(ctx, record):
        stats = StatsBuffer(NOBS, "cuda:0")
        per = torch.zeros(NOBS, 2 * npad, dtype=torch.float64, device="cuda:0")

        def sink(i, refs, obs, t):
            observable(refs, obs, something=0.0, time=t, stats_dev=stats.slot_ptr(i), sync=False, per_traj_dev=per[i].data_ptr())
        if record:
            with ctx.record() as rec:
                dispose(schedule(CudaBatchRepr(ctx, "qo"), bg, sink))
            return rec.program, stats, per
        dispose(schedule(CudaBatchRepr(ctx, "qo"), bg, sink))
        ctx.sync()
        return None, stats, per

    def close(a, b, what):
        a, b = a.cpu().numpy(), b.cpu().numpy()
        scale = max(np.abs(a).max(), 1.0)      # observables and statistics-per-trajectory are O(1): 1e-12 absolute below that
        assert np.abs(a - b).max() <= RTOL * scale, (what, np.abs(a - b).max(), scale)

    ctx = BatchContext(ntraj, seed=77)
    _, s_direct, p_direct = run(ctx, False)
    ctx._draw = 0
    prog, s_prog, p_prog = run(ctx, True)
    info = prog.info()
    assert info["ops"] > 40 and info["stats_rows"] == NOBS and info["table_bytes"] > 0
    prog.run(); ctx.sync()
    for i in range(NOBS):
        close(p_direct[i], p_prog[i], f"observable {i}")
    assert torch.equal(s_direct.t[:, 0], s_prog.t[:, 0])                      # identical alive counts: identical outcomes
    close(s_direct.t[:, 1:], s_prog.t[:, 1:] , "statistics") if ntraj < 2000 else None
    assert np.allclose(s_direct.t.cpu().numpy(), s_prog.t.cpu().numpy(), rtol=1e-9, atol=1e-9)
    assert 0 < s_prog.t[1, 0].item() < ntraj                                   # some purifications failed, some succeeded
    first = p_prog.clone()
    prog.run(); ctx.sync()                                                     # a second run recomputes the same thing
    assert torch.equal(first, p_prog)
    ctx2 = BatchContext(ntraj, seed=77, traj_offset=999 * 32)
    _, s_shift, p_shift = run(ctx2, False)
    prog.run(traj_shift=999 * 32); ctx.sync()
    for i in range(NOBS):
        close(p_shift[i], p_prog[i], f"shifted observable {i}")
    assert not torch.equal(p_shift, p_direct)
    prog.free(); ctx.close(); ctx2.close()


def test_dense_program_refuses_large_states_and_reports_errors():
    from qsb200 import ops
    from qsb200._lib import MethodError
    from qsb200.backend import BatchContext, CudaBatchRepr, CudaBatchState
    from qsb200._lib import KET
    ctx = BatchContext(64)
    with pytest.raises(MethodError):
        with ctx.record():
            CudaBatchState.zeros(ctx, KET, 6)          # 4^6 elements per trajectory do not fit on chip
    ctx.close()
