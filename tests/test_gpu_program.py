"""GPU: a recorded program (one fused launch, state resident in shared memory) must reproduce the per-op path
bit for bit -- same draws, same per-trajectory observables, same statistics -- and `traj_shift` must equal
running the schedule on a context whose trajectory ids are shifted."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def schedule(rep, bg, sink):
    """swap + purification + noisy waiting on a 3-node chain; sink(i, refs, obs) records observable i."""
    from qsb200 import ops
    from qsb200.circuits import EntanglementSwap, Purify2to1
    from qsb200.registers import Register, apply, initialize, project_traceout, traceout, uptotime
    regs = [Register(3, rep, bg) for _ in range(3)]
    pair = ops.Mixture([(0.9, ops.StabilizerState("XX ZZ")), (0.1, ops.MixedState(2))])
    initialize((regs[0][1], regs[1][1]), pair, time=0.0)
    initialize((regs[2][1], regs[1][2]), pair, time=0.1)
    initialize((regs[0][2], regs[2][2]), pair, time=0.3)
    initialize((regs[0][3], regs[2][3]), ops.StabilizerState("XX ZZ"), time=0.3)
    uptotime((regs[1][1], regs[0][1], regs[1][2], regs[2][1]), 0.5)
    EntanglementSwap()(regs[1][1], regs[0][1], regs[1][2], regs[2][1])
    sink(0, (regs[0][1], regs[2][1]), ops.tensor(ops.Z, ops.Z), 0.8)
    Purify2to1("X")(regs[0][1], regs[2][1], regs[0][2], regs[2][2])
    sink(1, (regs[0][1], regs[2][1]), ops.tensor(ops.X, ops.X), 1.0)
    sink(2, (regs[0][1], regs[2][1]), ops.SProjector(ops.StabilizerState("XX ZZ")), 1.0)
    apply((regs[0][3], regs[0][1]), ops.XCZ)
    m = project_traceout(regs[0][3], ops.Y)
    apply(regs[2][3], ops.H, mask=(m == 2))
    traceout(regs[2][3])
    sink(3, (regs[0][1], regs[2][1]), ops.tensor(ops.Y, ops.Y), 2.0)
    return regs


@pytest.mark.parametrize("ntraj", [1000, 40000])
def test_program_matches_per_op_path(ntraj):
    import torch
    from qsb200 import symbols as S
    from qsb200.backend import BatchContext, CudaBatchRepr
    from qsb200.registers import dispose, observable
    from qsb200.sharding import StatsBuffer
    bg = S.Depolarization(3.0)
    npad = (ntraj + 31) // 32 * 32

    def run(ctx, record):
        stats = StatsBuffer(4, "cuda:0")
        per = torch.zeros(4, 2 * npad, dtype=torch.float64, device="cuda:0")

        def sink(i, refs, obs, t):
            observable(refs, obs, something=0.0, time=t, stats_dev=stats.slot_ptr(i), sync=False,
                       per_traj_dev=per[i].data_ptr())
        if record:
            with ctx.record() as rec:
                dispose(schedule(CudaBatchRepr(ctx, "clifford"), bg, sink))
            return rec.program, stats, per
        dispose(schedule(CudaBatchRepr(ctx, "clifford"), bg, sink))
        ctx.sync()
        return None, stats, per

    ctx = BatchContext(ntraj, seed=99)
    _, s_direct, p_direct = run(ctx, False)
    ctx._draw = 0
    prog, s_prog, p_prog = run(ctx, True)
    info = prog.info()
    assert info["ops"] > 40 and info["stats_rows"] == 4 and not info["global_arena"]
    prog.run()
    ctx.sync()
    assert torch.equal(p_direct, p_prog)
    assert torch.equal(s_direct.t[:, :3], s_prog.t[:, :3])
    assert 0 < s_prog.t[1, 0].item() <= ntraj
    # a second run recomputes the same thing (statistics are zeroed by the program)
    prog.run(); ctx.sync()
    assert torch.equal(s_direct.t[:, :3], s_prog.t[:, :3])
    # traj_shift == the same schedule on a context whose global trajectory ids are shifted
    ctx2 = BatchContext(ntraj, seed=99, traj_offset=12345 * 32)
    _, s_shift, p_shift = run(ctx2, False)
    prog.run(traj_shift=12345 * 32); ctx.sync()
    assert torch.equal(p_shift, p_prog) and torch.equal(s_shift.t[:, :3], s_prog.t[:, :3])
    assert not torch.equal(p_shift, p_direct)
    # draw_base offsets every draw id: equal to the per-op path whose host counter starts there.  An odd offset changes
    # which draws share a Philox block (the draw pass then runs its per-spec kernel instead of the block-item list).
    for base in (6, 7):
        ctx._draw = base
        _, s_off, p_off = run(ctx, False)
        prog.run(draw_base=base); ctx.sync()
        assert torch.equal(p_off, p_prog) and torch.equal(s_off.t[:, :3], s_prog.t[:, :3]), base
        assert not torch.equal(p_off, p_direct)
    prog.free(); ctx.close(); ctx2.close()


def test_recording_refuses_host_visible_calls_and_trajectory_kets():
    from qsb200 import ops
    from qsb200._lib import MethodError, QsbError
    from qsb200.backend import BatchContext, CudaBatchRepr
    ctx = BatchContext(64)
    outside = CudaBatchRepr(ctx, "clifford").newstate()
    with pytest.raises(QsbError):
        with ctx.record():
            outside.apply([1], ops.X)               # handle from outside the recording
    with pytest.raises(MethodError):
        with ctx.record():
            CudaBatchRepr(ctx, "mc").newstate()     # QuantumMCRepr trajectory kets are not recordable
    with pytest.raises(MethodError):
        with ctx.record():
            CudaBatchRepr(ctx, "clifford").newstate()
            CudaBatchRepr(ctx, "qo").newstate()     # one program = one family of states (tableau or dense)
    with pytest.raises(MethodError):
        with ctx.record():
            s = CudaBatchRepr(ctx, "clifford").newstate()
            s.observable([1], ops.Z)                # stats_host is host-visible
    inside = None
    with ctx.record() as rec:
        inside = CudaBatchRepr(ctx, "clifford").newstate()
    with pytest.raises(QsbError):
        inside.apply([1], ops.X)                    # virtual handle used after the recording ended
    ctx.close()
