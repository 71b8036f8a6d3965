"""GPU: the launch shapes that produce the benchmark numbers, pinned to the per-op path AND to oracle twins.

bench.py's headline replays the firstgenrepeater Clifford schedule as a recorded program whose executor block
size is chosen at run time (k_tab_exec<1>: a non-multiple of 32 threads, partial last warp, ragged last block).
The small program tests (test_gpu_program.py) fit one wave and never reach that variant, so this file records
the schedule with bench.py's own parameters at ensemble sizes that force it and compares

  * every statistics row with the per-op path (bit-exact: integer tallies), and
  * the per-trajectory observables of followed lanes with oracle twins that follow their own control flow
    (whole run when no purification of the twin failed, else up to the twin's first failure -- the lockstep
    "masks" ensemble shares the success-branch schedule; tests/test_gpu_cohorts.py covers the exact mode).

Same for the repeatergrid program (42-thread blocks) and the dense executor on the full firstgenrepeater_dm
schedule (1e-12 relative)."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _bench():
    import bench
    return bench


def _run_fgr(ctx, formalism, mode, want_per_traj, nrows=256):
    """One pass of bench.py's firstgenrepeater workload -> (program or None, StatsBuffer, per-trajectory tensor, keys, log)."""
    import torch
    from qsb200.backend import CudaBatchRepr
    from qsb200.registers import observable_options
    from qsb200.sharding import StatsBuffer
    b = _bench()
    npad = (ctx.ntraj + 31) // 32 * 32
    stats = StatsBuffer(nrows, "cuda:0")
    per = torch.zeros(nrows if want_per_traj else 1, 2 * npad, dtype=torch.float64, device="cuda:0")
    slot, keys = [0], []

    def next_row():
        i = slot[0]; slot[0] += 1
        return stats.slot_ptr(i)

    def next_per():
        return per[slot[0] - 1].data_ptr()     # called after next_row for the same observable (dict order)

    def record(kind, tm, val):
        keys.append((kind, round(tm, 9), hasattr(val, "n_alive")))

    opts = dict(stats_dev=next_row, sync=False)
    if want_per_traj:
        opts["per_traj_dev"] = next_per
    ctx._draw = 0
    if mode == "program":
        with observable_options(**opts), ctx.record() as rec:
            log = b.wl_fgr(lambda: CudaBatchRepr(ctx, formalism), formalism, record)
        return rec.program, stats, per, keys, log, slot[0]
    with observable_options(**opts):
        log = b.wl_fgr(lambda: CudaBatchRepr(ctx, formalism), formalism, record)
    ctx.sync()
    return None, stats, per, keys, log, slot[0]


def _twin(formalism, t):
    from tests.oracle_backend import OracleContext, OracleRepr
    b = _bench()
    ctx = OracleContext(b.SEED, t)
    rec = []
    log = b.wl_fgr(lambda: OracleRepr(ctx, formalism), formalism, lambda k, tm, v: rec.append((k, round(tm, 9), complex(v))))
    return rec, log


def _compare_with_twins(formalism, follow, keys, per, nobs, ntraj, rtol=0.0):
    """per: [rows][2*npad] device tensor of the backend-reached observables, in call order (keys with flag True)."""
    import torch
    idx = torch.tensor(sorted(follow), device=per.device)
    re = per[:nobs].view(nobs, -1, 2)[:, idx, 0].cpu().numpy()       # [row][followed lane]; row = index of the observable() call
    assert len(keys) == nobs
    whole = 0
    for j, t in enumerate(sorted(follow)):
        orec, olog = _twin(formalism, t)
        t_fail = olog.get("first_failure_time", float("inf"))
        whole += t_fail == float("inf")
        assert [(k, tm) for k, tm, _ in orec] == [(k, tm) for k, tm, _ in keys]
        checked = 0
        for i, (k, tm, reached) in enumerate(keys):
            if tm < t_fail:                  # an unassigned slot gives `something` = 0.0 on both sides (the row stays zero)
                ov = orec[i][2].real
                gv = re[i, j]
                assert abs(gv - ov) <= rtol * max(1.0, abs(ov)), (formalism, t, k, tm, gv, ov)
                checked += reached
        assert checked > 0
    return whole


@pytest.mark.parametrize("ntraj,force_threads", [(1_000_000, None), (999_983, None), (40_009, "77")])
def test_headline_clifford_program_at_the_benchmarked_shape(ntraj, force_threads):
    import torch
    from qsb200.backend import BatchContext
    old = os.environ.pop("QSB_EXEC_THREADS", None)
    if force_threads:
        os.environ["QSB_EXEC_THREADS"] = force_threads
    try:
        ctx = BatchContext(ntraj, seed=_bench().SEED)
        prog, s_prog, p_prog, keys, log, nobs = _run_fgr(ctx, "clifford", "program", True)
    finally:
        os.environ.pop("QSB_EXEC_THREADS", None)
        if old is not None:
            os.environ["QSB_EXEC_THREADS"] = old
    info = prog.info()
    # the run-time sized executor with a partial last warp -- the variant BENCH times; never the fixed-shape fallback
    assert info["threads"] % 32 != 0 and not info["global_arena"], info
    if force_threads:
        assert info["threads"] == int(force_threads) and ((ntraj + 31) // 32) % info["threads"] != 0
    else:
        assert (ntraj + 31) // 32 > info["threads"] * 148 // 2          # more than one block per SM pair: a real multi-block grid
    assert info["ops"] > 1000 and nobs >= info["stats_rows"] > 200   # a sampling point before the first end-to-end pair reaches no state
    prog.run(); ctx.sync()
    rows_prog = s_prog.t[:nobs].clone()
    _, s_op, p_op, keys2, log2, nobs2 = _run_fgr(ctx, "clifford", "per_op", True)
    assert keys2 == keys and nobs2 == nobs and log2 == log
    assert torch.equal(rows_prog[:, :3], s_op.t[:nobs, :3])            # n_alive, sum, sum of squares: integer tallies
    assert torch.equal(p_prog[:nobs], p_op[:nobs])                     # every lane, every observable
    assert 0 < rows_prog[-1, 0].item() < ntraj                          # some lanes lost their end-to-end pair, some kept it
    # followed lanes, including the last valid lane, a lane of the partial last warp of block 0 and of the last block
    th = info["threads"]
    follow = {0, 33, 32 * (th - 1) + 5, 32 * th + 1, ntraj // 2, ntraj - 1, ntraj - 40, 32 * th * 147 + 7 if 32 * th * 147 + 7 < ntraj else 77}
    whole = _compare_with_twins("clifford", follow, keys, p_prog, nobs, ntraj)
    # a second run on shifted ids must equal a context created at that offset
    prog.run(traj_shift=5 * 32); ctx.sync()
    ctx2 = BatchContext(ntraj, seed=_bench().SEED, traj_offset=5 * 32)
    _, s_shift, _, _, _, _ = _run_fgr(ctx2, "clifford", "per_op", False)
    assert torch.equal(s_prog.t[:nobs, :3], s_shift.t[:nobs, :3])
    assert not torch.equal(s_prog.t[:nobs, :3], rows_prog[:, :3])
    prog.free(); ctx.close(); ctx2.close()
    assert whole >= 0


def test_headline_dense_program_on_the_full_firstgenrepeater_dm_schedule():
    """k_dense_exec on bench.py's firstgenrepeater_dm schedule vs the per-op kernels (1e-12 relative per lane) and oracle twins."""
    import torch
    from qsb200.backend import BatchContext
    ntraj = 20_011
    ctx = BatchContext(ntraj, seed=_bench().SEED)
    prog, s_prog, p_prog, keys, log, nobs = _run_fgr(ctx, "qo", "program", True)
    info = prog.info()
    assert info["ops"] > 1000 and nobs >= info["stats_rows"] > 200
    prog.run(); ctx.sync()
    _, s_op, p_op, keys2, log2, nobs2 = _run_fgr(ctx, "qo", "per_op", True)
    assert keys2 == keys and nobs2 == nobs and log2 == log
    a, b = p_prog[:nobs].cpu().numpy(), p_op[:nobs].cpu().numpy()
    assert np.abs(a - b).max() <= 1e-12 * max(1.0, np.abs(b).max()), np.abs(a - b).max()
    assert torch.equal(s_prog.t[:nobs, 0], s_op.t[:nobs, 0])           # identical alive counts <=> identical outcomes on every lane
    sa, sb = s_prog.t[:nobs].cpu().numpy(), s_op.t[:nobs].cpu().numpy()
    assert np.abs(sa - sb).max() <= 1e-12 * ntraj
    _compare_with_twins("qo", {0, 31, 32, 4097, ntraj - 1}, keys, p_prog, nobs, ntraj, rtol=1e-12)
    prog.free(); ctx.close()


@pytest.mark.parametrize("waves", [2, 6])
def test_repeatergrid_program_at_the_benchmarked_shape(waves):
    """bench.py's repeatergrid config: 1341 planes -> 42-thread executor blocks in shared memory; past 4 waves of them the same
    executor runs over a plane-major arena in global memory (the shape bench.py times at 10^7 lanes).  Statistics rows of the
    program vs per-op, both placements."""
    import torch
    from qsb200.backend import BatchContext, CudaBatchRepr
    from qsb200.registers import observable_options
    from qsb200.sharding import StatsBuffer
    b = _bench()
    ntraj = waves * 148 * 42 * 32 - 5     # whole waves of 148 blocks x 42 threads x 32 lanes, ragged last word
    ctx = BatchContext(ntraj, seed=b.SEED)

    def one(mode):
        stats = StatsBuffer(4096, "cuda:0")
        slot, keys = [0], []

        def next_row():
            i = slot[0]; slot[0] += 1
            return stats.slot_ptr(i)
        ctx._draw = 0
        rec = lambda k, tm, v: keys.append((k, round(tm, 9)))
        if mode == "program":
            with observable_options(stats_dev=next_row, sync=False), ctx.record() as r:
                log = b.wl_grid(lambda: CudaBatchRepr(ctx, "clifford"), "clifford", rec)
            return r.program, stats, keys, log, slot[0]
        with observable_options(stats_dev=next_row, sync=False):
            log = b.wl_grid(lambda: CudaBatchRepr(ctx, "clifford"), "clifford", rec)
        ctx.sync()
        return None, stats, keys, log, slot[0]

    prog, s_prog, keys, log, nobs = one("program")
    info = prog.info()
    assert info["planes"] > 1000 and bool(info["global_arena"]) == (waves > 4), info
    assert info["global_arena"] or info["threads"] == 42, info
    prog.run(); ctx.sync()
    _, s_op, keys2, log2, nobs2 = one("per_op")
    assert keys2 == keys and log2 == log and nobs2 == nobs > 0
    assert torch.equal(s_prog.t[:nobs, :3], s_op.t[:nobs, :3])
    assert (s_prog.t[:nobs, 0] == ntraj).all()                          # the grid never loses lanes (no purification)
    prog.free(); ctx.close()
