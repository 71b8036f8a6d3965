"""GPU: the workload drivers of BASELINE.json configs[2] (distributed_ghz_sensors) and configs[3] (repeatergrid)
on the CUDA backend, checked against oracle twins of followed trajectories fed the same Philox draws, and
against the known answers of tests/test_workloads_oracle.py on EVERY lane of the ensemble."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu
SEED = 0x6A5


def _ghz_gpu(formalism, ntraj, **kw):
    from qsb200.backend import BatchContext, CudaBatchRepr
    from qsb200.registers import dispose, observable_options
    from qsb200.workloads import ghz_sensors as G
    ctx = BatchContext(ntraj, seed=SEED)
    with observable_options(per_trajectory=True):
        k, fid, net, outcomes = G.f_tmbl_trial(lambda: CudaBatchRepr(ctx, formalism), **kw)
    vals = fid.values().copy()
    outs = np.stack([o.numpy() for o in outcomes]) if outcomes else np.zeros((0, ntraj), np.int64)
    stats = (fid.n_alive, fid.sum)
    dispose(net)
    ctx.close()
    return k, vals, outs, stats


def _ghz_oracle(formalism, t, **kw):
    from qsb200.workloads import ghz_sensors as G
    from tests.oracle_backend import OracleContext, OracleRepr
    ctx = OracleContext(SEED, t)
    k, fid, _, outcomes = G.f_tmbl_trial(lambda: OracleRepr(ctx, formalism), **kw)
    return k, complex(fid), [int(o) for o in outcomes]


@pytest.mark.parametrize("formalism,S", [("qo", 2), ("qo", 4), ("clifford", 4), ("clifford", 5)])
def test_ghz_perfect_pairs_give_unit_fidelity_on_every_lane(formalism, S):
    ntraj = 500
    k, vals, outs, stats = _ghz_gpu(formalism, ntraj, S=S, F=1.0, tau=1e15)
    assert k == S and outs.shape == (S, ntraj)
    assert np.abs(vals - 1.0).max() < 1e-12
    assert len({tuple(c) for c in outs.T}) > 1          # different outcome patterns, all corrected
    assert stats[0] == ntraj and abs(stats[1] - ntraj) < 1e-9


@pytest.mark.parametrize("formalism,S,links", [
    ("qo", 3, None), ("qo", 3, {1: 0.03, 3: 0.11}), ("clifford", 5, None), ("clifford", 4, {2: 0.05, 3: 0.06, 4: 0.2})])
def test_ghz_noisy_trial_matches_oracle_twins(formalism, S, links):
    ntraj, follow = 200, (0, 1, 37, 64, 199)
    kw = dict(S=S, F=0.93, tau=0.7, links=links)
    k, vals, outs, _ = _ghz_gpu(formalism, ntraj, **kw)
    for t in follow:
        ko, fo, oo = _ghz_oracle(formalism, t, **kw)
        assert ko == k and [int(x) for x in outs[:, t]] == oo, (t, outs[:, t], oo)
        if formalism == "clifford":
            assert vals[t] == fo
        else:
            assert abs(vals[t] - fo) <= 1e-12 * max(1.0, abs(fo)), (t, vals[t], fo)


def _grid_gpu(ntraj, mode, **kw):
    """-> list of (kind, time, per-trajectory values or None), statistics rows, classical log."""
    import torch
    from qsb200.backend import BatchContext, CudaBatchRepr
    from qsb200.registers import dispose, observable_options
    from qsb200.sharding import StatsBuffer
    from qsb200.workloads import repeatergrid as R
    ctx = BatchContext(ntraj, seed=SEED)
    rec = []
    if mode == "per_op":
        with observable_options(per_trajectory=True):
            net = R.run(lambda: CudaBatchRepr(ctx, "clifford"), record=lambda k, tm, v: rec.append((k, round(tm, 9), v.values().real.copy(), v.n_alive)), **kw)
        log = dict(net.log)
        dispose(net.registers)
        ctx.close()
        return rec, None, log
    stats = StatsBuffer(4096, "cuda:0")
    slot = [0]

    def opts():
        i = slot[0]; slot[0] += 1
        return stats.slot_ptr(i)

    with observable_options(stats_dev=opts, sync=False):
        with ctx.record() as r:
            net = R.run(lambda: CudaBatchRepr(ctx, "clifford"), record=lambda k, tm, v: rec.append((k, round(tm, 9), None, None)), **kw)
            dispose(net.registers)
    prog = r.program
    info = prog.info()
    prog.run(); ctx.sync()
    rows = stats.t[: slot[0]].cpu().numpy().copy()
    prog.free(); ctx.close()
    return rec, rows, dict(net.log, **{"program_" + k: v for k, v in info.items()})


GRID = dict(n=3, t_end=12.0, success_prob=0.01, tau=4.0)


def test_grid_per_op_matches_oracle_twins():
    from qsb200.workloads import repeatergrid as R
    from tests.oracle_backend import OracleContext, OracleRepr
    ntraj, follow = 128, (0, 5, 77, 127)
    rec, _, log = _grid_gpu(ntraj, "per_op", **GRID)
    assert log["swaps"] > 0 and log["consumed"] > 0 and len(rec) == 2 * log["consumed"]
    for t in follow:
        ctx = OracleContext(SEED, t)
        orec = []
        net = R.run(lambda: OracleRepr(ctx, "clifford"), record=lambda k, tm, v: orec.append((k, round(tm, 9), float(v))), **GRID)
        assert dict(net.log) == log
        assert [(k, tm) for k, tm, _ in orec] == [(k, tm) for k, tm, _, _ in rec]
        for (_, _, ov), (_, _, gv, _) in zip(orec, rec):
            assert gv[t] == ov
    allv = np.concatenate([gv for _, _, gv, _ in rec])
    assert set(np.unique(allv)) <= {-1.0, 1.0} and (allv == -1.0).any()


@pytest.mark.parametrize("kw", [GRID, dict(n=5, t_end=12.0, success_prob=0.01, tau=4.0)])
def test_grid_recorded_program_matches_per_op(kw):
    ntraj = 4096
    rec, _, log = _grid_gpu(ntraj, "per_op", **kw)
    rec2, rows, log2 = _grid_gpu(ntraj, "program", **kw)
    assert [(k, tm) for k, tm, _, _ in rec] == [(k, tm) for k, tm, _, _ in rec2]
    assert {k: v for k, v in log2.items() if not k.startswith("program_")} == log
    assert len(rows) == len(rec) > 0
    for (_, _, gv, na), row in zip(rec, rows):
        assert row[0] == na and row[1] == gv.sum() and row[2] == (gv * gv).sum()
    assert log2["program_global_arena"] in (0, 1)


def test_grid_without_noise_delivers_perfect_pairs_on_every_lane():
    rec, _, log = _grid_gpu(2000, "per_op", n=5, t_end=12.0, success_prob=0.01, tau=1e15)
    assert log["consumed"] > 0
    for _, _, gv, na in rec:
        assert na == 2000 and (gv == 1.0).all()
