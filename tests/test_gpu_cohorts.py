"""GPU: cohort forking (quantumsavory.jl_b200/cohorts.py) -- an ensemble whose host control flow depends on
measurement outcomes must reproduce, lane by lane, N independent runs of the reference's control flow.

The firstgenrepeater purifier frees the purified pair when a purification fails and the entangler refills the
slots later (/root/reference/examples/firstgenrepeater_lowlevel/setup.jl:209-237), so every trajectory has its own
classical history.  The oracle twin of lane t is a single-trajectory run that follows its own branches
(workloads/firstgenrepeater.py with ntraj == 1, exactly like the reference); the cohort tree must give lane t the
same observable at EVERY sampling point of the whole run -- also after any number of failed purifications -- and
the statistics rows must be the sums over the twins."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu
SEED = 0x1F6
KW = dict(sizes=(2, 3, 4, 3, 2), F=0.97, t_end=12.0, dt_sample=0.2)
NROWS = 2 * 61


def _bg():
    from qsb200 import symbols as S
    return S.Depolarization(20.0)


def _tree(ntraj, max_depth=None, traj_offset=0):
    import torch
    from qsb200.backend import BatchContext, CudaBatchRepr
    from qsb200.cohorts import CohortTree
    from qsb200.registers import dispose, observable_options
    from qsb200.sharding import StatsBuffer
    from qsb200.workloads import firstgenrepeater as fgr
    ctx = BatchContext(ntraj, seed=SEED, traj_offset=traj_offset)
    npad = (ntraj + 31) // 32 * 32
    stats = StatsBuffer(NROWS, "cuda:0")
    per = torch.zeros(NROWS, 2 * npad, dtype=torch.float64, device="cuda:0")
    slot = [0]

    def next_row():
        i = slot[0]; slot[0] += 1
        return stats.slot_ptr(i)

    def schedule():
        slot[0] = 0
        with observable_options(stats_dev=next_row, sync=False, per_traj_dev=lambda: per[slot[0] - 1].data_ptr()):
            _, _, net = fgr.monte_carlo_trajectory(lambda: CudaBatchRepr(ctx, "clifford"), background=_bg(), **KW)
        dispose(net.registers)
        assert slot[0] == NROWS

    tree = CohortTree(ctx, schedule, max_depth=max_depth)
    return ctx, tree, stats, per


def _twin(t):
    from qsb200.workloads import firstgenrepeater as fgr
    from tests.oracle_backend import OracleContext, OracleRepr
    ctx = OracleContext(SEED, t)
    vals = []
    _, _, net = fgr.monte_carlo_trajectory(lambda: OracleRepr(ctx, "clifford"), background=_bg(),
                                           record=lambda k, tm, v: vals.append(float(np.real(complex(v)))), **KW)   # unassigned slot: something = 0.0
    return np.array(vals), net


def _twins(ids):
    out, fails = [], []
    for t in ids:
        v, net = _twin(t)
        out.append(v); fails.append(net.log.get("failures", 0))
    return np.stack(out, 1), np.array(fails)      # [row][lane], failures per lane


def test_every_lane_follows_its_own_control_flow():
    import torch
    ntraj = 400
    ctx, tree, stats, per = _tree(ntraj)
    stats.t.zero_(); per.zero_()
    torch.cuda.synchronize()
    info = tree.run()
    ctx.sync(); torch.cuda.synchronize()
    got = per.view(NROWS, -1, 2)[:, :ntraj, 0].cpu().numpy()
    want, fails = _twins(range(ntraj))
    assert (fails > 0).mean() > 0.3 and fails.max() >= 2           # the run is dominated by trajectories with failed purifications
    assert info["censored"] == 0 and info["dropped"] == 0
    assert info["levels"] == fails.max() + 1                        # one tree level per fork of the deepest lane
    assert info["lanes_rerun"] == fails.sum()                       # a lane is re-run once per failed purification
    bad = np.argwhere(got != want)
    assert bad.size == 0, (bad[:5], got[tuple(bad[0])], want[tuple(bad[0])])
    rows = stats.t.cpu().numpy()
    assert np.array_equal(rows[:, 1], want.sum(1)) and np.array_equal(rows[:, 2], (want * want).sum(1))
    assert (rows[:, 0] <= ntraj).all() and rows[:, 0].max() > 0
    # the tree is cached: a second batch of the same ids records nothing new and gives the same rows
    nprog = info["programs"]
    stats.t.zero_(); per.zero_(); torch.cuda.synchronize()
    info2 = tree.run()
    ctx.sync(); torch.cuda.synchronize()
    assert info2["programs"] == nprog and info2["programs_recorded_this_run"] == 0
    assert np.array_equal(stats.t.cpu().numpy(), rows)
    # a shifted batch equals twins with shifted ids
    stats.t.zero_(); per.zero_(); torch.cuda.synchronize()
    tree.run(traj_shift=1000)
    ctx.sync(); torch.cuda.synchronize()
    got = per.view(NROWS, -1, 2)[:, :64, 0].cpu().numpy()
    want2, _ = _twins(range(1000, 1064))
    assert np.array_equal(got, want2)
    tree.free(); ctx.close()


def test_depth_cap_censors_instead_of_continuing_on_a_wrong_history():
    import torch
    ntraj = 300
    ctx, tree, stats, per = _tree(ntraj, max_depth=1)
    stats.t.zero_(); per.zero_(); torch.cuda.synchronize()
    info = tree.run()
    ctx.sync(); torch.cuda.synchronize()
    got = per.view(NROWS, -1, 2)[:, :ntraj, 0].cpu().numpy()
    want, fails = _twins(range(ntraj))
    assert info["levels"] == 2 and info["censored"] == int((fails >= 2).sum()) > 0
    ok = fails <= 1
    assert np.array_equal(got[:, ok], want[:, ok])                  # lanes within the cap are exact over the whole run
    rows = stats.t.cpu().numpy()
    full = want.copy(); full[:, ~ok] = 0.0
    assert (np.abs(rows[:, 1]) <= np.abs(want).sum(1) + 1e-9).all() and rows[:, 0].max() <= ntraj
    tree.free(); ctx.close()


def test_wave_of_lane_subsets_equals_separate_runs():
    """qsb_wave_run over several (program, lane subset) jobs == the same jobs launched one by one."""
    import torch
    ntraj = 5000
    ctx, tree, stats, per = _tree(ntraj)
    root = tree.root
    g = torch.Generator(device="cpu"); g.manual_seed(3)
    perm = torch.randperm(ntraj, generator=g).to(torch.int32).cuda()
    parts = [(0, 1), (1, 34), (34, 1500), (1500, 4097), (4097, 5000)]
    rows = []
    for mode in ("wave", "separate"):
        stats.t.zero_(); per.zero_(); tree.codes.fill_(-1); torch.cuda.synchronize()
        if mode == "wave":
            tree._launch([(root, perm.data_ptr() + 4 * a, b - a) for a, b in parts], 0)
        else:
            for a, b in parts:
                root.program.run_lanes(perm.data_ptr() + 4 * a, b - a, keep_stats=True)
        ctx.sync(); torch.cuda.synchronize()
        rows.append((stats.t.clone(), per.clone(), tree.codes.clone()))
    assert torch.equal(rows[0][0], rows[1][0]) and torch.equal(rows[0][1], rows[1][1]) and torch.equal(rows[0][2], rows[1][2])
    # and both equal the identity run of the whole ensemble (lanes are independent; rows are indexed by lane)
    stats.t.zero_(); per.zero_(); tree.codes.fill_(-1); torch.cuda.synchronize()
    root.program.run_lanes(None, ntraj, keep_stats=True)
    ctx.sync(); torch.cuda.synchronize()
    assert torch.equal(rows[0][0], stats.t) and torch.equal(rows[0][1], per) and torch.equal(rows[0][2], tree.codes)
    assert (tree.codes >= 0).any()
    tree.free(); ctx.close()


def test_dense_programs_fork_too():
    """The same lane-by-lane check in the density-matrix formalism (BASELINE configs[0]): cohort forking of DENSE recorded
    programs -- k_dense_exec with lane subsets, a live-lanes word and TOP_RETIRE.  Every lane against its oracle twin at every
    sampling point (1e-12), statistics rows against the sums over the twins."""
    import torch
    from qsb200 import symbols as S
    from qsb200.backend import BatchContext, CudaBatchRepr
    from qsb200.cohorts import CohortTree
    from qsb200.registers import dispose, observable_options
    from qsb200.sharding import StatsBuffer
    from qsb200.workloads import firstgenrepeater as fgr
    from tests.oracle_backend import OracleContext, OracleRepr
    ntraj, bg = 150, S.T1T2Noise(60.0, 40.0)
    kw = dict(sizes=(2, 3, 3, 2), F=0.95, t_end=8.0, dt_sample=0.2)
    nrows = 2 * 41
    ctx = BatchContext(ntraj, seed=SEED)
    npad = (ntraj + 31) // 32 * 32
    stats = StatsBuffer(nrows, "cuda:0")
    per = torch.zeros(nrows, 2 * npad, dtype=torch.float64, device="cuda:0")
    slot = [0]

    def next_row():
        i = slot[0]; slot[0] += 1
        return stats.slot_ptr(i)

    def schedule():
        slot[0] = 0
        with observable_options(stats_dev=next_row, sync=False, per_traj_dev=lambda: per[slot[0] - 1].data_ptr()):
            _, _, net = fgr.monte_carlo_trajectory(lambda: CudaBatchRepr(ctx, "qo"), background=bg, **kw)
        dispose(net.registers)
        assert slot[0] == nrows

    tree = CohortTree(ctx, schedule)
    stats.t.zero_(); per.zero_(); torch.cuda.synchronize()
    info = tree.run()
    ctx.sync(); torch.cuda.synchronize()
    got = per.view(nrows, -1, 2)[:, :ntraj, 0].cpu().numpy()
    want, fails = [], []
    for t in range(ntraj):
        octx = OracleContext(SEED, t)
        vals = []
        _, _, net = fgr.monte_carlo_trajectory(lambda: OracleRepr(octx, "qo"), background=bg,
                                               record=lambda k, tm, v: vals.append(float(np.real(complex(v)))), **kw)
        want.append(vals); fails.append(net.log.get("failures", 0))
    want, fails = np.array(want).T, np.array(fails)
    assert (fails > 0).mean() > 0.2 and info["censored"] == 0 and info["dropped"] == 0
    assert info["levels"] == fails.max() + 1 and info["lanes_rerun"] == fails.sum()
    assert np.abs(got - want).max() < 1e-12, np.abs(got - want).max()
    rows = stats.t.cpu().numpy()
    assert np.abs(rows[:, 1] - want.sum(1)).max() < 1e-10 and np.abs(rows[:, 2] - (want * want).sum(1)).max() < 1e-10
    tree.free(); ctx.close()


def test_program_budget_of_one_drops_every_fork_and_repeats():
    """max_programs = 1: every cohort key of level 1 is dropped, so run() leaves through the `no jobs` exit right after the
    scatter.  Two runs back to back must report the same dropped lanes and leave identical statistics (the second run's reset
    of the routing tables on torch's stream must not overtake the first run's scatter on the context's stream)."""
    import torch
    from qsb200.cohorts import CohortTree
    ntraj = 2000
    ctx, tree, stats, per = _tree(ntraj)
    tree.free()
    tree = CohortTree(ctx, tree.schedule, max_programs=1)
    rows = []
    for _ in range(3):
        stats.t.zero_(); per.zero_()
        torch.cuda.synchronize()
        info = tree.run()
        ctx.sync()
        assert info["levels"] == 1 and info["programs"] == 1 and info["dropped"] > 0 and info["lanes_rerun"] == 0
        rows.append((info["dropped"], stats.t.cpu().numpy().copy()))
    assert rows[0][0] == rows[1][0] == rows[2][0]
    assert np.array_equal(rows[0][1], rows[1][1]) and np.array_equal(rows[0][1], rows[2][1])
    # the lanes that were dropped are exactly the lanes with at least one failed purification
    full_ctx, full, fstats, fper = _tree(ntraj)
    fstats.t.zero_(); fper.zero_(); torch.cuda.synchronize()
    finfo = full.run(); full_ctx.sync()
    assert rows[0][0] == finfo["lanes_per_level"][1]
    full.free(); full_ctx.close()
    tree.free(); ctx.close()


def test_levels_with_global_arena_programs_run_job_by_job(monkeypatch):
    """A program too large for a shared-memory tile keeps its arena in global memory, and that executor instantiation takes one
    job per launch.  A tree level with several such jobs used to fail ("does not fit in shared memory"); now each goes out as its
    own launch pair.  QSB_EXEC_GLOBAL=1 (read at qsb_record_end) forces the placement for every program of a small tree; the
    statistics must equal the shared-memory tree's bit for bit."""
    import torch
    ntraj = 3000
    ctx, tree, stats, per = _tree(ntraj)
    stats.t.zero_(); per.zero_(); torch.cuda.synchronize()
    info = tree.run(); ctx.sync()
    want, want_per = stats.t.cpu().numpy().copy(), per.cpu().numpy().copy()
    assert tree.root.program.info()["global_arena"] == 0 and max(info["jobs_per_level"]) > 1
    tree.free(); ctx.close()
    monkeypatch.setenv("QSB_EXEC_GLOBAL", "1")
    ctx, tree, stats, per = _tree(ntraj)
    stats.t.zero_(); per.zero_(); torch.cuda.synchronize()
    info2 = tree.run(); ctx.sync()
    assert tree.root.program.info()["global_arena"] == 1
    assert all(n.program.info()["global_arena"] == 1 for n in tree.nodes)
    assert info2["jobs_per_level"] == info["jobs_per_level"] and info2["lanes_per_level"] == info["lanes_per_level"]
    assert np.array_equal(stats.t.cpu().numpy(), want)
    assert np.array_equal(per.cpu().numpy(), want_per)
    tree.free(); ctx.close()
