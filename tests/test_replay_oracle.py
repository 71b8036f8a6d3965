"""The compiled CPU trajectory loop (oracle/qsb_replay.c + oracle/trace.py: bench.py's CPU arm) must reproduce the
Python-driven oracle backend (tests/oracle_backend.py) trajectory by trajectory, bit for bit: same oracle functions,
same draws, same order -- for every formalism, with every trajectory following its own control flow."""
import numpy as np
import pytest

from oracle import trace
from qsb200 import ops
from qsb200 import symbols as S
from qsb200.registers import dispose, observable_options
from qsb200.workloads import firstgenrepeater as fgr
from tests.oracle_backend import OracleContext, OracleRepr

SEED = 0x1F6
KW = dict(sizes=(2, 3, 4, 3, 2), F=0.97, t_end=8.0, dt_sample=0.2)
NROWS = 2 * 41


def _setup(formalism):
    kw = dict(KW)
    if formalism == "mc":
        kw["noisy_pair"] = ops.StabilizerState("XX ZZ")
    bg = S.Depolarization(15.0) if formalism == "clifford" else S.T1T2Noise(150.0, 15.0)
    return kw, bg


def _tree(formalism, max_depth=None):
    kw, bg = _setup(formalism)
    slot = [0]

    def next_row():
        i = slot[0]; slot[0] += 1
        return i

    def schedule(make_repr):
        slot[0] = 0
        with observable_options(stats_dev=next_row):
            _, _, net = fgr.monte_carlo_trajectory(make_repr, background=bg, **kw)
        dispose(net.registers)
        assert slot[0] == NROWS
    return trace.TraceTree(schedule, formalism, NROWS, max_depth=max_depth)


def _twin(formalism, t):
    kw, bg = _setup(formalism)
    ctx = OracleContext(SEED, t)
    vals = []
    _, _, net = fgr.monte_carlo_trajectory(lambda: OracleRepr(ctx, formalism), background=bg,
                                           record=lambda k, tm, v: vals.append(float(np.real(complex(v)))), **kw)
    return np.array(vals), net.log.get("failures", 0)


@pytest.mark.parametrize("formalism,n", [("clifford", 96), ("qo", 24), ("mc", 24)])
def test_compiled_replay_equals_the_python_driven_oracle(formalism, n):
    tree = _tree(formalism)
    lo = 100
    res = tree.run(SEED, lo, lo + n, threads=4, per_traj=True)
    assert res["complete"] == n and res["censored"] == 0
    want, fails = zip(*[_twin(formalism, t) for t in range(lo, lo + n)])
    want = np.stack(want, 1)
    assert max(fails) >= 1 and res["nodes"] > 1                   # some trajectory left the all-success history
    assert np.array_equal(res["per_traj"], want), np.argwhere(res["per_traj"] != want)[:5]
    assert np.array_equal(res["rows"][:, 1], want.sum(1)) or np.allclose(res["rows"][:, 1], want.sum(1), rtol=0, atol=1e-12)
    # thread count does not change anything (per-thread accumulators, exact for the integer-valued Clifford rows)
    res1 = tree.run(SEED, lo, lo + n, threads=1, per_traj=True)
    assert np.array_equal(res1["per_traj"], res["per_traj"]) and res1["passes"] == 1      # the tree is complete now: one pass


def test_compiled_replay_censors_like_the_cohort_tree():
    n = 200
    tree = _tree("clifford", max_depth=1)
    res = tree.run(SEED, 0, n, threads=4, per_traj=True)
    want, fails = zip(*[_twin("clifford", t) for t in range(n)])
    want, fails = np.stack(want, 1), np.array(fails)
    assert res["censored"] == int((fails >= 2).sum()) > 0 and res["complete"] == int((fails < 2).sum())
    ok = fails <= 1
    assert np.array_equal(res["per_traj"][:, ok], want[:, ok])
