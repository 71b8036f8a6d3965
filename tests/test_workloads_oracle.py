"""CPU: the workload drivers of BASELINE.json configs[2] and [3] on the oracle backend (single trajectories).

Known answers: with perfect pairs and no decoherence the hub circuit + Pauli-frame corrections must leave the
sensors in |GHZ_k> for EVERY combination of measurement outcomes (examples/distributed_ghz_sensors/setup.jl:38-71),
and every pair the grid's corner consumer sees must be a perfect |Phi+> (<ZZ> = <XX> = 1, the assumption written
at src/ProtocolZoo/ProtocolZoo.jl:866-869) whatever the swap outcomes were.
"""
import numpy as np
import pytest

import qsb200  # noqa: F401
from qsb200.workloads import ghz_sensors as G
from qsb200.workloads import repeatergrid as R
from tests.oracle_backend import OracleContext, OracleRepr


@pytest.mark.parametrize("formalism", ["qo", "clifford"])
@pytest.mark.parametrize("S", [1, 2, 3, 4, 5])
def test_ghz_projection_with_corrections_is_exact_for_perfect_pairs(formalism, S):
    if formalism == "qo" and S == 5:
        pytest.skip("the oracle's reference-faithful dense products on a 10-qubit dm take minutes; S=5 is covered by Clifford")
    seen = set()
    for t in range(12 if (formalism == "clifford" or S < 4) else 4):
        ctx = OracleContext(0x6A, t)
        k, fid, net, outcomes = G.f_tmbl_trial(lambda: OracleRepr(ctx, formalism), S=S, F=1.0, tau=1e15)
        assert k == S
        assert abs(complex(fid).real - 1.0) < 1e-12 and abs(complex(fid).imag) < 1e-12
        seen.add(tuple(int(o) for o in outcomes))
    assert len(seen) > 1 or S == 1   # the corrections were exercised on different outcome patterns


def test_ghz_partial_link_patterns_and_noise():
    ctx = OracleContext(3, 0)
    k, fid, _, _ = G.f_tmbl_trial(lambda: OracleRepr(ctx, "qo"), S=5, links={2: 0.05, 5: 0.11})
    assert k == 2 and 0.8 < complex(fid).real < 0.99
    k, fid, _, _ = G.f_tmbl_trial(lambda: OracleRepr(ctx, "qo"), S=3, links={})
    assert k == 0 and np.isnan(fid)
    # dm and Clifford-ensemble agree on the noisy fidelity (Clifford: mean of 0/1 values over trajectories)
    links = {1: 0.02, 2: 0.07, 3: 0.15}
    ctx = OracleContext(3, 0)
    f_dm = complex(G.f_tmbl_trial(lambda: OracleRepr(ctx, "qo"), S=3, links=links, F=0.9)[1]).real
    vals = []
    for t in range(1500):
        c = OracleContext(11, t)
        vals.append(float(G.f_tmbl_trial(lambda: OracleRepr(c, "clifford"), S=3, links=links, F=0.9)[1]))
    assert abs(np.mean(vals) - f_dm) < 4 * np.std(vals) / np.sqrt(len(vals)) + 1e-9


def test_ghz_cohorts_and_link_sampling():
    co = list(G.cohorts(4, 0.2, 0.001, 0.01))
    assert len(co) == 16 and abs(sum(p for _, p in co) - 1) < 1e-12
    q = 1 - 0.99 ** 200
    assert abs(dict(co)[(1, 2, 3, 4)] - q ** 4) < 1e-15
    rng = np.random.default_rng(0)
    ups = [len(G.sample_links(rng, 4, 0.2, 0.001, 0.01)) for _ in range(4000)]
    assert abs(np.mean(ups) - 4 * q) < 0.08     # f_tmbl_expected_sensors, plots.jl:52-53
    links = G.sample_links(rng, 5, 0.2, 0.001, 0.5)
    assert all(0 < t <= 0.2 for t in links.values())


def test_grid_geometry_matches_reference_helpers():
    n = 5
    assert len(R.grid_edges(n)) == 2 * n * (n - 1)
    assert R.check_nodes(n, 13, 1) and not R.check_nodes(n, 13, 25) and R.check_nodes(n, 13, 25, low=False)
    assert R.check_nodes(n, 13, 13) and R.check_nodes(n, 13, 13, low=False)
    assert R.distance(n, 13, 1) == 4 and R.distance(n, 13, 25) == -4 and R.distance(n, 7, 6) == 1
    assert R._xy(n, 5) == (1, 5) and R._xy(n, 6) == (2, 1)


@pytest.mark.parametrize("n,kw", [(3, dict(t_end=10.0, success_prob=0.01)), (5, dict(t_end=12.0, success_prob=0.01))])
def test_grid_delivers_perfect_pairs_without_noise(n, kw):
    for t in range(3):
        ctx = OracleContext(0x77, t)
        rec = []
        net = R.run(lambda: OracleRepr(ctx, "clifford"), record=lambda k, tm, v: rec.append((k, tm, v)), n=n, tau=1e15, **kw)
        assert net.log["swaps"] > 0 and net.log["consumed"] > 0 and len(rec) == 2 * net.log["consumed"]
        assert all(v == 1.0 for _, _, v in rec)


def test_grid_schedule_is_classical_and_noise_flips_signs():
    logs, recs = [], []
    for t in range(4):
        ctx = OracleContext(0x77, t)
        rec = []
        net = R.run(lambda: OracleRepr(ctx, "clifford"), record=lambda k, tm, v: rec.append((k, round(tm, 9), v)),
                    n=5, t_end=20.0, success_prob=0.01, tau=4.0)
        logs.append(dict(net.log)); recs.append(rec)
    assert all(l == logs[0] for l in logs)                                  # one schedule shared by every trajectory
    assert all([(k, tm) for k, tm, _ in r] == [(k, tm) for k, tm, _ in recs[0]] for r in recs)
    vals = {v for r in recs for _, _, v in r}
    assert vals <= {1.0, -1.0} and -1.0 in vals                             # Pauli noise only ever flips a sign
