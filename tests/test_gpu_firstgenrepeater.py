"""GPU: the firstgenrepeater workload (BASELINE.json configs[0] and [1]; examples/firstgenrepeater_lowlevel of the
reference, 6.1_compare_formalisms_noplot.jl:10-57) on the CUDA backend, every sampled observable of followed lanes
compared with its oracle twin fed the same Philox draws.

The ensemble shares the classical schedule of the success branch (DESIGN.md "Lockstep ensemble"), a single
trajectory follows its own purification outcomes like the reference does: a twin whose purifications all succeeded
must agree on the WHOLE run, a twin with a failed purification up to the time of its first failure."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu
SEED = 0x1F6
KW = dict(sizes=(2, 3, 4, 3, 2), F=0.97, t_end=12.0, dt_sample=0.2)


def _setup(formalism):
    from qsb200 import ops
    from qsb200 import symbols as S
    kw = dict(KW)
    if formalism == "mc":
        kw["noisy_pair"] = ops.StabilizerState("XX ZZ")      # a mixture would be lowered to a density matrix
    bg = S.Depolarization(20.0) if formalism == "clifford" else S.T1T2Noise(200.0, 20.0)
    return kw, bg


def _gpu(formalism, ntraj):
    from qsb200.backend import BatchContext, CudaBatchRepr
    from qsb200.registers import dispose, observable_options
    from qsb200.workloads import firstgenrepeater as fgr
    kw, bg = _setup(formalism)
    ctx = BatchContext(ntraj, seed=SEED)
    rec = []

    def record(k, tm, v):      # an unassigned slot gives the plain `something` (no state, no lanes)
        if hasattr(v, "values"):
            rec.append((k, round(tm, 9), v.values().copy(), v.n_alive))
        else:
            rec.append((k, round(tm, 9), np.full(ntraj, complex(v)), 0))

    with observable_options(per_trajectory=True):
        _, _, net = fgr.monte_carlo_trajectory(lambda: CudaBatchRepr(ctx, formalism), background=bg, record=record, **kw)
    log = dict(net.log)
    dispose(net.registers)
    ctx.close()
    return rec, log


def _twin(formalism, t):
    from qsb200.workloads import firstgenrepeater as fgr
    from tests.oracle_backend import OracleContext, OracleRepr
    kw, bg = _setup(formalism)
    ctx = OracleContext(SEED, t)
    rec = []
    _, _, net = fgr.monte_carlo_trajectory(lambda: OracleRepr(ctx, formalism), background=bg,
                                           record=lambda k, tm, v: rec.append((k, round(tm, 9), complex(v))), **kw)
    return rec, dict(net.log)


@pytest.mark.parametrize("formalism", ["qo", "mc", "clifford"])
def test_firstgenrepeater_lanes_match_oracle_twins(formalism):
    ntraj, follow = 96, tuple(range(0, 96, 8))
    rec, log = _gpu(formalism, ntraj)
    assert log["pairs"] > 0 and log["swaps"] > 0 and log["purifications"] > 0 and len(rec) == 2 * 61
    whole = 0
    for t in follow:
        orec, olog = _twin(formalism, t)
        t_fail = olog.get("first_failure_time", float("inf"))
        if t_fail == float("inf"):
            whole += 1
            assert {k: olog[k] for k in ("pairs", "swaps", "purifications")} == {k: log[k] for k in ("pairs", "swaps", "purifications")}
        assert [(k, tm) for k, tm, _ in orec] == [(k, tm) for k, tm, _, _ in rec]
        checked = 0
        for (_, tm, ov), (_, _, gv, _) in zip(orec, rec):
            if tm >= t_fail:
                break
            checked += 1
            if formalism == "clifford":
                assert gv[t] == ov, (t, tm, gv[t], ov)
            else:
                assert abs(gv[t] - ov) <= 1e-12 * max(1.0, abs(ov)), (t, tm, gv[t], ov)
        assert checked > 0
    assert whole >= 1          # at least one followed lane is compared over the whole run
    if formalism != "mc":
        allv = np.concatenate([gv.real for _, _, gv, _ in rec])
        assert (np.abs(allv) <= 1.0 + 1e-12).all() and (np.abs(allv) > 0).any()
