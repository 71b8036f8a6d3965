"""Workload driver: examples/distributed_ghz_sensors of the reference (BASELINE.json configs[2]).

Restates the F-TMBL trial (`f_tmbl_trial`, examples/distributed_ghz_sensors/plots.jl:24-34; protocol
f_tmbl.jl:5-44) on top of registers.py: S one-slot sensors and one S-slot hub (setup.jl:27-28), one noisy
Bell pair per link that succeeded inside the fixed window, the hub circuit `ghz_project` (setup.jl:38-64),
the Pauli-frame corrections the sensors' EntanglementTracker applies on receipt of the hub's messages
(src/ProtocolZoo/ProtocolZoo.jl:505,569-573: EntanglementUpdateX -> Z gate, EntanglementUpdateZ -> X gate,
only when the reported outcome is 2) and `ghz_fidelity` (setup.jl:67-71).

What stays classical (host side, shared by the lockstep ensemble): WHICH links succeeded and WHEN
(`EntanglerProt` draws a Geometric number of attempts per link; that is control-plane randomness, so the
ensemble is partitioned into cohorts by link pattern -- `sample_links` draws one pattern, `cohorts`
enumerates them with their probabilities).  Everything quantum is per trajectory: the hub's S measurement
outcomes, and hence which corrections fire, differ lane by lane and are handled with masks.

Time bookkeeping follows the reference exactly: `ghz_project` calls `apply!` without `time`, so the operands
are only brought up to the latest access time among themselves (src/baseops/apply.jl:11-16) and
`ghz_fidelity` observes without `time` (no further decay).
"""
from __future__ import annotations

import numpy as np

from .. import ops
from ..registers import Register, apply, initialize, observable, project_traceout
from ..symbols import Depolarization

DEFAULTS = dict(S=5, F=0.99, success_prob=0.01, attempt_time=0.001, fixed_time=0.2, tau=1.0)  # setup.jl:9-13, f_tmbl.jl:49


def sample_links(rng: np.random.Generator, S, fixed_time, attempt_time, success_prob):
    """One classical draw of the entanglement window: {sensor: success time} for the links that came up.
    EntanglerProt(attempts=round(fixed_time/attempt_time), rounds=1) succeeds at attempt k ~ Geometric(p)
    if k <= attempts, and the pair is initialised at k*attempt_time (f_tmbl.jl:9-13)."""
    attempts = int(round(fixed_time / attempt_time))
    out = {}
    for i in range(1, S + 1):
        k = int(rng.geometric(success_prob))
        if k <= attempts:
            out[i] = k * attempt_time
    return out


def cohorts(S, fixed_time, attempt_time, success_prob):
    """Every link pattern (as a tuple of sensor ids) with its probability: the 2^S cohorts of SURVEY.md 8d row 3."""
    q = 1.0 - (1.0 - success_prob) ** int(round(fixed_time / attempt_time))
    for bits in range(1 << S):
        ent = tuple(i + 1 for i in range(S) if bits >> i & 1)
        yield ent, q ** len(ent) * (1 - q) ** (S - len(ent))


def build_sensor_net(S, make_repr, background):
    """setup.jl:27-28 -- S single-slot sensors plus an S-slot hub (vertex S+1)."""
    return [Register(1, make_repr(), background) for _ in range(S)] + [Register(S, make_repr(), background)]


def ghz_project(net, S, ent):
    """setup.jl:38-64 with the trackers' zero-delay corrections folded in (ProtocolZoo.jl:569-573)."""
    hub = net[S]
    first = ent[0]
    for i in ent[1:]:
        apply((hub[first], hub[i]), ops.CNOT)
    apply(hub[first], ops.H)
    m1 = project_traceout(hub[first], ops.Z)
    apply(net[first - 1][1], ops.Z, mask=(m1 == 2))       # EntanglementUpdateX -> updategate Z
    outcomes = [m1]
    for i in ent[1:]:
        m = project_traceout(hub[i], ops.Z)
        apply(net[i - 1][1], ops.X, mask=(m == 2))        # EntanglementUpdateZ -> updategate X
        outcomes.append(m)
    return outcomes


def ghz_fidelity(net, ent):
    """setup.jl:67-71 -- <GHZ_k| rho |GHZ_k> on the entangled sensors (NaN when no link came up)."""
    if not ent:
        return float("nan")
    if len(ent) == 1:
        # QuantumClifford.ghz(1) is the single generator X: the +1 eigenstate of X
        return observable([net[ent[0] - 1][1]], ops.SProjector(ops.StabilizerState("X")))
    return observable([net[i - 1][1] for i in ent], ops.SProjector(ops.ghz(len(ent))))


def f_tmbl_trial(make_repr, links=None, record=None, **kw):
    """One F-TMBL trial for a GIVEN classical link pattern `links` = {sensor: success time} (all S links at
    evenly spread times when omitted).  Returns (number of entangled sensors, fidelity, net, outcomes)."""
    p = dict(DEFAULTS); p.update(kw)
    S = p["S"]
    bg = p.get("background") or Depolarization(p["tau"])
    if links is None:
        links = {i: round(p["fixed_time"] * i / (S + 1), 9) for i in range(1, S + 1)}
    net = build_sensor_net(S, make_repr, bg)
    pair = ops.DepolarizedBellPair(F=p["F"])
    for i, t in sorted(links.items(), key=lambda kv: (kv[1], kv[0])):
        initialize((net[i - 1][1], net[S][i]), pair, time=t)   # EntanglerProt: chooseslotA=1, chooseslotB=i
    ent = sorted(links)
    outcomes = ghz_project(net, S, ent) if ent else []
    fid = ghz_fidelity(net, ent)
    if record is not None and ent:
        record("GHZ", p["fixed_time"], fid)
    return len(ent), fid, net, outcomes
