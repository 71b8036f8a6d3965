"""CircuitZoo circuits that generate the hot path's op sequences.

Restated from /root/reference/src/CircuitZoo/CircuitZoo.jl with the outcome-dependent branches
expressed through `mask=` so that the same code drives one trajectory (plain Int outcomes, bool
masks) or a lockstep ensemble (BatchOutcome / BatchMask):

    EntanglementSwap       CircuitZoo.jl:22-33
    LocalEntanglementSwap  CircuitZoo.jl:40-45
    Purify2to1             CircuitZoo.jl:144-161
    Purify2to1Node         CircuitZoo.jl:206-217
    Purify3to1             CircuitZoo.jl:279-305
"""
from __future__ import annotations

from . import ops
from .registers import apply, project_traceout, traceout


def _is2(meas):
    return meas == 2


def _branches(m) -> bool:
    """Can `if m:` be evaluated -- a plain bool, one trajectory, or an ensemble whose host loop forks (cohorts.CohortTree)?"""
    ctx = getattr(m, "ctx", None)
    return ctx is None or ctx.ntraj == 1 or getattr(ctx, "fork", None) is not None


class EntanglementSwap:
    def __call__(self, localL, remoteL, localR, remoteR):
        apply((localL, localR), ops.CNOT)
        xmeas = project_traceout(localL, ops.σˣ)
        zmeas = project_traceout(localR, ops.σᶻ)
        apply(remoteL, ops.Z, mask=_is2(xmeas))  # if xmeas==2 apply!(remoteL, Z)
        apply(remoteR, ops.X, mask=_is2(zmeas))  # if zmeas==2 apply!(remoteR, X)
        return xmeas, zmeas


class LocalEntanglementSwap:
    def __call__(self, localL, localR):
        apply((localL, localR), ops.CNOT)
        xmeas = project_traceout(localL, ops.σˣ)
        zmeas = project_traceout(localR, ops.σᶻ)
        return xmeas, zmeas


def _check_leaveout(name, *leaveouts):
    for l in leaveouts:
        if l not in ("X", "Y", "Z"):
            raise ValueError(f"`{name}` is parameterized by `leaveout` which has to be one of :X, :Y, or :Z; got {l!r}")


_P2_GATE_BASIS = {"X": (ops.CNOT, ops.σˣ), "Z": (ops.XCZ, ops.σᶻ), "Y": (ops.ZCY, ops.σʸ)}


class Purify2to1:
    """Returns `success` (bool, or a per-trajectory mask).  On failure the purified pair is traced
    out -- on the failing trajectories only when run on an ensemble."""

    def __init__(self, leaveout="X"):
        _check_leaveout("Purify2to1", leaveout)
        self.leaveout = leaveout

    def __call__(self, purifiedL, purifiedR, sacrificedL, sacrificedR):
        gate, basis = _P2_GATE_BASIS[self.leaveout]
        apply((sacrificedL, purifiedL), gate)
        apply((sacrificedR, purifiedR), gate)
        measa = project_traceout(sacrificedL, basis)
        measb = project_traceout(sacrificedR, basis)
        success = measa == measb
        if _branches(success):
            if not success:                            # the reference's own control flow (CircuitZoo.jl:156-158)
                traceout(purifiedL, purifiedR)
        else:                                          # lockstep ensemble without forking: per-trajectory mask
            traceout(purifiedL, purifiedR, mask=(measa != measb))
        return success


class Purify2to1Node:
    def __init__(self, leaveout="X"):
        _check_leaveout("Purify2to1Node", leaveout)
        self.leaveout = leaveout

    def __call__(self, purified, sacrificed):
        gate, basis = _P2_GATE_BASIS[self.leaveout]
        apply((sacrificed, purified), gate)
        return project_traceout(sacrificed, basis)


class Purify3to1:
    def __init__(self, leaveout1="X", leaveout2="Z"):
        _check_leaveout("Purify3to1", leaveout1, leaveout2)
        self.leaveout1, self.leaveout2 = leaveout1, leaveout2

    def __call__(self, purifiedL, purifiedR, sacrificedL1, sacrificedL2, sacrificedR1, sacrificedR2):
        meas = {"X": ops.σˣ, "Y": ops.σʸ, "Z": ops.σᶻ}
        if self.leaveout1 != "Y" or self.leaveout2 != "X":
            gates = {"X": ops.CNOT, "Y": ops.ZCY, "Z": ops.XCZ}
        else:
            gates = {"X": ops.CNOT, "Y": ops.XCY, "Z": ops.XCZ}
        gate1, gate2 = gates[self.leaveout1], gates[self.leaveout2]
        basis1, basis2 = meas[self.leaveout1], meas[self.leaveout2]
        apply((sacrificedL1, purifiedL), gate1)
        apply((sacrificedR1, purifiedR), gate1)
        apply((sacrificedR2, sacrificedR1), gate2)
        apply((sacrificedL2, sacrificedL1), gate2)
        measa1 = project_traceout(sacrificedL1, basis1)
        measb1 = project_traceout(sacrificedR1, basis1)
        measa2 = project_traceout(sacrificedL2, basis2)
        measb2 = project_traceout(sacrificedR2, basis2)
        s1, s2 = measa1 == measb1, measa2 == measb2
        if isinstance(s1, bool):
            success = s1 and s2
            failure = not success
        else:
            success = s1 & s2
            failure = ~success & (measa1.alive & measa2.alive & measb1.alive & measb2.alive)
        traceout(purifiedL, purifiedR, mask=failure)
        return success
