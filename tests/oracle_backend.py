"""A single-trajectory backend built on the CPU oracle, with the same method surface as
qsb200.backend.CudaBatchState.  TESTS ONLY: it lets the register layer + circuits (registers.py,
circuits.py) run against the oracle so that (a) the reference's known-answer tests can be re-expressed
on CPU, and (b) a GPU ensemble can be compared trajectory by trajectory with an oracle twin that
consumes the same (seed, trajectory, draw_id) draws.
"""
from __future__ import annotations

import numpy as np

from oracle import oracle as orc
from qsb200 import ops
from qsb200 import symbols as S
from qsb200._lib import ArgumentError, MethodError, QsbError


class OracleContext:
    """Mirrors BatchContext's draw numbering for ONE trajectory id."""

    ntraj = 1

    def __init__(self, seed=0x5153, traj=0):
        self.seed, self.traj = seed, traj
        self._draw = 0

    def next_draw(self):
        d = self._draw
        self._draw += 1
        return d

    def u(self):
        return orc.draw(self.seed, self.traj, self.next_draw())


class OracleRepr:
    def __init__(self, ctx: OracleContext, formalism="qo"):
        self.ctx, self.formalism = ctx, formalism

    def __eq__(self, o):
        return isinstance(o, OracleRepr) and o.ctx is self.ctx and o.formalism == self.formalism

    def __hash__(self):
        return hash((id(self.ctx), self.formalism))

    def newstate(self):
        if self.formalism == "clifford":
            return OracleState(self.ctx, orc.Stab.newstate(), "tab")
        return OracleState(self.ctx, orc.Dense.newstate(), "mc" if self.formalism == "mc" else "qo")

    def express(self, spec):
        ctx = self.ctx
        if isinstance(spec, OracleState):
            return spec
        if self.formalism == "clifford":
            if isinstance(spec, (ops.StabilizerState, ops.MixedStabilizer)):
                return OracleState(ctx, orc.Stab.from_generators(spec.n, spec.xs, spec.zs, spec.signs), "tab")
            if isinstance(spec, ops.Mixture):
                terms = [self.express(s) for _, s in spec.terms]
                w = np.array([w for w, _ in spec.terms]); cum = np.cumsum(w / w.sum())
                u = ctx.u()
                sel = next((i for i, c in enumerate(cum) if c > u), len(terms) - 1)
                return terms[sel]
            raise MethodError(3, "express to CliffordRepr not defined")
        flavor = "mc" if self.formalism == "mc" else "qo"
        if isinstance(spec, ops.StabilizerState):
            return OracleState(ctx, orc.Dense.from_array(spec.ket), flavor)
        if isinstance(spec, (ops.Mixture, ops.MixedStabilizer)):
            return OracleState(ctx, orc.Dense.from_array(spec.dm), "qo")
        a = np.asarray(spec, dtype=np.complex128)
        return OracleState(ctx, orc.Dense.from_array(a), flavor if a.ndim == 1 else "qo")


class OracleState:
    ispadded = False

    def __init__(self, ctx, s, flavor):
        self.ctx, self.s, self.flavor = ctx, s, flavor  # flavor: "qo" | "mc" | "tab"

    @property
    def nsubsystems(self):
        return self.s.n

    @property
    def is_mc(self):
        return self.flavor == "mc" and self.s.kind == 0

    def copy(self):
        return OracleState(self.ctx, self.s.copy(), self.flavor)

    def kill(self, lanes):
        raise NotImplementedError("single-trajectory backend: masks are plain bools")

    def subsystemcompose(self, *others):
        acc, flavor = self.s, self.flavor
        for o in others:
            if (flavor == "tab") != (o.flavor == "tab"):
                raise MethodError(3, "subsystemcompose of a tableau and a dense state")
            acc = acc.tensor(o.s)
            if flavor == "mc" and not (o.flavor == "mc" and o.s.kind == 0):
                flavor = "qo"
        if flavor == "mc" and acc.kind != 0:
            flavor = "qo"
        return OracleState(self.ctx, acc, flavor)

    def apply(self, indices, op, mask=None):
        idx = [i - 1 for i in indices]
        if self.flavor == "tab":
            if not isinstance(op, ops.Gate):
                raise MethodError(3, "dense operator on a tableau")
            self.s.apply(idx, op.table)
        else:
            self.s.apply(idx, op.mat if isinstance(op, ops.Gate) else np.asarray(op))
        return self

    def _evolve(self, indices, H, lind, duration):
        idx = [i - 1 for i in indices]
        if self.flavor == "tab":
            raise MethodError(3, "apply_noninstant! has no method for a Clifford state")
        if self.is_mc:
            d = self.ctx.next_draw() if lind else 0
            self.s.evolve_mcwf(idx, H, lind, duration, self.ctx.seed, self.ctx.traj, d)
        else:
            self.s.evolve_master(idx, H, lind, duration)
            self.flavor = "qo"
        return self

    def apply_noninstant(self, indices, operation, backgrounds):
        H = operation.hamiltonian
        H = H.mat if isinstance(H, ops.Gate) else np.asarray(H, dtype=np.complex128)
        lind = [(j, L) for j, bg in enumerate(backgrounds) if bg is not None for L in S.lindbladop(bg)]
        return self._evolve(indices, H, lind, operation.duration)

    def uptotime(self, idx, bg, dt):
        if isinstance(bg, S.LindbladOnly):
            return self._evolve([idx], None, [(0, L) for L in S.lindbladop(bg)], dt)
        d = self.ctx.next_draw()
        u = orc.draw(self.ctx.seed, self.ctx.traj, d)
        try:
            if self.flavor == "tab":
                self.s.uptotime(idx - 1, bg, dt, u)
            elif self.is_mc:
                self.s.uptotime_mc(idx - 1, bg, dt, u)
            else:
                self.s.uptotime(idx - 1, bg, dt)
                self.flavor = "qo"
        except orc.OracleError as e:
            if e.code == 3:
                raise MethodError(3, "no method") from e
            raise
        return self

    def project_traceout(self, idx, basis):
        u = self.ctx.u()
        if self.flavor == "tab":
            if not (isinstance(basis, ops.Gate) and basis.pauli in ("X", "Y", "Z")):
                raise MethodError(3, "Clifford measurements support X, Y, Z only")
            j, wr = self.s.project_traceout(idx - 1, basis.pauli, u)
            self.last_was_random = wr
        else:
            psis = S.EIGVECS[basis.pauli] if isinstance(basis, ops.Gate) else basis
            try:
                j = self.s.project_traceout(idx - 1, psis, u)
            except orc.OracleError as e:
                raise QsbError(e.code, str(e)) from e
        return j, (self if self.s.n > 0 else None)

    def traceout(self, idx):
        if self.is_mc:
            if self.s.n == 1:
                raise ArgumentError(7, "Partial trace can't be used to trace out all subsystems - use tr() instead.")
            self.s.project_traceout(idx - 1, S.EIGVECS["Z"], self.ctx.u())
            return self
        if self.s.n == 1:
            raise ArgumentError(7, "Partial trace can't be used to trace out all subsystems - use tr() instead.")
        self.s.ptrace(idx - 1)
        if self.flavor != "tab":
            self.flavor = "qo"
        return self

    def observable(self, indices, obs, **kw):
        idx = [i - 1 for i in indices]
        if self.flavor == "tab":
            if isinstance(obs, ops.Gate) and obs.pauli is not None:
                return float(self.s.expect_pauli(idx, obs.pauli))
            if isinstance(obs, ops.SProjector) and isinstance(obs.state, ops.StabilizerState):
                return float(self.s.expect_proj(idx, obs.state.spec))
            raise MethodError(3, "dense observable on a tableau")
        if isinstance(obs, ops.SProjector):
            return self.s.expect_proj(idx, obs.ket)
        return self.s.expect_op(idx, obs.mat if isinstance(obs, ops.Gate) else np.asarray(obs))
