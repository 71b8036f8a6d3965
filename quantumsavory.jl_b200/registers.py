"""Register / RegRef / StateRef bookkeeping above the backend boundary.

A Python restatement of the reference's register interface (the CALLER of the hot path; it stays in
Julia in a real deployment and is mirrored here only so that workloads and parity tests can be
driven without a Julia install):

    src/states_registers.jl:3-67      StateRef, Register, RegRef
    src/baseops/initialize.jl         initialize!      -> initialize
    src/baseops/subsystemcompose.jl   subsystemcompose, swap!
    src/baseops/apply.jl              apply!           -> apply
    src/baseops/uptotime.jl           uptotime!, overwritetime!
    src/baseops/observable.jl         observable
    src/baseops/traceout.jl           traceout!, project_traceout!, removebackref!
    src/noninstant.jl:9-13            NonInstantGate
    src/noninstant.jl:16-35           ConstantHamiltonianEvolution

It is backend agnostic: any state object with the methods listed in backend.py works (the CUDA
batched state in production, the oracle-backed single-trajectory state in tests), exactly like the
reference's `ThrowingTraceoutState` conformance example (test/general/traceout_tests.jl:8-16).
Slot indices are 1-BASED and slices are inclusive (`reg[1:2]` is slots 1 and 2), like the reference.

Batched extension (BASELINE.json north_star: "measurement-dependent branches handled by
per-trajectory masks"): `apply(..., mask=m)` and `traceout(..., mask=m)` take a per-trajectory mask
(or a plain bool for single-trajectory backends), so `if xmeas==2 apply!(r, Z) end` is written
`apply(r, Z, mask=(xmeas == 2))`.
"""
from __future__ import annotations

import contextlib
import itertools

from . import ops


class Qubit:
    pass


class StateRef:
    """state: the native (batched) state; registers / registerindices: back-references."""

    __slots__ = ("state", "registers", "registerindices")

    def __init__(self, state, registers, registerindices, copy=True):
        self.state = state.copy() if copy else state  # StateRef(...) copies: states_registers.jl:10
        self.registers = list(registers)
        self.registerindices = list(registerindices)


class Register:
    def __init__(self, traits, reprs=None, backgrounds=None):
        if isinstance(traits, int):
            traits = [Qubit() for _ in range(traits)]
        n = len(traits)
        if reprs is not None and not isinstance(reprs, (list, tuple)):
            reprs = [reprs] * n
        if backgrounds is not None and not isinstance(backgrounds, (list, tuple)):
            backgrounds = [backgrounds] * n
        self.traits = list(traits)
        self.reprs = list(reprs) if reprs is not None else [None] * n
        self.backgrounds = list(backgrounds) if backgrounds is not None else [None] * n
        self.staterefs = [None] * n
        self.stateindices = [0] * n
        self.accesstimes = [0.0] * n

    def __len__(self):
        return len(self.staterefs)

    def __getitem__(self, i):
        if isinstance(i, slice):
            start = 1 if i.start is None else i.start
            stop = len(self) if i.stop is None else i.stop
            step = 1 if i.step is None else i.step
            return tuple(RegRef(self, j) for j in range(start, stop + 1, step))
        if isinstance(i, (list, tuple)):
            return tuple(RegRef(self, j) for j in i)
        if not 1 <= i <= len(self):
            raise IndexError(f"slot {i} out of range 1:{len(self)}")
        return RegRef(self, i)


class RegRef:
    __slots__ = ("reg", "idx")

    def __init__(self, reg, idx):
        self.reg, self.idx = reg, idx


def _refs(refs):
    if isinstance(refs, RegRef):
        return [refs]
    return list(refs)


def _split(refs):
    r = _refs(refs)
    return [x.reg for x in r], [x.idx for x in r]


def isassigned(reg, i=None) -> bool:
    if isinstance(reg, RegRef):
        reg, i = reg.reg, reg.idx
    return reg.staterefs[i - 1] is not None


def nsubsystems(x) -> int:
    if isinstance(x, StateRef):
        return len(x.registers)
    if isinstance(x, Register):
        return len(x.staterefs)
    if isinstance(x, RegRef):
        return 1
    return ops.nsubsystems(x)


def stateof(ref: RegRef):
    return ref.reg.staterefs[ref.idx - 1]


def consistent_representation(regs, idx):
    reprs = {r.reprs[i - 1] for r, i in zip(regs, idx)}
    if len(reprs) != 1:
        raise ValueError("the slots being initialized use different representations; no common lowering is defined")
    rep = next(iter(reprs))
    if rep is None:
        raise ValueError("register slot has no representation")
    return rep


# ---- initialize! --------------------------------------------------------------------------------
def initialize(refs, state=None, time=None):
    """initialize!(refs, state; time) -- src/baseops/initialize.jl:17-58."""
    regs, indices = _split(refs)
    if state is None:  # initialize!(reg, i): newstate(trait, repr)
        if len(regs) != 1:
            raise ValueError("initialize(ref) without a state works on one slot")
        state = regs[0].reprs[indices[0] - 1].newstate()
    if isinstance(state, (tuple, list)) and not hasattr(state, "nsubsystems"):
        # STensor-like product given as a tuple of factors: initialize.jl:50-58
        total = sum(ops.nsubsystems(t) for t in state)
        if not (len(regs) == len(indices) == total):
            raise ValueError("Attempting to initialize a set of registers with a state that does not have the correct number of subsystems.")
        i = 0
        out = None
        for t in state:
            s = ops.nsubsystems(t)
            out = initialize([RegRef(r, j) for r, j in zip(regs[i:i + s], indices[i:i + s])], t, time=time)
            i += s
        return out
    if not (len(regs) == len(indices) == ops.nsubsystems(state)):
        raise ValueError("Attempting to initialize a set of registers with a state that does not have the correct number of subsystems.")
    rep = consistent_representation(regs, indices)
    native = rep.express(state)
    stateref = StateRef(native, regs, indices, copy=native is state)  # freshly expressed states are already private
    for si, (reg, ri) in enumerate(zip(regs, indices), start=1):
        if isassigned(reg, ri):
            raise RuntimeError("error")  # initialize.jl:38-40 throws "error"
        reg.staterefs[ri - 1] = stateref
        reg.stateindices[ri - 1] = si
        if time is not None:
            reg.accesstimes[ri - 1] = time
    return stateref


# ---- subsystemcompose / swap! ------------------------------------------------------------------
def unique_staterefs_with_offsets(staterefs):
    unique, seen = [], set()
    for ref in staterefs:
        if ref is None:
            raise ValueError("Cannot compose unassigned register slots.")
        if id(ref) not in seen:
            seen.add(id(ref))
            unique.append(ref)
    offsets, off = {}, 0
    for ref in unique:
        offsets[id(ref)] = off
        off += ref.state.nsubsystems
    return unique, offsets


def subsystemcompose(refs):
    """Ensure all given slots live in one state object -- src/baseops/subsystemcompose.jl:56-76."""
    regs, indices = _split(refs)
    staterefs, offsets = unique_staterefs_with_offsets([r.staterefs[i - 1] for r, i in zip(regs, indices)])
    if len(staterefs) == 1:
        return staterefs[0]  # the reference re-wraps (and copies) even here; nothing observable changes
    newstate = staterefs[0].state.subsystemcompose(*[s.state for s in staterefs[1:]])
    newregisters = [r for s in staterefs for r in s.registers]
    newindices = [i for s in staterefs for i in s.registerindices]
    newref = StateRef(newstate, newregisters, newindices, copy=False)
    for r, i in zip(newregisters, newindices):
        if r is None:
            continue
        oldref = r.staterefs[i - 1]
        r.staterefs[i - 1] = newref
        r.stateindices[i - 1] += offsets[id(oldref)]
    return newref


def swap(r1: RegRef, r2: RegRef, time=None):
    """swap!(r1, r2; time) -- src/baseops/subsystemcompose.jl:6-31."""
    reg1, i1, reg2, i2 = r1.reg, r1.idx, r2.reg, r2.idx
    if reg1 is reg2 and i1 == i2:
        return
    if reg1.accesstimes[i1 - 1] != reg2.accesstimes[i2 - 1]:
        maxtime = max(reg1.accesstimes[i1 - 1], reg2.accesstimes[i2 - 1])
        maxtime = maxtime if time is None else max(maxtime, time)
        uptotime(r1, maxtime)
        uptotime(r2, maxtime)
    s1, s2 = reg1.staterefs[i1 - 1], reg2.staterefs[i2 - 1]
    k1, k2 = reg1.stateindices[i1 - 1], reg2.stateindices[i2 - 1]
    reg1.staterefs[i1 - 1], reg2.staterefs[i2 - 1] = s2, s1
    reg1.stateindices[i1 - 1], reg2.stateindices[i2 - 1] = k2, k1
    if s1 is not None:
        s1.registers[k1 - 1] = reg2
        s1.registerindices[k1 - 1] = i2
    if s2 is not None:
        s2.registers[k2 - 1] = reg1
        s2.registerindices[k2 - 1] = i1


# ---- uptotime! ------------------------------------------------------------------------------------
def uptotime(refs, now):
    """Evolve the slots' states under their backgrounds up to `now` -- src/baseops/uptotime.jl:33-57."""
    regs, indices = _split(refs)
    records = [(r.staterefs[i - 1], r.stateindices[i - 1], r.backgrounds[i - 1], r.accesstimes[i - 1])
               for r, i in zip(regs, indices) if isassigned(r, i)]
    for _, group in itertools.groupby(records, key=lambda x: id(x[0])):  # consecutive grouping, like IterTools.groupby
        group = list(group)
        stateref = group[0][0]
        timegroups = [list(g) for _, g in itertools.groupby(group, key=lambda x: x[3])]
        timegroups.sort(key=lambda g: g[0][3])
        times = [g[0][3] for g in timegroups] + [now]
        for i in range(len(timegroups)):
            dt = times[i + 1] - times[i]
            if dt == 0:
                continue
            for g in timegroups[: i + 1]:
                for (_, idx, bg, _) in g:
                    if bg is None:
                        continue
                    stateref.state = stateref.state.uptotime(idx, bg, dt)  # uptotime.jl:22-24
    for r, i in zip(regs, indices):
        if r.accesstimes[i - 1] > now:
            raise RuntimeError(f"The simulation was commanded to apply an operation at time t={now} although the current "
                               f"simulation time is higher at t={r.accesstimes[i - 1]}. Consider using locks around the offending operations.")
        r.accesstimes[i - 1] = now


def overwritetime(refs, now):
    regs, indices = _split(refs)
    for r, i in zip(regs, indices):
        r.accesstimes[i - 1] = now
    return now


# ---- apply! ---------------------------------------------------------------------------------------
class NonInstantGate:
    """apply! then uptotime!(..., t + duration) -- src/noninstant.jl:9-13."""

    def __init__(self, gate, duration):
        self.gate, self.duration = gate, duration


class ConstantHamiltonianEvolution:
    """A Hamiltonian applied for `duration` -- src/noninstant.jl:16-19."""

    def __init__(self, hamiltonian, duration):
        self.hamiltonian, self.duration = hamiltonian, duration

    def __repr__(self):
        return f"ConstantHamiltonianEvolution(duration={self.duration})"


def _normalize_mask(mask):
    """A mask over a single trajectory is a plain Bool (the reference's scalar control flow)."""
    if mask is None or isinstance(mask, bool):
        return mask
    if getattr(getattr(mask, "ctx", None), "ntraj", 0) == 1:
        return bool(mask)
    return mask


def apply(refs, operation, time=None, mask=None):
    """apply!(refs, operation; time) -- src/baseops/apply.jl:10-23.  mask: per-trajectory condition."""
    mask = _normalize_mask(mask)
    if isinstance(operation, NonInstantGate):
        _, new_time = apply(refs, operation.gate, time=time, mask=mask)
        uptotime(refs, new_time + operation.duration)
        return refs, new_time + operation.duration
    regs, indices = _split(refs)
    max_time = max(r.accesstimes[i - 1] for r, i in zip(regs, indices))
    if time is not None:
        if time < max_time:
            raise RuntimeError(f"The simulation was commanded to apply {operation} at time t={time} although the current "
                               f"simulation time is higher at t={max_time}. Consider using locks around the offending operations.")
        max_time = time
    uptotime(refs, max_time)
    if isinstance(operation, ConstantHamiltonianEvolution):   # src/noninstant.jl:21-35
        if mask is not None:
            raise TypeError("ConstantHamiltonianEvolution takes no per-trajectory mask")
        stateref = subsystemcompose(refs)
        state_indices = [r.stateindices[i - 1] for r, i in zip(regs, indices)]
        stateref.state = stateref.state.apply_noninstant(state_indices, operation, [r.backgrounds[i - 1] for r, i in zip(regs, indices)])
        end_time = overwritetime(refs, max_time + operation.duration)
        return refs, end_time
    if mask is False:  # single-trajectory condition not met: the time bookkeeping above still happens
        return refs, max_time
    stateref = subsystemcompose(refs)
    state_indices = [r.stateindices[i - 1] for r, i in zip(regs, indices)]
    m = None if mask is None or mask is True else mask
    stateref.state = stateref.state.apply(state_indices, operation, mask=m) if m is not None else stateref.state.apply(state_indices, operation)
    return refs, max_time


# ---- observable -----------------------------------------------------------------------------------
_OBSERVABLE_KW: dict = {}


@contextlib.contextmanager
def observable_options(**kw):
    """Backend keywords (per_trajectory=True, ...) added to every observable() call inside the block, so a
    workload driver written against the reference's signature can be asked for per-trajectory values.
    A callable value is called once per observable() call, whether or not it reaches the backend (e.g. stats_dev=next_row
    hands every sampling point its own device statistics row; the row of an unassigned slot stays zero)."""
    global _OBSERVABLE_KW
    old, _OBSERVABLE_KW = _OBSERVABLE_KW, {**_OBSERVABLE_KW, **kw}
    try:
        yield
    finally:
        _OBSERVABLE_KW = old


def observable(refs, obs, something=None, time=None, **kw):
    """observable(refs, obs; something, time) -- src/baseops/observable.jl:12-41."""
    if _OBSERVABLE_KW:
        kw = {**_OBSERVABLE_KW, **kw}
    # callables are evaluated for EVERY call, also one that ends at an unassigned slot below: a row handed out per call
    # keeps its meaning (call index = sampling point) whatever the classical history of the run is
    kw = {k: (v() if callable(v) else v) for k, v in kw.items()}
    regs, indices = _split(refs)
    if len(regs) != len(indices):
        raise ValueError("The number of registers does not match the number of slot indices.")
    slots = set()
    for r, i in zip(regs, indices):
        if not 1 <= i <= len(r):
            raise IndexError(i)
        key = (id(r), i)
        if key in slots:
            raise ValueError("Each physical register slot can be observed at most once.")
        slots.add(key)
    staterefs = []
    for r, i in zip(regs, indices):
        ref = r.staterefs[i - 1]
        if ref is None:
            return something
        staterefs.append(ref)
    if time is not None:
        uptotime(refs, time)
    unique, offsets = unique_staterefs_with_offsets(staterefs)
    state = unique[0].state if len(unique) == 1 else unique[0].state.subsystemcompose(*[s.state for s in unique[1:]])
    state_indices = [r.stateindices[i - 1] + offsets[id(r.staterefs[i - 1])] for r, i in zip(regs, indices)]
    return state.observable(state_indices, obs, **kw)


# ---- traceout! / project_traceout! ------------------------------------------------------------------
def removebackref(s: StateRef, i: int):
    """src/baseops/traceout.jl:3-22 (unpadded storage)."""
    for r, ri in zip(s.registers, s.registerindices):
        if r is None:
            continue
        if r.stateindices[ri - 1] == i:
            r.staterefs[ri - 1] = None
            r.stateindices[ri - 1] = 0
        elif r.stateindices[ri - 1] > i:
            r.stateindices[ri - 1] -= 1
    del s.registerindices[i - 1]
    del s.registers[i - 1]
    return s


def _traceout_stateref(s: StateRef, i: int):
    s.state = s.state.traceout(i)
    removebackref(s, i)
    return s


def _traceout_slot(r: Register, i: int):
    stateref = r.staterefs[i - 1]
    if stateref is not None:
        if nsubsystems(stateref) > 1:
            _traceout_stateref(stateref, r.stateindices[i - 1])
        else:
            r.staterefs[i - 1] = None
            r.stateindices[i - 1] = 0
    return r


def traceout(*refs, mask=None):
    """traceout!(refs...) -- src/baseops/traceout.jl:72-119.  Complete StateRef groups are deleted
    without calling the backend (and without consuming a draw).

    mask (batched extension): delete only on the trajectories of the mask.  Only whole groups can be
    deleted per trajectory (that is the failed-purification case, CircuitZoo.jl:156-160); the slots
    stay assigned on the host and the state is marked dead on those lanes."""
    refs = [x for r in refs for x in (_refs(r))]
    mask = _normalize_mask(mask)
    if mask is False:
        return tuple(r.reg for r in refs)
    requested = {(id(r.reg), r.idx) for r in refs}
    candidates = {}
    for r in refs:
        s = r.reg.staterefs[r.idx - 1]
        if s is not None and (nsubsystems(s) > 1 or (mask is not None and mask is not True)):
            candidates.setdefault(id(s), s)
    if mask is not None and mask is not True:
        for s in candidates.values():
            if not all(reg is None or (id(reg), idx) in requested for reg, idx in zip(s.registers, s.registerindices)):
                raise ValueError("a masked traceout must cover every slot of the state (whole-group deletion)")
            s.state.kill(mask)
        return tuple(r.reg for r in refs)
    for s in candidates.values():
        if all(reg is None or (id(reg), idx) in requested for reg, idx in zip(s.registers, s.registerindices)):
            for stateindex in range(len(s.registers), 0, -1):
                if s.registers[stateindex - 1] is None:
                    continue
                removebackref(s, stateindex)
    return tuple(r.reg if r.reg.staterefs[r.idx - 1] is None else _traceout_slot(r.reg, r.idx) for r in refs)


def project_traceout(ref: RegRef, basis, time=None, f=None):
    """project_traceout!(ref, basis; time) -- src/baseops/traceout.jl:143-153."""
    reg, i = ref.reg, ref.idx
    stateref = reg.staterefs[i - 1]
    if stateref is None:
        raise ValueError("Cannot project and trace out an unassigned register slot.")
    if time is not None:
        uptotime(ref, time)
    stateindex = reg.stateindices[i - 1]
    j, stateref.state = stateref.state.project_traceout(stateindex, basis)
    removebackref(stateref, stateindex)
    return f(j) if f is not None else j


def dispose(registers):
    """Drop every state the given registers still hold and release its backend memory at once."""
    seen = set()
    for reg in registers:
        for i, s in enumerate(reg.staterefs):
            if s is not None and id(s) not in seen:
                seen.add(id(s))
                if hasattr(s.state, "free"):
                    s.state.free()
            reg.staterefs[i] = None
            reg.stateindices[i] = 0
