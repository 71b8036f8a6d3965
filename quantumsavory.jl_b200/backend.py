"""CudaBatchRepr / CudaBatchState: the backend generics of QuantumSavory on batched CUDA states.

One `CudaBatchState` stands where the reference keeps one native state (`StateRef.state[]`,
/root/reference/src/states_registers.jl:3-10) but carries N Monte-Carlo trajectories advanced in
lockstep.  Method <-> reference generic (all reached only through the register layer, registers.py):

    newstate(repr)                    newstate(::Qubit, repr)           quantumoptics.jl:66-72, clifford.jl:10-13
    express(spec, repr)               express(state, repr)              src/baseops/initialize.jl:49
    state.copy()                      Base.copy                          states_registers.jl:10
    state.nsubsystems                 nsubsystems(state)
    subsystemcompose(*states)         subsystemcompose(states...)       quantumoptics.jl:3-6, clifford.jl:1
    state.apply(indices, op, mask)    apply!(state, indices, op)        src/representations.jl:4-7
    state.uptotime(idx, bg, dt)       uptotime!(state, idx, bg, dt)     quantumoptics/uptotime.jl, clifford/uptotime.jl
    state.observable(indices, obs)    observable(state, indices, obs)   quantumoptics.jl:14-45, clifford/express.jl:16-52
    state.project_traceout(idx, b)    project_traceout!(state, idx, b)  quantumoptics.jl:47-64, clifford/express.jl:2-14
    state.traceout(idx)               _traceout_state(state, idx)       src/baseops/traceout.jl:46, clifford.jl:7-8

Indices are 1-BASED here, like the reference's `stateindex`; the C-ABI is 0-based.
Measurement results are `BatchOutcome`s: `outcome == 2` is a per-trajectory mask that conditional
`apply`/`traceout` calls consume (the "per-trajectory masks" of BASELINE.json's north_star).  With
ntraj == 1 they convert to the plain `Int` the reference returns.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib, ops
from . import symbols as S
from ._lib import DM, KET, MCKET, TAB, check


class BatchContext:
    """One trajectory shard on one GPU (qsb_ctx).  Owns the per-call draw numbering."""

    def __init__(self, ntraj: int, seed: int = 0x5153, device: int = 0, traj_offset: int = 0, stream: int | None = None):
        self.L = _lib.lib()
        p = C.c_void_p()
        check(self.L.qsb_ctx_create(device, ntraj, traj_offset, seed, C.c_void_p(stream) if stream else None, C.byref(p)))
        self.p = p
        self.ntraj = ntraj
        self.seed = seed
        self.traj_offset = traj_offset
        self.device = device
        self._draw = 0
        self.recording = False
        self.fork = None      # cohorts.Fork while a program of a cohort tree is being recorded (bool(mask) then forks)

    def next_draw(self) -> int:
        d = self._draw
        self._draw += 1
        return d

    def sync(self):
        check(self.L.qsb_ctx_sync(self.p))

    def set_traj_offset(self, traj_offset: int):
        """The next batch of the same schedule: global trajectory ids [traj_offset, traj_offset + ntraj)."""
        check(self.L.qsb_ctx_set_traj_offset(self.p, int(traj_offset)))
        self.traj_offset = int(traj_offset)

    def counters(self):
        c = (C.c_uint64 * 4)()
        check(self.L.qsb_ctx_counters(self.p, c))
        return {"launches": c[0], "algo_bytes": c[1], "ops": c[2], "draws": c[3]}

    def profile(self, enable: bool):
        check(self.L.qsb_ctx_profile(self.p, int(enable)))

    def profile_report(self):
        import json
        buf = C.create_string_buffer(1 << 16)
        check(self.L.qsb_ctx_profile_report(self.p, buf, len(buf)))
        return json.loads(buf.value.decode())

    def record(self):
        """with ctx.record() as rec: ...issue the schedule once...; rec.program.run() replays it fused."""
        return _Recording(self)

    def reset_counters(self):
        check(self.L.qsb_ctx_reset_counters(self.p))

    def close(self):
        """Destroys the context; every state / mask handle still alive is released with it."""
        if self.p:
            self.L.qsb_ctx_destroy(self.p)
            self.p = None

    def on_decision(self, mask, ordinal: int, take: bool, scripted: bool, fork):
        """Called by cohorts.Fork when the host evaluates `if mask:` inside a recording.  The lanes routed to a program
        agree with its script by construction; at an unscripted decision the lanes whose outcome disagrees with the
        branch taken leave the program (qsb_ctx_retire), marked with the ordinal for the tree driver."""
        if scripted:
            return
        from .cohorts import CENSORED
        away = (~mask) if take else mask
        check(self.L.qsb_ctx_retire(self.p, away.p, ordinal + (CENSORED if fork.leaf else 0), C.c_void_p(fork.tree.codes_ptr)))

    def mask(self, fill: bool) -> "BatchMask":
        p = C.c_void_p()
        check(self.L.qsb_mask_create(self.p, int(fill), C.byref(p)))
        return BatchMask(self, p)

    def random_mask(self, prob: float) -> "BatchMask":
        p = C.c_void_p()
        check(self.L.qsb_mask_random(self.p, float(prob), self.next_draw(), C.byref(p)))
        return BatchMask(self, p)


class Program:
    """A recorded op list (qsb_program): one fused launch per run()."""

    def __init__(self, ctx, p):
        self.ctx, self.p = ctx, p

    def run(self, traj_shift: int = 0, draw_base: int = 0):
        check(self.ctx.L.qsb_program_run(self.p, traj_shift, draw_base))

    def run_lanes(self, lanes_dev: int | None, count: int, traj_shift: int = 0, draw_base: int = 0, keep_stats: bool = False):
        """Runs the program on a subset of the ensemble: lanes_dev = device pointer to `count` uint32 lane indices
        (None: every lane).  keep_stats accumulates into the statistics rows instead of zeroing them."""
        check(self.ctx.L.qsb_program_run_lanes(self.p, C.c_void_p(lanes_dev) if lanes_dev else None, int(count), traj_shift, draw_base,
                                               _lib.RUN_KEEP_STATS if keep_stats else 0))

    def upload(self) -> int:
        """Re-sends the host op list to the device (async H2D); returns the bytes copied."""
        b = C.c_int64()
        check(self.ctx.L.qsb_program_upload(self.p, C.byref(b)))
        return b.value

    def info(self):
        a = (C.c_int64 * 8)()
        check(self.ctx.L.qsb_program_info(self.p, a))
        d = dict(zip(("ops", "planes", "stats_rows", "threads", "smem_bytes", "global_arena", "state_ops", "draws"), list(a)))
        b = C.c_uint64()
        check(self.ctx.L.qsb_program_table_bytes(self.p, C.byref(b)))
        d["table_bytes"] = b.value
        return d

    def export(self):
        """Raw records of a tableau program (qsb_program_export): {'ops','specs','obs','aux','stats','meta'} as bytes / int arrays."""
        out = {}
        for what, name in enumerate(("ops", "specs", "obs", "aux", "stats", "meta")):
            need = C.c_size_t(0)
            check(self.ctx.L.qsb_program_export(self.p, what, None, 0, C.byref(need)))
            buf = (C.c_uint8 * max(1, need.value))()
            check(self.ctx.L.qsb_program_export(self.p, what, buf, need.value, C.byref(need)))
            out[name] = bytes(buf[: need.value])
        out["meta"] = np.frombuffer(out["meta"], np.int32)
        out["stats"] = np.frombuffer(out["stats"], np.uint64)
        return out

    def free(self):
        if self.p and self.ctx.p:
            self.ctx.L.qsb_program_free(self.p)
        self.p = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


class _Recording:
    def __init__(self, ctx):
        self.ctx, self.program = ctx, None

    def __enter__(self):
        check(self.ctx.L.qsb_record_begin(self.ctx.p))
        self.ctx.recording = True
        return self

    def __exit__(self, et, ev, tb):
        p = C.c_void_p()
        self.ctx.recording = False
        rc = self.ctx.L.qsb_record_end(self.ctx.p, C.byref(p))
        if et is None:
            check(rc)
            self.program = Program(self.ctx, p)
        elif rc == 0:
            self.ctx.L.qsb_program_free(p)
        return False


class BatchMask:
    """One bit per trajectory."""

    def __init__(self, ctx: BatchContext, p):
        self.ctx, self.p = ctx, p

    def __del__(self):
        try:
            if self.p and self.ctx.p:
                self.ctx.L.qsb_mask_free(self.p)
        except Exception:
            pass
        self.p = None

    def _logic(self, other, op):
        out = self.ctx.mask(False)
        check(self.ctx.L.qsb_mask_logic(out.p, self.p, other.p if other is not None else None, op))
        return out

    def __and__(self, o): return self._logic(o, _lib.MASK_AND)
    def __or__(self, o): return self._logic(o, _lib.MASK_OR)
    def __xor__(self, o): return self._logic(o, _lib.MASK_XOR)
    def __invert__(self): return self._logic(None, _lib.MASK_NOT)
    def andnot(self, o): return self._logic(o, _lib.MASK_ANDNOT)

    def count(self) -> int:
        n = C.c_int64()
        check(self.ctx.L.qsb_mask_count(self.p, C.byref(n)))
        return n.value

    def numpy(self) -> np.ndarray:
        nw = (self.ctx.ntraj + 31) // 32
        w = np.zeros(nw, np.uint32)
        check(self.ctx.L.qsb_mask_to_host(self.p, w.ctypes.data_as(C.POINTER(C.c_uint32))))
        bits = np.unpackbits(w.view(np.uint8), bitorder="little")[: self.ctx.ntraj]
        return bits.astype(bool)

    def __bool__(self):
        """`if mask:` -- one trajectory: its bit.  While a cohort tree records a program (ctx.fork): the branch this
        program's classical history takes; the lanes that disagree are retired to the program of the other branch."""
        if self.ctx.ntraj == 1 and self.ctx.fork is None:
            return bool(self.count())
        if self.ctx.fork is not None:
            return self.ctx.fork.decide(self)
        raise TypeError("a BatchMask over more than one trajectory has no single truth value; pass it as mask= "
                        "(or run the schedule under cohorts.CohortTree, where `if mask:` forks the ensemble)")


class BatchOutcome:
    """Result of project_traceout! for all trajectories: 1-based outcomes in {1, 2}."""

    def __init__(self, ctx, is2: BatchMask, alive: BatchMask, was_random: BatchMask | None = None):
        self.ctx, self.is2, self.alive, self.was_random = ctx, is2, alive, was_random

    def __eq__(self, other):
        if isinstance(other, BatchOutcome):
            return ~(self.is2 ^ other.is2) & (self.alive & other.alive)
        if other == 2:
            return self.is2
        if other == 1:
            return self.alive.andnot(self.is2)
        return self.ctx.mask(False)

    def __ne__(self, other):
        if isinstance(other, BatchOutcome):
            return (self.is2 ^ other.is2) & (self.alive & other.alive)
        return self.alive.andnot(self == other)

    __hash__ = None

    def __int__(self):
        if self.ctx.ntraj != 1:
            raise TypeError("a BatchOutcome over more than one trajectory is not a single Int")
        return 2 if self.is2.count() else 1

    __index__ = __int__

    def numpy(self) -> np.ndarray:
        return 1 + self.is2.numpy().astype(np.int64)

    def counts(self):
        n2 = self.is2.count()
        na = self.alive.count()
        return {"alive": na, 1: na - n2, 2: n2}


class BatchValue:
    """observable() over all trajectories: ensemble statistics (+ per-trajectory values on request)."""

    def __init__(self, ctx, stats, per_traj=None):
        self.ctx = ctx
        self.n_alive, self.sum, self.sumsq, self.sum_im = (float(x) for x in stats)
        self._per_traj = per_traj

    @property
    def mean(self):
        return self.sum / self.n_alive if self.n_alive else float("nan")

    @property
    def mean_complex(self):
        return complex(self.sum, self.sum_im) / self.n_alive if self.n_alive else complex("nan")

    @property
    def var(self):
        return self.sumsq / self.n_alive - self.mean**2 if self.n_alive else float("nan")

    def __float__(self):
        return self.mean

    def __complex__(self):
        return self.mean_complex

    def values(self) -> np.ndarray:
        if self._per_traj is None:
            raise ValueError("per-trajectory values were not requested")
        return self._per_traj


def _ri(a) -> np.ndarray:
    return np.ascontiguousarray(np.asarray(a, dtype=np.complex128)).view(np.float64).reshape(-1)


def _colmajor_ri(m) -> np.ndarray:
    return _ri(np.asarray(m, dtype=np.complex128).T.copy())


def _dp(a):
    return a.ctypes.data_as(C.POINTER(C.c_double))


def _idx0(indices):
    return (C.c_int * len(indices))(*[int(i) - 1 for i in indices])


class CudaBatchRepr:
    """The representation object that selects this backend (`reprs` of `Register(traits, reprs, bgs)`,
    states_registers.jl:32-52).  formalism: "qo" (QuantumOpticsRepr), "mc" (QuantumMCRepr), "clifford"."""

    def __init__(self, ctx: BatchContext, formalism: str = "qo"):
        if formalism not in ("qo", "mc", "clifford"):
            raise ValueError(formalism)
        self.ctx, self.formalism = ctx, formalism

    def __eq__(self, o):
        return isinstance(o, CudaBatchRepr) and o.ctx is self.ctx and o.formalism == self.formalism

    def __hash__(self):
        return hash((id(self.ctx), self.formalism))

    # newstate(::Qubit, repr)
    def newstate(self):
        kind = {"qo": KET, "mc": MCKET, "clifford": TAB}[self.formalism]
        p = C.c_void_p()
        check(self.ctx.L.qsb_state_new(self.ctx.p, kind, 1, C.byref(p)))
        return CudaBatchState(self.ctx, p)

    # express(state, repr): lower a host-side state description to a native batched state
    def express(self, spec):
        ctx = self.ctx
        if isinstance(spec, CudaBatchState):
            return spec
        if self.formalism == "clifford":
            if isinstance(spec, ops.StabilizerState) or isinstance(spec, ops.MixedStabilizer):
                return CudaBatchState.from_generators(ctx, spec.n, spec.xs, spec.zs, spec.signs)
            if isinstance(spec, ops.Mixture):
                terms = [self.express(s) for _, s in spec.terms]
                w = np.array([w for w, _ in spec.terms], dtype=np.float64)
                arr = (C.c_void_p * len(terms))(*[t.p for t in terms])
                p = C.c_void_p()
                check(ctx.L.qsb_state_select(ctx.p, arr, _dp(w), len(terms), ctx.next_draw(), C.byref(p)))
                return CudaBatchState(ctx, p)
            raise _lib.MethodError(_lib.QSB_ERR_UNSUPPORTED, f"express({type(spec).__name__}, CliffordRepr) is not defined")
        ketkind = KET if self.formalism == "qo" else MCKET
        if isinstance(spec, ops.StabilizerState):
            return CudaBatchState.from_array(ctx, spec.ket, ketkind)
        if isinstance(spec, (ops.Mixture, ops.MixedStabilizer)):
            return CudaBatchState.from_array(ctx, spec.dm, DM)
        a = np.asarray(spec, dtype=np.complex128)
        return CudaBatchState.from_array(ctx, a, ketkind if a.ndim == 1 else DM)


class CudaBatchState:
    """Opaque handle to N trajectories' copies of one native state (qsb_state)."""

    ispadded = False  # ispadded(state) = false: removal physically drops the subsystem (traceout.jl:1-22)

    def __init__(self, ctx: BatchContext, p):
        self.ctx, self.p = ctx, p

    def free(self):
        """Release the device memory now (the Julia finalizer's job; explicit here because register
        <-> StateRef reference cycles would otherwise wait for Python's cycle collector)."""
        if self.p and self.ctx.p:
            self.ctx.L.qsb_state_free(self.p)
        self.p = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass

    # -- construction ----------------------------------------------------------------------------
    @classmethod
    def from_array(cls, ctx, a, kind):
        a = np.asarray(a, dtype=np.complex128)
        n = int(round(np.log2(a.shape[0])))
        data = _ri(a) if a.ndim == 1 else _colmajor_ri(a)
        p = C.c_void_p()
        check(ctx.L.qsb_state_from_host(ctx.p, kind, n, _dp(data), C.byref(p)))
        return cls(ctx, p)

    @classmethod
    def from_generators(cls, ctx, n, xs, zs, signs):
        xs = np.ascontiguousarray(xs, dtype=np.uint8); zs = np.ascontiguousarray(zs, dtype=np.uint8)
        signs = np.ascontiguousarray(signs, dtype=np.uint8)
        u8 = lambda a: a.ctypes.data_as(C.POINTER(C.c_uint8))
        p = C.c_void_p()
        check(ctx.L.qsb_tab_from_host(ctx.p, int(n), int(xs.shape[0]), u8(xs), u8(zs), u8(signs), C.byref(p)))
        return cls(ctx, p)

    @classmethod
    def zeros(cls, ctx, kind, n):
        p = C.c_void_p()
        check(ctx.L.qsb_state_new(ctx.p, kind, n, C.byref(p)))
        return cls(ctx, p)

    # -- trait queries -----------------------------------------------------------------------------
    @property
    def kind(self):
        return self.ctx.L.qsb_state_kind(self.p)

    @property
    def nsubsystems(self):
        return self.ctx.L.qsb_state_nsubsystems(self.p)

    @property
    def default_repr(self):
        return CudaBatchRepr(self.ctx, {KET: "qo", DM: "qo", MCKET: "mc", TAB: "clifford"}[self.kind])

    def copy(self):
        p = C.c_void_p()
        check(self.ctx.L.qsb_state_copy(self.p, C.byref(p)))
        return CudaBatchState(self.ctx, p)

    def alive(self) -> BatchMask:
        p = C.c_void_p()
        check(self.ctx.L.qsb_state_alive(self.p, C.byref(p)))
        return BatchMask(self.ctx, p)

    def kill(self, lanes: BatchMask):
        """Masked whole-group traceout!: the state stops existing on `lanes` only."""
        check(self.ctx.L.qsb_state_kill(self.p, lanes.p))
        return self

    # -- read-back (tests) -------------------------------------------------------------------------
    def trajectory(self, t: int) -> np.ndarray:
        n, kind = self.nsubsystems, self.kind
        L = 2**n if kind in (KET, MCKET) else 4**n
        out = np.empty(2 * L)
        check(self.ctx.L.qsb_state_to_host(self.p, t, _dp(out)))
        v = out.view(np.complex128)
        return v.copy() if kind in (KET, MCKET) else v.reshape(2**n, 2**n).T.copy()

    def tableau(self, t: int):
        n = self.nsubsystems
        xs = np.zeros((2 * n, n), np.uint8); zs = np.zeros_like(xs); sg = np.zeros(2 * n, np.uint8); ac = np.zeros(n, np.uint8)
        u8 = lambda a: a.ctypes.data_as(C.POINTER(C.c_uint8))
        check(self.ctx.L.qsb_tab_to_host(self.p, t, u8(xs), u8(zs), u8(sg), u8(ac)))
        return xs, zs, sg, ac

    # -- backend generics ----------------------------------------------------------------------------
    def subsystemcompose(self, *others):
        states = (self,) + others
        arr = (C.c_void_p * len(states))(*[s.p for s in states])
        p = C.c_void_p()
        check(self.ctx.L.qsb_tensor(self.ctx.p, arr, len(states), C.byref(p)))
        return CudaBatchState(self.ctx, p)

    def apply(self, indices, op, mask: BatchMask | None = None):
        m = mask.p if mask is not None else None
        if self.kind == TAB:
            if not isinstance(op, ops.Gate):
                raise _lib.MethodError(_lib.QSB_ERR_UNSUPPORTED, "apply! of a dense operator on a Clifford state")
            try:
                table = op.table
            except ValueError as e:
                raise _lib.MethodError(_lib.QSB_ERR_UNSUPPORTED, str(e))
            check(self.ctx.L.qsb_apply_clifford(self.p, _idx0(indices), len(indices), table.ctypes.data_as(C.POINTER(C.c_uint8)), m))
        else:
            mat = op.mat if isinstance(op, ops.Gate) else np.asarray(op, dtype=np.complex128)
            if mat.shape[0] != 2 ** len(indices):
                raise _lib.ArgumentError(_lib.QSB_ERR_BAD_ARG, "operator size does not match the number of indices")
            check(self.ctx.L.qsb_apply(self.p, _idx0(indices), len(indices), _dp(_colmajor_ri(mat)), m))
        return self

    def _evolve(self, indices, H, lind, duration):
        """lind: [(position in `indices`, 2x2 collapse operator)] -> qsb_evolve."""
        k = len(indices)
        Hri = None if H is None else _colmajor_ri(H)
        pos = (C.c_int * max(1, len(lind)))(*[p for p, _ in lind])
        Lri = np.concatenate([_colmajor_ri(L) for _, L in lind]) if lind else None
        draw = self.ctx.next_draw() if (self.kind == MCKET and lind) else 0
        check(self.ctx.L.qsb_evolve(self.p, _idx0(indices), k, None if Hri is None else _dp(Hri), len(lind),
                                    pos if lind else None, None if Lri is None else _dp(Lri), float(duration), draw))
        return self

    def apply_noninstant(self, indices, operation, backgrounds):
        """apply_noninstant!(state, state_indices, ConstantHamiltonianEvolution, backgrounds) --
        src/backends/quantumoptics/noninstant.jl:1-37 (collapse operators: _embedded_lindblad_operators, uptotime.jl:6-18)."""
        if self.kind == TAB:
            raise _lib.MethodError(_lib.QSB_ERR_UNSUPPORTED, "apply_noninstant! has no method for a Clifford state")
        H = operation.hamiltonian
        H = H.mat if isinstance(H, ops.Gate) else np.asarray(H, dtype=np.complex128)
        if H.shape[0] != 2 ** len(indices):
            raise _lib.ArgumentError(_lib.QSB_ERR_BAD_ARG, "Hamiltonian size does not match the number of indices")
        lind = [(j, L) for j, bg in enumerate(backgrounds) if bg is not None for L in S.lindbladop(bg)]
        return self._evolve(indices, H, lind, operation.duration)

    def uptotime(self, idx, background, dt):
        if isinstance(background, S.LindbladOnly):   # krausops === nothing: uptotime.jl:24-29 (mcwf), :51-54 (master)
            if self.kind == TAB:
                raise _lib.MethodError(_lib.QSB_ERR_UNSUPPORTED, "no Lindblad fallback for a Clifford state")
            return self._evolve([idx], None, [(0, L) for L in S.lindbladop(background)], dt)
        prm = (C.c_double * 2)(*background.params)
        check(self.ctx.L.qsb_channel(self.p, int(idx) - 1, background.kind, prm, float(dt), self.ctx.next_draw()))
        return self

    def project_traceout(self, idx, basis):
        """-> (BatchOutcome, self or None).  basis: ops.X/Y/Z (sigma^x..) or a pair of kets."""
        ctx = self.ctx
        alive = self.alive()
        o, r = C.c_void_p(), C.c_void_p()
        if isinstance(basis, ops.Gate) and basis.pauli in ("X", "Y", "Z"):
            code, kets = {"X": 1, "Y": 2, "Z": 3}[basis.pauli], None
        else:
            code = _lib.BASIS_KETS
            kets = _dp(_ri(np.stack([np.asarray(b, dtype=np.complex128) for b in basis])))
        want_r = self.kind == TAB
        check(ctx.L.qsb_measure(self.p, int(idx) - 1, code, kets, ctx.next_draw(), C.byref(o), C.byref(r) if want_r else None, None))
        out = BatchOutcome(ctx, BatchMask(ctx, o), alive, BatchMask(ctx, r) if want_r and r else None)
        return out, (self if self.nsubsystems > 0 else None)

    def traceout(self, idx):
        check(self.ctx.L.qsb_ptrace(self.p, int(idx) - 1, self.ctx.next_draw() if self.kind == MCKET else 0))
        return self

    def observable(self, indices, obs, per_trajectory: bool = False, stats_dev: int | None = None, sync: bool = True,
                   per_traj_dev: int | None = None):
        """-> BatchValue.  stats_dev: optional device pointer (double[4]) that receives the statistics
        without a host sync (sync=False), e.g. a torch tensor that is NCCL-all-reduced later."""
        ctx = self.ctx
        if ctx.fork is not None and not ctx.fork.tally:
            return BatchValue(ctx, np.zeros(4))     # prefix of a cohort program: the parent program tallied this observable
        k = len(indices)
        sign = 0
        keep = None
        if isinstance(obs, ops.Gate) and obs.pauli is not None:
            codes = np.array([{"I": 0, "X": 1, "Y": 2, "Z": 3}[c] for c in obs.pauli], np.uint8)
            kind, ptr, keep = _lib.OBS_PAULI, codes.ctypes.data_as(C.c_void_p), codes
        elif isinstance(obs, ops.SProjector):
            if self.kind == TAB:
                st = obs.state
                if not isinstance(st, ops.StabilizerState):
                    raise _lib.MethodError(_lib.QSB_ERR_UNSUPPORTED, "projector on a non-stabilizer ket for a Clifford state")
                buf = np.concatenate([st.xs.reshape(-1), st.zs.reshape(-1), st.signs.reshape(-1)]).astype(np.uint8)
                kind, ptr, keep = _lib.OBS_PROJ_STAB, buf.ctypes.data_as(C.c_void_p), buf
            else:
                buf = _ri(obs.ket)
                kind, ptr, keep = _lib.OBS_PROJ_KET, buf.ctypes.data_as(C.c_void_p), buf
        elif self.kind == TAB:
            return self._tab_dense_observable(indices, obs.mat if isinstance(obs, ops.Gate) else np.asarray(obs, dtype=np.complex128), stats_dev)
        else:
            mat = obs.mat if isinstance(obs, ops.Gate) else np.asarray(obs, dtype=np.complex128)
            buf = _colmajor_ri(mat)
            kind, ptr, keep = _lib.OBS_OP, buf.ctypes.data_as(C.c_void_p), buf
        per_dev = C.c_void_p(per_traj_dev) if per_traj_dev else None
        per_t = None
        if per_trajectory:
            import torch
            npad = (ctx.ntraj + 31) // 32 * 32
            per_t = torch.empty(2 * npad, dtype=torch.float64, device=f"cuda:{ctx.device}")
            per_dev = C.c_void_p(per_t.data_ptr())
        stats = np.zeros(4)
        check(ctx.L.qsb_expect(self.p, _idx0(indices), k, kind, ptr, sign, per_dev,
                               C.c_void_p(stats_dev) if stats_dev else None, _dp(stats) if sync else None))
        del keep
        vals = None
        if per_trajectory:
            ctx.sync()
            vals = per_t.cpu().numpy().view(np.complex128)[: ctx.ntraj]
        return BatchValue(ctx, stats, vals)


def _pauli_decomposition(mat, k):
    """O = sum_P c_P P over the 4^k Pauli strings (factor 0 least significant); returns [(codes, c_P)] with c_P != 0."""
    import itertools
    out = []
    for codes in itertools.product(range(4), repeat=k):
        P = S.tensor(*[(S.Id, S.X, S.Y, S.Z)[c] for c in codes])
        c = np.trace(P @ mat) / 2 ** k
        if abs(c) > 1e-15:
            out.append((np.array(codes, np.uint8), complex(c)))
    return out


_WARNED_DENSE_ON_TAB = [False]


def _tab_dense_observable(self, indices, mat, stats_dev):
    """observable(::MixedDestabilizer, indices, ::Operator) -- src/backends/clifford/express.jl:22-46.  The reference converts the
    tableau to a ket (pure states only, one `:stabilizer_to_ket` warning); here the k-qubit operator is expanded in the Pauli basis
    and every term is one exact tableau expectation, so nothing dense is built and mixed states work too."""
    import warnings
    import torch
    ctx = self.ctx
    if stats_dev:
        raise _lib.MethodError(_lib.QSB_ERR_UNSUPPORTED, "a dense observable on a Clifford state is host-visible (no stats_dev / recording)")
    k = len(indices)
    if mat.shape != (2 ** k, 2 ** k):
        raise _lib.ArgumentError(_lib.QSB_ERR_BAD_ARG, "operator size does not match the number of indices")
    if not _WARNED_DENSE_ON_TAB[0]:
        _WARNED_DENSE_ON_TAB[0] = True
        warnings.warn("dense observable on a Clifford stabilizer state: evaluated as %d Pauli expectations (event :stabilizer_to_ket of the reference)" % 4 ** k)
    npad = (ctx.ntraj + 31) // 32 * 32
    buf = torch.empty(2 * npad, dtype=torch.float64, device=f"cuda:{ctx.device}")
    torch.cuda.current_stream(ctx.device).synchronize()   # the block may carry pending work of torch's stream; the library writes it on its own
    acc = np.zeros(npad, np.complex128)
    stats = np.zeros(4)
    n_alive = 0.0
    for codes, c in _pauli_decomposition(mat, k):
        check(ctx.L.qsb_expect(self.p, _idx0(indices), k, _lib.OBS_PAULI, codes.ctypes.data_as(C.c_void_p), 0, C.c_void_p(buf.data_ptr()), None, _dp(stats)))
        n_alive = stats[0]
        # qsb_expect has synchronised (stats_host): the row is complete.  Accumulate on the host from a blocking copy, so that the next
        # term's write of `buf` (context stream) cannot overtake a read still queued on torch's stream.  Dead lanes hold 0.
        acc += c * buf.cpu().numpy()[0::2]
    vals = acc[: ctx.ntraj]
    out = np.array([n_alive, vals.real.sum(), (vals.real ** 2).sum(), vals.imag.sum()])
    return BatchValue(ctx, out, vals)


CudaBatchState._tab_dense_observable = _tab_dense_observable


def subsystemcompose(*states):
    return states[0].subsystemcompose(*states[1:])
