"""Tracing backend + trace-tree driver for the compiled CPU trajectory loop (oracle/qsb_replay.c).

TEST / BENCH INFRASTRUCTURE ONLY -- never imported by the product package.

`TraceRepr` has the method surface of tests/oracle_backend.OracleRepr / qsb200.backend.CudaBatchRepr, but instead of
computing it writes the backend calls the host schedule makes (registers.py, circuits.py, workloads/) into an op list.
`TraceTree` keeps one op list per classical history (script of `if mask:` decisions, exactly like
qsb200.cohorts.CohortTree keeps one recorded program per history) and hands the tree to qsr_run, which replays it for
every trajectory through the oracle's own C functions with OpenMP over trajectories; a trajectory that reaches a
history nobody traced yet is reported back, its history is traced, and it runs again.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

(OP_NEW, OP_FROM_DATA, OP_FROM_GENERATORS, OP_SELECT, OP_COPY, OP_TENSOR, OP_APPLY, OP_APPLY_CLIFFORD, OP_CHANNEL_DM, OP_CHANNEL_MC,
 OP_CHANNEL_TAB, OP_MEASURE_DENSE, OP_MEASURE_TAB, OP_PTRACE, OP_EXPECT_OP, OP_EXPECT_PROJ_KET, OP_EXPECT_PAULI, OP_EXPECT_PROJ_STAB,
 OP_MASK_CONST, OP_MASK_LOGIC, OP_MASK_RANDOM, OP_DECIDE, OP_FREE) = range(23)
M_AND, M_OR, M_XOR, M_ANDNOT, M_NOT, M_XNOR = range(6)

OP_DTYPE = np.dtype([("type", "i4"), ("n", "i4"), ("k", "i4"), ("nsrc", "i4"), ("dst", "i4"), ("src", "i4", 5), ("idx", "i4", 6),
                     ("mask", "i4", 3), ("iv", "i4", 4), ("tab", "i4"), ("row", "i4"), ("draw", "u8"), ("dpay", "i8"), ("bpay", "i8")], align=True)


class _Node(C.Structure):
    _fields_ = [("ops", C.c_void_p), ("nops", C.c_int64), ("dpay", C.c_void_p), ("bpay", C.c_void_p),
                ("nstates", C.c_int32), ("nmasks", C.c_int32), ("leaf", C.c_int32), ("child", C.c_void_p), ("nchild", C.c_int32)]


def build(force: bool = False) -> str:
    so = os.path.join(_HERE, "libqsb_replay.so")
    srcs = [os.path.join(_HERE, f) for f in ("qsb_replay.c", "qsb_oracle.c")]
    if force or not os.path.exists(so) or any(os.path.getmtime(so) < os.path.getmtime(s) for s in srcs):
        subprocess.check_call(["make", "-C", _HERE, "-s", "libqsb_replay.so"] + (["-B"] if force else []))
    return so


def lib():
    global _LIB
    if _LIB is None:
        L = C.CDLL(build())
        L.qsr_run.restype = C.c_int64
        L.qsr_run.argtypes = [C.c_void_p, C.c_int, C.c_uint64, C.c_int64, C.c_int64, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p,
                              C.c_void_p, C.c_void_p]
        assert L.qsr_sizeof_op() == OP_DTYPE.itemsize and L.qsr_sizeof_node() == C.sizeof(_Node), (L.qsr_sizeof_op(), OP_DTYPE.itemsize)
        _LIB = L
    return _LIB


# ------------------------------------------------------------------------------------------------
# the tracing backend
# ------------------------------------------------------------------------------------------------
class TraceContext:
    """Stands where a BatchContext / OracleContext stands; ntraj = 2 keeps outcomes symbolic (masks, not bools)."""

    ntraj = 2

    def __init__(self):
        self.fork = None
        self.reset()

    def reset(self):
        self._draw = 0
        self.ops, self.dpay, self.bpay = [], [], bytearray()
        self.nstates = self.nmasks = 0

    def next_draw(self):
        d = self._draw
        self._draw += 1
        return d

    def emit(self, type_, **kw):
        self.ops.append((type_, kw))

    def d(self, values) -> int:
        off = len(self.dpay)
        self.dpay.extend(float(v) for v in values)
        return off

    def b(self, values) -> int:
        off = len(self.bpay)
        self.bpay.extend(bytes(int(v) & 0xFF for v in values))
        return off

    def new_state(self):
        self.nstates += 1
        return self.nstates - 1

    def new_mask(self):
        self.nmasks += 1
        return TraceMask(self, self.nmasks - 1)

    def mask(self, fill: bool):
        m = self.new_mask()
        self.emit(OP_MASK_CONST, mask=(m.id, -1, -1), iv=(int(bool(fill)),))
        return m

    def random_mask(self, prob: float):
        m = self.new_mask()
        self.emit(OP_MASK_RANDOM, mask=(m.id, -1, -1), draw=self.next_draw(), dpay=self.d([prob]))
        return m

    def on_decision(self, mask, ordinal, take, scripted, fork):
        self.emit(OP_DECIDE, mask=(-1, mask.id, -1), iv=(ordinal, int(take)))

    def pack(self):
        a = np.zeros(len(self.ops), OP_DTYPE)
        for f in ("src", "idx", "mask"):
            a[f] = -1
        a["dst"] = a["row"] = a["dpay"] = a["bpay"] = -1
        for i, (t, kw) in enumerate(self.ops):
            r = a[i]
            r["type"] = t
            for k, v in kw.items():
                if k in ("src", "idx", "mask", "iv"):
                    r[k][: len(v)] = v
                else:
                    r[k] = v
        return a, np.asarray(self.dpay + [0.0], np.float64), np.frombuffer(bytes(self.bpay) + b"\0", np.uint8).copy()


def _ri(a) -> np.ndarray:
    return np.ascontiguousarray(np.asarray(a, dtype=np.complex128)).view(np.float64).reshape(-1)


class TraceMask:
    def __init__(self, ctx, id_):
        self.ctx, self.id = ctx, id_

    def _logic(self, other, op):
        out = self.ctx.new_mask()
        self.ctx.emit(OP_MASK_LOGIC, mask=(out.id, self.id, other.id if other is not None else -1), iv=(op,))
        return out

    def __and__(self, o): return self._logic(o, M_AND)
    def __or__(self, o): return self._logic(o, M_OR)
    def __xor__(self, o): return self._logic(o, M_XOR)
    def __invert__(self): return self._logic(None, M_NOT)
    def andnot(self, o): return self._logic(o, M_ANDNOT)

    def __bool__(self):
        if self.ctx.fork is None:
            raise TypeError("`if mask:` needs a trace tree (a script of decisions)")
        return self.ctx.fork.decide(self)


class TraceOutcome:
    """project_traceout! result: 1-based outcome, kept symbolic."""

    def __init__(self, ctx, is2):
        self.ctx, self.is2 = ctx, is2

    def __eq__(self, other):
        if isinstance(other, TraceOutcome):
            return self.is2._logic(other.is2, M_XNOR)
        if other == 2:
            return self.is2
        if other == 1:
            return ~self.is2
        return self.ctx.mask(False)

    def __ne__(self, other):
        if isinstance(other, TraceOutcome):
            return self.is2 ^ other.is2
        return ~(self == other)

    __hash__ = None


class TraceValue:
    n_alive = 1

    def __float__(self):
        return float("nan")

    def __complex__(self):
        return complex("nan")


class TraceRepr:
    def __init__(self, ctx: TraceContext, formalism="qo"):
        self.ctx, self.formalism = ctx, formalism

    def __eq__(self, o):
        return isinstance(o, TraceRepr) and o.ctx is self.ctx and o.formalism == self.formalism

    def __hash__(self):
        return hash((id(self.ctx), self.formalism))

    def newstate(self):
        ctx = self.ctx
        tab = self.formalism == "clifford"
        s = TraceState(ctx, ctx.new_state(), 1, "tab" if tab else "ket", "tab" if tab else ("mc" if self.formalism == "mc" else "qo"))
        ctx.emit(OP_NEW, dst=s.id, tab=int(tab))
        return s

    def _from_generators(self, spec):
        ctx = self.ctx
        s = TraceState(ctx, ctx.new_state(), spec.n, "tab", "tab")
        ngen = int(np.asarray(spec.signs).shape[0])
        pay = list(np.asarray(spec.xs, np.uint8).reshape(-1)) + list(np.asarray(spec.zs, np.uint8).reshape(-1)) + list(np.asarray(spec.signs, np.uint8).reshape(-1))
        ctx.emit(OP_FROM_GENERATORS, dst=s.id, n=spec.n, k=ngen, tab=1, bpay=ctx.b(pay))
        return s

    def _from_array(self, a, flavor):
        ctx = self.ctx
        a = np.asarray(a, dtype=np.complex128)
        n = int(round(np.log2(a.shape[0])))
        kind = "ket" if a.ndim == 1 else "dm"
        s = TraceState(ctx, ctx.new_state(), n, kind, flavor if kind == "ket" else "qo")
        data = _ri(a) if a.ndim == 1 else _ri(np.asfortranarray(a).T.copy())
        ctx.emit(OP_FROM_DATA, dst=s.id, n=n, iv=(0 if kind == "ket" else 1,), dpay=ctx.d(data))
        return s

    def express(self, spec):
        from qsb200 import ops
        ctx = self.ctx
        if isinstance(spec, TraceState):
            return spec
        if self.formalism == "clifford":
            if isinstance(spec, (ops.StabilizerState, ops.MixedStabilizer)):
                return self._from_generators(spec)
            if isinstance(spec, ops.Mixture):
                terms = [self.express(s) for _, s in spec.terms]
                w = np.array([w for w, _ in spec.terms]); cum = np.cumsum(w / w.sum())
                out = TraceState(ctx, ctx.new_state(), terms[0].n, "tab", "tab")
                ctx.emit(OP_SELECT, dst=out.id, src=[t.id for t in terms], nsrc=len(terms), draw=ctx.next_draw(), dpay=ctx.d(cum), tab=1)
                for t in terms:
                    t.freed = True            # consumed by the SELECT
                return out
            raise TypeError("express to CliffordRepr not defined")
        flavor = "mc" if self.formalism == "mc" else "qo"
        if isinstance(spec, ops.StabilizerState):
            return self._from_array(spec.ket, flavor)
        if isinstance(spec, (ops.Mixture, ops.MixedStabilizer)):
            return self._from_array(spec.dm, "qo")
        return self._from_array(np.asarray(spec, dtype=np.complex128), flavor)


class TraceState:
    ispadded = False

    def __init__(self, ctx, id_, n, kind, flavor):
        self.ctx, self.id, self.n, self.kind, self.flavor, self.freed = ctx, id_, n, kind, flavor, False

    @property
    def nsubsystems(self):
        return self.n

    @property
    def tab(self):
        return self.kind == "tab"

    @property
    def is_mc(self):
        return self.flavor == "mc" and self.kind == "ket"

    def free(self):
        if not self.freed:
            self.freed = True
            self.ctx.emit(OP_FREE, dst=self.id)

    def copy(self):
        c = TraceState(self.ctx, self.ctx.new_state(), self.n, self.kind, self.flavor)
        self.ctx.emit(OP_COPY, dst=c.id, src=[self.id], tab=int(self.tab))
        return c

    def kill(self, lanes):
        raise NotImplementedError("per-lane deletion is the lockstep-masks extension; a traced trajectory follows its own control flow")

    def subsystemcompose(self, *others):
        ctx = self.ctx
        acc = self
        for o in others:
            if acc.tab != o.tab:
                raise TypeError("subsystemcompose of a tableau and a dense state")
            kind = "tab" if acc.tab else ("ket" if acc.kind == "ket" and o.kind == "ket" else "dm")
            flavor = acc.flavor
            if flavor == "mc" and not (o.flavor == "mc" and o.kind == "ket"):
                flavor = "qo"
            if flavor == "mc" and kind != "ket":
                flavor = "qo"
            new = TraceState(ctx, ctx.new_state(), acc.n + o.n, kind, flavor)
            ctx.emit(OP_TENSOR, dst=new.id, src=[acc.id, o.id], tab=int(acc.tab))
            if acc is not self:
                acc.free()
            acc = new
        return acc

    def apply(self, indices, op, mask=None):
        from qsb200 import ops
        ctx = self.ctx
        idx = [i - 1 for i in indices]
        m = (-1, mask.id, -1) if mask is not None else (-1, -1, -1)
        if self.tab:
            if not isinstance(op, ops.Gate):
                raise TypeError("dense operator on a tableau")
            ctx.emit(OP_APPLY_CLIFFORD, dst=self.id, idx=idx, k=len(idx), mask=m, bpay=ctx.b(np.asarray(op.table, np.uint8).reshape(-1)), tab=1)
        else:
            U = op.mat if isinstance(op, ops.Gate) else np.asarray(op)
            ctx.emit(OP_APPLY, dst=self.id, idx=idx, k=len(idx), mask=m, dpay=ctx.d(_ri(np.asarray(U, np.complex128).T.copy())))
        return self

    def uptotime(self, idx, bg, dt):
        ctx = self.ctx
        d = ctx.next_draw()                      # every backend consumes one draw id per uptotime! call (BatchContext does too)
        pay = ctx.d(list(bg.params) + [0.0] * (2 - len(bg.params)) + [dt])
        if self.tab:
            if bg.kind not in (1, 2):
                raise TypeError("no paulinoise method for this background on a Clifford state")
            ctx.emit(OP_CHANNEL_TAB, dst=self.id, idx=[idx - 1], iv=(bg.kind,), draw=d, dpay=pay, tab=1)
        elif self.is_mc:
            ctx.emit(OP_CHANNEL_MC, dst=self.id, idx=[idx - 1], iv=(bg.kind,), draw=d, dpay=pay)
        else:
            ctx.emit(OP_CHANNEL_DM, dst=self.id, idx=[idx - 1], iv=(bg.kind,), dpay=pay)
            self.kind, self.flavor = "dm", "qo"
        return self

    def _measure_dense(self, idx, psis):
        ctx = self.ctx
        m = ctx.new_mask()
        ps = np.stack([np.asarray(p, dtype=np.complex128) for p in psis])
        ctx.emit(OP_MEASURE_DENSE, dst=self.id, idx=[idx - 1], mask=(m.id, -1, -1), draw=ctx.next_draw(), dpay=ctx.d(_ri(ps)))
        self.n -= 1
        return m

    def project_traceout(self, idx, basis):
        from qsb200 import ops
        from qsb200 import symbols as S
        ctx = self.ctx
        if self.tab:
            if not (isinstance(basis, ops.Gate) and basis.pauli in ("X", "Y", "Z")):
                raise TypeError("Clifford measurements support X, Y, Z only")
            m = ctx.new_mask()
            ctx.emit(OP_MEASURE_TAB, dst=self.id, idx=[idx - 1], iv=({"X": 1, "Y": 2, "Z": 3}[basis.pauli],), mask=(m.id, -1, -1), draw=ctx.next_draw(), tab=1)
            self.n -= 1
        else:
            m = self._measure_dense(idx, S.EIGVECS[basis.pauli] if isinstance(basis, ops.Gate) else basis)
        return TraceOutcome(ctx, m), (self if self.n > 0 else None)

    def traceout(self, idx):
        from qsb200 import symbols as S
        if self.n == 1:
            raise ValueError("Partial trace can't be used to trace out all subsystems - use tr() instead.")
        if self.is_mc:
            self._measure_dense(idx, S.EIGVECS["Z"])
            return self
        self.ctx.emit(OP_PTRACE, dst=self.id, idx=[idx - 1], tab=int(self.tab))
        self.n -= 1
        if not self.tab:
            self.kind, self.flavor = "dm", "qo"
        return self

    def observable(self, indices, obs, stats_dev=None, **kw):
        """stats_dev: the statistics ROW (an int) this observable is tallied into."""
        from qsb200 import ops
        from qsb200 import symbols as S
        ctx = self.ctx
        idx = [i - 1 for i in indices]
        row = -1 if stats_dev is None else int(stats_dev)
        common = dict(dst=self.id, idx=idx, k=len(idx), row=row, tab=int(self.tab))
        if self.tab:
            if isinstance(obs, ops.Gate) and obs.pauli is not None:
                ctx.emit(OP_EXPECT_PAULI, bpay=ctx.b([{"I": 0, "X": 1, "Y": 2, "Z": 3}[c] for c in obs.pauli]), **common)
            elif isinstance(obs, ops.SProjector) and isinstance(obs.state, ops.StabilizerState):
                xs, zs, signs = S.stabilizer_generators(obs.state.spec)
                ctx.emit(OP_EXPECT_PROJ_STAB, bpay=ctx.b(list(xs.reshape(-1)) + list(zs.reshape(-1)) + list(signs.reshape(-1))), **common)
            else:
                raise TypeError("dense observable on a tableau")
        elif isinstance(obs, ops.SProjector):
            ctx.emit(OP_EXPECT_PROJ_KET, dpay=ctx.d(_ri(obs.ket)), **common)
        else:
            O = obs.mat if isinstance(obs, ops.Gate) else np.asarray(obs)
            ctx.emit(OP_EXPECT_OP, dpay=ctx.d(_ri(np.asarray(O, np.complex128).T.copy())), **common)
        return TraceValue()


# ------------------------------------------------------------------------------------------------
# the tree of op lists and the compiled replay
# ------------------------------------------------------------------------------------------------
class TraceTree:
    """schedule(make_repr): runs the host schedule once against the tracing backend (make_repr() -> TraceRepr).
    max_depth: forks allowed per trajectory before it is censored (None: unlimited) -- the same rule as
    qsb200.cohorts.CohortTree, so that the two arms do the same work."""

    default = True
    codes_ptr = 0

    def __init__(self, schedule, formalism, nrows, max_depth=None):
        self.schedule, self.formalism, self.nrows, self.max_depth = schedule, formalism, nrows, max_depth
        self.ctx = TraceContext()
        self.nodes = []          # dicts: script, taken, ops/dpay/bpay arrays, children {ordinal: index}, leaf
        self.trace_s = 0.0
        self._record(())
        self._c = None

    def _record(self, script, parent=None, ordinal=None):
        import time
        from qsb200.cohorts import Fork
        t0 = time.perf_counter()
        ctx = self.ctx
        leaf = self.max_depth is not None and sum(1 for b in script if b != self.default) >= self.max_depth
        fork = Fork(self, script, leaf)
        ctx.reset()
        ctx.fork = fork
        try:
            self.schedule(lambda: TraceRepr(ctx, self.formalism))
        finally:
            ctx.fork = None
        ops, dpay, bpay = ctx.pack()
        node = dict(script=tuple(script), taken=tuple(fork.taken), ops=ops, dpay=dpay, bpay=bpay, children={}, leaf=leaf,
                    nstates=ctx.nstates, nmasks=ctx.nmasks, index=len(self.nodes))
        if parent is not None:
            # the child's op list must agree with its parent's up to and including the DECIDE it forks at (the replay
            # continues at the same op index): any non-determinism of the host schedule would show up here
            pos = [i for i in range(len(parent["ops"])) if parent["ops"]["type"][i] == OP_DECIDE and parent["ops"]["iv"][i][0] == ordinal][0]
            a, b = parent["ops"][: pos + 1].copy(), ops[: pos + 1].copy()
            a["iv"][pos] = b["iv"][pos] = 0
            if not all(np.array_equal(a[f], b[f]) for f in OP_DTYPE.names):   # field by field: struct padding is not copied
                raise RuntimeError(f"host schedule is not deterministic: history {script} diverges from its parent before decision {ordinal}")
        self.nodes.append(node)
        self._c = None
        self.trace_s += time.perf_counter() - t0
        return node

    def child(self, node, ordinal):
        c = node["children"].get(ordinal)
        if c is None:
            t = node["taken"]
            c = self._record(t[:ordinal] + (not t[ordinal],), parent=node, ordinal=ordinal)["index"]
            node["children"][ordinal] = c
        return c

    def _carray(self):
        if self._c is None:
            arr = (_Node * len(self.nodes))()
            keep = []
            for i, n in enumerate(self.nodes):
                nchild = len(n["taken"])
                ch = np.full(max(nchild, 1), -1, np.int32)
                for o, c in n["children"].items():
                    ch[o] = c
                keep.append(ch)
                arr[i] = _Node(n["ops"].ctypes.data, len(n["ops"]), n["dpay"].ctypes.data, n["bpay"].ctypes.data, n["nstates"], n["nmasks"],
                               int(n["leaf"]), ch.ctypes.data, nchild)
            self._c = (arr, keep)
        return self._c[0]

    def run(self, seed, traj_lo, traj_hi, threads=0, per_traj=False):
        """-> dict(rows [nrows,3], per_traj [nrows, n] or None, complete, censored, passes, seconds of the last full pass)."""
        import time
        L = lib()
        n = traj_hi - traj_lo
        passes = 0
        while True:
            rows = np.zeros((self.nrows, 3))
            per = np.zeros((self.nrows, n)) if per_traj else None
            status = np.zeros(n, np.int32)
            need = np.zeros((n, 2), np.int32)
            counters = np.zeros(4, np.int64)
            t0 = time.perf_counter()
            rc = L.qsr_run(C.cast(self._carray(), C.c_void_p), len(self.nodes), seed, traj_lo, traj_hi, self.nrows, threads,
                           rows.ctypes.data, per.ctypes.data if per is not None else None, status.ctypes.data, need.ctypes.data, counters.ctypes.data)
            dt = time.perf_counter() - t0
            passes += 1
            if rc < 0:
                raise RuntimeError(f"oracle error in {-rc} trajectories (status codes {sorted(set(status[status < 0]))})")
            if rc == 0:
                return dict(rows=rows, per_traj=per, complete=int(counters[0]), censored=int(counters[1]), passes=passes, seconds=dt, nodes=len(self.nodes))
            for node_i, ordinal in sorted({(int(a), int(b)) for a, b in need[status == 2]}):
                self.child(self.nodes[node_i], ordinal)
