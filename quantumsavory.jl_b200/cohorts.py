"""Cohort forking: outcome-dependent HOST control flow for a lockstep ensemble.

The reference runs one event loop per trajectory, so host code may branch on a measurement outcome and do
something different afterwards -- a failed purification frees the purified pair and the entangler later
refills those slots (/root/reference/src/CircuitZoo/CircuitZoo.jl:156-160,
examples/firstgenrepeater_lowlevel/setup.jl:209-237).  A batched backend sits under ONE host loop, which can follow
one classical history at a time.  This module makes `if mask:` on a per-trajectory mask mean what it means in the
reference:

* the host schedule is recorded as a program (backend.BatchContext.record) under a *script* = the branches taken at
  the first len(script) decisions; every later `if mask:` takes the default branch;
* at each unscripted decision the lanes whose outcome disagrees are RETIRED from the program (qsb_ctx_retire): they
  stop counting in its statistics and are marked with the ordinal of the decision;
* the tree driver groups the retired lanes by (program, ordinal), records -- once, cached: the tree depends on the
  schedule only, not on the trajectories -- the program whose script is the parent's decisions up to that ordinal
  with the last one flipped, and re-runs those lanes under it FROM t = 0 with the same global trajectory ids (hence
  the same draws: the re-run is bit-identical up to the fork, then continues on the lane's own branch).  Observables
  before the fork are not tallied again (the parent counted them).  All cohorts of one tree level run as ONE fused
  launch pair (qsb_wave_run).

Every (trajectory, observable) is therefore counted exactly once, in the program whose classical history is that
trajectory's own: the ensemble statistics and the per-trajectory series equal those of N independent runs of the
reference's control flow (tests/test_gpu_cohorts.py compares every lane with an oracle twin that follows its own
branches).  `max_depth` caps the number of forks per trajectory; a lane that disagrees once more is *censored*
(dropped from later statistics, counted in `last["censored"]`) -- never silently continued on a wrong history.
"""
from __future__ import annotations

import ctypes as C

from . import _lib
from ._lib import check

CENSORED = 1 << 20   # added to the ordinal of a decision taken in a leaf program: the lane leaves the tree there


class Fork:
    """State of one recording of a cohort tree: which branch `if mask:` takes at each decision."""

    def __init__(self, tree, script, leaf):
        self.tree, self.script, self.leaf = tree, tuple(bool(b) for b in script), leaf
        self.k = 0
        self.taken = []

    @property
    def tally(self) -> bool:
        """Observables before the last scripted decision were already counted by the parent program."""
        return self.k >= len(self.script)

    def decide(self, mask) -> bool:
        cached = getattr(mask, "_decided", None)
        if cached is not None:          # the same mask used in a second `if` (circuit, then protocol): same decision
            return cached
        k = self.k
        self.k += 1
        scripted = k < len(self.script)
        take = self.script[k] if scripted else self.tree.default
        mask.ctx.on_decision(mask, k, take, scripted, self)   # the backend's part: retire the lanes that disagree
        mask._decided = take
        self.taken.append(take)
        return take


class _Node:
    __slots__ = ("script", "program", "taken", "children", "index")

    def __init__(self, script, program, taken, index):
        self.script, self.program, self.taken, self.children, self.index = script, program, tuple(taken), {}, index


class CohortTree:
    """schedule(): issues the host schedule once against `ctx` (called under ctx.record() with ctx.fork set).

    run(traj_shift) executes the whole tree for one batch of ctx.ntraj trajectories; statistics rows must be zeroed by
    the caller before (every program accumulates into the rows captured at record time)."""

    def __init__(self, ctx, schedule, max_depth=None, default=True, max_programs=None):
        import torch
        self.torch = torch
        self.ctx, self.schedule, self.max_depth, self.default, self.max_programs = ctx, schedule, max_depth, default, max_programs
        self.device = torch.device("cuda", ctx.device)
        self.codes = torch.full((ctx.ntraj,), -1, dtype=torch.int32, device=self.device)
        self.codes_ptr = self.codes.data_ptr()
        self.lane_node = torch.zeros(ctx.ntraj, dtype=torch.int32, device=self.device)
        self.lanes_buf = [torch.zeros(ctx.ntraj, dtype=torch.int32, device=self.device) for _ in range(2)]   # lane subsets of the next level
        self.maxdec = 1          # decisions per program seen so far (key = program * maxdec + ordinal)
        self.nodes = []
        self.record_ms = 0.0
        self.last = {}
        self.root = self._record(())

    # -- recording -----------------------------------------------------------------------------------
    def _record(self, script):
        import time
        t0 = time.perf_counter()
        ctx = self.ctx
        leaf = self.max_depth is not None and sum(1 for i, b in enumerate(script) if b != self.default) >= self.max_depth
        fork = Fork(self, script, leaf)
        ctx.fork, ctx._draw = fork, 0
        try:
            with ctx.record() as rec:
                self.schedule()
        finally:
            ctx.fork = None
        node = _Node(tuple(script), rec.program, fork.taken, len(self.nodes))
        self.nodes.append(node)
        self.maxdec = max(self.maxdec, len(fork.taken))
        self.record_ms += 1e3 * (time.perf_counter() - t0)
        return node

    def child(self, node, ordinal):
        """The program of the lanes that disagreed with `node` at decision `ordinal` (recorded on first use)."""
        c = node.children.get(ordinal)
        if c is None:
            if self.max_programs is not None and len(self.nodes) >= self.max_programs:
                return None
            c = self._record(node.taken[:ordinal] + (not node.taken[ordinal],))
            node.children[ordinal] = c
        return c

    # -- execution -----------------------------------------------------------------------------------
    def _launch(self, jobs, traj_shift):
        """jobs: list of (node, lanes device pointer or None, count) -> ONE fused launch pair."""
        L, n = self.ctx.L, len(jobs)
        progs = (C.c_void_p * n)(*[j[0].program.p.value for j in jobs])
        lanes = (C.c_void_p * n)(*[j[1] for j in jobs])
        counts = (C.c_int64 * n)(*[j[2] for j in jobs])
        check(L.qsb_wave_run(self.ctx.p, progs, lanes, counts, n, traj_shift, 0, _lib.RUN_KEEP_STATS))

    def run(self, traj_shift: int = 0, nlanes: int | None = None):
        """One batch: every lane of the context (or the first `nlanes` lanes) through the tree of its own history.

        Per level: ONE fused launch pair over all cohorts (qsb_wave_run), then the routing of the retired lanes to the next
        level on the device (qsb_cohort_count / qsb_cohort_scatter); the only host-visible step is the read-back of the
        cohort sizes, because the host has to record the programs of histories it meets for the first time."""
        import numpy as np
        torch = self.torch
        ctx, L = self.ctx, self.ctx.L
        tstream = torch.cuda.current_stream(self.device)   # codes / lane tables are torch tensors used only as device buffers:
        self.codes.fill_(-1)                                # fence once unless the context was created on torch's stream
        self.lane_node.zero_()
        first = None
        if nlanes is not None and nlanes < ctx.ntraj:
            first = torch.arange(nlanes, dtype=torch.int32, device=self.device)
        self._first = first                                   # read by the launch: keep it alive
        tstream.synchronize()
        self._launch([(self.root, first.data_ptr() if first is not None else None, nlanes if first is not None else ctx.ntraj)], traj_shift)
        levels, rerun, censored, dropped = 1, 0, 0, 0
        jobs_per_level, lanes_per_level = [1], [nlanes if first is not None else ctx.ntraj]
        h2d, d2h = 88, 0        # bytes that really cross PCIe/NVLink-C2C per batch: job tables and routing tables down, cohort sizes up
        nprog0 = len(self.nodes)
        codes_p, node_p = C.c_void_p(self.codes_ptr), C.c_void_p(self.lane_node.data_ptr())
        while True:
            nn, md = len(self.nodes), self.maxdec
            counts = np.zeros(nn * md + 1, np.int32)
            check(L.qsb_cohort_count(ctx.p, codes_p, node_p, nn, md, CENSORED, counts.ctypes.data_as(C.c_void_p)))   # waits for the level
            d2h += counts.nbytes
            censored += int(counts[-1])
            keys = np.flatnonzero(counts[:-1])
            if keys.size == 0:
                break
            offsets = np.zeros(nn * md, np.int32)
            child = np.full(nn * md, -1, np.int32)
            jobs, off = [], 0
            base = self.lanes_buf[levels & 1].data_ptr()       # double buffered: the wave in flight reads the other half
            for k in keys.tolist():
                cnt = int(counts[k])
                node = self.child(self.nodes[k // md], k % md)
                if node is None:                               # program budget exhausted: these lanes are dropped, and counted
                    dropped += cnt
                    continue
                offsets[k], child[k] = off, node.index
                jobs.append((node, base + 4 * off, cnt))
                off += cnt
            check(L.qsb_cohort_scatter(ctx.p, codes_p, node_p, nn, md, CENSORED, offsets.ctypes.data_as(C.c_void_p), child.ctypes.data_as(C.c_void_p), C.c_void_p(base)))
            h2d += offsets.nbytes + child.nbytes
            if not jobs:
                ctx.sync()      # the scatter is still writing codes / lane_node on the context's stream; the next run() resets them on torch's
                break
            self._launch(jobs, traj_shift)
            rerun += off
            levels += 1
            jobs_per_level.append(len(jobs)); lanes_per_level.append(off)
            h2d += 88 * len(jobs)
        self.last = {"levels": levels, "lanes_rerun": rerun, "censored": censored, "dropped": dropped, "programs": len(self.nodes),
                     "programs_recorded_this_run": len(self.nodes) - nprog0, "jobs_per_level": jobs_per_level, "lanes_per_level": lanes_per_level,
                     "h2d_bytes": h2d, "d2h_bytes": d2h}
        return self.last

    def free(self):
        for n in self.nodes:
            n.program.free()
        self.nodes = []
