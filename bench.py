#!/usr/bin/env python
"""bench.py -- trajectories/sec of the register-state hot path on B200 (contract in DESIGN.md "Measurement").

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--workload NAME] [--ntraj M]

One "step" = one complete pass of the workload's event schedule over one ensemble of `ntraj` trajectories per GPU
(weak scaling: every GPU owns its own block of global trajectory ids; `--scaling strong` divides ntraj by the GPU
count).  Default workload = BASELINE.json configs[1]: firstgenrepeater, CliffordRepr tableaux, depolarizing noise,
10^6 trajectories per GPU, every trajectory following its OWN control flow (cohort tree).
Rank 0 prints ONE JSON line.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

# The CPU arms (cpu_baseline*, --impl reference) run on rank 0 with ALL the host cores this process may use, whatever the
# launcher exported: torch.distributed.run sets OMP_NUM_THREADS=1, which made the round-2 reference arm a one-thread run at
# N > 1.  Set before any OpenMP runtime is loaded, and passed to qsr_run explicitly as well.
HOST_THREADS = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
if int(os.environ.get("RANK", 0)) == 0:
    os.environ["OMP_NUM_THREADS"] = str(HOST_THREADS)

import numpy as np  # noqa: E402

SEED = 0x5153
FGR = dict(sizes=(2, 3, 4, 3, 2), F=0.97, t_end=25.0, dt_sample=0.2)  # 6.1_compare_formalisms_noplot.jl:11-21
FGR_SETS = {   # the two parameter sets of BASELINE.json configs[0]
    "ref5": dict(sizes=(2, 3, 4, 3, 2), F=0.9, t_end=30.0, dt_sample=0.2, bg=("T2", 10.0)),          # (i) examples/firstgenrepeater/3_purifier_example.jl:10-18,63
    "t1t2_4": dict(sizes=(2, 3, 3, 2), F=0.9, t_end=30.0, dt_sample=0.2, bg=("T1T2", 100.0, 10.0)),  # (ii) 4 nodes, T1T2Noise(100, 10)
}


# ------------------------------------------------------------------------------------------------
# workloads (shared by the GPU arm and the CPU arm: the same host code drives either backend)
# ------------------------------------------------------------------------------------------------
def wl_fgr(make_repr, formalism, record, params=None):
    from qsb200 import symbols as S
    from qsb200.workloads import firstgenrepeater as fgr
    from qsb200.registers import dispose
    kw = dict(FGR)
    bg = S.Depolarization(100.0) if formalism == "clifford" else S.T1T2Noise(1000.0, 100.0)
    if params:
        kw = {k: v for k, v in FGR_SETS[params].items() if k != "bg"}
        b = FGR_SETS[params]["bg"]
        bg = S.T2Dephasing(b[1]) if b[0] == "T2" else S.T1T2Noise(b[1], b[2])
    if formalism == "mc":   # a mixture would be lowered to a density matrix (express of a mixed state): pure pairs keep the states kets
        from qsb200 import ops
        kw["noisy_pair"] = ops.StabilizerState("XX ZZ")
    _, _, net = fgr.monte_carlo_trajectory(make_repr, background=bg, record=record, **kw)
    dispose(net.registers)
    return net.log


def wl_fgr_ref5(make_repr, formalism, record):
    return wl_fgr(make_repr, formalism, record, params="ref5")


def wl_fgr_t1t2_4(make_repr, formalism, record):
    return wl_fgr(make_repr, formalism, record, params="t1t2_4")


def wl_sweep12(make_repr, formalism, record, n=12, rounds=2, seed=11):
    """BASELINE.json configs[4]: composed n-qubit dm register; random 1q/2q gates, T1T2 noise on every
    qubit, project_traceout! of the last qubit, re-compose a fresh |0>."""
    from qsb200 import ops
    from qsb200 import symbols as S
    from qsb200.registers import Register, apply, dispose, initialize, observable, project_traceout, uptotime
    rng = np.random.default_rng(seed)
    reg = Register(n, make_repr(), S.T1T2Noise(100.0, 10.0))
    for i in range(1, n, 2):
        initialize((reg[i], reg[i + 1]), ops.DepolarizedBellPair(F=0.95))
    for i in range(1, n - 1, 2):
        apply((reg[i], reg[i + 2]), ops.CNOT)
    t = 0.0
    for r in range(rounds):
        for _ in range(3):
            apply(reg[int(rng.integers(n)) + 1], [ops.H, ops.X, ops.Z][rng.integers(3)])
            a, b = (int(x) + 1 for x in rng.choice(n, 2, replace=False))
            apply((reg[a], reg[b]), ops.CNOT)
        t += float(rng.uniform(0.01, 0.1))
        uptotime(reg[1:n], t)
        m = project_traceout(reg[n], ops.Z)
        initialize(reg[n], time=t)
        apply((reg[n - 1], reg[n]), ops.CNOT)
        record("ZZ", t, observable((reg[1], reg[2]), ops.tensor(ops.Z, ops.Z)))
    dispose([reg])
    return {"rounds": rounds, "n": n}


def wl_ghz(make_repr, formalism, record, S=4):
    """BASELINE.json configs[2]: one F-TMBL trial of distributed_ghz_sensors with all S links up (S=4: four pairs
    composed to an 8-qubit dm at the hub, 3 CNOT + H, 4 measurements, masked Pauli corrections, GHZ_4 fidelity)."""
    from qsb200.registers import dispose
    from qsb200.workloads import ghz_sensors as G
    k, _, net, _ = G.f_tmbl_trial(make_repr, record=record, S=S)
    dispose(net)
    return {"S": S, "entangled": k}


def wl_grid(make_repr, formalism, record):
    """BASELINE.json configs[3]: repeatergrid 5x5, 4 slots per node, CliffordRepr, Depolarization(10), t_end=30."""
    from qsb200.registers import dispose
    from qsb200.workloads import repeatergrid as R
    net = R.run(make_repr, record=record, n=5, regsize=4, tau=10.0, seed=1)
    dispose(net.registers)
    return net.log


FGR_DESC = "examples/firstgenrepeater_lowlevel, sizes [2,3,4,3,2], F=0.97"
WORKLOADS = {
    "firstgenrepeater_clifford": dict(fn=wl_fgr, formalism="clifford", ntraj=1_000_000, mode="cohort",
                                      desc=FGR_DESC + ", CliffordRepr, Depolarization(100), t=0:0.2:25"),
    "firstgenrepeater_clifford_masks": dict(fn=wl_fgr, formalism="clifford", ntraj=1_000_000, mode="program",
                                            desc=FGR_DESC + ", CliffordRepr, Depolarization(100), t=0:0.2:25; lockstep ensemble on the success-branch schedule "
                                                 "(failed purifications = per-trajectory masks; NOT the reference's per-trajectory control flow: round-1 headline, kept for continuity)"),
    "firstgenrepeater_dm": dict(fn=wl_fgr, formalism="qo", ntraj=1_000_000, mode="program",
                                desc=FGR_DESC + ", QuantumOpticsRepr dm, T1T2Noise(1000,100), t=0:0.2:25; lockstep masks ensemble"),
    "firstgenrepeater_dm_cohort": dict(fn=wl_fgr, formalism="qo", ntraj=1_000_000, mode="cohort",
                                       desc=FGR_DESC + ", QuantumOpticsRepr dm, T1T2Noise(1000,100), t=0:0.2:25; cohort tree of DENSE recorded programs: "
                                            "every trajectory follows its own control flow (the reference's semantics)"),
    "firstgenrepeater_dm_ref5": dict(fn=wl_fgr_ref5, formalism="qo", ntraj=1_000_000, mode="program",
                                     desc="examples/firstgenrepeater/3_purifier_example.jl parameters: sizes [2,3,4,3,2], F=0.9, T2Dephasing(10), t=0:0.2:30, dm; lockstep masks ensemble"),
    "firstgenrepeater_dm_t1t2_4": dict(fn=wl_fgr_t1t2_4, formalism="qo", ntraj=1_000_000, mode="program",
                                       desc="BASELINE configs[0] as worded: 4-node chain sizes [2,3,3,2], F=0.9, T1T2Noise(100,10), t=0:0.2:30, dm; lockstep masks ensemble"),
    "firstgenrepeater_mcket": dict(fn=wl_fgr, formalism="mc", ntraj=1_000_000, mode="per_op",
                                   desc=FGR_DESC.replace("F=0.97", "perfect pairs") + ", QuantumMCRepr kets, T1T2Noise(1000,100) as sampled Kraus jumps, t=0:0.2:25; lockstep masks ensemble"),
    "ghz_sensors_dm": dict(fn=wl_ghz, formalism="qo", ntraj=16384, mode="per_op",
                           desc="examples/distributed_ghz_sensors F-TMBL trial, S=4 all links up (peak 8-qubit dm = 1 MiB per trajectory), F=0.99, Depolarization(1.0); "
                                "resident chunk of 16384 trajectories per step (10^6 trajectories = 62 chunks: --workload ghz_sensors_dm --chunks 62)"),
    "repeatergrid_clifford": dict(fn=wl_grid, formalism="clifford", ntraj=10_000_000, mode="program",
                                  desc="examples/repeatergrid 5x5 grid, 4 slots per node, CliffordRepr, Depolarization(10) (T1Decay has no Clifford method), success_prob=0.001, t_end=30"),
    "sweep12_dm": dict(fn=wl_sweep12, formalism="qo", ntraj=256, mode="per_op",
                       desc="12-qubit composed dm register: random H/X/Z + CNOT, T1T2Noise(100,10) on every qubit, project_traceout!, recompose"),
}
DEFAULT_EXTRAS = "firstgenrepeater_clifford_masks,firstgenrepeater_dm,firstgenrepeater_dm_cohort,firstgenrepeater_dm_ref5,firstgenrepeater_dm_t1t2_4,firstgenrepeater_mcket,ghz_sensors_dm,repeatergrid_clifford,sweep12_dm"


# ------------------------------------------------------------------------------------------------
# CPU arm.  (1) compiled: the host schedule traced once per classical history (oracle/trace.py) and replayed through
# the oracle's C functions with OpenMP over trajectories (oracle/qsb_replay.c) -- what `cpu_baseline` and
# `--impl reference` report; (2) python: the same oracle called op by op from the Python host (round-1 number,
# interpreter bound, kept beside it as cpu_baseline_python).
# ------------------------------------------------------------------------------------------------
CPU_SAMPLE = {   # trajectories per timed CPU pass (about 5-20 s of work on 16 cores), Python-driven trajectories per core
    "firstgenrepeater_clifford": (100_000, 40), "firstgenrepeater_clifford_masks": (0, 0), "firstgenrepeater_dm": (4_000, 0), "firstgenrepeater_dm_cohort": (4_000, 0),
    "firstgenrepeater_dm_ref5": (4_000, 0), "firstgenrepeater_dm_t1t2_4": (4_000, 0), "firstgenrepeater_mcket": (40_000, 0),
    "ghz_sensors_dm": (16, 0), "repeatergrid_clifford": (20_000, 0), "sweep12_dm": (0, 0)}


def make_trace_tree(workload, max_depth):
    from oracle import trace
    from qsb200.registers import observable_options
    w = WORKLOADS[workload]
    slot = [0]

    def next_row():
        i = slot[0]; slot[0] += 1
        return i

    def schedule(make_repr):
        slot[0] = 0
        with observable_options(stats_dev=next_row):
            w["fn"](make_repr, w["formalism"], lambda *a: None)
    return trace.TraceTree(schedule, w["formalism"], 4096, max_depth=max_depth)


def cpu_compiled(workload, ntraj, max_depth, traj_lo=0, passes=1, tree=None):
    """-> dict(value traj/s, seconds, cores, nodes, trace_s, censored, complete, rows).  The op lists are traced untimed
    (the GPU arm records its programs untimed too); the timed region is the compiled replay of `ntraj` trajectories."""
    from oracle import trace
    tree = tree or make_trace_tree(workload, max_depth)
    r = tree.run(SEED, traj_lo, traj_lo + ntraj, threads=HOST_THREADS)   # grows the tree until every history of the sample is traced
    secs = [r["seconds"]] if r["passes"] == 1 else []
    while len(secs) < passes:
        r = tree.run(SEED, traj_lo, traj_lo + ntraj, threads=HOST_THREADS)
        secs.append(r["seconds"])
    s = float(np.median(secs))
    cores = int(trace.lib().qsr_max_threads())             # what omp_set_num_threads(HOST_THREADS) left in force
    if cores == 1 and (os.cpu_count() or 1) > 1 and HOST_THREADS > 1:
        raise RuntimeError("the compiled CPU replay ran on one thread of a multi-core box: refusing to report it as the CPU arm")
    return dict(value=(r["complete"]) / s, seconds=s, cores=cores, nodes=r["nodes"], trace_s=tree.trace_s,
                censored=r["censored"], complete=r["complete"], rows=r["rows"], tree=tree)


def _cpu_worker(args):
    workload, traj_lo, traj_hi = args
    sys.path.insert(0, ROOT)
    import qsb200  # noqa: F401
    from tests.oracle_backend import OracleContext, OracleRepr
    w = WORKLOADS[workload]
    t0 = time.perf_counter()
    for t in range(traj_lo, traj_hi):
        ctx = OracleContext(SEED, t)
        w["fn"](lambda: OracleRepr(ctx, w["formalism"]), w["formalism"], lambda *a: None)
    return traj_hi - traj_lo, time.perf_counter() - t0


def cpu_python(workload, per_core, cores=None):
    """Times `per_core` trajectories on each host core, oracle driven op by op from Python (processes, not threads)."""
    import multiprocessing as mp
    cores = cores or HOST_THREADS
    jobs = [(workload, c * per_core, (c + 1) * per_core) for c in range(cores)]
    with mp.get_context("fork").Pool(cores) as pool:
        pool.map(_cpu_worker, [(workload, 0, 1)] * cores)     # pool start-up and imports are not timed
        t0 = time.perf_counter()
        res = pool.map(_cpu_worker, jobs)
        wall = time.perf_counter() - t0
    return dict(ntraj=sum(r[0] for r in res), wall=wall, cores=cores)


def cpu_baseline_for(workload, max_depth, want_python=True):
    n, per_core = CPU_SAMPLE.get(workload, (0, 0))
    if not n:
        return None, None, None
    c = cpu_compiled(workload, n, max_depth)
    what = "complete trajectories" if WORKLOADS[workload]["mode"] == "cohort" else "trajectories"
    base = {"value": c["value"], "unit": "trajectories/s", "cores": c["cores"], "kind": "port-c-openmp",
            "sample": f"{n} trajectories replayed by oracle/qsb_replay.c on {c['cores']} OpenMP threads in {c['seconds']:.2f} s ({c['complete']} {what}, {c['censored']} censored at "
                      f"fork budget {max_depth}); {c['nodes']} op lists traced untimed in {c['trace_s']:.1f} s",
            "note": "every trajectory follows its own control flow through the oracle's C functions (embed + dense products / generator-list stabilizers, the reference's algorithm)"}
    py = None
    if want_python and per_core:
        r = cpu_python(workload, per_core)
        py = {"value": r["ntraj"] / r["wall"], "unit": "trajectories/s", "cores": r["cores"], "kind": "port-python",
              "sample": f"{per_core} trajectories per core on {r['cores']} processes ({r['wall']:.1f} s): the same oracle called op by op from the Python host layer (interpreter bound)"}
    return base, py, c


# ------------------------------------------------------------------------------------------------
# GPU arm
# ------------------------------------------------------------------------------------------------
class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100", "-i", str(self.index)],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        self.thread.join(timeout=2)
        sm = [float(r[0]) for r in self.rows if len(r) >= 7 and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) >= 7 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[i] for r in self.rows if len(r) >= 7 for i in range(4) if r[3 + i].lower().startswith("active")})
        pw = [float(r[2]) for r in self.rows if len(r) >= 7 and r[2].replace(".", "").isdigit()]
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons,
                "power_w_max": max(pw) if pw else None, "samples": len(sm)}


def ncu_facts(workload, kernel):
    """Per-launch facts of `kernel` from the committed ncu --set full captures of this workload (profiles/ncu_kernels.json):
    dram bytes, executed warp instructions, grid x block -- or None when there is no capture for it."""
    try:
        with open(os.path.join(ROOT, "profiles", "ncu_kernels.json")) as f:
            return json.load(f).get(f"{workload}|{kernel}")
    except Exception:
        return None


def ncu_facts_provenance(kernels):
    """-> (commit the captures were taken at, [kernel source files that changed since]).  The facts are only as good as the
    kernels they were captured from: profiles/ncu_kernels.json records a hash of every source file per kernel."""
    import hashlib
    try:
        with open(os.path.join(ROOT, "profiles", "ncu_kernels.json")) as f:
            about = json.load(f).get("_about", {})
        changed = set()
        for k in kernels:
            for src in about.get("kernel_sources", {}).get(k, []):
                with open(os.path.join(ROOT, "quantumsavory.jl_b200", "csrc", src), "rb") as f:
                    if hashlib.sha256(f.read()).hexdigest()[:16] != about.get("sources_sha256_16", {}).get(src):
                        changed.add(src)
        return about.get("captured_at_commit"), sorted(changed)
    except Exception:
        return None, ["(provenance unreadable)"]


def measured_peak():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class GpuRun:
    """One workload on this rank's GPU: ensemble behind the C-ABI, statistics kept on the device."""

    def __init__(self, workload, ntraj, rank, world, local_rank, cohort_depth=3):
        import torch
        import qsb200  # noqa: F401
        from qsb200.backend import BatchContext, CudaBatchRepr
        from qsb200.sharding import StatsBuffer
        self.torch = torch
        self.workload = workload
        self.w = WORKLOADS[workload]
        self.ntraj, self.rank, self.world, self.cohort_depth = ntraj, rank, world, cohort_depth
        torch.cuda.set_device(local_rank)
        self.device = torch.device("cuda", local_rank)
        self.stream = torch.cuda.Stream(device=self.device)
        self.ctx = BatchContext(ntraj, seed=SEED, device=local_rank, traj_offset=rank * ntraj, stream=self.stream.cuda_stream)
        self.CudaBatchRepr = CudaBatchRepr
        self.stats = StatsBuffer(4096, self.device)
        self.flush = torch.empty(256 << 20, dtype=torch.uint8, device=self.device)  # > 126 MB L2
        self.h2d_bytes = self.d2h_extra = 0
        self.per_op_batch = 0
        self.batch = 0
        self.tree = self.program = None
        self.nobs = 0

    # -- the host schedule against this context (per-op: launches; inside ctx.record(): captured) --------------
    def step(self):
        ctx, stats = self.ctx, self.stats
        ctx._draw = 0
        if not ctx.recording:   # every per-op pass simulates a FRESH batch of global trajectory ids (like run_program does)
            ctx.set_traj_offset(self.rank * self.ntraj + self.per_op_batch * self.ntraj * self.world)
            self.per_op_batch += 1
        slot = [0]
        keys = []

        def record(kind, tm, val):
            keys.append((kind, tm, None if hasattr(val, "n_alive") else val))

        def make_repr():
            return self.CudaBatchRepr(ctx, self.w["formalism"])

        # every observable() call gets its own device statistics row (no host sync)
        from qsb200.registers import observable_options

        def next_row():
            i = slot[0]; slot[0] += 1
            return stats.slot_ptr(i)
        with observable_options(stats_dev=next_row, sync=False), self.torch.cuda.stream(self.stream):
            log = self.w["fn"](make_repr, self.w["formalism"], record)
        self.nobs = slot[0]
        self.keys = keys
        self.log = log
        return log

    def record_program(self):
        """Runs the host schedule ONCE in record mode (nothing launches) and keeps the fused program."""
        t0 = time.perf_counter()
        with self.ctx.record() as rec:
            log = self.step()
        self.program = rec.program
        self.program_log = log
        self.record_ms = 1e3 * (time.perf_counter() - t0)
        return log

    def record_tree(self):
        """Cohort tree: one recorded program per classical history (cohorts.CohortTree); grown during the warm-up batches."""
        from qsb200.cohorts import CohortTree
        t0 = time.perf_counter()
        with self.torch.cuda.stream(self.stream):
            self.tree = CohortTree(self.ctx, self.step, max_depth=self.cohort_depth)
        self.program = self.tree.root.program
        self.program_log = self.log
        self.record_ms = 1e3 * (time.perf_counter() - t0)

    def upload_program(self):
        with self.torch.cuda.stream(self.stream):
            self.h2d_bytes = self.program.upload()

    def run_program(self):
        """One fused launch over a fresh batch of ntraj trajectories per GPU (global ids never repeat)."""
        self.batch += 1
        self.program.run(traj_shift=self.batch * self.ntraj * self.world)
        return self.program_log

    def run_tree(self):
        self.batch += 1
        with self.torch.cuda.stream(self.stream):
            self.stats.t.zero_()
            self.tree_info = self.tree.run(traj_shift=self.batch * self.ntraj * self.world)
        self.h2d_bytes, self.d2h_extra = self.tree_info["h2d_bytes"], self.tree_info["d2h_bytes"]
        return self.program_log

    def finish_step(self, reduce=True):
        """The step's result crosses GPUs (one all-reduce) and comes back to the host (D2H).
        reduce=False: rank-local read-back (the profiled extra step runs on rank 0 only -- no collective there)."""
        with self.torch.cuda.stream(self.stream):
            if reduce:
                self.stats.allreduce()
            host = self.stats.t[: max(self.nobs, 1)].cpu()
        return host.numpy()


def parity_check(g, workload, ntraj, rank, world, local_rank):
    """Untimed: replays the FIRST timed batch of trajectory ids through the fused program once more and through the per-op
    path (one launch per backend call, a context created at the same global ids) and compares every statistics row.
    Clifford rows are integer tallies and must be identical; dense rows agree to 1e-12 x trajectories (summation order)."""
    import torch
    shift = 1 * ntraj * world
    g.stats.t.zero_()
    g.program.run(traj_shift=shift)
    g.stream.synchronize()
    torch.cuda.synchronize()
    rows_prog = g.stats.t[: g.nobs].cpu().numpy().copy()
    ref = GpuRun(workload, ntraj, rank, world, local_rank)
    ref.per_op_batch = 1                 # step() re-bases the context on rank * ntraj + 1 * ntraj * world == the ids of `shift`
    ref.step()
    ref.stream.synchronize()
    torch.cuda.synchronize()
    rows_op = ref.stats.t[: ref.nobs].cpu().numpy().copy()
    ref.ctx.close()
    exact = g.w["formalism"] == "clifford"
    same_shape = rows_op.shape == rows_prog.shape
    if same_shape:
        diff = float(np.abs(rows_prog[:, :3] - rows_op[:, :3]).max()) if len(rows_op) else 0.0
        alive_equal = bool((rows_prog[:, 0] == rows_op[:, 0]).all())
        equal = alive_equal and (diff == 0.0 if exact else diff <= 1e-12 * ntraj)
    else:
        diff, alive_equal, equal = float("nan"), False, False
    res = {"rows": int(len(rows_prog)), "equal": bool(equal), "against": "per-op path (one launch per backend call) on the same global trajectory ids",
           "tolerance": "bit-exact (integer tallies)" if exact else "1e-12 x trajectories on the sums, identical alive counts",
           "max_abs_diff": diff, "alive_counts_equal": alive_equal, "trajectory_ids": [rank * ntraj + shift, rank * ntraj + shift + ntraj]}
    if not equal:
        raise RuntimeError(f"parity check failed for {workload}: {res}")
    return res


def cpu_bitsliced(g, ntraj, rank, world, no_cpu=False):
    """The OPTIMISED CPU baseline: the exported op records of the recorded tableau program replayed on
    the host cores with the GPU's own bit-sliced tableau core compiled by g++ (oracle/qsb_bitsliced.cpp, 32 trajectories per word,
    OpenMP over words).  Same global trajectory ids as the first timed batch; the integer statistics rows must equal the GPU's."""
    import time
    import torch
    from oracle import bitsliced
    ex = g.program.export()
    shift = 1 * ntraj * world
    g.stats.t.zero_()
    g.program.run(traj_shift=shift)
    g.stream.synchronize(); torch.cuda.synchronize()
    rows_gpu = g.stats.t[: g.nobs].cpu().numpy()[:, :3].copy()
    n = ntraj if ntraj <= (1 << 21) else (1 << 21)            # bounded sample for the very large ensembles (the rows are compared on the full one only)
    traj0 = rank * ntraj + shift
    bitsliced.run(ex, SEED, traj0, min(n, 1 << 14))           # build / page in
    t0 = time.perf_counter()
    rows_cpu, threads = bitsliced.run(ex, SEED, traj0, n)
    dt = time.perf_counter() - t0
    # row r of the replay belongs to the statistics slot its device pointer names (slots of never-assigned observables stay zero)
    slots = ((ex["stats"].astype(np.int64) - int(g.stats.t.data_ptr())) // 32).astype(np.int64)
    placed = np.zeros_like(rows_gpu)
    ok_slots = bool(((slots >= 0) & (slots < len(rows_gpu))).all())
    if ok_slots:
        placed[slots] = rows_cpu[:, :3]
    equal = bool(n == ntraj and ok_slots and np.array_equal(rows_gpu, placed))
    res = {"value": n / dt, "unit": "trajectories/s", "cores": threads, "kind": "bit-sliced-c++-openmp",
           "sample": f"{n} trajectories of the same recorded program (lockstep masks) in {dt:.2f} s on {threads} OpenMP threads",
           "rows_equal_gpu": equal if n == ntraj else None,
           "note": "the GPU design on the CPU: exported op records, csrc/qsb_tab_core.cuh compiled by g++ -O3, one 32-trajectory word column per task"}
    if n == ntraj and not equal:
        raise RuntimeError(f"bit-sliced CPU replay disagrees with the GPU program: {res}")
    return res


def parity_check_tree(g, cpu, n):
    """Untimed: the first n trajectory ids through the cohort tree vs the compiled CPU replay of the same ids under the same
    fork budget (cpu["rows"], oracle/qsb_replay.c): every trajectory follows its own control flow on both sides, so the
    statistics rows -- alive counts, sums, sums of squares at every sampling point -- must be IDENTICAL."""
    import torch
    with torch.cuda.stream(g.stream):
        g.stats.t.zero_()
        info = g.tree.run(traj_shift=0, nlanes=n)
    g.stream.synchronize(); torch.cuda.synchronize()
    rows_gpu = g.stats.t[: g.nobs].cpu().numpy()[:, :3]
    rows_cpu = cpu["rows"][: g.nobs]
    exact = g.w["formalism"] == "clifford"
    # tableau rows are integer tallies (identical); dense rows are floating-point sums over the same per-trajectory values in a
    # different order: identical alive counts, sums within 1e-12 x trajectories
    equal = bool(np.array_equal(rows_gpu, rows_cpu)) if exact else bool(np.array_equal(rows_gpu[:, 0], rows_cpu[:, 0]) and np.abs(rows_gpu - rows_cpu).max() <= 1e-12 * n)
    res = {"rows": int(g.nobs), "equal": equal, "trajectories": n, "against": "oracle/qsb_replay.c (compiled CPU replay, every trajectory on its own control flow) on the same global trajectory ids and fork budget",
           "tolerance": "bit-exact (integer tallies)" if exact else "identical alive counts, sums within 1e-12 x trajectories",
           "max_abs_diff": float(np.abs(rows_gpu - rows_cpu).max()), "censored_gpu": info["censored"], "censored_cpu": cpu["censored"]}
    if not equal or info["censored"] != cpu["censored"]:
        raise RuntimeError(f"cohort-tree parity check failed: {res}")
    return res


def censor_bias_check(g, workload, n):
    """Untimed: what the fork budget costs in the STATISTICS.  The first n trajectory ids run through the GPU tree under the
    benchmarked fork budget and through the compiled CPU replay with NO budget (every trajectory to the end of its own history,
    the reference's semantics, examples/firstgenrepeater_lowlevel/setup.jl:209-237).  A censored lane leaves the GPU's rows at its
    next disagreement, so later sampling points average over the survivors only; the difference of the ensemble means is that bias."""
    import torch
    free = cpu_compiled(workload, n, None)
    assert free["censored"] == 0 and free["complete"] == n
    with torch.cuda.stream(g.stream):
        g.stats.t.zero_()
        info = g.tree.run(traj_shift=0, nlanes=n)
    g.stream.synchronize(); torch.cuda.synchronize()
    rg, rf = g.stats.t[: g.nobs].cpu().numpy()[:, :3], free["rows"][: g.nobs]
    both = (rg[:, 0] > 0) & (rf[:, 0] > 0)
    mg, mf = rg[both, 1] / rg[both, 0], rf[both, 1] / rf[both, 0]
    d = np.abs(mg - mf)
    i = int(d.argmax())
    return {"trajectories": n, "censored_gpu": info["censored"], "uncensored_op_lists": free["nodes"],
            "against": "oracle/qsb_replay.c without a fork budget (every trajectory to the end of its own history) on the same trajectory ids",
            "max_abs_diff_of_ensemble_means": float(d.max()), "rms_diff_of_ensemble_means": float(np.sqrt((d ** 2).mean())), "at_row": int(np.flatnonzero(both)[i]),
            "mean_gpu_censored": float(mg[i]), "mean_uncensored": float(mf[i]),
            "max_alive_count_diff": float(np.abs(rg[:, 0] - rf[:, 0]).max()),
            "standard_error_of_a_mean_at_this_n": float(1.0 / np.sqrt(n))}


def run_gpu_workload(workload, ntraj, steps, warmup, rank, world, local_rank, profile=True, mode="per_op", cohort_depth=3, chunks=1, check=True, keep=None):
    """mode "per_op": every backend call is one launch (the drop-in path, host schedule re-run each step).
    mode "program": the schedule is recorded once (untimed, reported as record_ms) and each step replays it as ONE fused
    launch pair on a fresh batch of trajectories; e2e additionally re-sends the op list from the host.
    mode "cohort": a tree of recorded programs, one per classical history; each step runs every trajectory of a fresh batch
    through the program of its own history (one fused launch pair per tree level).
    chunks > 1: a step = `chunks` consecutive batches of ntraj resident trajectories (states larger than HBM / chunks)."""
    import torch
    import torch.distributed as dist
    g = GpuRun(workload, ntraj, rank, world, local_rank, cohort_depth=cohort_depth)
    ctx = g.ctx
    program, cohort = mode == "program", mode == "cohort"
    if program:
        g.record_program()
    if cohort:
        g.record_tree()
    one = g.run_tree if cohort else (g.run_program if program else g.step)

    def do_step():
        for _ in range(chunks):
            log = one()
        return log

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(warmup):
        do_step(); g.finish_step()
    tree_record_ms = g.tree.record_ms if cohort else 0.0      # the tree is grown by the warm-up batches (untimed, like record_ms)
    barrier()
    ctx.reset_counters()
    sampler = ClockSampler(local_rank)
    sampler.start()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    dev_ms = 0.0
    censored = rerun = 0
    new_programs_timed = 0
    barrier()
    t0 = time.perf_counter()
    for _ in range(steps):
        with torch.cuda.stream(g.stream):
            g.flush.zero_()                       # flush L2 between timed iterations (not timed on the device clock)
        if program:
            g.upload_program()                    # host op list -> device: inside e2e, before the device-timed region
        with torch.cuda.stream(g.stream):
            ev0.record(g.stream)
        log = do_step()
        with torch.cuda.stream(g.stream):
            ev1.record(g.stream)
        host_stats = g.finish_step()              # e2e: all-reduce + D2H of the step's result
        g.stream.synchronize()
        dev_ms += ev0.elapsed_time(ev1)
        if cohort:
            censored += g.tree_info["censored"] + g.tree_info["dropped"]; rerun += g.tree_info["lanes_rerun"]
            new_programs_timed += g.tree_info["programs_recorded_this_run"]
    barrier()
    wall = time.perf_counter() - t0
    counters = ctx.counters()
    # nvidia-smi answers every ~100 ms (slower with 8 GPUs polled at once) and a timed region of a few fused launches can be
    # shorter than that: keep the same step running, untimed, until the sampler holds readings taken under this load
    probe_steps, tp = 0, time.perf_counter()
    while len(sampler.rows) < 3 and time.perf_counter() - tp < 4.0 and sampler.proc is not None:
        do_step(); g.finish_step(reduce=False); g.stream.synchronize()
        probe_steps += 1
    clocks = sampler.stop()
    clocks["timed_region_ms"] = wall * 1e3
    if probe_steps:
        clocks["note"] = f"{probe_steps} untimed steps of the same workload appended so that nvidia-smi sampled the clocks under this load"
    barrier()
    t = torch.tensor([dev_ms, wall * 1e3], dtype=torch.float64, device=g.device)
    c = torch.tensor([float(censored)], dtype=torch.float64, device=g.device)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(c, op=dist.ReduceOp.SUM)
    dev_ms, wall_ms, censored_all = float(t[0]), float(t[1]), float(c[0])
    total = ntraj * chunks * world * steps
    done = total - censored_all                   # trajectories that ran their whole schedule on their own control flow
    out = dict(workload=workload, ntraj_per_gpu=ntraj * chunks, dev_ms_per_step=dev_ms / steps, wall_ms_per_step=wall_ms / steps,
               value=done / (dev_ms / 1e3), e2e=done / (wall_ms / 1e3),
               launches_per_step=counters["launches"] / steps, algo_bytes_per_step=counters["algo_bytes"] / steps,
               ops_per_step=counters["ops"] / steps, nobs=g.nobs, d2h_bytes=g.nobs * 32 + g.d2h_extra, clocks=clocks, log=log,
               sample_stats=host_stats[: min(4, len(host_stats))].tolist(), mode=mode, h2d_bytes=g.h2d_bytes)
    # what a "trajectory" of this number is: lanes still holding an end-to-end pair at the last sampling point
    if len(host_stats):
        out["alive_at_end_fraction"] = float(host_stats[-1][0]) / (ntraj * world)
    if program or cohort:
        out["program"] = dict(g.program.info(), record_ms=g.record_ms)
    if program:
        # amortised view: the one-off recording of the schedule (host time) spread over the timed steps
        out["e2e_incl_record"] = total / ((wall_ms + g.record_ms) / 1e3)
        if rank == 0 and check:
            out["parity_check"] = parity_check(g, workload, ntraj, rank, world, local_rank)
            if g.w["formalism"] == "clifford" and world == 1 and not os.environ.get("QSB_PROFILING"):
                try:
                    out["cpu_baseline_optimised"] = cpu_bitsliced(g, ntraj, rank, world)
                except RuntimeError:
                    raise
                except Exception as e:      # the optimised arm is a report, never a reason to lose the GPU line
                    out["cpu_baseline_optimised"] = {"error": repr(e)[:200]}
    if cohort:
        info = g.tree_info
        out["ensemble"] = {"mode": "cohort tree: every trajectory runs under the recorded program of its OWN classical history (cohorts.py)",
                           "max_forks_per_trajectory": cohort_depth, "programs": info["programs"], "levels": info["levels"],
                           "jobs_per_level": info["jobs_per_level"], "lanes_per_level": info["lanes_per_level"],
                           "lanes_with_any_failure_fraction": info["lanes_per_level"][1] / ntraj if len(info["lanes_per_level"]) > 1 else 0.0,
                           "complete_fraction": done / total, "censored_fraction": censored_all / total,
                           "lane_runs_per_trajectory": 1.0 + rerun / (ntraj * steps),
                           "tree_record_ms": tree_record_ms, "programs_recorded_in_timed_steps": new_programs_timed,
                           "note": "value/e2e count COMPLETE trajectories only; a trajectory whose fork budget is spent is censored at its next disagreement "
                                   "(dropped from later statistics and from the numerator), never continued on a wrong history"}
        out["e2e_incl_record"] = done / ((wall_ms + tree_record_ms) / 1e3)
    if profile and rank == 0:
        ctx.profile(True)
        do_step(); g.finish_step(reduce=False)
        rows = ctx.profile_report()
        ctx.profile(False)
        rows = [r for r in rows if r["launches"]]
        tot = sum(r["ms"] for r in rows) or 1.0
        rows.sort(key=lambda r: -r["ms"])
        peak, which = measured_peak()
        for r in rows:
            r["share"] = r["ms"] / tot
            r["gbs"] = r["algo_bytes"] / (r["ms"] * 1e-3) / 1e9 if r["ms"] > 0 else 0.0
            r["us_per_launch"] = 1e3 * r["ms"] / r["launches"]
        top = rows[0]
        out["kernels"] = rows[:8]
        facts = ncu_facts(workload, top["kernel"])
        out["roofline"] = {"bound": "hbm", "kernel": top["kernel"], "achieved": top["gbs"], "peak": peak, "unit": "GB/s",
                           "frac": top["gbs"] / peak, "traffic": facts["dram_bytes"] if facts else None, "peak_source": which,
                           "algo_bytes_per_launch": top["algo_bytes"] / top["launches"], "us_per_launch": top["us_per_launch"],
                           "share_of_step": top["share"]}
        if top["gbs"] > peak:
            out["roofline"]["note"] = (f"algorithmic bytes / event time = {top['gbs']:.0f} GB/s exceeds the measured copy peak: launches of a few microseconds whose operands are "
                                       "partly served by the 126 MB L2 (or over-counted algorithmic bytes); reported as measured, not capped")
        if program or cohort:
            out["roofline"] = fused_roofline(out, rows, g, workload, ntraj * chunks, peak, which, clocks)
    if keep is not None:
        keep.append(g)
    else:
        ctx.close()
    return out


def fused_roofline(out, rows, g, workload, lanes, peak, which, clocks):
    """Roofline block of a fused launch (pair).  The ensemble state never leaves the SM, so HBM is NOT what bounds these
    kernels: the block names the real bound -- instruction issue -- and gives
        frac = executed warp instructions / (SMs x 4 schedulers x SM clock x kernel time)
    with the instruction counts taken from the committed ncu --set full captures of the same launch shapes
    (profiles/ncu_kernels.json).  SURVEY 8d's table-byte figure (what the same op list would account op by op) is kept as
    hbm_equivalent next to the DRAM bytes the launches really move (traffic)."""
    info = out["program"]
    fused = [r for r in rows if r["kernel"] in ("k_tab_exec", "k_draw_bits", "k_dense_exec")]
    sec = sum(r["ms"] for r in fused) * 1e-3
    lane_runs = out["ensemble"]["lane_runs_per_trajectory"] * lanes if "ensemble" in out else lanes
    table_bytes = info["table_bytes"] * lane_runs / g.ntraj
    kname = "+".join(sorted(r["kernel"] for r in fused))
    sm_hz = (clocks.get("sm_mhz") or 1965.0) * 1e6
    sms = g.torch.cuda.get_device_properties(g.device).multi_processor_count
    inst = dram = 0.0
    have = True
    per_kernel = {}
    base = workload.replace("_masks", "").replace("_cohort", "")
    wave_facts = True
    for r in fused:
        # a cohort tree has captures of its own wave launches (summed over the levels of one step: the re-runs are inside); otherwise
        # the single-program capture is scaled by the lane-runs of the step
        f = ncu_facts(base + "_cohort", r["kernel"]) if "ensemble" in out else None
        if f:
            scale = lanes / f["trajectories"]
        else:
            wave_facts = False
            f = ncu_facts(base, r["kernel"])
            if not f:
                have = False
                continue
            scale = lane_runs / f["trajectories"]      # instructions and bytes scale with the lanes processed
        inst += f["warp_inst"] * scale; dram += f["dram_bytes"] * scale
        per_kernel[r["kernel"]] = {"ms": r["ms"], "issue_frac": f["warp_inst"] * scale / (sms * 4 * sm_hz * r["ms"] * 1e-3),
                                   "ncu": {k: f[k] for k in ("grid", "block", "issue_active", "source") if k in f}}
    issue_peak = sms * 4 * sm_hz                       # warp instructions per second the device can issue
    facts_commit, facts_changed = ncu_facts_provenance([r["kernel"] for r in fused])
    blk = {"modelled": True,
           "source": f"kernel time: CUDA events in this run; warp instructions and DRAM bytes: ncu captures at commit {facts_commit} (profiles/ncu_kernels.json) of "
                     + ("the tree's own wave launches, summed over the levels of one 10^6-lane step" if ("ensemble" in out and wave_facts and have)
                        else "the single-program launch, scaled linearly by the lanes processed"),
           "facts_stale": facts_changed or None,       # kernel source files edited since the capture: the counts may no longer describe this build
           "bound": "issue", "kernel": kname, "unit": "warp-inst/s", "peak": issue_peak, "peak_source": f"{sms} SMs x 4 schedulers x {sm_hz / 1e6:.0f} MHz (SM clock sampled under load)",
           "achieved": inst / sec if have else None, "frac": inst / sec / issue_peak if have else None,
           "traffic": dram if have else None, "traffic_source": "profiles/ncu_kernels.json (dram__bytes_read.sum + dram__bytes_write.sum of the ncu --set full captures, scaled to the lanes of this step)",
           "us_per_step": sec * 1e6, "share_of_step": sum(r["share"] for r in fused), "per_kernel": per_kernel,
           "hbm_equivalent": {"bytes": table_bytes, "gbs": table_bytes / sec / 1e9, "frac_of_hbm_peak": table_bytes / sec / 1e9 / peak, "hbm_peak_gbs": peak, "peak_source": which,
                              "note": "SURVEY 8d per-op algorithmic bytes of the recorded ops x lanes processed: what the per-op path would move through HBM for the same work; "
                                      "the fused launches keep the states in shared memory and move only `traffic`"},
           "hbm_frac_real": dram / sec / 1e9 / peak if have else None,
           "alu": {"philox_draws_per_s": info["draws"] * lane_runs / sec, "state_ops_per_s": info["state_ops"] * lane_runs / g.ntraj / sec,
                   "draws_per_trajectory": info["draws"], "recorded_ops": info["ops"]}}
    return blk


def _emit(line: dict):
    """The ONE JSON line goes to the process's real stdout; everything else printed while the bench runs (NCCL's
    version banner, library chatter) was redirected to stderr so that stdout stays machine-readable."""
    os.write(_REAL_STDOUT, (json.dumps(line) + "\n").encode())


HEADLINE_MAX_BYTES = 4000      # the driver keeps a tail of stdout: round 2's 34.6 KB line could not be parsed from it


def _short(x, n=160):
    return x if not isinstance(x, str) or len(x) <= n else x[: n - 3] + "..."


def _headline(full: dict) -> dict:
    """The compact record for stdout (<= HEADLINE_MAX_BYTES).  The full record (kernel tables, per-workload extras, logs) goes to
    the side file named in `details`.  Long prose is cut, then optional keys are dropped, until the line fits."""
    keys = ("metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling", "vs_baseline", "dtype", "data",
            "config", "clocks", "e2e", "gpu_launches", "roofline", "cpu_baseline", "parity_check", "censor_bias", "ensemble", "state_ops_per_sec",
            "useful_state_ops_per_sec", "cpu_baseline_optimised", "alive_at_end_fraction", "extra", "details")
    line = {k: full[k] for k in keys if k in full}
    if isinstance(line.get("config"), dict):
        line["config"] = {k: _short(v) for k, v in line["config"].items()}
    if isinstance(line.get("clocks"), dict):
        line["clocks"] = {k: line["clocks"][k] for k in ("sm_mhz", "sm_max_mhz", "reasons", "power_w_max", "samples") if k in line["clocks"]}
    if isinstance(line.get("e2e"), dict):
        line["e2e"] = {k: _short(v, 120) for k, v in line["e2e"].items()}
    if isinstance(line.get("roofline"), dict):
        r = line["roofline"]
        line["roofline"] = {k: _short(r[k], 120) for k in ("bound", "kernel", "achieved", "peak", "unit", "frac", "traffic", "peak_source", "modelled", "facts_stale", "note", "source") if k in r}
    for k in ("cpu_baseline", "cpu_baseline_optimised"):
        if isinstance(line.get(k), dict):
            line[k] = {kk: _short(v, 220) for kk, v in line[k].items() if kk != "note"}
    if isinstance(line.get("parity_check"), dict):
        line["parity_check"] = {k: _short(v, 100) for k, v in line["parity_check"].items()}
    if isinstance(line.get("censor_bias"), dict):
        line["censor_bias"] = {k: _short(v, 60) for k, v in line["censor_bias"].items()
                               if k in ("trajectories", "censored_gpu", "max_abs_diff_of_ensemble_means", "rms_diff_of_ensemble_means", "uncensored_op_lists", "error")}
    if isinstance(line.get("ensemble"), dict):
        e = line["ensemble"]
        line["ensemble"] = {k: e[k] for k in ("complete_fraction", "censored_fraction", "lane_runs_per_trajectory", "programs", "levels", "max_forks_per_trajectory",
                                              "tree_record_ms") if k in e}
    if isinstance(line.get("extra"), dict):      # one number per extra workload (same unit as `value`); everything else is in the side file
        line["extra"] = {n: (float(f"{v['value']:.4g}") if isinstance(v.get("value"), (int, float)) else "error: " + _short(str(v.get("error")), 60))
                         for n, v in line["extra"].items()}
    for drop in (None, "extra", "cpu_baseline_optimised", "alive_at_end_fraction", "parity_check", "ensemble"):
        if drop:
            line.pop(drop, None)
        if len(json.dumps(line)) <= HEADLINE_MAX_BYTES:
            break
    return line


def _write_details(full: dict, path: str):
    """Full record -> side file (and nothing on stdout).  Returns the path written, or None."""
    try:
        os.makedirs(os.path.dirname(os.path.abspath(path)), exist_ok=True)
        with open(path, "w") as f:
            json.dump(full, f, indent=1)
        return path
    except OSError as e:
        print(f"[bench] could not write {path}: {e}", file=sys.stderr)
        return None


_REAL_STDOUT = 1


def main():
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="firstgenrepeater_clifford", choices=sorted(WORKLOADS))
    ap.add_argument("--ntraj", type=int, default=0, help="trajectories per GPU (default: the config's)")
    ap.add_argument("--extras", default=DEFAULT_EXTRAS, help="extra workloads reported under 'extra' (N=1 only, except sweep12_dm)")
    ap.add_argument("--extras-out", default="", help="side file for the full record (kernel tables, per-workload extras); default gpurun_out/bench_extra_n<N>.json")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--mode", default="auto", choices=["auto", "per_op", "program", "cohort"],
                    help="cohort: tree of recorded programs, every trajectory on its own control flow; program: one recorded op list "
                         "(lockstep masks) replayed as one fused launch; per_op: one launch per backend call")
    ap.add_argument("--cohort-depth", type=int, default=3, help="forks allowed per trajectory before it is censored")
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"], help="strong: the config's trajectory count is divided by the GPU count")
    ap.add_argument("--chunks", type=int, default=1, help="batches of ntraj resident trajectories per step")
    args = ap.parse_args()
    if args.impl == "b200" and not os.environ.get("QSB_PROFILING"):  # under ncu a short warm-up is fine; never for a reported number
        args.warmup = max(args.warmup, 3)
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local_rank = int(os.environ.get("LOCAL_RANK", 0))
    w = WORKLOADS[args.workload]
    ntraj = args.ntraj or w["ntraj"]
    if args.scaling == "strong":
        ntraj = (ntraj // max(world, args.gpus) + 31) // 32 * 32
    mode = args.mode if args.mode != "auto" else w["mode"]
    config = {"workload": f"{args.workload}: {w['desc']}", "trajectories_per_gpu": ntraj * args.chunks, "seed": SEED,
              "l2": "L2 flushed (256 MiB memset) between timed steps; per-step state footprint >> L2 for the dm workloads",
              "ensemble": ("cohort tree: every trajectory follows its own control flow (the reference's semantics), at most %d forks per trajectory" % args.cohort_depth) if mode == "cohort"
                          else "lockstep, shared classical schedule, per-trajectory masks (DESIGN.md)"}

    if args.impl == "reference":
        # the reference's own CPU path (compiled oracle port: Julia is absent, DESIGN.md "Oracle"), all host cores, rank 0 only
        if rank != 0:
            return
        n = CPU_SAMPLE.get(args.workload, (0, 0))[0] or 1000
        tree = make_trace_tree(args.workload, args.cohort_depth if mode == "cohort" else None)
        vals, c = [], None
        for _ in range(max(1, min(args.warmup, 1))):
            c = cpu_compiled(args.workload, n, None, tree=tree)
        for _ in range(args.steps):
            c = cpu_compiled(args.workload, n, None, tree=tree)
            vals.append(c["value"])
        v = float(np.median(vals))
        sample = (f"{n} trajectories per step replayed by oracle/qsb_replay.c on {c['cores']} OpenMP threads ({c['complete']} complete, {c['censored']} censored), "
                  f"{args.steps} steps (median); {c['nodes']} op lists traced untimed in {c['trace_s']:.1f} s")
        _emit(({"impl": "reference", "metric": "trajectories/sec", "value": v, "unit": "trajectories/s", "n_gpus": args.gpus,
                "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * c["seconds"], "higher_is_better": True,
                "scaling": args.scaling, "vs_baseline": None, "dtype": "u32 bit-sliced tableau" if w["formalism"] == "clifford" else "f64 complex",
                "data": "synthetic", "config": config,
                "cpu_baseline": {"value": v, "unit": "trajectories/s", "cores": c["cores"], "kind": "port-c-openmp", "sample": sample},
                "e2e": {"value": v, "unit": "trajectories/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))
        return

    import torch
    import torch.distributed as dist
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    config["mode"] = {"program": "recorded program: host schedule captured once (untimed, program.record_ms), every step = ONE fused launch pair over a fresh batch of trajectories",
                      "cohort": "cohort tree of recorded programs (grown during warm-up, untimed: ensemble.tree_record_ms); every step = one fused launch pair per tree level over a fresh batch",
                      "per_op": "per-op: host schedule re-run every step, one launch per backend call"}[mode]
    kept = []
    main_res = run_gpu_workload(args.workload, ntraj, args.steps, args.warmup, rank, world, local_rank, mode=mode, cohort_depth=args.cohort_depth, chunks=args.chunks,
                                keep=kept if (mode == "cohort" and rank == 0 and world == 1 and not args.no_cpu) else None)
    cpu = cpu_py = censor_bias = None
    parity = main_res.get("parity_check")
    if rank == 0 and not args.no_cpu and CPU_SAMPLE.get(args.workload, (0, 0))[0]:
        # rank 0 at every N (same sample, all host cores), so that each line carries its own same-run CPU denominator
        cpu, cpu_py, craw = cpu_baseline_for(args.workload, args.cohort_depth if mode == "cohort" else None, want_python=(world == 1))
        if kept:
            # cross-check of the two arms: the same trajectory ids through the GPU tree and the compiled CPU replay
            n = 4096
            c = cpu_compiled(args.workload, n, args.cohort_depth, tree=craw["tree"])
            parity = parity_check_tree(kept[0], c, n)
            try:
                censor_bias = censor_bias_check(kept[0], args.workload, 2048)
            except Exception as e:   # a diagnostic: it must not take the headline down
                censor_bias = {"error": repr(e)[:200]}
    for g in kept:
        g.tree.free(); g.ctx.close()
    extra = {}
    keep = ("cpu_baseline", "cpu_baseline_python", "ntraj_per_gpu", "dev_ms_per_step", "value", "e2e", "launches_per_step", "algo_bytes_per_step", "ops_per_step", "roofline", "kernels",
            "mode", "program", "log", "parity_check", "e2e_incl_record", "alive_at_end_fraction", "ensemble", "cpu_baseline_optimised")
    if world == 1 and mode in ("program", "cohort") and args.extras:
        try:  # the same workload through the per-call drop-in path (lockstep masks), for comparison
            r = run_gpu_workload(args.workload, ntraj, max(2, min(args.steps, 3)), 3, 0, 1, local_rank, mode="per_op")
            extra[args.workload + "_per_op"] = {k: r[k] for k in keep if k in r}
        except Exception as e:
            extra[args.workload + "_per_op"] = {"error": repr(e)[:300]}
    if args.extras:
        for name in [x for x in args.extras.split(",") if x and x != args.workload]:
            if world > 1 and name != "sweep12_dm":       # BASELINE configs[4] is asked for at 1/2/4/8 GPUs; the others are N=1 lines
                continue
            try:
                r = run_gpu_workload(name, WORKLOADS[name]["ntraj"], max(2, min(args.steps, 3)), 3, rank, world, local_rank, mode=WORKLOADS[name]["mode"], profile=(world == 1))
                extra[name] = {k: r[k] for k in keep if k in r}
                extra[name]["state_ops_per_sec"] = r["ops_per_step"] * world / (r["dev_ms_per_step"] * 1e-3)
                if rank == 0 and world == 1 and not args.no_cpu:   # a short compiled-CPU sample of the same workload
                    base, _, _ = cpu_baseline_for(name, 2, want_python=False)   # fork budget 2 bounds the number of op lists to trace
                    if base:
                        extra[name]["cpu_baseline"] = base
            except Exception as e:  # an extra must never take the headline down
                extra[name] = {"error": repr(e)[:300]}
    if rank == 0:
        line = {"metric": "trajectories/sec", "value": main_res["value"], "unit": "trajectories/s", "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": main_res["dev_ms_per_step"], "higher_is_better": True, "scaling": args.scaling,
                "vs_baseline": None, "dtype": "u32 bit-sliced tableau" if w["formalism"] == "clifford" else "f64 complex", "data": "synthetic",
                "config": config, "clocks": main_res["clocks"],
                "e2e": {"value": main_res["e2e"], "unit": "trajectories/s", "h2d_bytes_per_step": main_res["h2d_bytes"], "d2h_bytes_per_step": main_res["d2h_bytes"],
                        "incl_recording": main_res.get("e2e_incl_record"),
                        "note": {"program": "wall clock around upload of the host op list (pinned) + qsb_program_run + all-reduce + D2H of the step's statistics",
                                 "cohort": "wall clock around the whole tree run: launches, per-level cohort routing (D2H of cohort keys and sizes, H2D of the job tables; the recorded "
                                           "programs stay resident on the device), all-reduce + D2H of the step's statistics",
                                 "per_op": "wall clock around the public API calls incl. the all-reduce + D2H of the step's statistics; inputs are host-side symbols/matrices passed by value through the C-ABI"}[mode]},
                "program": main_res.get("program"), "parity_check": parity, "censor_bias": censor_bias, "ensemble": main_res.get("ensemble"),
                "alive_at_end_fraction": main_res.get("alive_at_end_fraction"),
                "gpu_launches": int(main_res["launches_per_step"] * args.steps),
                "state_ops_per_sec": main_res["ops_per_step"] * world / (main_res["dev_ms_per_step"] * 1e-3),
                "roofline": main_res.get("roofline"), "kernels": main_res.get("kernels"), "cpu_baseline": cpu, "cpu_baseline_python": cpu_py,
                "cpu_baseline_optimised": main_res.get("cpu_baseline_optimised"),
                "workload_log": main_res["log"], "extra": extra}
        ens = main_res.get("ensemble") or {}
        # ops of the lanes' FIRST run only: a cohort tree re-runs a forked lane from t = 0, state_ops_per_sec counts that work too
        line["useful_state_ops_per_sec"] = line["state_ops_per_sec"] / ens["lane_runs_per_trajectory"] if ens.get("lane_runs_per_trajectory") else line["state_ops_per_sec"]
        line["details"] = _write_details(line, args.extras_out or os.path.join(ROOT, "gpurun_out", f"bench_extra_n{world}.json"))
        head = _headline(line)
        assert len(json.dumps(head)) <= HEADLINE_MAX_BYTES, len(json.dumps(head))
        _emit(head)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
