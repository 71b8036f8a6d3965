#!/usr/bin/env python
"""bench.py -- trajectories/sec of the register-state hot path on B200 (contract in DESIGN.md "Measurement").

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--workload NAME] [--ntraj M]

One "step" = one complete pass of the workload's event schedule over one lockstep ensemble of
`ntraj` trajectories per GPU (weak scaling: every GPU owns its own block of global trajectory ids).
Default workload = BASELINE.json configs[1]: firstgenrepeater, CliffordRepr tableaux, depolarizing
noise, 10^6 trajectories per GPU.  Rank 0 prints ONE JSON line.
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

import numpy as np  # noqa: E402

SEED = 0x5153
FGR = dict(sizes=(2, 3, 4, 3, 2), F=0.97, t_end=25.0, dt_sample=0.2)  # 6.1_compare_formalisms_noplot.jl:11-21


# ------------------------------------------------------------------------------------------------
# workloads (shared by the GPU arm and the CPU arm: the same host code drives either backend)
# ------------------------------------------------------------------------------------------------
def wl_fgr(make_repr, formalism, record):
    from qsb200 import symbols as S
    from qsb200.workloads import firstgenrepeater as fgr
    bg = S.Depolarization(100.0) if formalism == "clifford" else S.T1T2Noise(1000.0, 100.0)
    from qsb200.registers import dispose
    kw = dict(FGR)
    if formalism == "mc":   # a mixture would be lowered to a density matrix (express of a mixed state): pure pairs keep the states kets
        from qsb200 import ops
        kw["noisy_pair"] = ops.StabilizerState("XX ZZ")
    _, _, net = fgr.monte_carlo_trajectory(make_repr, background=bg, record=record, **kw)
    dispose(net.registers)
    return net.log


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


WORKLOADS = {
    "firstgenrepeater_clifford": dict(fn=wl_fgr, formalism="clifford", ntraj=1_000_000,
                                      desc="examples/firstgenrepeater_lowlevel, sizes [2,3,4,3,2], F=0.97, CliffordRepr, Depolarization(100), t=0:0.2:25"),
    "firstgenrepeater_dm": dict(fn=wl_fgr, formalism="qo", ntraj=1_000_000, program=True,
                                desc="examples/firstgenrepeater_lowlevel, sizes [2,3,4,3,2], F=0.97, QuantumOpticsRepr dm, T1T2Noise(1000,100), t=0:0.2:25"),
    "firstgenrepeater_mcket": dict(fn=wl_fgr, formalism="mc", ntraj=1_000_000, program=False,
                                   desc="examples/firstgenrepeater_lowlevel, sizes [2,3,4,3,2], perfect pairs, QuantumMCRepr kets, T1T2Noise(1000,100) as sampled Kraus jumps, t=0:0.2:25"),
    "ghz_sensors_dm": dict(fn=wl_ghz, formalism="qo", ntraj=16384,
                           desc="examples/distributed_ghz_sensors F-TMBL trial, S=4 all links up (peak 8-qubit dm = 1 MiB per trajectory), F=0.99, Depolarization(1.0); "
                                "10^6 trajectories = 62 steps of 16384 resident trajectories"),
    "repeatergrid_clifford": dict(fn=wl_grid, formalism="clifford", ntraj=10_000_000,
                                  desc="examples/repeatergrid 5x5 grid, 4 slots per node, CliffordRepr, Depolarization(10) (T1Decay has no Clifford method), success_prob=0.001, t_end=30"),
    "sweep12_dm": dict(fn=wl_sweep12, formalism="qo", ntraj=256,
                       desc="12-qubit composed dm register: random H/X/Z + CNOT, T1T2Noise(100,10) on every qubit, project_traceout!, recompose"),
}


# ------------------------------------------------------------------------------------------------
# CPU arm: the oracle port of the reference's CPU path, one trajectory at a time on the host cores
# ------------------------------------------------------------------------------------------------
def _cpu_worker(args):
    workload, traj_lo, traj_hi = args
    sys.path.insert(0, ROOT)
    import qsb200  # noqa: F401
    from tests.oracle_backend import OracleContext, OracleRepr
    w = WORKLOADS[workload]
    acc = {}
    t0 = time.perf_counter()
    for t in range(traj_lo, traj_hi):
        ctx = OracleContext(SEED, t)

        def record(kind, tm, val, acc=acc):
            a = acc.setdefault((kind, round(tm, 9)), [0, 0.0])
            a[0] += 1; a[1] += float(np.real(complex(val)))
        w["fn"](lambda: OracleRepr(ctx, w["formalism"]), w["formalism"], record)
    return traj_hi - traj_lo, time.perf_counter() - t0, acc


def cpu_run(workload, per_core, cores=None):
    """Times `per_core` trajectories on each host core (processes, not threads: the driver is Python)."""
    import multiprocessing as mp
    cores = cores or os.cpu_count() or 1
    jobs = [(workload, c * per_core, (c + 1) * per_core) for c in range(cores)]
    t0 = time.perf_counter()
    with mp.get_context("fork").Pool(cores) as pool:
        res = pool.map(_cpu_worker, jobs)
    wall = time.perf_counter() - t0
    ntraj = sum(r[0] for r in res)
    acc = {}
    for r in res:
        for k, (c, s) in r[2].items():
            a = acc.setdefault(k, [0, 0.0]); a[0] += c; a[1] += s
    return dict(ntraj=ntraj, wall=wall, cores=cores, stats=acc, busy=max(r[1] for r in res))


def cpu_sample_size(workload):
    return {"firstgenrepeater_clifford": 60, "firstgenrepeater_dm": 100, "firstgenrepeater_mcket": 100, "sweep12_dm": 0, "ghz_sensors_dm": 2, "repeatergrid_clifford": 40}.get(workload, 20)


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


def ncu_traffic(workload, ntraj, kernel):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch from the committed ncu --set full capture of this exact
    workload / size / kernel (profiles/ncu_traffic.json), or None when there is no capture for it."""
    try:
        with open(os.path.join(ROOT, "profiles", "ncu_traffic.json")) as f:
            return json.load(f).get(f"{workload}|{ntraj}|{kernel}")
    except Exception:
        return None


def measured_peak():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class GpuRun:
    """One workload on this rank's GPU: lockstep ensemble behind the C-ABI, statistics kept on the device."""

    def __init__(self, workload, ntraj, rank, world, local_rank):
        import torch
        import qsb200  # noqa: F401
        from qsb200.backend import BatchContext, CudaBatchRepr
        from qsb200.sharding import StatsBuffer
        self.torch = torch
        self.w = WORKLOADS[workload]
        self.ntraj, self.rank, self.world = ntraj, rank, world
        torch.cuda.set_device(local_rank)
        self.device = torch.device("cuda", local_rank)
        self.stream = torch.cuda.Stream(device=self.device)
        self.ctx = BatchContext(ntraj, seed=SEED, device=local_rank, traj_offset=rank * ntraj, stream=self.stream.cuda_stream)
        self.CudaBatchRepr = CudaBatchRepr
        self.stats = StatsBuffer(1024, self.device)
        self.flush = torch.empty(256 << 20, dtype=torch.uint8, device=self.device)  # > 126 MB L2
        self.h2d_bytes = 0
        self.per_op_batch = 0

    def step(self):
        """One pass of the workload; returns (#observables recorded, host log).  No host sync inside."""
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

        # every observable() that reaches the backend leaves its statistics in its own device row (no host sync)
        from qsb200.registers import observable_options

        def next_row():
            i = slot[0]; slot[0] += 1
            return stats.slot_ptr(i)
        with observable_options(stats_dev=next_row, sync=False), self.torch.cuda.stream(self.stream):
            log = self.w["fn"](make_repr, self.w["formalism"], record)
        self.nobs = slot[0]
        self.keys = keys
        return log

    def record_program(self):
        """Runs the host schedule ONCE in record mode (nothing launches) and keeps the fused program.
        Later steps replay it on fresh trajectory ids: the classical schedule is shared by the ensemble."""
        t0 = time.perf_counter()
        with self.ctx.record() as rec:
            log = self.step()
        self.program = rec.program
        self.program_log = log
        self.record_ms = 1e3 * (time.perf_counter() - t0)
        self.batch = 0
        return log

    def upload_program(self):
        with self.torch.cuda.stream(self.stream):
            self.h2d_bytes = self.program.upload()

    def run_program(self):
        """One fused launch over a fresh batch of ntraj trajectories per GPU (global ids never repeat)."""
        self.batch += 1
        self.program.run(traj_shift=self.batch * self.ntraj * self.world)
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


def run_gpu_workload(workload, ntraj, steps, warmup, rank, world, local_rank, profile=True, mode="per_op"):
    """mode "per_op": every backend call is one launch (the drop-in path, host schedule re-run each step).
    mode "program": the schedule is recorded once (untimed, reported as record_ms) and each step replays it as
    ONE fused launch on a fresh batch of trajectories; e2e additionally re-sends the op list from the host."""
    import torch
    import torch.distributed as dist
    g = GpuRun(workload, ntraj, rank, world, local_rank)
    ctx = g.ctx
    program = mode == "program"
    if program:
        g.record_program()
    do_step = g.run_program if program else g.step

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(warmup):
        do_step(); g.finish_step()
    barrier()
    ctx.reset_counters()
    sampler = ClockSampler(local_rank)
    sampler.start()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    dev_ms = 0.0
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
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dev_ms, wall_ms = float(t[0]), float(t[1])
    out = dict(workload=workload, ntraj_per_gpu=ntraj, dev_ms_per_step=dev_ms / steps, wall_ms_per_step=wall_ms / steps,
               value=ntraj * world * steps / (dev_ms / 1e3), e2e=ntraj * world * steps / (wall_ms / 1e3),
               launches_per_step=counters["launches"] / steps, algo_bytes_per_step=counters["algo_bytes"] / steps,
               ops_per_step=counters["ops"] / steps, nobs=g.nobs, d2h_bytes=g.nobs * 32, clocks=clocks, log=log,
               sample_stats=host_stats[: min(4, len(host_stats))].tolist(), mode=mode, h2d_bytes=g.h2d_bytes)
    if program:
        out["program"] = dict(g.program.info(), record_ms=g.record_ms)
        # amortised view: the one-off recording of the schedule (host time) spread over the timed steps
        out["e2e_incl_record"] = ntraj * world * steps / ((wall_ms + g.record_ms) / 1e3)
        if rank == 0:
            out["parity_check"] = parity_check(g, workload, ntraj, rank, world, local_rank)
    if profile and rank == 0:
        ctx.profile(True)
        do_step(); g.finish_step(reduce=False)
        rows = ctx.profile_report()
        ctx.profile(False)
        rows = [r for r in rows if r["launches"]]
        total = sum(r["ms"] for r in rows) or 1.0
        rows.sort(key=lambda r: -r["ms"])
        peak, which = measured_peak()
        for r in rows:
            r["share"] = r["ms"] / total
            r["gbs"] = r["algo_bytes"] / (r["ms"] * 1e-3) / 1e9 if r["ms"] > 0 else 0.0
            r["us_per_launch"] = 1e3 * r["ms"] / r["launches"]
        top = rows[0]
        out["kernels"] = rows[:8]
        out["roofline"] = {"bound": "hbm", "kernel": top["kernel"], "achieved": top["gbs"], "peak": peak, "unit": "GB/s",
                           "frac": top["gbs"] / peak, "traffic": None, "peak_source": which,
                           "algo_bytes_per_launch": top["algo_bytes"] / top["launches"], "us_per_launch": top["us_per_launch"],
                           "share_of_step": top["share"]}
        if program:
            # The fused pair k_draw_bits + k_tab_exec executes the recorded ops of ALL trajectories in one pass with the
            # tableaux resident in shared memory.  roofline.achieved follows the task's definition: SURVEY 8d's per-op
            # algorithmic bytes x the ops one run processes (= what the same op list accounts on the per-op path),
            # divided by the run's device time (both kernels).  What the run really has to move through HBM is far
            # less (decision words, op list, statistics): reported as fused_hbm_*.
            info = out["program"]
            fused = [r for r in rows if r["kernel"] in ("k_tab_exec", "k_draw_bits", "k_dense_exec")]
            sec = sum(r["ms"] for r in fused) * 1e-3 / max(1, fused[0]["launches"])
            hbm = sum(r["algo_bytes"] for r in fused) / max(1, fused[0]["launches"])
            gbs = info["table_bytes"] / sec / 1e9
            kname = "+".join(sorted(r["kernel"] for r in fused))
            traffic = ncu_traffic(workload, ntraj, kname)
            out["roofline"] = {"bound": "hbm", "kernel": kname, "achieved": gbs, "peak": peak, "unit": "GB/s", "frac": gbs / peak,
                               "traffic": traffic["bytes"] if traffic else None, "traffic_source": traffic["source"] if traffic else None, "peak_source": which, "algo_bytes_per_launch": info["table_bytes"], "us_per_launch": sec * 1e6,
                               "share_of_step": sum(r["share"] for r in fused),
                               "fused_hbm_bytes_per_launch": hbm, "fused_hbm_gbs": hbm / sec / 1e9,
                               "note": ("algorithmic bytes = SURVEY 8d table bytes of the recorded ops (per-op accounting of the same op list); the fused run keeps "
                                        "the states in shared memory, so its compulsory HBM traffic is only fused_hbm_bytes_per_launch and the kernels are "
                                        "bound on chip (Clifford: Philox4x32-10 draws + bit-sliced row logic; dense: fp64 pipe + shared-memory bandwidth), see alu"),
                               "alu": {"philox_draws_per_s": info["draws"] * ntraj / sec, "state_ops_per_s": info["state_ops"] / sec,
                                       "draws_per_trajectory": info["draws"], "recorded_ops": info["ops"]}}
    ctx.close()
    return out


def _emit(line: dict):
    """The ONE JSON line goes to the process's real stdout; everything else printed while the bench runs (NCCL's
    version banner, library chatter) was redirected to stderr so that stdout stays machine-readable."""
    os.write(_REAL_STDOUT, (json.dumps(line) + "\n").encode())


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
    ap.add_argument("--extras", default="firstgenrepeater_dm,firstgenrepeater_mcket,ghz_sensors_dm,repeatergrid_clifford,sweep12_dm", help="extra workloads reported under 'extra' (N=1 only)")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--mode", default="auto", choices=["auto", "per_op", "program"],
                    help="program: recorded op list replayed as one fused launch (Clifford workloads); per_op: one launch per backend call")
    args = ap.parse_args()
    if args.impl == "b200" and not os.environ.get("QSB_PROFILING"):  # under ncu a short warm-up is fine; never for a reported number
        args.warmup = max(args.warmup, 3)
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local_rank = int(os.environ.get("LOCAL_RANK", 0))
    w = WORKLOADS[args.workload]
    ntraj = args.ntraj or w["ntraj"]
    config = {"workload": f"{args.workload}: {w['desc']}", "trajectories_per_gpu": ntraj, "seed": SEED,
              "l2": "L2 flushed (256 MiB memset) between timed steps; per-step state footprint >> L2 for the dm workloads",
              "ensemble": "lockstep, shared classical schedule, per-trajectory masks (DESIGN.md)"}

    if args.impl == "reference":
        # the reference's own CPU path (oracle port: Julia is absent, DESIGN.md "Oracle"), all host cores, rank 0 only
        if rank != 0:
            return
        per_core = max(1, cpu_sample_size(args.workload) // 4)
        vals = []
        for _ in range(max(1, args.warmup > 0)):
            cpu_run(args.workload, 2)
        for _ in range(args.steps):
            r = cpu_run(args.workload, per_core)
            vals.append(r["ntraj"] / r["wall"])
        v = float(np.median(vals))
        sample = f"{per_core} trajectories per core x {r['cores']} cores per step, {args.steps} steps (median)"
        _emit(({"impl": "reference", "metric": "trajectories/sec", "value": v, "unit": "trajectories/s", "n_gpus": args.gpus,
                          "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * r["wall"], "higher_is_better": True,
                          "scaling": "weak", "vs_baseline": None, "dtype": "u32 bit-sliced tableau" if w["formalism"] == "clifford" else "f64 complex",
                          "data": "synthetic", "config": config,
                          "cpu_baseline": {"value": v, "unit": "trajectories/s", "cores": r["cores"], "kind": "port", "sample": sample},
                          "e2e": {"value": v, "unit": "trajectories/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))
        return

    import torch
    import torch.distributed as dist
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    mode = args.mode if args.mode != "auto" else ("program" if w.get("program", w["formalism"] == "clifford") else "per_op")
    config["mode"] = ("recorded program: host schedule captured once (untimed, program.record_ms), every step = ONE fused launch over a fresh batch of trajectories"
                      if mode == "program" else "per-op: host schedule re-run every step, one launch per backend call")
    main_res = run_gpu_workload(args.workload, ntraj, args.steps, args.warmup, rank, world, local_rank, mode=mode)
    extra = {}
    keep = ("cpu_baseline", "ntraj_per_gpu", "dev_ms_per_step", "value", "e2e", "launches_per_step", "algo_bytes_per_step", "ops_per_step", "roofline", "kernels", "mode", "program", "log",
            "parity_check", "e2e_incl_record")
    if world == 1 and mode == "program" and args.extras:
        try:  # the same workload through the per-call drop-in path, for comparison
            r = run_gpu_workload(args.workload, ntraj, max(2, min(args.steps, 3)), 3, 0, 1, local_rank, mode="per_op")
            extra[args.workload + "_per_op"] = {k: r[k] for k in keep if k in r}
        except Exception as e:
            extra[args.workload + "_per_op"] = {"error": repr(e)[:300]}
    if world == 1 and args.extras:
        for name in [x for x in args.extras.split(",") if x and x != args.workload]:
            try:
                r = run_gpu_workload(name, WORKLOADS[name]["ntraj"], max(2, min(args.steps, 3)), 3, 0, 1, local_rank,
                                     mode="program" if WORKLOADS[name].get("program", WORKLOADS[name]["formalism"] == "clifford") else "per_op")
                extra[name] = {k: r[k] for k in keep if k in r}
                extra[name]["state_ops_per_sec"] = r["ops_per_step"] / (r["dev_ms_per_step"] * 1e-3)
                per_core = max(2, cpu_sample_size(name) // 3) if cpu_sample_size(name) else 0
                if per_core and not args.no_cpu:   # a short CPU sample of the same workload (oracle port), a few seconds per extra
                    c = cpu_run(name, per_core)
                    extra[name]["cpu_baseline"] = {"value": c["ntraj"] / c["wall"], "unit": "trajectories/s", "cores": c["cores"], "kind": "port",
                                                   "sample": f"{per_core} full trajectories per core on {c['cores']} processes ({c['wall']:.1f} s wall)"}
            except Exception as e:  # an extra must never take the headline down
                extra[name] = {"error": repr(e)[:300]}
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu and cpu_sample_size(args.workload):
        per_core = cpu_sample_size(args.workload)
        r = cpu_run(args.workload, per_core)
        cpu = {"value": r["ntraj"] / r["wall"], "unit": "trajectories/s", "cores": r["cores"], "kind": "port",
               "sample": f"{per_core} full trajectories per core on {r['cores']} processes ({r['wall']:.1f} s wall): oracle port driven by the same host schedule code"}
    if rank == 0:
        line = {"metric": "trajectories/sec", "value": main_res["value"], "unit": "trajectories/s", "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": main_res["dev_ms_per_step"], "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "u32 bit-sliced tableau" if w["formalism"] == "clifford" else "f64 complex", "data": "synthetic",
                "config": config, "clocks": main_res["clocks"],
                "e2e": {"value": main_res["e2e"], "unit": "trajectories/s", "h2d_bytes_per_step": main_res["h2d_bytes"], "d2h_bytes_per_step": main_res["d2h_bytes"],
                        "note": ("wall clock around upload of the host op list (pinned) + qsb_program_run + all-reduce + D2H of the step's statistics" if mode == "program" else
                                 "wall clock around the public API calls incl. the all-reduce + D2H of the step's statistics; inputs are host-side symbols/matrices passed by value through the C-ABI")},
                "program": main_res.get("program"), "parity_check": main_res.get("parity_check"),
                "gpu_launches": int(main_res["launches_per_step"] * args.steps),
                "state_ops_per_sec": main_res["ops_per_step"] * world / (main_res["dev_ms_per_step"] * 1e-3),
                "roofline": main_res.get("roofline"), "kernels": main_res.get("kernels"), "cpu_baseline": cpu,
                "workload_log": main_res["log"], "extra": extra}
        _emit(line)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
