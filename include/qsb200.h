/*
 * qsb200.h -- C ABI of libqsb200.so: a batched (N Monte-Carlo trajectories in lockstep)
 * register-state backend for QuantumSavory's backend API, sm_100a only.
 *
 * Every entry point replaces one backend generic that QuantumSavory's register layer
 * (src/baseops/ *.jl files) calls on the native state object.  File:line citations are relative to the
 * reference checkout (QuantumSavory.jl v0.7.2).  The Julia `ccall` glue that binds these symbols
 * is shown in INTEGRATION.md; quantumsavory.jl_b200/_lib.py binds the same symbols with ctypes.
 *
 * Conventions
 *   - plain pointers and sizes only; all handles are opaque; no CUDA/torch types in signatures
 *     (a cudaStream_t is passed as void*).
 *   - subsystem indices are 0-BASED here (reference `stateindex` minus one).
 *   - complex data is interleaved (re, im) double; matrices column-major; subsystem 0 is the
 *     least-significant bit of a basis index (QuantumOpticsBase: tensor(a,b).data = kron(b,a)).
 *   - every function returns a qsb_status; qsb_last_error() has the message (thread-local).
 *   - a state handle is single-owner and is mutated in place, *including* its size and kind
 *     (project_traceout!/traceout! shrink it, uptotime! promotes ket -> dm): the Julia methods
 *     return the same wrapper object, which is what src/baseops/apply.jl:20-21, uptotime.jl:23 and
 *     traceout.jl:50-51,150 store back into the StateRef.
 *   - calls on one context are issued in order on its stream; only functions documented as
 *     "host-visible" synchronise.
 *   - randomness is counter based: u = Philox4x32-10(key=seed, ctr=(global trajectory id,
 *     draw_id)) -> 53-bit uniform in [0,1).  The caller numbers stochastic calls with draw_id
 *     (one id per call site execution, shared by all trajectories), so an oracle fed the same
 *     (seed, trajectory, draw_id) consumes the same draws ("RNG").
 */
#ifndef QSB200_H
#define QSB200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct qsb_ctx qsb_ctx;     /* one trajectory shard on one GPU */
typedef struct qsb_state qsb_state; /* one batched native state (what StateRef.state[] holds) */
typedef struct qsb_mask qsb_mask;   /* one bit per trajectory (outcome==2, alive, success, ...) */

enum qsb_status {
    QSB_OK = 0,
    QSB_ERR_CUDA = 1,
    QSB_ERR_BAD_INDEX = 2,      /* BoundsError-like */
    QSB_ERR_UNSUPPORTED = 3,    /* MethodError-like: no method for this (state kind, op) pair */
    QSB_ERR_NOT_NORMALIZED = 4, /* "State not normalized. Could be due to passing wrong state to `initialize!`"
                                   (quantumoptics.jl:52) or a failed `@assert sum(P) ~ 1` (:60) */
    QSB_ERR_BAD_ARG = 5,
    QSB_ERR_OOM = 6,
    QSB_ERR_TRACE_ALL = 7       /* "Partial trace can't be used to trace out all subsystems - use tr() instead." (mcket.jl:47-49) */
};

enum qsb_kind {
    QSB_KET = 0,   /* QuantumOpticsBase.Ket      (QuantumOpticsRepr, pure) */
    QSB_DM = 1,    /* QuantumOpticsBase.Operator (QuantumOpticsRepr, mixed) */
    QSB_MCKET = 2, /* MCKet                      (QuantumMCRepr, mcket.jl:2-4) */
    QSB_TAB = 3    /* QuantumClifford.MixedDestabilizer (CliffordRepr), bit-sliced over trajectories */
};

enum qsb_background { QSB_BG_T1 = 0, QSB_BG_T2 = 1, QSB_BG_DEPOL = 2, QSB_BG_T1T2 = 3 }; /* src/backgrounds.jl */
enum qsb_basis { QSB_BASIS_X = 1, QSB_BASIS_Y = 2, QSB_BASIS_Z = 3, QSB_BASIS_KETS = 4 };
enum qsb_obs {
    QSB_OBS_PAULI = 0,     /* obs: uint8_t[k] codes 0 I,1 X,2 Y,3 Z (tensor product over idx)          */
    QSB_OBS_OP = 1,        /* obs: double[2*4^k] dense k-qubit operator, column-major (dense kinds)      */
    QSB_OBS_PROJ_KET = 2,  /* obs: double[2*2^k] ket psi; value <psi|rho|psi>            (dense kinds)   */
    QSB_OBS_PROJ_STAB = 3  /* obs: uint8_t xs[k*k], zs[k*k], signs[k] stabilizer generators (QSB_TAB)    */
};
enum qsb_mask_op { QSB_MASK_AND = 0, QSB_MASK_OR = 1, QSB_MASK_XOR = 2, QSB_MASK_ANDNOT = 3, QSB_MASK_NOT = 4, QSB_MASK_XNOR = 5 };

const char* qsb_last_error(void);
int qsb_version(void);

/* ---- context ---------------------------------------------------------------------------------
 * A context owns `ntraj` trajectories with global ids [traj_offset, traj_offset+ntraj) (section 8e:
 * shard trajectories across GPUs, Philox keyed by the GLOBAL id so results do not depend on the
 * number of GPUs).  `cuda_stream` may be NULL (a private non-blocking stream is created). */
int qsb_ctx_create(int device, int64_t ntraj, int64_t traj_offset, uint64_t seed, void* cuda_stream, qsb_ctx** out);
int qsb_ctx_destroy(qsb_ctx* ctx);
int qsb_ctx_set_stream(qsb_ctx* ctx, void* cuda_stream);
int qsb_ctx_sync(qsb_ctx* ctx); /* host-visible: waits, then reports sticky device-side errors (QSB_ERR_NOT_NORMALIZED) */
int64_t qsb_ctx_ntraj(const qsb_ctx* ctx);
/* re-bases the context on the global trajectory ids [traj_offset, traj_offset+ntraj): the next batch of the same
 * schedule (the replicate loop of 6.1_compare_formalisms_noplot.jl:66-73) without re-creating the context */
int qsb_ctx_set_traj_offset(qsb_ctx* ctx, int64_t traj_offset);
/* counters[0]=kernel launches, [1]=algorithmic bytes (SURVEY 8d table x trajectories), [2]=state ops x trajectories, [3]=draw ids consumed */
int qsb_ctx_counters(const qsb_ctx* ctx, uint64_t counters[4]);
int qsb_ctx_reset_counters(qsb_ctx* ctx);
/* Measurement aid: when enabled every kernel launch is bracketed by CUDA events on the context's
 * stream; the report is a JSON array of {kernel, launches, ms, algo_bytes} rows (host-visible). */
int qsb_ctx_profile(qsb_ctx* ctx, int enable);
int qsb_ctx_profile_report(qsb_ctx* ctx, char* buf, size_t buflen);

/* ---- state lifetime --------------------------------------------------------------------------*/
/* newstate(::Qubit, repr): |0> ket / MCKet / tableau of Z1 -- quantumoptics.jl:66-72, clifford.jl:10-13;
 * nsub>1 gives |0..0> (used by tests and the synthetic sweep). */
int qsb_state_new(qsb_ctx* ctx, int kind, int nsub, qsb_state** out);
/* initialize!(regs, indices, state): broadcast one native ket (2^n) / operator (4^n) to all
 * trajectories -- src/baseops/initialize.jl:32-45 (StateRef copies, states_registers.jl:10). */
int qsb_state_from_host(qsb_ctx* ctx, int kind, int nsub, const double* data_ri, qsb_state** out);
/* Same for a stabilizer state given by `ngen` <= nsub generators (bytes xs[g*nsub+q], zs[...],
 * signs[g] 1 = minus); ngen < nsub gives a mixed MixedDestabilizer(S"...") state. */
int qsb_tab_from_host(qsb_ctx* ctx, int nsub, int ngen, const uint8_t* xs, const uint8_t* zs, const uint8_t* signs, qsb_state** out);
/* express(sum_i w_i state_i, repr) sampled per trajectory (one draw): trajectory takes states[i] for
 * the first i with cumsum(w)_i > u -- the stochastic-mixture initial states of CliffordRepr,
 * src/StatesZoo/depolarized.jl:39-43, firstgenrepeater_lowlevel/clifford_setup.jl:10. */
int qsb_state_select(qsb_ctx* ctx, const qsb_state* const* states, const double* weights, int m, uint64_t draw_id, qsb_state** out);
int qsb_state_copy(const qsb_state* st, qsb_state** out); /* Base.copy(state) */
int qsb_state_free(qsb_state* st);                         /* Julia finalizer */
int qsb_state_kind(const qsb_state* st);
int qsb_state_nsubsystems(const qsb_state* st);            /* nsubsystems(state) */
int qsb_state_promote_dm(qsb_state* st);                   /* dm(ket) in place (uptotime.jl:2) */
/* read one trajectory back (host-visible; tests / debugging) */
int qsb_state_to_host(const qsb_state* st, int64_t traj, double* out_ri);
/* 2n rows (n destabilizer/logical-X rows then n stabilizer/logical-Z rows): xs[r*n+q], zs[r*n+q],
 * signs[r] (only meaningful for the last n rows), active[i] = 1 when stabilizer row i is a real
 * stabilizer (rank = sum(active)). */
int qsb_tab_to_host(const qsb_state* st, int64_t traj, uint8_t* xs, uint8_t* zs, uint8_t* signs, uint8_t* active);

/* ---- backend generics ------------------------------------------------------------------------*/
/* subsystemcompose(states...): tensor, first factor least significant; Ket(x)Operator promotes with
 * dm() -- quantumoptics.jl:3-6, mcket.jl:35-44, clifford.jl:1.  Inputs are left untouched. */
int qsb_tensor(qsb_ctx* ctx, const qsb_state* const* states, int m, qsb_state** out);
/* apply!(state, indices, op) for the dense kinds: psi <- U psi, rho <- U rho U', U a k-qubit
 * (k = 1, 2) operator whose factor j acts on subsystem idx[j] -- QuantumOpticsBase.apply! reached
 * from src/baseops/apply.jl:20; mcket.jl:24-26.  mask (nullable): only trajectories whose bit is set
 * are touched (outcome-conditioned corrections, src/CircuitZoo/CircuitZoo.jl:26-31). */
int qsb_apply(qsb_state* st, const int* idx, int k, const double* U_ri, const qsb_mask* mask);
/* apply!(state::MixedDestabilizer, sGate(idx...)): conjugation table of a k-qubit Clifford,
 * entry t = x0|z0<<1|x1<<2|z1<<3 -> x0'|z0'<<1|x1'<<2|z1'<<3|flip<<(2k) (Hermitian convention, (1,1)=Y). */
int qsb_apply_clifford(qsb_state* st, const int* idx, int k, const uint8_t* table, const qsb_mask* mask);
/* uptotime!(state, idx, background, dt) -- DM: sum_k K rho K' in closed form (uptotime.jl:44-65,
 * tables :106-173); KET: promoted to DM first (:1-4); MCKET: one sampled Kraus branch (:20-42);
 * TAB: Pauli noise, one draw always consumed (clifford/uptotime.jl:1-12,26-39; T1/T1T2 -> UNSUPPORTED). */
int qsb_channel(qsb_state* st, int idx, int background, const double params[2], double dt, uint64_t draw_id);
/* apply_noninstant!(state, state_indices, ConstantHamiltonianEvolution(H, duration), backgrounds) --
 * src/backends/quantumoptics/noninstant.jl:10-37 (timeevolution.master / schroedinger / mcwf), and the
 * `krausops === nothing` branch of uptotime! (uptotime.jl:24-29,51-54; H_ri == NULL, k == 1):
 *   d rho/dt = -i[H, rho] + sum_j (L_j rho L_j' - 1/2 {L_j' L_j, rho}) for `duration`.
 * H_ri: 2^k x 2^k Hamiltonian on idx[] (column-major, nullable = 0); L_ri: nl 2x2 collapse operators (lindbladop,
 * noninstant.jl:44-97), operator j acting on subsystem idx[l_pos[j]].  The generator is constant, so the propagator
 * exp(L t) of the k-qubit block is computed once on the host and applied in one pass (no solver tolerance).
 * DM / KET (promoted with dm(), :21-23): exact.  MCKET: nl == 0 -> exp(-iHt) psi (:31-32); nl > 0 -> one sampled
 * branch of the canonical Kraus form of exp(L t) (an unravelling whose ensemble mean is the master-equation solution,
 * like mcwf :33-34; consumes draw_id).  TAB: QSB_ERR_UNSUPPORTED (the reference has no method either). */
int qsb_evolve(qsb_state* st, const int* idx, int k, const double* H_ri, int nl, const int* l_pos, const double* L_ri,
               double duration, uint64_t draw_id);
/* project_traceout!(state, stateindex, basis) -> (j, newstate): quantumoptics.jl:47-64 (+ express.jl:7-12
 * for symbolic X/Y/Z bases, mcket.jl:30-33) and clifford/express.jl:2-14.  The state shrinks in place
 * (nsub-1; 0 when the last subsystem was measured).  *outcome2 (new mask, caller frees) has bit t set
 * when trajectory t got outcome 2 (the -1 eigenstate); was_random (nullable; TAB only) marks the
 * trajectories whose outcome consumed the draw; counts_host (nullable, host-visible) = {alive, n1, n2}.
 * kets_ri: 2 kets x 2 amplitudes when basis == QSB_BASIS_KETS. */
int qsb_measure(qsb_state* st, int idx, int basis, const double* kets_ri, uint64_t draw_id,
                qsb_mask** outcome2, qsb_mask** was_random, int64_t* counts_host);
/* traceout!: _traceout_state(state, i) -- ptrace for Ket (-> DM) / Operator (src/baseops/traceout.jl:46),
 * QuantumClifford.ptrace (clifford.jl:7-8), sampled canonical-basis measurement for MCKET (mcket.jl:46-54;
 * consumes draw_id, QSB_ERR_TRACE_ALL when nsub == 1). */
int qsb_ptrace(qsb_state* st, int idx, uint64_t draw_id);
/* observable(state, indices, obs): quantumoptics.jl:14-45, mcket.jl:28, clifford/express.jl:16-20,49-52.
 * per_traj_dev (nullable): device double[2*Npad] (re, im per trajectory; 0 on dead lanes).
 * stats_dev / stats_host (nullable): {n_alive, sum re, sum re^2, sum im} over alive trajectories
 * -- the only data that ever crosses GPUs (one NCCL all-reduce of these doubles, section 8e).
 * stats_host != NULL makes the call host-visible.  obs_sign: 1 negates the observable. */
int qsb_expect(qsb_state* st, const int* idx, int k, int obs_kind, const void* obs, int obs_sign,
               double* per_traj_dev, double* stats_dev, double* stats_host);

/* The Pauli-string case of qsb_expect (obs_kind == QSB_OBS_PAULI, dense kinds) for callers that must not allocate.
 * No counterpart in the reference: QuantumSavory has no notion of who owns GPU memory.  qsb_expect keeps its reduction
 * partials in a per-context buffer that it frees and re-allocates when a larger observable arrives; that is a problem
 * for (i) a host whose HBM is already owned by another allocator (CUDA.jl's pool, torch's caching allocator), where an
 * allocation made behind its back can fail or force a pool trim that stalls the whole device, and (ii) a loop with a
 * latency bound per call, where the growth lands in the middle of the loop.  This variant calls no cudaMalloc / cudaFree
 * / stream-ordered allocator and uploads nothing: the partials live in a device array built into the library, of
 * QSB_STATIC_SCRATCH_BYTES, which every GPU that loads the library holds until the process exits.
 *   - limit: 2 * nsplit * Npad + 4 * ceil(Npad / 256) doubles must fit (Npad = trajectories of the context rounded up to 32);
 *     the call lowers its split count to fit and returns QSB_ERR_BAD_ARG with a message when the context has more than
 *     QSB_STATIC_MAX_TRAJ trajectories.  Fewer splits = a different summation order than qsb_expect: values agree to
 *     rounding (1e-12 in the tests), not bit for bit.
 *   - codes: uint8_t[k], 0 I, 1 X, 2 Y, 3 Z on subsystem idx[j]; per_traj_dev / stats_dev as in qsb_expect (device pointers
 *     owned by the caller, at least one non-NULL); there is no host-visible output and the call never synchronises.
 *   - the first call on each device looks up the array's device address (it may make the CUDA runtime load the library's
 *     code onto that device): make it once per device during set-up, before any stream capture is opened.
 *   - ONE array per device is shared by all contexts and calls and keeps the previous call's partials.  Nothing needs
 *     clearing (every slot read is written first by the same call), but two calls that can overlap on the GPU -- different
 *     streams, no event between them -- race on it.  Calls on one stream are safe.
 *   - refused: tableau states (their expectation uses no scratch), calls inside qsb_record_begin/_end. */
#define QSB_STATIC_SCRATCH_BYTES 4194304 /* 4 MiB */
#define QSB_STATIC_MAX_TRAJ 260096
int qsb_expect_pauli_static(qsb_state* st, const int* idx, int k, const uint8_t* codes, int obs_sign,
                            double* per_traj_dev, double* stats_dev);

/* ---- per-trajectory masks --------------------------------------------------------------------*/
int qsb_mask_create(qsb_ctx* ctx, int fill, qsb_mask** out);
int qsb_mask_free(qsb_mask* m);
int qsb_mask_logic(qsb_mask* dst, const qsb_mask* a, const qsb_mask* b, int op); /* dst may alias a/b */
int qsb_mask_count(const qsb_mask* m, int64_t* out);       /* host-visible */
int qsb_mask_to_host(const qsb_mask* m, uint32_t* words);  /* host-visible; ceil(ntraj/32) words */
int qsb_mask_from_host(qsb_mask* m, const uint32_t* words);
int qsb_mask_random(qsb_ctx* ctx, double p, uint64_t draw_id, qsb_mask** out); /* bit = (u < p) */
/* lanes where this state is assigned ("alive"); dead lanes model a slot that was traced out in
 * that trajectory only (failed purification, src/CircuitZoo/CircuitZoo.jl:156-160). */
int qsb_state_alive(const qsb_state* st, qsb_mask** out);
int qsb_state_kill(qsb_state* st, const qsb_mask* lanes);

/* ---- recorded programs (command buffer) ------------------------------------------------------------
 * Between qsb_record_begin and qsb_record_end every call on the context is CAPTURED instead of launched
 * (like a CUDA graph capture): handles created inside are virtual, host-visible calls (counts_host,
 * stats_host, *_to_host, qsb_mask_count) are refused, and device pointers passed as stats_dev /
 * per_traj_dev are captured by address.  A program holds tableau (QSB_TAB) states or dense (KET/DM) states, never both;
 * MCKET trajectory kets are refused.  qsb_program_run executes the whole op list as ONE fused pass: tableau programs with
 * each 32-trajectory word's tableaux resident in shared memory (a thread only ever touches its own word column, so no
 * grid sync is needed between ops), dense programs with every live density matrix of a 32-trajectory tile in shared memory.  traj_shift re-bases the ensemble on fresh global trajectory ids (a new batch of the same
 * schedule); draw_base offsets every draw id.  SURVEY.md 7.2 "command buffer", 8f-2. */
typedef struct qsb_program qsb_program;
int qsb_record_begin(qsb_ctx* ctx);
int qsb_record_end(qsb_ctx* ctx, qsb_program** out);
int qsb_program_run(qsb_program* prog, uint64_t traj_shift, uint64_t draw_base);
/* ---- cohort forking: outcome-dependent host control flow -----------------------------------
 * The host follows ONE classical history per recorded program.  Where the reference branches on a measurement outcome
 * and the branch changes what the host does next (`if !success ... end` after a purification,
 * src/CircuitZoo/CircuitZoo.jl:156-160 and examples/firstgenrepeater_lowlevel/setup.jl:209-237), the host picks one
 * branch for the program being recorded and calls qsb_ctx_retire with the lanes whose outcome disagrees: from that op on
 * they count in no statistic of this program, and codes_dev[lane] (nullable; int32 per trajectory of the context, lane =
 * index inside the context's ensemble) is set to `code`.  The host then records the program of the other history once
 * (cached: the tree of programs depends on the schedule only) and re-runs the retired lanes under it from t = 0 with
 * qsb_program_run_lanes / qsb_wave_run: lane i of a job is the trajectory with lane index lanes_dev[i] (same global id,
 * same draws: its run is identical up to the fork and continues on its own branch).  Statistics rows and per-trajectory
 * rows are indexed by lane index, so every program of the tree writes into the same buffers; flags & QSB_RUN_KEEP_STATS
 * accumulates into the statistics rows instead of zeroing them first.  qsb_wave_run executes any number of (program,
 * lane subset) jobs in ONE launch pair (tableau programs) or as one launch per job over a pool of streams (dense programs);
 * lanes_dev[j] == NULL runs job j on the whole ensemble. */
enum { QSB_RUN_KEEP_STATS = 1 };
int qsb_ctx_retire(qsb_ctx* ctx, const qsb_mask* lanes, int32_t code, int32_t* codes_dev);
int qsb_program_run_lanes(qsb_program* prog, const uint32_t* lanes_dev, int64_t count, uint64_t traj_shift, uint64_t draw_base, int flags);
int qsb_wave_run(qsb_ctx* ctx, qsb_program* const* progs, const uint32_t* const* lanes_dev, const int64_t* counts, int njobs,
                 uint64_t traj_shift, uint64_t draw_base, int flags);
/* Routing between two levels of the tree, on the device.  node_dev[lane] = index of the program the lane ran in last; a lane
 * retired with code c < censor_base has key node * maxdec + c.  qsb_cohort_count returns the number of lanes per key
 * (counts_host[nnodes * maxdec] = lanes retired with c >= censor_base: censored) -- the one host-visible step of a level;
 * the host records the programs of new keys and calls qsb_cohort_scatter with, per key, the offset of its lanes in
 * lanes_out_dev and the index of the program they continue in (-1: no program, the lanes are dropped): lanes_out_dev then
 * holds the lane indices grouped by key (the lane subsets of the next qsb_wave_run), node_dev is updated and codes_dev is
 * reset to -1. */
int qsb_cohort_count(qsb_ctx* ctx, const int32_t* codes_dev, const int32_t* node_dev, int nnodes, int maxdec, int32_t censor_base, int32_t* counts_host);
int qsb_cohort_scatter(qsb_ctx* ctx, int32_t* codes_dev, int32_t* node_dev, int nnodes, int maxdec, int32_t censor_base,
                       const int32_t* offsets_host, const int32_t* child_host, uint32_t* lanes_out_dev);
/* re-sends the recorded op list (the program's only host-side input) to the device, async on the stream */
int qsb_program_upload(qsb_program* prog, int64_t* bytes);
/* info: {ops, arena planes, statistics rows, threads per block, shared bytes per block, 1 if arena in global memory,
 *        register-state ops per run summed over trajectories, draws per run} */
int qsb_program_info(const qsb_program* prog, int64_t info[8]);
/* SURVEY.md 8d table bytes of the recorded ops summed over trajectories: the algorithmic traffic the SAME op list
 * accounts when every op is its own launch (the numerator of roofline.achieved for the fused launch). */
int qsb_program_table_bytes(const qsb_program* prog, uint64_t* bytes);
/* verification aid (host-visible): the raw records of a tableau program as the executor reads them -- what = 0 op records, 1 draw
 * specs, 2 observable tables, 3 constant tables, 4 statistics row pointers, 5 int32 {planes, slots, ops, specs, obs, aux, rows,
 * first slot}; layouts in csrc/qsb_exec_kernel.cuh, not a stable ABI.  bench.py's bit-sliced CPU baseline replays them. */
int qsb_program_export(const qsb_program* prog, int what, void* buf, size_t buflen, size_t* needed);
int qsb_program_free(qsb_program* prog);

#ifdef __cplusplus
}
#endif
#endif /* QSB200_H */
