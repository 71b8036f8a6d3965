/*
 * qsb_replay.c -- compiled trajectory loop over the CPU ORACLE (TEST / BENCH INFRASTRUCTURE ONLY).
 *
 * bench.py's CPU arm (`cpu_baseline`, `--impl reference`) must time the reference's algorithm, not a Python
 * interpreter.  The reference runs one event loop per trajectory and calls the backend generics op by op
 * (examples/firstgenrepeater_lowlevel/6.1_compare_formalisms_noplot.jl:10-57 and :66-73 for the replicate loop).
 * Here the host schedule is traced ONCE per classical history by oracle/trace.py (the same registers.py / circuits.py
 * / workloads code, driven with a tracing backend) into a tree of op lists, and this file replays the tree for every
 * trajectory through the oracle's own functions (qsb_oracle.c: qc_* / qo_*, the line-by-line restatement of
 * src/backends/{clifford,quantumoptics}/*.jl), `#pragma omp parallel for` over trajectories
 * (the reference's own idiom: Threads.@threads over replicates, examples/congestionchain/3_aggregate_of_multiple_simulations.jl:46).
 *
 * A trajectory follows ITS OWN control flow: at a DECIDE op (an `if` of the host on a measurement outcome, e.g.
 * `if !success` after Purify2to1, src/CircuitZoo/CircuitZoo.jl:156-160) it compares its outcome with the branch the
 * op list took and, when they differ, continues in the child op list of that history at the next op (the op lists agree
 * up to and including the DECIDE, so no work is repeated).  A missing child is reported back so that the host can trace
 * it; a DECIDE of a leaf list (fork budget spent) censors the trajectory exactly like the GPU cohort tree does.
 *
 * Nothing under quantumsavory.jl_b200/ links or loads this file.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

/* ---- the oracle (qsb_oracle.c, compiled into the same shared object) ------------------------------------------ */
typedef struct qo_state qo_state;
typedef struct qc_state qc_state;
double qso_draw(uint64_t seed, uint64_t traj, uint64_t draw_id);
qo_state* qo_copy(const qo_state* s);
void qo_free(qo_state* s);
qo_state* qo_newstate(void);
qo_state* qo_from_data(int kind, int n, const double* ri);
qo_state* qo_tensor(const qo_state* a, const qo_state* b);
int qo_kind(const qo_state* s);
int qo_apply(qo_state* s, const int* idx, int k, const double* U_ri);
int qo_uptotime_dm(qo_state* s, int idx, int bg, const double* prm, double dt);
int qo_uptotime_mc(qo_state* s, int idx, int bg, const double* prm, double dt, double u, int* branch_out);
int qo_project_traceout(qo_state* s, int idx, const double* psis_ri, double u, int* j_out);
int qo_ptrace(qo_state* s, int idx);
void qo_expect_op(const qo_state* s, const int* idx, int k, const double* O_ri, double* out_ri);
void qo_expect_proj(const qo_state* s, const int* idx, int k, const double* psi_ri, double* out_ri);
qc_state* qc_copy(const qc_state* s);
void qc_free(qc_state* s);
qc_state* qc_newstate(void);
qc_state* qc_from_generators(int n, int r, const uint8_t* xs, const uint8_t* zs, const uint8_t* signs);
qc_state* qc_tensor(const qc_state* a, const qc_state* b);
int qc_apply(qc_state* s, const int* idx, int k, const uint8_t* table);
int qc_uptotime(qc_state* s, int idx, int bg, const double* prm, double dt, double u);
int qc_ptrace(qc_state* s, int q);
int qc_project_traceout(qc_state* s, int idx, int basis, double u, int* was_random);
int qc_expect_pauli(const qc_state* s, const int* idx, int k, const uint8_t* paulis, int sign);
double qc_expect_proj(const qc_state* s, const int* idx, int k, const uint8_t* gx, const uint8_t* gz, const uint8_t* gsign);

/* ---- trace format (written by oracle/trace.py; keep in sync with OP_* there) ---------------------------------- */
enum { ROP_NEW = 0, ROP_FROM_DATA, ROP_FROM_GENERATORS, ROP_SELECT, ROP_COPY, ROP_TENSOR, ROP_APPLY, ROP_APPLY_CLIFFORD,
       ROP_CHANNEL_DM, ROP_CHANNEL_MC, ROP_CHANNEL_TAB, ROP_MEASURE_DENSE, ROP_MEASURE_TAB, ROP_PTRACE, ROP_EXPECT_OP, ROP_EXPECT_PROJ_KET,
       ROP_EXPECT_PAULI, ROP_EXPECT_PROJ_STAB, ROP_MASK_CONST, ROP_MASK_LOGIC, ROP_MASK_RANDOM, ROP_DECIDE, ROP_FREE };
enum { RM_AND = 0, RM_OR, RM_XOR, RM_ANDNOT, RM_NOT, RM_XNOR };

typedef struct {
    int32_t type, n, k, nsrc;
    int32_t dst, src[5];     /* state ids */
    int32_t idx[6];          /* 0-based subsystem indices */
    int32_t mask[3];         /* [0] result mask id, [1] [2] operands / condition (-1: none) */
    int32_t iv[4];           /* op specific small integers */
    int32_t tab;             /* 1: tableau state, 0: dense */
    int32_t row;             /* EXPECT: statistics row (-1: none) */
    uint64_t draw;
    int64_t dpay, bpay;      /* offsets into the double / byte payload tables (-1: none) */
} qsr_op;

typedef struct {
    const qsr_op* ops; int64_t nops;
    const double* dpay; const uint8_t* bpay;
    int32_t nstates, nmasks, leaf;
    const int32_t* child; int32_t nchild;     /* child[ordinal] = node index of the history that disagrees there, -1: not traced yet */
} qsr_node;

typedef struct { void* p; int tab; } slot_t;

static void slot_free(slot_t* s) { if (s->p) { if (s->tab) qc_free((qc_state*)s->p); else qo_free((qo_state*)s->p); } s->p = NULL; }

/* status per trajectory: 0 complete, 1 censored at a leaf, 2 needs a child (need_node / need_ordinal), <0 oracle error code */
typedef struct { int status, need_node, need_ordinal; int nrec; } traj_result;

/* one trajectory through the tree; rec_row / rec_val receive its (row, value) pairs in order */
static traj_result run_one(const qsr_node* nodes, uint64_t seed, uint64_t traj, int maxstates, int maxmasks,
                           slot_t* st, uint8_t* mk, int32_t* rec_row, double* rec_val, int rec_cap) {
    traj_result R = {0, -1, -1, 0};
    memset(st, 0, sizeof(slot_t) * (size_t)maxstates);
    memset(mk, 0, (size_t)maxmasks);
    int node = 0;
    const qsr_node* N = &nodes[0];
    for (int64_t i = 0; i < N->nops && R.status == 0; ++i) {
        const qsr_op* o = &N->ops[i];
        const double* dp = o->dpay >= 0 ? N->dpay + o->dpay : NULL;
        const uint8_t* bp = o->bpay >= 0 ? N->bpay + o->bpay : NULL;
        if (o->mask[1] >= 0 && (o->type == ROP_APPLY || o->type == ROP_APPLY_CLIFFORD) && !mk[o->mask[1]]) continue;   /* if cond ... end */
        int rc = 0;
        switch (o->type) {
        case ROP_NEW: st[o->dst].tab = o->tab; st[o->dst].p = o->tab ? (void*)qc_newstate() : (void*)qo_newstate(); break;
        case ROP_FROM_DATA: st[o->dst].tab = 0; st[o->dst].p = qo_from_data(o->iv[0], o->n, dp); break;
        case ROP_FROM_GENERATORS: st[o->dst].tab = 1; st[o->dst].p = qc_from_generators(o->n, o->k, bp, bp + o->k * o->n, bp + 2 * o->k * o->n); break;
        case ROP_SELECT: {   /* express(mixture, CliffordRepr()): one term sampled per call, first i with cum_i > u */
            const double u = qso_draw(seed, traj, o->draw);
            int sel = o->nsrc - 1;
            for (int j = 0; j < o->nsrc; ++j) if (dp[j] > u) { sel = j; break; }
            st[o->dst] = st[o->src[sel]]; st[o->src[sel]].p = NULL;
            for (int j = 0; j < o->nsrc; ++j) slot_free(&st[o->src[j]]);
            break; }
        case ROP_COPY: st[o->dst].tab = o->tab; st[o->dst].p = o->tab ? (void*)qc_copy((qc_state*)st[o->src[0]].p) : (void*)qo_copy((qo_state*)st[o->src[0]].p); break;
        case ROP_TENSOR:
            st[o->dst].tab = o->tab;
            st[o->dst].p = o->tab ? (void*)qc_tensor((qc_state*)st[o->src[0]].p, (qc_state*)st[o->src[1]].p) : (void*)qo_tensor((qo_state*)st[o->src[0]].p, (qo_state*)st[o->src[1]].p);
            break;
        case ROP_APPLY: rc = qo_apply((qo_state*)st[o->dst].p, o->idx, o->k, dp); break;
        case ROP_APPLY_CLIFFORD: rc = qc_apply((qc_state*)st[o->dst].p, o->idx, o->k, bp); break;
        case ROP_CHANNEL_DM: rc = qo_uptotime_dm((qo_state*)st[o->dst].p, o->idx[0], o->iv[0], dp, dp[2]); break;
        case ROP_CHANNEL_MC: { int br; rc = qo_uptotime_mc((qo_state*)st[o->dst].p, o->idx[0], o->iv[0], dp, dp[2], qso_draw(seed, traj, o->draw), &br); break; }
        case ROP_CHANNEL_TAB: { int r2 = qc_uptotime((qc_state*)st[o->dst].p, o->idx[0], o->iv[0], dp, dp[2], qso_draw(seed, traj, o->draw)); rc = r2 < 0 ? -r2 : 0; break; }
        case ROP_MEASURE_DENSE: { int j = 0; rc = qo_project_traceout((qo_state*)st[o->dst].p, o->idx[0], dp, qso_draw(seed, traj, o->draw), &j); mk[o->mask[0]] = (uint8_t)(j == 2); break; }
        case ROP_MEASURE_TAB: { int wr; const int j = qc_project_traceout((qc_state*)st[o->dst].p, o->idx[0], o->iv[0], qso_draw(seed, traj, o->draw), &wr); mk[o->mask[0]] = (uint8_t)(j == 2); break; }
        case ROP_PTRACE: rc = o->tab ? qc_ptrace((qc_state*)st[o->dst].p, o->idx[0]) : qo_ptrace((qo_state*)st[o->dst].p, o->idx[0]); break;
        case ROP_EXPECT_OP: case ROP_EXPECT_PROJ_KET: case ROP_EXPECT_PAULI: case ROP_EXPECT_PROJ_STAB: {
            double v[2] = {0, 0};
            if (o->type == ROP_EXPECT_OP) qo_expect_op((qo_state*)st[o->dst].p, o->idx, o->k, dp, v);
            else if (o->type == ROP_EXPECT_PROJ_KET) qo_expect_proj((qo_state*)st[o->dst].p, o->idx, o->k, dp, v);
            else if (o->type == ROP_EXPECT_PAULI) v[0] = (double)qc_expect_pauli((qc_state*)st[o->dst].p, o->idx, o->k, bp, 0);
            else v[0] = qc_expect_proj((qc_state*)st[o->dst].p, o->idx, o->k, bp, bp + o->k * o->k, bp + 2 * o->k * o->k);
            if (o->row >= 0 && R.nrec < rec_cap) { rec_row[R.nrec] = o->row; rec_val[R.nrec] = v[0]; R.nrec++; }
            break; }
        case ROP_MASK_CONST: mk[o->mask[0]] = (uint8_t)(o->iv[0] != 0); break;
        case ROP_MASK_LOGIC: {
            const uint8_t x = mk[o->mask[1]], y = o->mask[2] >= 0 ? mk[o->mask[2]] : 0;
            uint8_t r;
            switch (o->iv[0]) { case RM_AND: r = x & y; break; case RM_OR: r = x | y; break; case RM_XOR: r = x ^ y; break;
                                case RM_ANDNOT: r = x & (uint8_t)!y; break; case RM_NOT: r = (uint8_t)!x; break; default: r = (uint8_t)!(x ^ y); break; }
            mk[o->mask[0]] = r & 1;
            break; }
        case ROP_MASK_RANDOM: mk[o->mask[0]] = (uint8_t)(qso_draw(seed, traj, o->draw) < dp[0]); break;
        case ROP_DECIDE: {   /* iv: {ordinal, branch taken by this op list} */
            const int mine = mk[o->mask[1]] != 0;
            if (mine == (o->iv[1] != 0)) break;
            if (N->leaf) { R.status = 1; break; }
            const int c = o->iv[0] < N->nchild ? N->child[o->iv[0]] : -1;
            if (c < 0) { R.status = 2; R.need_node = node; R.need_ordinal = o->iv[0]; break; }
            node = c; N = &nodes[c];      /* same op index in the child's list: the lists agree up to and including this DECIDE */
            break; }
        case ROP_FREE: slot_free(&st[o->dst]); break;
        default: rc = 99; break;
        }
        if (rc) R.status = -rc;
    }
    for (int s = 0; s < maxstates; ++s) slot_free(&st[s]);
    return R;
}

/* Replays trajectories [traj_lo, traj_hi) (global ids; Philox keyed by (seed, id, draw id) like the GPU path).
 * rows: [nrows][3] = {count, sum, sum of squares} accumulated over the trajectories that completed or were censored;
 * per_traj (nullable): [nrows][traj_hi - traj_lo] values (0 where the trajectory recorded nothing);
 * status (nullable): per trajectory 0 / 1 / 2 / <0; need (nullable): [n][2] = (node, ordinal) for status 2.
 * Returns the number of trajectories with status 2 (missing child: trace it and call again), or <0 on an oracle error. */
int64_t qsr_run(const qsr_node* nodes, int nnodes, uint64_t seed, int64_t traj_lo, int64_t traj_hi, int nrows, int threads,
                double* rows, double* per_traj, int32_t* status, int32_t* need, int64_t* counters) {
    int maxstates = 1, maxmasks = 1;
    for (int i = 0; i < nnodes; ++i) { if (nodes[i].nstates > maxstates) maxstates = nodes[i].nstates; if (nodes[i].nmasks > maxmasks) maxmasks = nodes[i].nmasks; }
    const int64_t n = traj_hi - traj_lo;
    int64_t missing = 0, censored = 0, errors = 0, complete = 0;
#ifdef _OPENMP
    if (threads > 0) omp_set_num_threads(threads);
#endif
#pragma omp parallel reduction(+ : missing, censored, errors, complete)
    {
        slot_t* st = (slot_t*)malloc(sizeof(slot_t) * (size_t)maxstates);
        uint8_t* mk = (uint8_t*)malloc((size_t)maxmasks);
        int32_t* rrow = (int32_t*)malloc(sizeof(int32_t) * (size_t)(nrows + 1));
        double* rval = (double*)malloc(sizeof(double) * (size_t)(nrows + 1));
        double* acc = (double*)calloc((size_t)nrows * 3, sizeof(double));
#pragma omp for schedule(dynamic, 8)
        for (int64_t t = 0; t < n; ++t) {
            traj_result R = run_one(nodes, seed, (uint64_t)(traj_lo + t), maxstates, maxmasks, st, mk, rrow, rval, nrows);
            if (status) status[t] = R.status;
            if (need) { need[2 * t] = R.need_node; need[2 * t + 1] = R.need_ordinal; }
            if (R.status == 2) { missing++; continue; }      /* nothing of this trajectory is tallied: it runs again once its history is traced */
            if (R.status < 0) { errors++; continue; }
            if (R.status == 1) censored++; else complete++;
            for (int k = 0; k < R.nrec; ++k) {
                const int r = rrow[k];
                acc[3 * r] += 1.0; acc[3 * r + 1] += rval[k]; acc[3 * r + 2] += rval[k] * rval[k];
                if (per_traj) per_traj[(size_t)r * (size_t)n + (size_t)t] = rval[k];
            }
        }
#pragma omp critical
        { for (int i = 0; i < nrows * 3; ++i) rows[i] += acc[i]; }
        free(st); free(mk); free(rrow); free(rval); free(acc);
    }
    if (counters) { counters[0] = complete; counters[1] = censored; counters[2] = missing; counters[3] = errors; }
    return errors ? -errors : missing;
}

int qsr_sizeof_op(void) { return (int)sizeof(qsr_op); }
int qsr_sizeof_node(void) { return (int)sizeof(qsr_node); }
int qsr_max_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}
