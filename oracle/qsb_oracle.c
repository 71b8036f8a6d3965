/*
 * qsb_oracle.c -- CPU ORACLE (TEST INFRASTRUCTURE ONLY; never linked into the product).
 *
 * A plain, single-trajectory C restatement of the register-state evolution path of
 * QuantumSavory.jl (reference checkout /root/reference, v0.7.2), i.e. SURVEY.md section 8(a)
 * rows a1..a12.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
 * reference legs may load this library.  The product (libqsb200.so) must never call into it.
 *
 * PARITY STATUS
 *   Dense (QuantumOptics-family) part: restates the reference's L1 adapter line by line
 *   (embed + dense products, Kraus sums with temporaries, all-branches measurement).  The
 *   third-party numerics it stands on (QuantumOpticsBase tensor/embed/expect/ptrace) are not
 *   vendored under /root/reference and Julia is not installed, so the conventions (subsystem 1
 *   is the least-significant index, tensor(a,b)=kron(b,a)) are restated from the published
 *   package behaviour and pinned by the reference's own known-answer tests re-expressed in
 *   tests/test_oracle_kats.py (Bell/GHZ projectors, swap chains, purification truth tables,
 *   doctest values 0.9999546000702374 and 0.9999999999999998, Kraus closed forms).
 *   Clifford part: QuantumClifford (>=0.11.6) source is absent; "parity unpinned" for raw
 *   tableau bits.  What is pinned (and what the GPU path is compared on, bit-exactly) is every
 *   group-invariant quantity: measurement outcomes given the injected draw, the was-random
 *   flag, Pauli / stabilizer-projector expectation values, rank, and the canonical (RREF)
 *   generator set.  The stabilizer algorithms below are the textbook generator-list ones
 *   (no destabilizers) so that they are independent of the GPU's bit-sliced pair tableau.
 *
 * Conventions (reference file:line)
 *   - |0> = spinup = Z1 = index 0, |1> = spindown = index 1
 *       src/backends/quantumoptics/quantumoptics.jl:66, uptotime.jl:68-69
 *   - composite linear index (0-based) = sum_k i_k 2^k with k the 0-based subsystem number,
 *     i.e. subsystem 1 is the fastest index (QuantumOpticsBase tensor(a,b).data = kron(b,a))
 *   - density matrices dense column-major: d[r + c*D]
 *   - operator on `indices`: factor m of the operator acts on subsystem indices[m]; factor 0
 *     is the least significant bit of the operator's own index (CNOT: first index = control,
 *     uptotime.jl:83)
 *   - outcomes are 1-based (1 <-> +1 eigenstate), src/baseops/traceout.jl:150
 */
#include <complex.h>
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

typedef double _Complex cplx;

/* ------------------------------------------------------------------------------------------
 * Counter-based RNG: Philox4x32-10 (Salmon et al., SC'11).  draw(seed, traj, draw_id) is the
 * "injected draw" both the oracle and the GPU path consume; key=(seed lo, seed hi),
 * counter=(traj lo, traj hi, draw lo, draw hi).  u in [0,1) with 53 bits.
 * ---------------------------------------------------------------------------------------- */
static void philox4x32_10(const uint32_t ctr_in[4], const uint32_t key_in[2], uint32_t out[4]) {
    uint32_t c0 = ctr_in[0], c1 = ctr_in[1], c2 = ctr_in[2], c3 = ctr_in[3];
    uint32_t k0 = key_in[0], k1 = key_in[1];
    for (int r = 0; r < 10; ++r) {
        uint64_t p0 = (uint64_t)0xD2511F53u * c0;
        uint64_t p1 = (uint64_t)0xCD9E8D57u * c2;
        uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0;
        uint32_t n1 = (uint32_t)p1;
        uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1;
        uint32_t n3 = (uint32_t)p0;
        c0 = n0; c1 = n1; c2 = n2; c3 = n3;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
    out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

void qso_philox(const uint32_t* ctr, const uint32_t* key, uint32_t* out) { philox4x32_10(ctr, key, out); }

/* The injected draw: one Philox block (counter = trajectory, draw_id / 2) carries two 53-bit uniforms,
 * words (0,1) for even draw ids and (2,3) for odd ones. */
double qso_draw(uint64_t seed, uint64_t traj, uint64_t draw_id) {
    uint64_t block = draw_id >> 1;
    uint32_t ctr[4] = {(uint32_t)traj, (uint32_t)(traj >> 32), (uint32_t)block, (uint32_t)(block >> 32)};
    uint32_t key[2] = {(uint32_t)seed, (uint32_t)(seed >> 32)};
    uint32_t o[4];
    philox4x32_10(ctr, key, o);
    uint64_t bits = (draw_id & 1) ? (((uint64_t)o[3] << 32) | o[2]) : (((uint64_t)o[1] << 32) | o[0]);
    return (double)(bits >> 11) * (1.0 / 9007199254740992.0);
}

/* ==========================================================================================
 *                                   DENSE (ket / dm) PART
 * ======================================================================================== */
typedef struct {
    int kind; /* 0 = ket, 1 = density operator */
    int n;    /* number of qubit subsystems */
    cplx* d;
} qo_state;

static size_t qo_len(int kind, int n) { return kind == 0 ? ((size_t)1 << n) : ((size_t)1 << (2 * n)); }

qo_state* qo_alloc(int kind, int n) {
    qo_state* s = (qo_state*)malloc(sizeof(qo_state));
    s->kind = kind; s->n = n;
    s->d = (cplx*)calloc(qo_len(kind, n), sizeof(cplx));
    return s;
}
void qo_free(qo_state* s) { if (s) { free(s->d); free(s); } }
int qo_kind(const qo_state* s) { return s->kind; }
int qo_nsubsystems(const qo_state* s) { return s->n; }

/* Base.copy(state): StateRef copies on initialize!, src/states_registers.jl:10 */
qo_state* qo_copy(const qo_state* s) {
    qo_state* c = qo_alloc(s->kind, s->n);
    memcpy(c->d, s->d, qo_len(s->kind, s->n) * sizeof(cplx));
    return c;
}
/* newstate(::Qubit, ::QuantumOpticsRepr) = copy(_l), quantumoptics.jl:66-69 */
qo_state* qo_newstate(void) { qo_state* s = qo_alloc(0, 1); s->d[0] = 1.0; return s; }

qo_state* qo_from_data(int kind, int n, const double* ri) {
    qo_state* s = qo_alloc(kind, n);
    size_t L = qo_len(kind, n);
    for (size_t i = 0; i < L; ++i) s->d[i] = ri[2 * i] + I * ri[2 * i + 1];
    return s;
}
void qo_get(const qo_state* s, double* out) {
    size_t L = qo_len(s->kind, s->n);
    for (size_t i = 0; i < L; ++i) { out[2 * i] = creal(s->d[i]); out[2 * i + 1] = cimag(s->d[i]); }
}

/* dm(ket), used by subsystemcompose(k::Ket, op::Operator) quantumoptics.jl:5-6 and uptotime.jl:2 */
qo_state* qo_dm(const qo_state* k) {
    if (k->kind == 1) return qo_copy(k);
    size_t D = (size_t)1 << k->n;
    qo_state* s = qo_alloc(1, k->n);
    for (size_t c = 0; c < D; ++c)
        for (size_t r = 0; r < D; ++r) s->d[r + c * D] = k->d[r] * conj(k->d[c]);
    return s;
}

/* subsystemcompose(states...) = tensor(states...), quantumoptics.jl:3-6; first factor is the
 * least significant.  Ket (x) Operator promotes the ket with dm(). */
qo_state* qo_tensor(const qo_state* a, const qo_state* b) {
    if (a->kind != b->kind) {
        qo_state* pa = qo_dm(a); qo_state* pb = qo_dm(b);
        qo_state* r = qo_tensor(pa, pb);
        qo_free(pa); qo_free(pb);
        return r;
    }
    int n = a->n + b->n;
    qo_state* s = qo_alloc(a->kind, n);
    size_t Da = (size_t)1 << a->n, Db = (size_t)1 << b->n, D = Da * Db;
    if (a->kind == 0) {
        for (size_t ib = 0; ib < Db; ++ib)
            for (size_t ia = 0; ia < Da; ++ia) s->d[ia + ib * Da] = a->d[ia] * b->d[ib];
    } else {
        for (size_t cb = 0; cb < Db; ++cb) for (size_t ca = 0; ca < Da; ++ca)
        for (size_t rb = 0; rb < Db; ++rb) for (size_t ra = 0; ra < Da; ++ra)
            s->d[(ra + rb * Da) + (ca + cb * Da) * D] = a->d[ra + ca * Da] * b->d[rb + cb * Db];
    }
    return s;
}

/* embed(basis, indices, op): full 2^n x 2^n operator (column-major) from a k-qubit operator
 * (column-major 2^k x 2^k).  QuantumOpticsBase.embed as used at quantumoptics.jl:24 and
 * uptotime.jl:32,57.  idx are 0-based subsystem numbers, pairwise distinct. */
static void qo_embed(int n, const int* idx, int k, const cplx* op, cplx* out) {
    size_t D = (size_t)1 << n, K = (size_t)1 << k;
    size_t tmask = 0;
    for (int m = 0; m < k; ++m) tmask |= (size_t)1 << idx[m];
    memset(out, 0, D * D * sizeof(cplx));
    for (size_t c = 0; c < D; ++c) {
        size_t cj = 0;
        for (int m = 0; m < k; ++m) cj |= ((c >> idx[m]) & 1) << m;
        for (size_t i = 0; i < K; ++i) {
            size_t r = c & ~tmask;
            for (int m = 0; m < k; ++m) r |= ((i >> m) & 1) << idx[m];
            out[r + c * D] = op[i + cj * K];
        }
    }
}

static void matmul(size_t D, const cplx* A, const cplx* B, cplx* C) { /* C = A*B, column-major */
    for (size_t c = 0; c < D; ++c)
        for (size_t r = 0; r < D; ++r) {
            cplx acc = 0;
            for (size_t m = 0; m < D; ++m) acc += A[r + m * D] * B[m + c * D];
            C[r + c * D] = acc;
        }
}
static void matmul_adj(size_t D, const cplx* A, const cplx* B, cplx* C) { /* C = A*B' */
    for (size_t c = 0; c < D; ++c)
        for (size_t r = 0; r < D; ++r) {
            cplx acc = 0;
            for (size_t m = 0; m < D; ++m) acc += A[r + m * D] * conj(B[c + m * D]);
            C[r + c * D] = acc;
        }
}

/* apply!(state, indices, op): QuantumOpticsBase.apply! -- psi <- U psi ; rho <- U rho U'
 * with U = embed(basis, indices, op); reached from src/baseops/apply.jl:20. */
int qo_apply(qo_state* s, const int* idx, int k, const double* U_ri) {
    size_t D = (size_t)1 << s->n, K = (size_t)1 << k;
    cplx* op = (cplx*)malloc(K * K * sizeof(cplx));
    for (size_t i = 0; i < K * K; ++i) op[i] = U_ri[2 * i] + I * U_ri[2 * i + 1];
    cplx* E = (cplx*)malloc(D * D * sizeof(cplx));
    qo_embed(s->n, idx, k, op, E);
    if (s->kind == 0) {
        cplx* out = (cplx*)calloc(D, sizeof(cplx));
        for (size_t c = 0; c < D; ++c) for (size_t r = 0; r < D; ++r) out[r] += E[r + c * D] * s->d[c];
        memcpy(s->d, out, D * sizeof(cplx)); free(out);
    } else {
        cplx* t1 = (cplx*)malloc(D * D * sizeof(cplx));
        cplx* t2 = (cplx*)malloc(D * D * sizeof(cplx));
        matmul(D, E, s->d, t1);
        matmul_adj(D, t1, E, t2);
        memcpy(s->d, t2, D * D * sizeof(cplx));
        free(t1); free(t2);
    }
    free(E); free(op);
    return 0;
}

/* Background kinds (src/backgrounds.jl): */
enum { BG_T1 = 0, BG_T2 = 1, BG_DEPOL = 2, BG_T1T2 = 3 };

/* krausops(bg, dt): uptotime.jl:106-109 (T1), :126-129 (T2), :139-142 (Depolarization),
 * :157-173 (T1T2).  Each K is 2x2 column-major: K[r + 2c].  Returns the count. */
int qo_krausops(int bg, const double* prm, double dt, cplx K[][4]) {
    const cplx lh[4] = {0, 0, 1, 0};  /* sigmap = |0><1| : row 0, col 1 */
    const cplx ll[4] = {1, 0, 0, 0};
    const cplx hh[4] = {0, 0, 0, 1};
    const cplx id[4] = {1, 0, 0, 1};
    const cplx zz[4] = {1, 0, 0, -1};
    const cplx xx[4] = {0, 1, 1, 0};
    const cplx yy[4] = {0, I, -I, 0}; /* sigmay = [[0,-i],[i,0]] column-major */
    if (bg == BG_T1) {
        double p = exp(-dt / prm[0]);
        for (int i = 0; i < 4; ++i) { K[0][i] = sqrt(1 - p) * lh[i]; K[1][i] = sqrt(p) * hh[i] + ll[i]; }
        return 2;
    } else if (bg == BG_T2) {
        double p = 1 - exp(-dt / prm[0]);
        for (int i = 0; i < 4; ++i) { K[0][i] = sqrt(1 - p / 2) * id[i]; K[1][i] = sqrt(p / 2) * zz[i]; }
        return 2;
    } else if (bg == BG_DEPOL) {
        double p = 1 - exp(-dt / prm[0]);
        for (int i = 0; i < 4; ++i) {
            K[0][i] = sqrt(1 - 3 * p / 4) * id[i];
            K[1][i] = sqrt(p) / 2 * xx[i]; K[2][i] = sqrt(p) / 2 * yy[i]; K[3][i] = sqrt(p) / 2 * zz[i];
        }
        return 4;
    } else if (bg == BG_T1T2) {
        double t1 = prm[0], t2 = prm[1];
        double p = exp(-dt / t1);
        cplx kt1[2][4];
        for (int i = 0; i < 4; ++i) { kt1[0][i] = sqrt(1 - p) * lh[i]; kt1[1][i] = sqrt(p) * hh[i] + ll[i]; }
        double tphi_inv = 1 / t2 - 1 / (2 * t1);
        if (tphi_inv <= 0) { memcpy(K[0], kt1[0], sizeof(kt1[0])); memcpy(K[1], kt1[1], sizeof(kt1[1])); return 2; }
        double tphi = 1 / tphi_inv;
        double pphi = 1 - exp(-dt / tphi);
        cplx kd[2][4];
        for (int i = 0; i < 4; ++i) { kd[0][i] = sqrt(1 - pphi / 2) * id[i]; kd[1][i] = sqrt(pphi / 2) * zz[i]; }
        int cnt = 0; /* [F*E for F in kraus_dephase for E in kraus_T1] */
        for (int f = 0; f < 2; ++f) for (int e = 0; e < 2; ++e) {
            for (int c = 0; c < 2; ++c) for (int r = 0; r < 2; ++r) {
                cplx acc = 0;
                for (int m = 0; m < 2; ++m) acc += kd[f][r + 2 * m] * kt1[e][m + 2 * c];
                K[cnt][r + 2 * c] = acc;
            }
            ++cnt;
        }
        return 4;
    }
    return -1;
}

/* uptotime!(state::Operator, idx, bg, dt): nstate = sum_k K rho K' with K=embed(...), three
 * temporaries and two products per Kraus operator -- uptotime.jl:44-65.  A Ket is promoted
 * with dm() first (uptotime.jl:1-4).  Returns 0, or 4 when |tr| is not ~1 (the @assert :63). */
int qo_uptotime_dm(qo_state* s, int idx, int bg, const double* prm, double dt) {
    if (s->kind == 0) {
        qo_state* p = qo_dm(s);
        free(s->d); s->d = p->d; s->kind = 1; free(p);
    }
    cplx K[4][4];
    int nk = qo_krausops(bg, prm, dt, K);
    if (nk < 0) return 3;
    size_t D = (size_t)1 << s->n;
    cplx* nstate = (cplx*)calloc(D * D, sizeof(cplx));
    cplx* tmpl = (cplx*)malloc(D * D * sizeof(cplx));
    cplx* tmpr = (cplx*)malloc(D * D * sizeof(cplx));
    cplx* E = (cplx*)malloc(D * D * sizeof(cplx));
    for (int k = 0; k < nk; ++k) {
        qo_embed(s->n, &idx, 1, K[k], E);
        matmul(D, E, s->d, tmpl);
        matmul_adj(D, tmpl, E, tmpr);
        for (size_t i = 0; i < D * D; ++i) nstate[i] += tmpr[i];
    }
    cplx tr = 0;
    for (size_t i = 0; i < D; ++i) tr += nstate[i + i * D];
    memcpy(s->d, nstate, D * D * sizeof(cplx));
    free(nstate); free(tmpl); free(tmpr); free(E);
    return fabs(cabs(tr) - 1.0) < 1.5e-8 ? 0 : 4;
}

/* uptotime!(state::MCKet, idx, bg, dt): branches K_k psi, probabilities, threshold=rand()*total,
 * first k with cumsum > threshold (else last), normalize -- uptotime.jl:20-42. */
int qo_uptotime_mc(qo_state* s, int idx, int bg, const double* prm, double dt, double u, int* branch_out) {
    if (s->kind != 0) return 3;
    cplx K[4][4];
    int nk = qo_krausops(bg, prm, dt, K);
    if (nk < 0) return 3;
    size_t D = (size_t)1 << s->n;
    cplx* E = (cplx*)malloc(D * D * sizeof(cplx));
    cplx* br[4]; double P[4]; double total = 0;
    for (int k = 0; k < nk; ++k) {
        br[k] = (cplx*)calloc(D, sizeof(cplx));
        qo_embed(s->n, &idx, 1, K[k], E);
        for (size_t c = 0; c < D; ++c) for (size_t r = 0; r < D; ++r) br[k][r] += E[r + c * D] * s->d[c];
        double nn = 0;
        for (size_t r = 0; r < D; ++r) nn += creal(br[k][r]) * creal(br[k][r]) + cimag(br[k][r]) * cimag(br[k][r]);
        double nrm = sqrt(nn);
        P[k] = nrm * nrm; total += P[k];
    }
    double thr = u * total, cum = 0; int sel = nk - 1;
    for (int k = 0; k < nk; ++k) { cum += P[k]; if (cum > thr) { sel = k; break; } }
    double nn = 0;
    for (size_t r = 0; r < D; ++r) nn += creal(br[sel][r]) * creal(br[sel][r]) + cimag(br[sel][r]) * cimag(br[sel][r]);
    double nrm = sqrt(nn);
    for (size_t r = 0; r < D; ++r) s->d[r] = br[sel][r] / nrm;
    for (int k = 0; k < nk; ++k) free(br[k]);
    free(E);
    if (branch_out) *branch_out = sel;
    return fabs(total - 1.0) < 1.5e-8 ? 0 : 4;
}

/* _project_and_drop(state, project_on, basis_index): (<psi|_idx (x) I) state (|psi>_idx (x) I),
 * should_upstream.jl:15-36; remaining subsystems keep their order. */
static void project_and_drop(const qo_state* s, int idx, const cplx psi[2], cplx* out) {
    int n = s->n;
    size_t Dn = (size_t)1 << (n - 1);
    size_t lowmask = ((size_t)1 << idx) - 1;
    if (s->kind == 0) {
        for (size_t a = 0; a < Dn; ++a) {
            size_t base = (a & lowmask) | ((a & ~lowmask) << 1);
            out[a] = conj(psi[0]) * s->d[base] + conj(psi[1]) * s->d[base | ((size_t)1 << idx)];
        }
    } else {
        size_t D = (size_t)1 << n;
        for (size_t b = 0; b < Dn; ++b) for (size_t a = 0; a < Dn; ++a) {
            size_t ra = (a & lowmask) | ((a & ~lowmask) << 1);
            size_t cb = (b & lowmask) | ((b & ~lowmask) << 1);
            cplx acc = 0;
            for (int sr = 0; sr < 2; ++sr) for (int sc = 0; sc < 2; ++sc)
                acc += conj(psi[sr]) * s->d[(ra | ((size_t)sr << idx)) + (cb | ((size_t)sc << idx)) * D] * psi[sc];
            out[a + b * Dn] = acc;
        }
    }
}

/* project_traceout!(state, stateindex, psis): quantumoptics.jl:47-64.
 *  n==1 : overlaps -> cumsum -> error 4 ("State not normalized...") if sum !~ 1 -> j=first>=u ; state gone
 *  n>1  : all branches, probs=_branch_prob (norm^2 | real trace), cumsum, j = first >= u, normalize.
 * psis_ri: 2 kets x 2 amplitudes x (re,im).  *j_out is 1-based.  If findfirst finds nothing
 * (u above the cumulative sum through rounding) the last branch is taken. */
int qo_project_traceout(qo_state* s, int idx, const double* psis_ri, double u, int* j_out) {
    cplx psi[2][2];
    for (int j = 0; j < 2; ++j) for (int a = 0; a < 2; ++a) psi[j][a] = psis_ri[4 * j + 2 * a] + I * psis_ri[4 * j + 2 * a + 1];
    if (s->n == 1) {
        double ov[2];
        for (int j = 0; j < 2; ++j) {
            if (s->kind == 0) { cplx o = conj(psi[j][0]) * s->d[0] + conj(psi[j][1]) * s->d[1]; ov[j] = creal(o) * creal(o) + cimag(o) * cimag(o); }
            else {
                cplx acc = 0;
                for (int r = 0; r < 2; ++r) for (int c = 0; c < 2; ++c) acc += conj(psi[j][r]) * s->d[r + 2 * c] * psi[j][c];
                ov[j] = creal(acc);
            }
        }
        double c1 = ov[0], c2 = ov[0] + ov[1];
        if (fabs(c2 - 1.0) > 1.5e-8) return 4;
        *j_out = (c1 >= u) ? 1 : 2;
        s->n = 0; s->d[0] = 1.0; /* caller drops the state (returns `nothing`) */
        return 0;
    }
    size_t Ln = qo_len(s->kind, s->n - 1);
    cplx* res[2]; double P[2];
    size_t Dn = (size_t)1 << (s->n - 1);
    for (int j = 0; j < 2; ++j) {
        res[j] = (cplx*)malloc(Ln * sizeof(cplx));
        project_and_drop(s, idx, psi[j], res[j]);
        if (s->kind == 0) {
            double nn = 0;
            for (size_t a = 0; a < Dn; ++a) nn += creal(res[j][a]) * creal(res[j][a]) + cimag(res[j][a]) * cimag(res[j][a]);
            double nrm = sqrt(nn); P[j] = nrm * nrm;
        } else {
            double tr = 0;
            for (size_t a = 0; a < Dn; ++a) tr += creal(res[j][a + a * Dn]);
            P[j] = tr;
        }
    }
    int rc = fabs(P[0] + P[1] - 1.0) < 1.5e-8 ? 0 : 4;
    int j = (P[0] >= u) ? 0 : 1;
    /* normalize(results[j]): ket / norm ; operator / trace */
    double scale;
    if (s->kind == 0) {
        double nn = 0;
        for (size_t a = 0; a < Dn; ++a) nn += creal(res[j][a]) * creal(res[j][a]) + cimag(res[j][a]) * cimag(res[j][a]);
        scale = sqrt(nn);
    } else {
        cplx tr = 0;
        for (size_t a = 0; a < Dn; ++a) tr += res[j][a + a * Dn];
        scale = creal(tr);
    }
    free(s->d);
    s->d = res[j]; s->n -= 1;
    for (size_t i = 0; i < Ln; ++i) s->d[i] /= scale;
    free(res[1 - j]);
    *j_out = j + 1;
    return rc;
}

/* traceout!: _traceout_state default -> ptrace(state, idx), src/baseops/traceout.jl:46.
 * ptrace of a Ket returns a density operator (QuantumOpticsBase). */
int qo_ptrace(qo_state* s, int idx) {
    if (s->n == 1) return 7; /* "Partial trace can't be used to trace out all subsystems" */
    if (s->kind == 0) { qo_state* p = qo_dm(s); free(s->d); s->d = p->d; s->kind = 1; free(p); }
    int n = s->n;
    size_t D = (size_t)1 << n, Dn = D >> 1, lowmask = ((size_t)1 << idx) - 1;
    cplx* out = (cplx*)malloc(Dn * Dn * sizeof(cplx));
    for (size_t b = 0; b < Dn; ++b) for (size_t a = 0; a < Dn; ++a) {
        size_t ra = (a & lowmask) | ((a & ~lowmask) << 1);
        size_t cb = (b & lowmask) | ((b & ~lowmask) << 1);
        cplx acc = 0;
        for (int sb = 0; sb < 2; ++sb) acc += s->d[(ra | ((size_t)sb << idx)) + (cb | ((size_t)sb << idx)) * D];
        out[a + b * Dn] = acc;
    }
    free(s->d); s->d = out; s->n = n - 1;
    return 0;
}

/* observable(state, indices, op) = expect(embed(basis, indices, op), state): quantumoptics.jl:14-28.
 * expect(O, rho) = tr(O rho); expect(O, psi) = <psi|O|psi>. */
void qo_expect_op(const qo_state* s, const int* idx, int k, const double* O_ri, double* out_ri) {
    size_t D = (size_t)1 << s->n, K = (size_t)1 << k;
    cplx* op = (cplx*)malloc(K * K * sizeof(cplx));
    for (size_t i = 0; i < K * K; ++i) op[i] = O_ri[2 * i] + I * O_ri[2 * i + 1];
    cplx* E = (cplx*)malloc(D * D * sizeof(cplx));
    qo_embed(s->n, idx, k, op, E);
    cplx acc = 0;
    if (s->kind == 0) {
        for (size_t c = 0; c < D; ++c) for (size_t r = 0; r < D; ++r) acc += conj(s->d[r]) * E[r + c * D] * s->d[c];
    } else {
        for (size_t r = 0; r < D; ++r) for (size_t m = 0; m < D; ++m) acc += E[r + m * D] * s->d[m + r * D];
    }
    out_ri[0] = creal(acc); out_ri[1] = cimag(acc);
    free(E); free(op);
}

/* observable(state, indices, proj::SProjector): quantumoptics.jl:31-41.  When the projector
 * covers the whole state: permutesystems(state, indices) then |<a|b>|^2 or <psi|rho|psi>;
 * otherwise falls back to the outer-product operator path above. */
void qo_expect_proj(const qo_state* s, const int* idx, int k, const double* psi_ri, double* out_ri) {
    size_t K = (size_t)1 << k;
    cplx* psi = (cplx*)malloc(K * sizeof(cplx));
    for (size_t i = 0; i < K; ++i) psi[i] = psi_ri[2 * i] + I * psi_ri[2 * i + 1];
    if (k == s->n) {
        /* permuted state index p (bit m = subsystem idx[m]) <-> original index */
        size_t D = K;
        size_t* map = (size_t*)malloc(D * sizeof(size_t));
        for (size_t p = 0; p < D; ++p) { size_t o = 0; for (int m = 0; m < k; ++m) o |= ((p >> m) & 1) << idx[m]; map[p] = o; }
        cplx acc = 0;
        if (s->kind == 0) {
            for (size_t p = 0; p < D; ++p) acc += conj(s->d[map[p]]) * psi[p]; /* a' * b */
            double v = creal(acc) * creal(acc) + cimag(acc) * cimag(acc);
            out_ri[0] = v; out_ri[1] = 0;
        } else {
            for (size_t r = 0; r < D; ++r) for (size_t c = 0; c < D; ++c) acc += conj(psi[r]) * s->d[map[r] + map[c] * D] * psi[c];
            out_ri[0] = creal(acc); out_ri[1] = cimag(acc);
        }
        free(map);
    } else {
        double* O = (double*)malloc(K * K * 2 * sizeof(double));
        for (size_t c = 0; c < K; ++c) for (size_t r = 0; r < K; ++r) {
            cplx v = psi[r] * conj(psi[c]);
            O[2 * (r + c * K)] = creal(v); O[2 * (r + c * K) + 1] = cimag(v);
        }
        qo_expect_op(s, idx, k, O, out_ri);
        free(O);
    }
    free(psi);
}

/* ------------------------------------------------------------------------------------------
 * Non-instant evolution: ConstantHamiltonianEvolution and the Lindblad fallback of uptotime!
 *   src/noninstant.jl:21-35, src/backends/quantumoptics/noninstant.jl:1-37, uptotime.jl:24-29,51-54.
 * The reference hands embedded operators to QuantumOptics' timeevolution.master / schroedinger / mcwf, which
 * integrate with OrdinaryDiffEq's DP5 (Dormand-Prince 5(4), adaptive, reltol 1e-6 / abstol 1e-8 -- the package
 * defaults; the packages are not vendored under /root/reference, so the published algorithm is restated):
 *     master:        d rho/dt = -i (H rho - rho H) + sum_j (L_j rho L_j' - 1/2 L_j'L_j rho - 1/2 rho L_j'L_j)
 *     schroedinger:  d psi/dt = -i H psi
 *     mcwf:          d psi/dt = -i (H - i/2 sum L_j'L_j) psi until |psi|^2 drops to a uniform draw, then a jump
 *                    psi <- L_j psi / |L_j psi| with probability ~ |L_j psi|^2, new draw, continue.
 * Results carry the solver's tolerance: comparisons against the exact propagator use 1e-5.
 * ---------------------------------------------------------------------------------------- */
typedef struct { int n; size_t D; const cplx* H; int nl; cplx* const* L; cplx* const* G; int is_dm; int nonherm; cplx* tmp; } evo_rhs;
static void evo_f(const evo_rhs* e, const cplx* y, cplx* dy) {
    size_t D = e->D;
    if (e->is_dm) {
        memset(dy, 0, D * D * sizeof(cplx));
        if (e->H) {
            matmul(D, e->H, y, e->tmp);
            for (size_t i = 0; i < D * D; ++i) dy[i] += -I * e->tmp[i];
            matmul(D, y, e->H, e->tmp);
            for (size_t i = 0; i < D * D; ++i) dy[i] += I * e->tmp[i];
        }
        for (int j = 0; j < e->nl; ++j) {
            cplx* t2 = e->tmp + D * D;
            matmul(D, e->L[j], y, e->tmp); matmul_adj(D, e->tmp, e->L[j], t2);
            for (size_t i = 0; i < D * D; ++i) dy[i] += t2[i];
            matmul(D, e->G[j], y, e->tmp);
            for (size_t i = 0; i < D * D; ++i) dy[i] -= 0.5 * e->tmp[i];
            matmul(D, y, e->G[j], e->tmp);
            for (size_t i = 0; i < D * D; ++i) dy[i] -= 0.5 * e->tmp[i];
        }
    } else {
        for (size_t r = 0; r < D; ++r) {
            cplx acc = 0;
            if (e->H) for (size_t c = 0; c < D; ++c) acc += -I * e->H[r + c * D] * y[c];
            if (e->nonherm) for (int j = 0; j < e->nl; ++j) for (size_t c = 0; c < D; ++c) acc += -0.5 * e->G[j][r + c * D] * y[c];
            dy[r] = acc;
        }
    }
}
/* one Dormand-Prince step of size h from y (k1 = f(y) given); writes the 5th-order solution and returns the scaled
 * error norm sqrt(mean(|err_i / (abstol + reltol max(|y_i|, |ynew_i|))|^2)) */
static double dp5_step(const evo_rhs* e, size_t N, const cplx* y, const cplx* k1, double h, cplx* ynew, cplx* k7, cplx* work) {
    static const double a21 = 1.0 / 5, a31 = 3.0 / 40, a32 = 9.0 / 40, a41 = 44.0 / 45, a42 = -56.0 / 15, a43 = 32.0 / 9,
        a51 = 19372.0 / 6561, a52 = -25360.0 / 2187, a53 = 64448.0 / 6561, a54 = -212.0 / 729,
        a61 = 9017.0 / 3168, a62 = -355.0 / 33, a63 = 46732.0 / 5247, a64 = 49.0 / 176, a65 = -5103.0 / 18656,
        b1 = 35.0 / 384, b3 = 500.0 / 1113, b4 = 125.0 / 192, b5 = -2187.0 / 6784, b6 = 11.0 / 84,
        e1 = 71.0 / 57600, e3 = -71.0 / 16695, e4 = 71.0 / 1920, e5 = -17253.0 / 339200, e6 = 22.0 / 525, e7 = -1.0 / 40;
    cplx *k2 = work, *k3 = work + N, *k4 = work + 2 * N, *k5 = work + 3 * N, *k6 = work + 4 * N, *yt = work + 5 * N;
    for (size_t i = 0; i < N; ++i) yt[i] = y[i] + h * a21 * k1[i];
    evo_f(e, yt, k2);
    for (size_t i = 0; i < N; ++i) yt[i] = y[i] + h * (a31 * k1[i] + a32 * k2[i]);
    evo_f(e, yt, k3);
    for (size_t i = 0; i < N; ++i) yt[i] = y[i] + h * (a41 * k1[i] + a42 * k2[i] + a43 * k3[i]);
    evo_f(e, yt, k4);
    for (size_t i = 0; i < N; ++i) yt[i] = y[i] + h * (a51 * k1[i] + a52 * k2[i] + a53 * k3[i] + a54 * k4[i]);
    evo_f(e, yt, k5);
    for (size_t i = 0; i < N; ++i) yt[i] = y[i] + h * (a61 * k1[i] + a62 * k2[i] + a63 * k3[i] + a64 * k4[i] + a65 * k5[i]);
    evo_f(e, yt, k6);
    for (size_t i = 0; i < N; ++i) ynew[i] = y[i] + h * (b1 * k1[i] + b3 * k3[i] + b4 * k4[i] + b5 * k5[i] + b6 * k6[i]);
    evo_f(e, ynew, k7);
    double acc = 0;
    for (size_t i = 0; i < N; ++i) {
        cplx err = h * (e1 * k1[i] + e3 * k3[i] + e4 * k4[i] + e5 * k5[i] + e6 * k6[i] + e7 * k7[i]);
        double sc = 1e-8 + 1e-6 * fmax(cabs(y[i]), cabs(ynew[i]));
        acc += (creal(err) * creal(err) + cimag(err) * cimag(err)) / (sc * sc);
    }
    return sqrt(acc / (double)N);
}
/* adaptive integration over [0, T]; stop_norm2 >= 0 (mcwf): return at the first accepted step after which |y|^2 <=
 * stop_norm2, with *t_out the time reached.  Returns 0, or 4 when the step size underflows. */
static int dp5_integrate(const evo_rhs* e, size_t N, cplx* y, double t0, double T, double stop_norm2, double* t_out) {
    cplx* k1 = (cplx*)malloc(N * sizeof(cplx)); cplx* k7 = (cplx*)malloc(N * sizeof(cplx));
    cplx* yn = (cplx*)malloc(N * sizeof(cplx)); cplx* work = (cplx*)malloc(6 * N * sizeof(cplx));
    double t = t0, h = (T - t0) * 1e-2;
    int rc = 0;
    evo_f(e, y, k1);
    while (t < T) {
        if (t + h > T) h = T - t;
        if (h < 1e-14 * fmax(1.0, fabs(T))) { rc = 4; break; }
        double err = dp5_step(e, N, y, k1, h, yn, k7, work);
        if (err <= 1.0) {
            if (stop_norm2 >= 0) {
                double n2 = 0; for (size_t i = 0; i < N; ++i) n2 += creal(yn[i]) * creal(yn[i]) + cimag(yn[i]) * cimag(yn[i]);
                if (n2 <= stop_norm2) {
                    /* the jump lies inside this step: bisect the step size on |y(t + s)|^2 = stop_norm2 (each trial is one more
                     * DP5 step from the same start, the continuous-callback root search of the reference restated) */
                    double lo = 0, hi = h;
                    for (int it = 0; it < 40; ++it) {
                        double mid = 0.5 * (lo + hi);
                        dp5_step(e, N, y, k1, mid, yn, k7, work);
                        double m2 = 0; for (size_t i = 0; i < N; ++i) m2 += creal(yn[i]) * creal(yn[i]) + cimag(yn[i]) * cimag(yn[i]);
                        if (m2 <= stop_norm2) hi = mid; else lo = mid;
                    }
                    dp5_step(e, N, y, k1, hi, yn, k7, work);
                    memcpy(y, yn, N * sizeof(cplx)); t += hi;
                    break;
                }
            }
            memcpy(y, yn, N * sizeof(cplx)); memcpy(k1, k7, N * sizeof(cplx));   /* first-same-as-last */
            t += h;
        }
        double q = err > 0 ? 0.9 * pow(err, -0.2) : 5.0;
        h *= fmin(5.0, fmax(0.2, q));
    }
    if (t_out) *t_out = t;
    free(k1); free(k7); free(yn); free(work);
    return rc;
}
/* builds the embedded operators exactly as _noninstant_operators / _embedded_lindblad_operators do (noninstant.jl:1-8,
 * uptotime.jl:6-18): H = embed(basis, state_indices, H_k), L_j = embed(basis, [idx[l_pos[j]]], L_j) */
static void evo_build(const qo_state* s, const int* idx, int k, const double* H_ri, int nl, const int* l_pos, const double* L_ri,
                      cplx** H_out, cplx** L, cplx** G) {
    size_t D = (size_t)1 << s->n, K = (size_t)1 << k;
    *H_out = NULL;
    if (H_ri) {
        cplx* hk = (cplx*)malloc(K * K * sizeof(cplx));
        for (size_t i = 0; i < K * K; ++i) hk[i] = H_ri[2 * i] + I * H_ri[2 * i + 1];
        *H_out = (cplx*)malloc(D * D * sizeof(cplx));
        qo_embed(s->n, idx, k, hk, *H_out);
        free(hk);
    }
    for (int j = 0; j < nl; ++j) {
        cplx l2[4];
        for (int i = 0; i < 4; ++i) l2[i] = L_ri[8 * j + 2 * i] + I * L_ri[8 * j + 2 * i + 1];
        L[j] = (cplx*)malloc(D * D * sizeof(cplx)); G[j] = (cplx*)malloc(D * D * sizeof(cplx));
        int target = idx[l_pos[j]];
        qo_embed(s->n, &target, 1, l2, L[j]);
        for (size_t c = 0; c < D; ++c) for (size_t r = 0; r < D; ++r) { cplx acc = 0; for (size_t m = 0; m < D; ++m) acc += conj(L[j][m + r * D]) * L[j][m + c * D]; G[j][r + c * D] = acc; }
    }
}
/* apply_noninstant!(state::Operator | Ket, ...) -> timeevolution.master (a Ket is promoted with dm(), noninstant.jl:21-23) */
int qo_evolve_master(qo_state* s, const int* idx, int k, const double* H_ri, int nl, const int* l_pos, const double* L_ri, double duration) {
    if (nl > 8) return 3;
    if (s->kind == 0) { qo_state* p = qo_dm(s); free(s->d); s->d = p->d; s->kind = 1; free(p); }
    size_t D = (size_t)1 << s->n;
    cplx *H, *L[8], *G[8];
    evo_build(s, idx, k, H_ri, nl, l_pos, L_ri, &H, L, G);
    evo_rhs e = {s->n, D, H, nl, L, G, 1, 0, (cplx*)malloc(2 * D * D * sizeof(cplx))};
    int rc = dp5_integrate(&e, D * D, s->d, 0.0, duration, -1.0, NULL);
    free(e.tmp); free(H);
    for (int j = 0; j < nl; ++j) { free(L[j]); free(G[j]); }
    return rc;
}
/* apply_noninstant!(state::MCKet, ...): schroedinger when there are no collapse operators, else mcwf (noninstant.jl:25-37).
 * Draws: u_j = qso_draw(seed, traj, draw_id + j * 2^32), j = 0, 1, ... (jump thresholds and jump choices alternate). */
int qo_evolve_mcwf(qo_state* s, const int* idx, int k, const double* H_ri, int nl, const int* l_pos, const double* L_ri, double duration,
                   uint64_t seed, uint64_t traj, uint64_t draw_id, int* njumps_out) {
    if (s->kind != 0 || nl > 8) return 3;
    size_t D = (size_t)1 << s->n;
    cplx *H, *L[8], *G[8];
    evo_build(s, idx, k, H_ri, nl, l_pos, L_ri, &H, L, G);
    evo_rhs e = {s->n, D, H, nl, L, G, 0, 1, NULL};
    int rc = 0, jumps = 0;
    uint64_t sub = 0;
    double t = 0;
    if (nl == 0) rc = dp5_integrate(&e, D, s->d, 0.0, duration, -1.0, NULL);
    else while (t < duration && rc == 0) {
        double r = qso_draw(seed, traj, draw_id + ((sub++) << 32));
        rc = dp5_integrate(&e, D, s->d, t, duration, r, &t);
        if (rc || t >= duration) break;
        double P[8] = {0}, total = 0;
        cplx* br = (cplx*)malloc(D * sizeof(cplx));
        for (int j = 0; j < nl; ++j) {
            double nn = 0;
            for (size_t r2 = 0; r2 < D; ++r2) { cplx acc = 0; for (size_t c = 0; c < D; ++c) acc += L[j][r2 + c * D] * s->d[c]; nn += creal(acc) * creal(acc) + cimag(acc) * cimag(acc); }
            P[j] = nn; total += nn;
        }
        double thr = qso_draw(seed, traj, draw_id + ((sub++) << 32)) * total, cum = 0;
        int sel = nl - 1;
        for (int j = 0; j < nl; ++j) { cum += P[j]; if (cum > thr) { sel = j; break; } }
        for (size_t r2 = 0; r2 < D; ++r2) { cplx acc = 0; for (size_t c = 0; c < D; ++c) acc += L[sel][r2 + c * D] * s->d[c]; br[r2] = acc; }
        double nrm = sqrt(P[sel]);
        for (size_t r2 = 0; r2 < D; ++r2) s->d[r2] = br[r2] / nrm;
        free(br);
        ++jumps;
    }
    double nn = 0;
    for (size_t r2 = 0; r2 < D; ++r2) nn += creal(s->d[r2]) * creal(s->d[r2]) + cimag(s->d[r2]) * cimag(s->d[r2]);
    nn = sqrt(nn);
    for (size_t r2 = 0; r2 < D; ++r2) s->d[r2] /= nn;   /* mcwf returns normalised kets */
    if (njumps_out) *njumps_out = jumps;
    free(H);
    for (int j = 0; j < nl; ++j) { free(L[j]); free(G[j]); }
    return rc;
}

/* ==========================================================================================
 *                           STABILIZER (CliffordRepr) PART
 *  Generator-list representation of a (possibly mixed) stabilizer state: r <= n independent
 *  commuting Hermitian Paulis with signs.  x/z are n bytes per generator, sign in {0,1}.
 *  Replaces QuantumClifford.MixedDestabilizer at the level of group-invariant behaviour
 *  (src/backends/clifford/*.jl call sites cited per function).
 * ======================================================================================== */
#define QC_MAXN 64
typedef struct {
    int n, r;
    uint8_t x[QC_MAXN][QC_MAXN], z[QC_MAXN][QC_MAXN];
    uint8_t ph[QC_MAXN]; /* exponent of i, always 0 or 2 for stored generators */
} qc_state;

typedef struct { uint8_t x[QC_MAXN], z[QC_MAXN]; uint8_t ph; } qc_pauli;

/* exponent of i picked up when multiplying single-qubit Paulis (x1,z1)*(x2,z2), Aaronson-Gottesman g */
static int ag_g(int x1, int z1, int x2, int z2) {
    if (!x1 && !z1) return 0;
    if (x1 && z1) return z2 - x2;
    if (x1 && !z1) return z2 * (2 * x2 - 1);
    return x2 * (1 - 2 * z2);
}
/* dst <- src * dst (left multiplication; they commute whenever it matters) */
static void pauli_mul_left(int n, uint8_t* dx, uint8_t* dz, uint8_t* dph, const uint8_t* sx, const uint8_t* sz, uint8_t sph) {
    int e = *dph + sph;
    for (int q = 0; q < n; ++q) { e += ag_g(sx[q], sz[q], dx[q], dz[q]); dx[q] ^= sx[q]; dz[q] ^= sz[q]; }
    *dph = (uint8_t)(((e % 4) + 4) % 4);
}
static int pauli_anticomm(int n, const uint8_t* ax, const uint8_t* az, const uint8_t* bx, const uint8_t* bz) {
    int c = 0;
    for (int q = 0; q < n; ++q) c ^= (ax[q] & bz[q]) ^ (az[q] & bx[q]);
    return c;
}

qc_state* qc_alloc(int n) { qc_state* s = (qc_state*)calloc(1, sizeof(qc_state)); s->n = n; return s; }
void qc_free(qc_state* s) { free(s); }
qc_state* qc_copy(const qc_state* s) { qc_state* c = (qc_state*)malloc(sizeof(qc_state)); memcpy(c, s, sizeof(qc_state)); return c; }
int qc_nqubits(const qc_state* s) { return s->n; }
int qc_rank(const qc_state* s) { return s->r; }
/* newstate(::Qubit, ::CliffordRepr) = express(Z1, CliffordRepr()): stabilizer +Z, clifford.jl:10-13 */
qc_state* qc_newstate(void) { qc_state* s = qc_alloc(1); s->r = 1; s->z[0][0] = 1; return s; }
/* generators given as bytes: xs[g*n+q], zs[g*n+q], signs[g] in {0,1} (1 = minus) */
qc_state* qc_from_generators(int n, int r, const uint8_t* xs, const uint8_t* zs, const uint8_t* signs) {
    qc_state* s = qc_alloc(n); s->r = r;
    for (int g = 0; g < r; ++g) { for (int q = 0; q < n; ++q) { s->x[g][q] = xs[g * n + q]; s->z[g][q] = zs[g * n + q]; } s->ph[g] = signs[g] ? 2 : 0; }
    return s;
}
void qc_get_generators(const qc_state* s, uint8_t* xs, uint8_t* zs, uint8_t* signs) {
    for (int g = 0; g < s->r; ++g) { for (int q = 0; q < s->n; ++q) { xs[g * s->n + q] = s->x[g][q]; zs[g * s->n + q] = s->z[g][q]; } signs[g] = s->ph[g] == 2; }
}

/* subsystemcompose(states::MixedDestabilizer...) = QuantumClifford.tensor, clifford.jl:1 */
qc_state* qc_tensor(const qc_state* a, const qc_state* b) {
    qc_state* s = qc_alloc(a->n + b->n); s->r = a->r + b->r;
    for (int g = 0; g < a->r; ++g) { memcpy(s->x[g], a->x[g], a->n); memcpy(s->z[g], a->z[g], a->n); s->ph[g] = a->ph[g]; }
    for (int g = 0; g < b->r; ++g) { memcpy(s->x[a->r + g] + a->n, b->x[g], b->n); memcpy(s->z[a->r + g] + a->n, b->z[g], b->n); s->ph[a->r + g] = b->ph[g]; }
    return s;
}

/* apply!(state, sGate(idx...)): conjugate every generator by a k-qubit (k=1,2) Clifford given
 * as a lookup table over the generator's restriction to the target qubits: entry
 * t = x0 | z0<<1 | x1<<2 | z1<<3  ->  table[t] = x0' | z0'<<1 | x1'<<2 | z1'<<3 | signflip<<4
 * where (x,z) uses the Hermitian convention (1,1)=Y.  Call sites: src/baseops/apply.jl:20 via
 * express(op, CliffordRepr(), UseAsOperation()); uptotime.jl:7 (sX/sY/sZ). */
int qc_apply(qc_state* s, const int* idx, int k, const uint8_t* table) {
    for (int g = 0; g < s->r; ++g) {
        int t = 0;
        for (int m = 0; m < k; ++m) t |= (s->x[g][idx[m]] << (2 * m)) | (s->z[g][idx[m]] << (2 * m + 1));
        int o = table[t];
        for (int m = 0; m < k; ++m) { s->x[g][idx[m]] = (o >> (2 * m)) & 1; s->z[g][idx[m]] = (o >> (2 * m + 1)) & 1; }
        if ((o >> (2 * k)) & 1) s->ph[g] ^= 2;
    }
    return 0;
}

/* uptotime!(state::MixedDestabilizer, idx, bg, dt): r=rand() (always consumed); cum=1; for each
 * (prob, P): cum -= prob; if r>cum apply P and stop -- clifford/uptotime.jl:1-12 with paulinoise
 * T2 -> ((p/2,Z),) :26-29 and Depolarization -> ((p/4,X),(p/4,Y),(p/4,Z)) :36-39.
 * Returns the Pauli applied (0 none, 1 X, 2 Y, 3 Z) or -3 for backgrounds with no method. */
int qc_uptotime(qc_state* s, int idx, int bg, const double* prm, double dt, double u) {
    double probs[3]; int ops[3]; int np;
    if (bg == BG_T2) { double p = 1 - exp(-dt / prm[0]); probs[0] = p / 2; ops[0] = 3; np = 1; }
    else if (bg == BG_DEPOL) { double p = 1 - exp(-dt / prm[0]); probs[0] = probs[1] = probs[2] = p / 4; ops[0] = 1; ops[1] = 2; ops[2] = 3; np = 3; }
    else return -3;
    double cum = 1.0;
    for (int i = 0; i < np; ++i) {
        cum -= probs[i];
        if (u > cum) {
            int P = ops[i];
            int px = (P == 1 || P == 2), pz = (P == 3 || P == 2);
            for (int g = 0; g < s->r; ++g) if ((s->x[g][idx] & pz) ^ (s->z[g][idx] & px)) s->ph[g] ^= 2;
            return P;
        }
    }
    return 0;
}

/* Bring the generator list to reduced row-echelon form (columns: x_0..x_{n-1}, z_0..z_{n-1}).
 * This is the canonical representative of the stabilizer group used for group equality. */
void qc_canonicalize(qc_state* s) {
    int n = s->n, row = 0;
    for (int col = 0; col < 2 * n && row < s->r; ++col) {
        int q = col % n, isz = col >= n, piv = -1;
        for (int g = row; g < s->r; ++g) if ((isz ? s->z[g][q] : s->x[g][q])) { piv = g; break; }
        if (piv < 0) continue;
        if (piv != row) {
            uint8_t tx[QC_MAXN], tz[QC_MAXN], tp;
            memcpy(tx, s->x[piv], n); memcpy(tz, s->z[piv], n); tp = s->ph[piv];
            memcpy(s->x[piv], s->x[row], n); memcpy(s->z[piv], s->z[row], n); s->ph[piv] = s->ph[row];
            memcpy(s->x[row], tx, n); memcpy(s->z[row], tz, n); s->ph[row] = tp;
        }
        for (int g = 0; g < s->r; ++g)
            if (g != row && (isz ? s->z[g][q] : s->x[g][q])) pauli_mul_left(n, s->x[g], s->z[g], &s->ph[g], s->x[row], s->z[row], s->ph[row]);
        ++row;
    }
}

/* Is the (commuting) Pauli P = +/- a group element?  Returns 1 and the exponent of i of the
 * group element equal to P's bits in *ph_out (0 -> +P in group, 2 -> -P), else 0. */
static int group_membership(const qc_state* s, const qc_pauli* P, uint8_t* ph_out) {
    qc_state* c = qc_copy(s);
    qc_canonicalize(c);
    int n = s->n;
    uint8_t ax[QC_MAXN] = {0}, az[QC_MAXN] = {0}, aph = 0; /* accumulated product */
    uint8_t rx[QC_MAXN], rz[QC_MAXN];
    memcpy(rx, P->x, n); memcpy(rz, P->z, n);
    /* reduce by each RREF row's pivot (its first set bit in column order x_0..x_{n-1},z_0..) */
    for (int row = 0; row < c->r; ++row) {
        int first = -1;
        for (int cc = 0; cc < 2 * n; ++cc) { int qq = cc % n; if (cc >= n ? c->z[row][qq] : c->x[row][qq]) { first = cc; break; } }
        if (first < 0) continue;
        int q = first % n, isz = first >= n;
        if (isz ? rz[q] : rx[q]) {
            for (int k = 0; k < n; ++k) { rx[k] ^= c->x[row][k]; rz[k] ^= c->z[row][k]; }
            /* acc <- gen * acc (generators commute, order irrelevant) */
            pauli_mul_left(n, ax, az, &aph, c->x[row], c->z[row], c->ph[row]);
        }
    }
    int zero = 1;
    for (int k = 0; k < n; ++k) if (rx[k] | rz[k]) zero = 0;
    free(c);
    if (!zero) return 0;
    *ph_out = aph;
    return 1;
}

/* project!(state, P): measure Hermitian Pauli P (ph in {0,2}).  Outcome random iff P is not
 * +/- in the group: either it anticommutes with a generator, or it commutes with all of them
 * but is independent (mixed state; rank grows).  u is the injected draw; the random result is
 * 0x2 (outcome 2) when u >= 0.5 else 0x0.  Mirrors projectX!/Y!/Z! + the
 * `isnothing(res) && (res = phases[anticom] = rand((0x0,0x2)))` step of projectremoverand!
 * as used at src/backends/clifford/express.jl:6.  Returns result phase (0 or 2). */
int qc_project(qc_state* s, const qc_pauli* P, double u, int* was_random) {
    int n = s->n, piv = -1;
    for (int g = 0; g < s->r; ++g) if (pauli_anticomm(n, s->x[g], s->z[g], P->x, P->z)) { piv = g; break; }
    if (piv >= 0) {
        for (int g = 0; g < s->r; ++g)
            if (g != piv && pauli_anticomm(n, s->x[g], s->z[g], P->x, P->z))
                pauli_mul_left(n, s->x[g], s->z[g], &s->ph[g], s->x[piv], s->z[piv], s->ph[piv]);
        uint8_t res = (u >= 0.5) ? 2 : 0;
        memcpy(s->x[piv], P->x, n); memcpy(s->z[piv], P->z, n); s->ph[piv] = (uint8_t)((P->ph + res) & 3);
        *was_random = 1;
        return res;
    }
    uint8_t gph;
    if (group_membership(s, P, &gph)) {
        *was_random = 0;
        /* group holds i^gph * bits; measured operator is i^P.ph * bits -> eigenvalue i^(gph - P.ph) */
        return ((gph - P->ph) & 3);
    }
    /* independent commuting Pauli on a mixed state: random, rank += 1 */
    uint8_t res = (u >= 0.5) ? 2 : 0;
    memcpy(s->x[s->r], P->x, n); memcpy(s->z[s->r], P->z, n); s->ph[s->r] = (uint8_t)((P->ph + res) & 3);
    s->r += 1;
    *was_random = 1;
    return res;
}

/* Remove qubit q from every generator, dropping the generators that act non-trivially on it
 * after Gaussian elimination on its (x,z) columns: the stabilizer group of ptrace_q(rho) is
 * { g in G : g|_q = I }.  QuantumClifford.ptrace(state,(i,)) at clifford.jl:7-8. */
int qc_ptrace(qc_state* s, int q) {
    int n = s->n;
    int used[2] = {-1, -1};
    for (int pass = 0; pass < 2; ++pass) { /* pass 0: x column, pass 1: z column */
        int piv = -1;
        for (int g = 0; g < s->r; ++g) {
            if (g == used[0]) continue;
            if (pass == 0 ? s->x[g][q] : s->z[g][q]) { piv = g; break; }
        }
        if (piv < 0) continue;
        for (int g = 0; g < s->r; ++g) {
            if (g == piv || g == used[0]) continue;
            if (pass == 0 ? s->x[g][q] : s->z[g][q]) pauli_mul_left(n, s->x[g], s->z[g], &s->ph[g], s->x[piv], s->z[piv], s->ph[piv]);
        }
        used[pass] = piv;
    }
    int w = 0;
    for (int g = 0; g < s->r; ++g) {
        if (g == used[0] || g == used[1]) continue;
        int c = 0;
        for (int k = 0; k < n; ++k) { if (k == q) continue; s->x[w][c] = s->x[g][k]; s->z[w][c] = s->z[g][k]; ++c; }
        s->x[w][n - 1] = 0; s->z[w][n - 1] = 0;
        s->ph[w] = s->ph[g];
        ++w;
    }
    s->r = w; s->n = n - 1;
    return 0;
}

/* project_traceout!(state::MixedDestabilizer, idx, basis): projectremoverand!(state,
 * projectX!/Y!/Z!, idx); returns Int(res)÷2+1 -- clifford/express.jl:2-14.
 * basis: 1 X, 2 Y, 3 Z.  After the projection qubit idx is in a product state, so removing it
 * is the partial trace. */
int qc_project_traceout(qc_state* s, int idx, int basis, double u, int* was_random) {
    qc_pauli P; memset(&P, 0, sizeof(P));
    P.x[idx] = (basis == 1 || basis == 2); P.z[idx] = (basis == 3 || basis == 2);
    int res = qc_project(s, &P, u, was_random);
    qc_ptrace(s, idx);
    return res / 2 + 1;
}

/* observable(state::MixedDestabilizer, indices, Pauli): QuantumClifford.expect(embed(n,indices,P),
 * state) in {-1,0,+1} -- clifford/express.jl:16-20.  paulis[m] in {0 I,1 X,2 Y,3 Z}; sign: 1 = -P. */
int qc_expect_pauli(const qc_state* s, const int* idx, int k, const uint8_t* paulis, int sign) {
    qc_pauli P; memset(&P, 0, sizeof(P));
    for (int m = 0; m < k; ++m) { P.x[idx[m]] = (paulis[m] == 1 || paulis[m] == 2); P.z[idx[m]] = (paulis[m] == 3 || paulis[m] == 2); }
    P.ph = sign ? 2 : 0;
    for (int g = 0; g < s->r; ++g) if (pauli_anticomm(s->n, s->x[g], s->z[g], P.x, P.z)) return 0;
    uint8_t gph;
    if (!group_membership(s, &P, &gph)) return 0;
    return ((gph - P.ph) & 3) == 0 ? 1 : -1;
}

/* observable(state, indices, SProjector(stabilizer state)): QuantumClifford.expect(indices,
 * pstate, state) -- clifford/express.jl:49-52.  tr(Pi rho) with Pi = 2^-k sum_{h in <g_1..g_k>} h:
 * the h with non-zero expectation form a subgroup H on which the sign is a character, so the
 * value is |H|/2^k when every sign is +, else 0.  Generators of the projector state are k Paulis
 * on k qubits: gx[g*k+m], gz[g*k+m], gsign[g]; qubit m of the projector sits on subsystem idx[m]. */
double qc_expect_proj(const qc_state* s, const int* idx, int k, const uint8_t* gx, const uint8_t* gz, const uint8_t* gsign) {
    int n = s->n;
    long total = 0;
    for (unsigned sub = 0; sub < (1u << k); ++sub) {
        uint8_t hx[QC_MAXN] = {0}, hz[QC_MAXN] = {0}, hph = 0;
        for (int g = 0; g < k; ++g) if ((sub >> g) & 1) {
            uint8_t ex[QC_MAXN] = {0}, ez[QC_MAXN] = {0};
            for (int m = 0; m < k; ++m) { ex[idx[m]] = gx[g * k + m]; ez[idx[m]] = gz[g * k + m]; }
            pauli_mul_left(n, hx, hz, &hph, ex, ez, gsign[g] ? 2 : 0);
        }
        qc_pauli P; memset(&P, 0, sizeof(P));
        memcpy(P.x, hx, n); memcpy(P.z, hz, n); P.ph = hph;
        int anti = 0;
        for (int g = 0; g < s->r; ++g) if (pauli_anticomm(n, s->x[g], s->z[g], P.x, P.z)) { anti = 1; break; }
        if (anti) continue;
        uint8_t gph;
        if (!group_membership(s, &P, &gph)) continue;
        total += (((gph - P.ph) & 3) == 0) ? 1 : -1;
    }
    return (double)total / (double)(1u << k);
}
