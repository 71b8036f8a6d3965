// qsb_tab.cu -- bit-sliced stabilizer tableaux (QSB_TAB, the CliffordRepr path), sm_100a.
//
// One 32-bit word holds the same tableau bit of 32 consecutive trajectories; one thread owns one
// word column of every plane, so all row/column operations are word-wide boolean algebra with
// per-lane masks for data-dependent pivots and nothing is ever exchanged between threads.
// The tableau is a list of n symplectic pairs (D_i, S_i) with a per-lane `active` bit (see
// qsb_internal.cuh): rank, generator order and destabilizer signs are not observable, so they are
// chosen to keep the ensemble in lockstep instead of mirroring QuantumClifford's row order.
#include <math.h>

#include "qsb_internal.cuh"
#include "qsb_tab_core.cuh"

#define TTPB 128

static inline unsigned tgrid(const qsb_ctx* ctx) { return (unsigned)((ctx->nwords + TTPB - 1) / TTPB); }

// ---------------------------------------------------------------------------------------------
// construction / copies
// ---------------------------------------------------------------------------------------------
__global__ void k_tab_init(uint32_t* p, uint32_t* alive, const uint32_t* valid, size_t W, int n, int active) {
    size_t w = blockIdx.x * (size_t)TTPB + threadIdx.x;
    if (w >= W) return;
    tab_init_word(p, W, w, n, active);
    alive[w] = valid[w];
}
int qsbt_new(qsb_ctx* ctx, int n, int active, qsb_state** out) {
    qsb_state* st;
    QSB_TRY(qsb_state_alloc(ctx, QSB_TAB, n, &st));
    if (ctx->rec) { TabOp* o = qsb_rec_push(ctx, TOP_INIT); o->n = n; o->s[0] = st->vbase; o->al[0] = st->valive; o->iv[0] = active; *out = st; return QSB_OK; }
    QSB_PRE(ctx, "k_tab_init");
    k_tab_init<<<tgrid(ctx), TTPB, 0, ctx->stream>>>((uint32_t*)st->data, st->alive, ctx->valid, (size_t)ctx->nwords, n, active);
    QSB_LAUNCH_CHECK(ctx);
    *out = st;
    return QSB_OK;
}

__global__ void k_tab_tensor(uint32_t* out, uint32_t* alive_out, const uint32_t* a, const uint32_t* b, const uint32_t* alive_a, const uint32_t* alive_b,
                             size_t W, int na, int nb) {
    size_t w = blockIdx.x * (size_t)TTPB + threadIdx.x;
    if (w >= W) return;
    tab_tensor_word(out, a, b, W, w, na, nb);
    alive_out[w] = alive_a[w] & alive_b[w];
}
int qsbt_tensor2(qsb_ctx* ctx, const qsb_state* a, const qsb_state* b, qsb_state** out) {
    if (a->n + b->n > MAXN) QSB_FAIL(QSB_ERR_UNSUPPORTED, "composite tableau would exceed %d qubits", MAXN);
    qsb_state* st;
    QSB_TRY(qsb_state_alloc(ctx, QSB_TAB, a->n + b->n, &st));
    if (ctx->rec) {
        TabOp* o = qsb_rec_push(ctx, TOP_TENSOR); o->n = a->n; o->n2 = b->n;
        o->s[0] = st->vbase; o->s[1] = a->vbase; o->s[2] = b->vbase; o->al[0] = st->valive; o->al[1] = a->valive; o->al[2] = b->valive;
        qsb_account(ctx, "k_tab_tensor", (uint64_t)(qsb_tab_planes(a->n) + qsb_tab_planes(b->n) + qsb_tab_planes(st->n)) * ctx->ntraj / 8); ctx->ops += ctx->ntraj; *out = st; return QSB_OK;
    }
    QSB_PRE(ctx, "k_tab_tensor");
    k_tab_tensor<<<tgrid(ctx), TTPB, 0, ctx->stream>>>((uint32_t*)st->data, st->alive, (const uint32_t*)a->data, (const uint32_t*)b->data, a->alive, b->alive,
                                                       (size_t)ctx->nwords, a->n, b->n);
    QSB_LAUNCH_CHECK(ctx);
    qsb_account(ctx, "k_tab_tensor", (uint64_t)(qsb_tab_planes(a->n) + qsb_tab_planes(b->n) + qsb_tab_planes(st->n)) * ctx->ntraj / 8); ctx->ops += ctx->ntraj;
    *out = st;
    return QSB_OK;
}

struct TabSelArgs { const uint32_t* src[8]; const uint32_t* alive[8]; double cum[8]; int m; };
__global__ void k_tab_select(uint32_t* out, uint32_t* alive_out, TabSelArgs a, size_t W, int planes, uint64_t seed, uint64_t traj0, uint64_t draw_id) {
    size_t w = blockIdx.x * (size_t)TTPB + threadIdx.x;
    if (w >= W) return;
    uint32_t m[8];
    for (int i = 0; i < a.m; ++i) m[i] = 0;
    for (int l = 0; l < 32; ++l) {
        double u = qsb_draw(seed, traj0 + (uint64_t)(w * 32 + l), draw_id);
        int sel = a.m - 1;
        for (int i = 0; i < a.m; ++i) if (a.cum[i] > u) { sel = i; break; }
        m[sel] |= 1u << l;
    }
    for (int p = 0; p < planes; ++p) {
        uint32_t v = 0;
        for (int i = 0; i < a.m; ++i) v |= m[i] & a.src[i][(size_t)p * W + w];
        out[(size_t)p * W + w] = v;
    }
    uint32_t al = 0;
    for (int i = 0; i < a.m; ++i) al |= m[i] & a.alive[i][w];
    alive_out[w] = al;
}
int qsbt_select(qsb_ctx* ctx, const qsb_state* const* states, const double* w, int m, uint64_t draw_id, qsb_state** out) {
    if (m > 8) QSB_FAIL(QSB_ERR_UNSUPPORTED, "mixtures of more than 8 states are not implemented");
    qsb_state* st;
    QSB_TRY(qsb_state_alloc(ctx, QSB_TAB, states[0]->n, &st));
    TabSelArgs a; a.m = m;
    double total = 0, cum = 0;
    for (int i = 0; i < m; ++i) total += w[i];
    for (int i = 0; i < m; ++i) { cum += w[i] / total; a.cum[i] = cum; a.src[i] = (const uint32_t*)states[i]->data; a.alive[i] = states[i]->alive; }
    if (ctx->rec) {
        if (m > 5) { qsb_state_free(st); QSB_FAIL(QSB_ERR_UNSUPPORTED, "recorded mixtures are limited to 5 states"); }
        TabOp* o = qsb_rec_push(ctx, TOP_SELECT); o->n = st->n; o->n2 = m; o->draw = draw_id;
        o->s[0] = st->vbase; o->al[0] = st->valive;
        for (int i = 0; i < m; ++i) { o->s[1 + i] = states[i]->vbase; o->al[1 + i] = states[i]->valive; o->f[i] = a.cum[i]; }
        qsb_account(ctx, "k_tab_select", (uint64_t)2 * qsb_tab_planes(st->n) * ctx->ntraj / 8); ctx->ops += ctx->ntraj; *out = st; return QSB_OK;
    }
    QSB_PRE(ctx, "k_tab_select");
    k_tab_select<<<tgrid(ctx), TTPB, 0, ctx->stream>>>((uint32_t*)st->data, st->alive, a, (size_t)ctx->nwords, qsb_tab_planes(st->n), ctx->seed, (uint64_t)ctx->traj_offset, draw_id);
    QSB_LAUNCH_CHECK(ctx);
    qsb_account(ctx, "k_tab_select", (uint64_t)2 * qsb_tab_planes(st->n) * ctx->ntraj / 8); ctx->ops += ctx->ntraj;
    *out = st;
    return QSB_OK;
}

int qsbt_to_host(const qsb_state* st, int64_t traj, uint8_t* xs, uint8_t* zs, uint8_t* signs, uint8_t* active) {
    qsb_ctx* ctx = st->ctx;
    int n = st->n, planes = qsb_tab_planes(n), S = 4 * n + 2;
    std::vector<uint32_t> h(planes);
    QSB_CUDA(cudaMemcpy2DAsync(h.data(), sizeof(uint32_t), (const uint32_t*)st->data + traj / 32, ctx->nwords * sizeof(uint32_t),
                               sizeof(uint32_t), planes, cudaMemcpyDeviceToHost, ctx->stream));
    QSB_CUDA(cudaStreamSynchronize(ctx->stream));
    int l = (int)(traj % 32);
    auto bit = [&](int plane) { return (uint8_t)((h[plane] >> l) & 1u); };
    for (int i = 0; i < n; ++i) {
        for (int c = 0; c < n; ++c) {
            xs[i * n + c] = bit(i * S + c); zs[i * n + c] = bit(i * S + n + c);
            xs[(n + i) * n + c] = bit(i * S + 2 * n + c); zs[(n + i) * n + c] = bit(i * S + 3 * n + c);
        }
        if (signs) { signs[i] = 0; signs[n + i] = bit(i * S + 4 * n); }
        if (active) active[i] = bit(i * S + 4 * n + 1);
    }
    return QSB_OK;
}

// ---------------------------------------------------------------------------------------------
// apply!: Clifford gate as a GF(2)-linear map of (x,z) bits plus a sign truth table (a4)
// ---------------------------------------------------------------------------------------------
__global__ void k_tab_gate(uint32_t* p, size_t W, TabGateArgs a, const uint32_t* __restrict__ mask) {
    size_t w = blockIdx.x * (size_t)TTPB + threadIdx.x;
    if (w >= W) return;
    tab_gate_word(p, W, w, a, mask ? mask[w] : 0xFFFFFFFFu);
}
int qsbt_apply(qsb_state* st, const int* idx, int k, const uint8_t* table, const qsb_mask* mask) {
    qsb_ctx* ctx = st->ctx;
    TabGateArgs a; memset(&a, 0, sizeof(a));
    a.n = st->n; a.k = k;
    for (int j = 0; j < k; ++j) a.q[j] = idx[j];
    int nb = 2 * k, full = (1 << nb) - 1;
    int cols[4];
    for (int j = 0; j < nb; ++j) cols[j] = table[1 << j] & full;
    for (int t = 0; t < (1 << nb); ++t) {
        int img = 0;
        for (int j = 0; j < nb; ++j) if ((t >> j) & 1) img ^= cols[j];
        if (img != (table[t] & full)) QSB_FAIL(QSB_ERR_BAD_ARG, "conjugation table is not a Clifford gate (not linear at entry %d)", t);
        if ((table[t] >> nb) & 1) a.lut |= 1u << t;
    }
    if (table[0] != 0) QSB_FAIL(QSB_ERR_BAD_ARG, "conjugation table maps the identity to a non-identity");
    for (int b = 0; b < nb; ++b) for (int j = 0; j < nb; ++j) if ((cols[j] >> b) & 1) a.outmask[b] |= 1 << j;
    a.anf = tab_lut_to_anf(a.lut, nb);
    if (mask) QSB_TRY(qsb_handle_ok(ctx, mask->vplane >= 0, mask->gen));
    if (ctx->rec) {
        TabOp* o = qsb_rec_push(ctx, TOP_GATE); o->n = a.n; o->k = a.k; o->q[0] = a.q[0]; o->q[1] = a.q[1]; o->lut = a.lut; o->u[0] = a.anf;
        for (int j = 0; j < 4; ++j) o->outmask[j] = a.outmask[j];
        o->s[0] = st->vbase; o->m[0] = mask ? mask->vplane : -1;
        qsb_account(ctx, "k_tab_gate", (uint64_t)2 * (2 * st->n) * (2 * k + 2) * ctx->ntraj / 8); ctx->ops += ctx->ntraj; return QSB_OK;
    }
    QSB_PRE(ctx, "k_tab_gate");
    k_tab_gate<<<tgrid(ctx), TTPB, 0, ctx->stream>>>((uint32_t*)st->data, (size_t)ctx->nwords, a, mask ? mask->w : nullptr);
    QSB_LAUNCH_CHECK(ctx);
    qsb_account(ctx, "k_tab_gate", (uint64_t)2 * (2 * st->n) * (2 * k + 2) * ctx->ntraj / 8); ctx->ops += ctx->ntraj;
    return QSB_OK;
}

// ---------------------------------------------------------------------------------------------
// uptotime!: Pauli noise (a7).  r=rand() always; cum=1; cum-=prob; if r>cum apply P, stop.
// ---------------------------------------------------------------------------------------------
// one thread per trajectory draws and decides, the warp ballots the word's error masks, then lane i updates the sign of S_i
// (a thread per word would run its 32 Philox draws serially: 18 -> 13 us per launch at 10^6 trajectories)
#define NZPB 256
__global__ void __launch_bounds__(NZPB) k_tab_noise(uint32_t* p, size_t W, TabNoiseArgs a, uint64_t seed, uint64_t traj0, uint64_t draw_id) {
    const size_t t = blockIdx.x * (size_t)NZPB + threadIdx.x, w = t >> 5;
    if (w >= W) return;   // whole warps leave together
    const uint32_t d = tab_noise_decide(a, qsb_draw(seed, traj0 + (uint64_t)t, draw_id));
    const uint32_t fx = __ballot_sync(0xffffffffu, d & 1u), fz = __ballot_sync(0xffffffffu, d & 2u);
    if ((fx | fz) == 0) return;
    const int i = (int)(threadIdx.x & 31u);
    if (i < a.n) {   // sign of S_i flips where it anticommutes with the applied Pauli (tab_noise_apply, one row per lane; n <= 32)
        TabView v = make_view(p, W, w, a.n);
        const uint32_t flip = (v.ld(v.sx(i, a.q)) & fz) ^ (v.ld(v.sz(i, a.q)) & fx);
        if (flip) v.st(v.sg(i), v.ld(v.sg(i)) ^ flip);
    }
}
int qsbt_channel(qsb_state* st, int idx, int bg, const double* prm, double dt, uint64_t draw_id) {
    qsb_ctx* ctx = st->ctx;
    TabNoiseArgs a; memset(&a, 0, sizeof(a));
    a.n = st->n; a.q = idx;
    // paulinoise(): clifford/uptotime.jl:26-39
    double probs[3];
    if (bg == QSB_BG_T2) { double p = 1 - exp(-dt / prm[0]); probs[0] = p / 2; a.op[0] = 3; a.np = 1; }
    else if (bg == QSB_BG_DEPOL) { double p = 1 - exp(-dt / prm[0]); probs[0] = probs[1] = probs[2] = p / 4; a.op[0] = 1; a.op[1] = 2; a.op[2] = 3; a.np = 3; }
    else QSB_FAIL(QSB_ERR_UNSUPPORTED, "no paulinoise method for this background on a Clifford state (reference raises MethodError)");
    double cum = 1.0;
    for (int i = 0; i < a.np; ++i) { cum -= probs[i]; a.cum[i] = cum; }
    if (ctx->rec) {
        TabOp* o = qsb_rec_push(ctx, TOP_NOISE); o->n = a.n; o->q[0] = a.q; o->k = a.np; o->draw = draw_id; o->s[0] = st->vbase;
        for (int j = 0; j < 3; ++j) { o->f[j] = a.cum[j]; o->iv[j] = a.op[j]; }
        qsb_account(ctx, "k_tab_noise", (uint64_t)(2 * st->n) * 3 * ctx->ntraj / 8); ctx->ops += ctx->ntraj; ctx->draws++; return QSB_OK;
    }
    QSB_PRE(ctx, "k_tab_noise");
    k_tab_noise<<<(unsigned)((ctx->nwords * 32 + NZPB - 1) / NZPB), NZPB, 0, ctx->stream>>>((uint32_t*)st->data, (size_t)ctx->nwords, a, ctx->seed, (uint64_t)ctx->traj_offset, draw_id);
    QSB_LAUNCH_CHECK(ctx);
    qsb_account(ctx, "k_tab_noise", (uint64_t)(2 * st->n) * 3 * ctx->ntraj / 8); ctx->ops += ctx->ntraj; ctx->draws++;
    return QSB_OK;
}

// ---------------------------------------------------------------------------------------------
// Pauli projection machinery
// ---------------------------------------------------------------------------------------------
__global__ void k_tab_measure(uint32_t* in, uint32_t* out, size_t W, int n, int q, int tx, int tz, uint64_t seed, uint64_t traj0, uint64_t draw_id,
                              const uint32_t* __restrict__ alive, uint32_t* __restrict__ outcome2, uint32_t* __restrict__ was_random, long long* counts) {
    size_t w = blockIdx.x * (size_t)TTPB + threadIdx.x;
    if (w >= W) return;
    uint32_t rnd;
    uint32_t outcome = tab_measure_word(in, out, W, w, w, n, q, tx, tz, seed, traj0, draw_id, &rnd);
    uint32_t al = alive[w];
    outcome2[w] = outcome & al;
    if (was_random) was_random[w] = rnd & al;
    if (counts) {
        int na = __popc(al), n2 = __popc(outcome & al);
        if (na) atomicAdd((unsigned long long*)&counts[0], (unsigned long long)na);
        if (na - n2) atomicAdd((unsigned long long*)&counts[1], (unsigned long long)(na - n2));
        if (n2) atomicAdd((unsigned long long*)&counts[2], (unsigned long long)n2);
    }
}
__global__ void k_tab_ptrace(uint32_t* in, uint32_t* out, size_t W, int n, int q) {
    size_t w = blockIdx.x * (size_t)TTPB + threadIdx.x;
    if (w >= W) return;
    tab_ptrace_word(in, out, W, w, n, q);
}
__global__ void k_tab_force(uint32_t* p, size_t W, int n, uint32_t px, uint32_t pz, int sign) {
    size_t w = blockIdx.x * (size_t)TTPB + threadIdx.x;
    if (w >= W) return;
    tab_force_word(p, W, w, n, px, pz, sign);
}

static int swap_in(qsb_state* st, void* nd, int n_new) {
    qsb_ctx* ctx = st->ctx;
    QSB_TRY(qsb_dev_free(ctx, st->data));
    st->data = nd; st->n = n_new; st->bytes = qsb_state_bytes(ctx, QSB_TAB, n_new);
    return QSB_OK;
}

int qsbt_measure(qsb_state* st, int idx, int basis, uint64_t draw_id, qsb_mask* outcome2, qsb_mask* was_random, int64_t* counts_host) {
    qsb_ctx* ctx = st->ctx;
    int n = st->n;
    int tx = (basis == QSB_BASIS_X || basis == QSB_BASIS_Y), tz = (basis == QSB_BASIS_Z || basis == QSB_BASIS_Y);
    if (ctx->rec) {
        // the qubit is removed IN PLACE (remove_qubit tolerates aliasing): the state keeps its base, the tail is recycled
        TabOp* o = qsb_rec_push(ctx, TOP_MEASURE); o->n = n; o->q[0] = idx; o->iv[0] = tx; o->iv[1] = tz; o->draw = draw_id;
        o->s[0] = st->vbase; o->s[1] = st->vbase; o->al[1] = st->valive; o->m[0] = outcome2->vplane; o->m[1] = was_random ? was_random->vplane : -1;
        qsb_rec_shrink(ctx, st->vbase, qsb_tab_planes(n), qsb_tab_planes(n - 1));
        st->n = n - 1; st->bytes = qsb_state_bytes(ctx, QSB_TAB, n - 1);
        qsb_account(ctx, "k_tab_measure", (uint64_t)(qsb_tab_planes(n) + qsb_tab_planes(n - 1)) * ctx->ntraj / 8); ctx->ops += ctx->ntraj; return QSB_OK;
    }
    void* nd;
    QSB_TRY(qsb_dev_alloc(ctx, &nd, qsb_state_bytes(ctx, QSB_TAB, n - 1)));
    if (counts_host) QSB_CUDA(cudaMemsetAsync(ctx->counts_tmp, 0, 3 * sizeof(long long), ctx->stream));
    QSB_PRE(ctx, "k_tab_measure");
    k_tab_measure<<<tgrid(ctx), TTPB, 0, ctx->stream>>>((uint32_t*)st->data, (uint32_t*)nd, (size_t)ctx->nwords, n, idx, tx, tz, ctx->seed, (uint64_t)ctx->traj_offset, draw_id,
                                                        st->alive, outcome2->w, was_random ? was_random->w : nullptr, counts_host ? ctx->counts_tmp : nullptr);
    QSB_LAUNCH_CHECK(ctx);
    qsb_account(ctx, "k_tab_measure", (uint64_t)(qsb_tab_planes(n) + qsb_tab_planes(n - 1)) * ctx->ntraj / 8); ctx->ops += ctx->ntraj;
    QSB_TRY(swap_in(st, nd, n - 1));
    if (counts_host) {
        long long h[3];
        QSB_CUDA(cudaMemcpyAsync(h, ctx->counts_tmp, sizeof(h), cudaMemcpyDeviceToHost, ctx->stream));
        QSB_TRY(qsb_ctx_sync(ctx));
        counts_host[0] = h[0]; counts_host[1] = h[1]; counts_host[2] = h[2];
    }
    return QSB_OK;
}
int qsbt_ptrace(qsb_state* st, int idx) {
    qsb_ctx* ctx = st->ctx;
    int n = st->n;
    if (ctx->rec) {
        TabOp* o = qsb_rec_push(ctx, TOP_PTRACE); o->n = n; o->q[0] = idx; o->s[0] = st->vbase; o->s[1] = st->vbase;
        qsb_rec_shrink(ctx, st->vbase, qsb_tab_planes(n), qsb_tab_planes(n - 1));
        st->n = n - 1; st->bytes = qsb_state_bytes(ctx, QSB_TAB, n - 1);
        qsb_account(ctx, "k_tab_ptrace", (uint64_t)(qsb_tab_planes(n) + qsb_tab_planes(n - 1)) * ctx->ntraj / 8); ctx->ops += ctx->ntraj; return QSB_OK;
    }
    void* nd;
    QSB_TRY(qsb_dev_alloc(ctx, &nd, qsb_state_bytes(ctx, QSB_TAB, n - 1)));
    QSB_PRE(ctx, "k_tab_ptrace");
    k_tab_ptrace<<<tgrid(ctx), TTPB, 0, ctx->stream>>>((uint32_t*)st->data, (uint32_t*)nd, (size_t)ctx->nwords, n, idx);
    QSB_LAUNCH_CHECK(ctx);
    qsb_account(ctx, "k_tab_ptrace", (uint64_t)(qsb_tab_planes(n) + qsb_tab_planes(n - 1)) * ctx->ntraj / 8); ctx->ops += ctx->ntraj;
    return swap_in(st, nd, n - 1);
}
int qsbt_from_host(qsb_ctx* ctx, int n, int ngen, const uint8_t* xs, const uint8_t* zs, const uint8_t* signs, qsb_state** out) {
    // generators must commute and be independent (StabilizerState / MixedDestabilizer(S"...") semantics)
    for (int a = 0; a < ngen; ++a) for (int b = 0; b < a; ++b) {
        int c = 0;
        for (int q = 0; q < n; ++q) c ^= (xs[a * n + q] & zs[b * n + q]) ^ (zs[a * n + q] & xs[b * n + q]);
        if (c) QSB_FAIL(QSB_ERR_BAD_ARG, "stabilizer generators %d and %d anticommute", b + 1, a + 1);
    }
    qsb_state* st;
    QSB_TRY(qsbt_new(ctx, n, 0, &st));
    for (int g = 0; g < ngen; ++g) {
        uint32_t px = 0, pz = 0;
        for (int q = 0; q < n; ++q) { if (xs[g * n + q]) px |= 1u << q; if (zs[g * n + q]) pz |= 1u << q; }
        if ((px | pz) == 0) { qsb_state_free(st); QSB_FAIL(QSB_ERR_BAD_ARG, "stabilizer generator %d is the identity", g + 1); }
        if (ctx->rec) { TabOp* o = qsb_rec_push(ctx, TOP_FORCE); o->n = n; o->s[0] = st->vbase; o->u[0] = px; o->u[1] = pz; o->iv[0] = signs ? signs[g] : 0; continue; }
        QSB_PRE(ctx, "k_tab_force");
        k_tab_force<<<tgrid(ctx), TTPB, 0, ctx->stream>>>((uint32_t*)st->data, (size_t)ctx->nwords, n, px, pz, signs ? signs[g] : 0);
        QSB_LAUNCH_CHECK(ctx);
    }
    qsb_account(ctx, "k_tab_force", (uint64_t)qsb_tab_planes(n) * ctx->ntraj / 8); ctx->ops += ctx->ntraj;
    *out = st;
    return QSB_OK;
}

// ---------------------------------------------------------------------------------------------
// observable on a tableau (a12)
// ---------------------------------------------------------------------------------------------
// Sums are dyadic rationals: exact in fp64, so the atomics below are order independent.
__global__ void k_tab_expect(uint32_t* p, size_t W, TabObsArgs a, const uint32_t* __restrict__ alive, int64_t npad,
                             double* __restrict__ per_traj, double* __restrict__ stats) {
    size_t w = blockIdx.x * (size_t)TTPB + threadIdx.x;
    double s_alive = 0, s_v = 0, s_v2 = 0;
    if (w < W) {
        uint32_t cnt[8], cntm[8];
        tab_expect_word(p, W, w, a, cnt, cntm);
        uint32_t al = alive[w];
        double inv = 1.0 / (double)a.nelem;
        for (int l = 0; l < 32; ++l) {
            int plus = 0, minus = 0;
            for (int b = 0; b < 8; ++b) { plus |= ((cnt[b] >> l) & 1) << b; minus |= ((cntm[b] >> l) & 1) << b; }
            bool on = (al >> l) & 1u;
            double val = on ? (double)(plus - minus) * inv : 0.0;
            if (per_traj) { size_t t = w * 32 + l; per_traj[2 * t] = val; per_traj[2 * t + 1] = 0.0; }
            if (on) { s_alive += 1.0; s_v += val; s_v2 += val * val; }
        }
    }
    if (stats) {
        for (int o = 16; o; o >>= 1) {
            s_alive += __shfl_xor_sync(0xffffffffu, s_alive, o);
            s_v += __shfl_xor_sync(0xffffffffu, s_v, o);
            s_v2 += __shfl_xor_sync(0xffffffffu, s_v2, o);
        }
        if ((threadIdx.x & 31) == 0 && s_alive != 0.0) { atomicAdd(&stats[0], s_alive); atomicAdd(&stats[1], s_v); atomicAdd(&stats[2], s_v2); }
    }
}

// host-side product of Hermitian Paulis given as (x,z,sign) with the i-exponent tracked
static void host_pauli_mul(int n, uint32_t& ax, uint32_t& az, int& aph, uint32_t bx, uint32_t bz, int bph) {
    int e = aph + bph;
    for (int q = 0; q < n; ++q) {
        int x1 = (bx >> q) & 1, z1 = (bz >> q) & 1, x2 = (ax >> q) & 1, z2 = (az >> q) & 1;
        if (x1 && z1) e += z2 - x2;
        else if (x1) e += z2 * (2 * x2 - 1);
        else if (z1) e += x2 * (1 - 2 * z2);
    }
    ax ^= bx; az ^= bz; aph = ((e % 4) + 4) % 4;
}

int qsbt_expect(qsb_state* st, const int* idx, int k, int obs_kind, const void* obs, int obs_sign,
                double* per_traj, double* stats_dev, double* stats_host) {
    qsb_ctx* ctx = st->ctx;
    int n = st->n;
    TabObsArgs a; memset(&a, 0, sizeof(a));
    a.n = n; a.k = k;
    if (obs_kind == QSB_OBS_PAULI) {
        const uint8_t* codes = (const uint8_t*)obs;
        for (int m = 0; m < k; ++m) {
            if (codes[m] > 3) QSB_FAIL(QSB_ERR_BAD_ARG, "bad Pauli code %d", codes[m]);
            if (codes[m] == 1 || codes[m] == 2) a.px[0] |= 1u << idx[m];
            if (codes[m] == 3 || codes[m] == 2) a.pz[0] |= 1u << idx[m];
        }
        a.sign[0] = obs_sign ? 1 : 0;
        a.nelem = 1;
    } else if (obs_kind == QSB_OBS_PROJ_STAB) {
        if (k > 6) QSB_FAIL(QSB_ERR_UNSUPPORTED, "stabilizer projectors on more than 6 qubits are not implemented");
        if (obs_sign) QSB_FAIL(QSB_ERR_UNSUPPORTED, "negated projector observable");
        const uint8_t* xs = (const uint8_t*)obs; const uint8_t* zs = xs + k * k; const uint8_t* sg = zs + k * k;
        uint32_t gx[6], gz[6];
        for (int g = 0; g < k; ++g) { gx[g] = gz[g] = 0; for (int m = 0; m < k; ++m) { if (xs[g * k + m]) gx[g] |= 1u << idx[m]; if (zs[g * k + m]) gz[g] |= 1u << idx[m]; } }
        a.nelem = 1 << k;
        for (int sub = 0; sub < a.nelem; ++sub) {
            uint32_t hx = 0, hz = 0; int hph = 0;
            for (int g = 0; g < k; ++g) if ((sub >> g) & 1) host_pauli_mul(n, hx, hz, hph, gx[g], gz[g], sg[g] ? 2 : 0);
            if (hph & 1) QSB_FAIL(QSB_ERR_BAD_ARG, "projector generators do not commute");
            a.px[sub] = hx; a.pz[sub] = hz; a.sign[sub] = hph == 2;
        }
    } else {
        QSB_FAIL(QSB_ERR_UNSUPPORTED, "dense observables on a Clifford state are not implemented (the reference converts a pure tableau to a ket, clifford/express.jl:22-46)");
    }
    if (ctx->rec) {
        TabOp* o = qsb_rec_push(ctx, TOP_EXPECT); o->n = n; o->s[0] = st->vbase; o->al[0] = st->valive;
        o->obs = qsb_rec_push_obs(ctx, &a, sizeof(a)); o->stats = stats_dev; o->per_traj = per_traj;
        qsb_rec_note_stats(ctx, stats_dev);
        qsb_account(ctx, "k_tab_expect", (uint64_t)n * (2 * n + 2) * ctx->ntraj / 8); ctx->ops += ctx->ntraj; return QSB_OK;
    }
    double* sd = stats_dev ? stats_dev : ((stats_host) ? ctx->stats_tmp : nullptr);
    if (sd) QSB_CUDA(cudaMemsetAsync(sd, 0, 4 * sizeof(double), ctx->stream));
    QSB_PRE(ctx, "k_tab_expect");
    k_tab_expect<<<tgrid(ctx), TTPB, 0, ctx->stream>>>((uint32_t*)st->data, (size_t)ctx->nwords, a, st->alive, ctx->npad, per_traj, sd);
    QSB_LAUNCH_CHECK(ctx);
    qsb_account(ctx, "k_tab_expect", (uint64_t)n * (2 * n + 2) * ctx->ntraj / 8); ctx->ops += ctx->ntraj;
    if (stats_host) {
        QSB_CUDA(cudaMemcpyAsync(stats_host, sd, 4 * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
        QSB_TRY(qsb_ctx_sync(ctx));
    }
    return QSB_OK;
}
