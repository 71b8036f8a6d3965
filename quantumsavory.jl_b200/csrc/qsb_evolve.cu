// qsb_evolve.cu -- non-instant evolution under a constant Hamiltonian and Lindblad collapse operators.
//
// Replaces apply_noninstant!(state, state_indices, ConstantHamiltonianEvolution(H, t), backgrounds)
// (src/backends/quantumoptics/noninstant.jl:10-37) and the `krausops === nothing` branch of uptotime!
// (src/backends/quantumoptics/uptotime.jl:24-29,51-54).  The reference integrates
//     d rho / dt = -i [H, rho] + sum_j ( L_j rho L_j' - 1/2 { L_j' L_j, rho } )
// with an adaptive Dormand-Prince solver (timeevolution.master / schroedinger / mcwf) on the full 2^n space.
//
// B200-first restatement: H acts on k <= 2 subsystems and every L_j on one of them, so the generator is a constant
// 4^k x 4^k matrix acting on the k-qubit block of rho and the solution after `duration` is its matrix exponential.  The
// host computes exp(L t) once per call (scaling and squaring, <= 16 x 16) and ONE streaming pass applies it to every
// 4^k-element group of every trajectory: the same access pattern and the same HBM roofline as k_apply_dm<K>
// (2 * 16 * 4^n bytes per trajectory), no time stepping on the device, no solver tolerance.
//   Operator / Ket (promoted with dm(), noninstant.jl:21-23): k_apply_super<K>.
//   MCKet without collapse operators (schroedinger, noninstant.jl:31-32): U = exp(-i H t) through k_apply_ket<K>.
//   MCKet with collapse operators (mcwf, noninstant.jl:33-34): one sampled branch of the canonical Kraus form of exp(L t)
//   (eigenvectors of its Choi matrix).  Like mcwf this is an unravelling of the master equation into normalised pure
//   states whose ensemble mean is the master-equation solution; individual jump records are not comparable with the
//   reference's (its draws come from inside the ODE solver's callbacks).  k = 1: the sampled-Kraus kernels of uptotime!;
//   k = 2: k_rdm2_ket (reduced 4x4 density matrix) -> k_mc2_finalize (probabilities psi' K'K psi, draw) -> k_apply_kraus2_sel.
#include <math.h>

#include <complex>
#include <vector>

#include "qsb_internal.cuh"

typedef std::complex<double> cd;

#define TPB 128

// rho' = S vec(rho) on every 4^K-element block addressed by row bits pos[] and column bits pos[] + n.
// Block element (r, c) has vec index v = r + c 2^K; S is column-major S[v + w 4^K], staged in shared memory.
struct SuperArgs { int pos[2]; int sorted[4]; };
template <int K>
__global__ void __launch_bounds__(TPB) k_apply_super(double2* d, int64_t npad, size_t groups, int n, SuperArgs a, const double2* __restrict__ S) {
    constexpr int M = 1 << K, V = M * M;
    __shared__ double2 sS[V * V];
    for (int i = threadIdx.x; i < V * V; i += TPB) sS[i] = S[i];
    __syncthreads();
    int64_t t = blockIdx.x * (int64_t)TPB + threadIdx.x;
    if (t >= npad) return;
    for (size_t g = blockIdx.y; g < groups; g += gridDim.y) {
        size_t base = g;
#pragma unroll
        for (int i = 0; i < 2 * K; ++i) base = qsb_insert0(base, a.sorted[i]);
        auto elem = [&](int v) {
            const int r = v & (M - 1), c = v >> K;
            size_t e = base;
#pragma unroll
            for (int b = 0; b < K; ++b) e |= ((size_t)((r >> b) & 1) << a.pos[b]) | ((size_t)((c >> b) & 1) << (a.pos[b] + n));
            return e * npad + t;
        };
        double2 x[V];
#pragma unroll
        for (int v = 0; v < V; ++v) x[v] = d[elem(v)];
#pragma unroll 1
        for (int v = 0; v < V; ++v) {   // one output element per trip: the V inputs stay in registers, S comes from broadcast LDS
            double2 acc = make_double2(0, 0);
#pragma unroll
            for (int w = 0; w < V; ++w) cfma(acc, sS[v + w * V], x[w]);
            d[elem(v)] = acc;
        }
    }
}

// ---- two-subsystem sampled Kraus branch on a trajectory ket (the mcwf case of a 2-qubit ConstantHamiltonianEvolution) ----------
// table layout in device memory: E[m] = K_m' K_m (16 complex, column-major 4x4) for m < 16, then K[m] (16 complex each)
// reduced 4x4 density matrix of the two target qubits, split over the other 2^(n-2) indices: partial[(s*32 + 2 v (+1)) * npad + t]
__global__ void __launch_bounds__(TPB) k_rdm2_ket(const double2* __restrict__ d, int64_t npad, int q0, int q1, size_t range, int nsplit, double* __restrict__ partial) {
    int64_t t = blockIdx.x * (int64_t)TPB + threadIdx.x;
    if (t >= npad) return;
    const int s = blockIdx.y;
    const size_t lo = range * s / nsplit, hi = range * (s + 1) / nsplit;
    const int qa = q0 < q1 ? q0 : q1, qb = q0 < q1 ? q1 : q0;
    double2 R[16];   // R[a + 4 b] = sum psi_a conj(psi_b), a = bit(q0) | bit(q1) << 1
#pragma unroll
    for (int v = 0; v < 16; ++v) R[v] = make_double2(0, 0);
    for (size_t g = lo; g < hi; ++g) {
        const size_t base = qsb_insert0(qsb_insert0(g, qa), qb);
        double2 x[4];
#pragma unroll
        for (int a = 0; a < 4; ++a) x[a] = d[(base | ((size_t)(a & 1) << q0) | ((size_t)(a >> 1) << q1)) * npad + t];
#pragma unroll
        for (int b = 0; b < 4; ++b)
#pragma unroll
            for (int a = 0; a < 4; ++a) { const double2 c = cmulc(x[a], x[b]); R[a + 4 * b].x += c.x; R[a + 4 * b].y += c.y; }
    }
    const size_t o = (size_t)s * 32 * npad + t;
#pragma unroll
    for (int v = 0; v < 16; ++v) { partial[o + (size_t)(2 * v) * npad] = R[v].x; partial[o + (size_t)(2 * v + 1) * npad] = R[v].y; }
}
__global__ void k_mc2_finalize(const double* __restrict__ partial, int nsplit, int64_t npad, const double2* __restrict__ tab, int nk,
                               uint64_t seed, uint64_t traj0, uint64_t draw_id, double* __restrict__ scale, int* __restrict__ sel) {
    int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (t >= npad) return;
    double2 R[16];
    for (int v = 0; v < 16; ++v) {
        double re = 0, im = 0;
        for (int s = 0; s < nsplit; ++s) { const size_t o = (size_t)s * 32 * npad + t; re += partial[o + (size_t)(2 * v) * npad]; im += partial[o + (size_t)(2 * v + 1) * npad]; }
        R[v] = make_double2(re, im);
    }
    // P_m = psi' E_m psi = sum_{a,b} E_m[b,a] R[a,b]   (R[a + 4b] = sum psi_a conj(psi_b))
    double P[16], total = 0;
    for (int m = 0; m < nk; ++m) {
        double acc = 0;
        for (int b = 0; b < 4; ++b) for (int a = 0; a < 4; ++a) { const double2 e = tab[m * 16 + b + 4 * a]; acc += e.x * R[a + 4 * b].x - e.y * R[a + 4 * b].y; }
        P[m] = acc; total += acc;
    }
    const double thr = qsb_draw(seed, traj0 + (uint64_t)t, draw_id) * total;
    double cum = 0;
    int choice = nk - 1;
    for (int m = 0; m < nk; ++m) { cum += P[m]; if (cum > thr) { choice = m; break; } }
    sel[t] = choice;
    scale[t] = 1.0 / sqrt(P[choice] > 0 ? P[choice] : 1.0);
}
__global__ void __launch_bounds__(TPB) k_apply_kraus2_sel(double2* d, int64_t npad, size_t groups, int q0, int q1, const double2* __restrict__ tab,
                                                          const double* __restrict__ scale, const int* __restrict__ sel) {
    int64_t t = blockIdx.x * (int64_t)TPB + threadIdx.x;
    if (t >= npad) return;
    const double2* K = tab + 256 + (size_t)sel[t] * 16;   // column-major 4x4: K[r + 4 c]
    const double sc = scale[t];
    double2 k[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) k[i] = cscale(K[i], sc);
    const int qa = q0 < q1 ? q0 : q1, qb = q0 < q1 ? q1 : q0;
    for (size_t g = blockIdx.y; g < groups; g += gridDim.y) {
        const size_t base = qsb_insert0(qsb_insert0(g, qa), qb);
        double2 x[4];
#pragma unroll
        for (int a = 0; a < 4; ++a) x[a] = d[(base | ((size_t)(a & 1) << q0) | ((size_t)(a >> 1) << q1)) * npad + t];
#pragma unroll
        for (int r = 0; r < 4; ++r) {
            double2 acc = make_double2(0, 0);
#pragma unroll
            for (int c = 0; c < 4; ++c) cfma(acc, k[r + 4 * c], x[c]);
            d[(base | ((size_t)(r & 1) << q0) | ((size_t)(r >> 1) << q1)) * npad + t] = acc;
        }
    }
}

// ---- host: small dense complex matrices, column-major ------------------------------------------------
static void mm(int N, const cd* A, const cd* B, cd* C) {
    for (int c = 0; c < N; ++c) for (int r = 0; r < N; ++r) { cd acc = 0; for (int m = 0; m < N; ++m) acc += A[r + m * N] * B[m + c * N]; C[r + c * N] = acc; }
}
// exp(A): scaling and squaring around a Taylor series (|A / 2^s|_1 <= 1/4, 20 terms: truncation < 1e-30)
static void expm(int N, const cd* A, cd* out) {
    double norm = 0;
    for (int c = 0; c < N; ++c) { double s = 0; for (int r = 0; r < N; ++r) s += std::abs(A[r + c * N]); if (s > norm) norm = s; }
    int sq = 0;
    while (norm > 0.25 && sq < 60) { norm *= 0.5; ++sq; }
    std::vector<cd> X(N * N), term(N * N), tmp(N * N), acc(N * N);
    const double scale = ldexp(1.0, -sq);
    for (int i = 0; i < N * N; ++i) X[i] = A[i] * scale;
    for (int i = 0; i < N * N; ++i) { term[i] = 0; acc[i] = 0; }
    for (int i = 0; i < N; ++i) { term[i + i * N] = 1; acc[i + i * N] = 1; }
    for (int j = 1; j <= 20; ++j) {
        mm(N, term.data(), X.data(), tmp.data());
        for (int i = 0; i < N * N; ++i) { term[i] = tmp[i] / (double)j; acc[i] += term[i]; }
    }
    for (int s = 0; s < sq; ++s) { mm(N, acc.data(), acc.data(), tmp.data()); acc.swap(tmp); }
    for (int i = 0; i < N * N; ++i) out[i] = acc[i];
}
// 2x2 operator on factor `pos` of a k-qubit space (factor 0 least significant): embed(basis, [pos], op)
static void embed1(int k, int pos, const cd* op2, cd* out) {
    const int D = 1 << k;
    for (int c = 0; c < D; ++c) for (int r = 0; r < D; ++r) {
        const bool same = ((r ^ c) & ~(1 << pos)) == 0;
        out[r + c * D] = same ? op2[((r >> pos) & 1) + 2 * ((c >> pos) & 1)] : cd(0);
    }
}
// Liouvillian in the column-stacking convention vec(A rho B)[r + c D] = sum A[r,r'] rho[r',c'] B[c',c]
static void liouvillian(int k, const cd* H, int nl, const int* l_pos, const cd* L2, std::vector<cd>& M) {
    const int D = 1 << k, V = D * D;
    M.assign((size_t)V * V, cd(0));
    auto at = [&](int r, int c, int r2, int c2) -> cd& { return M[(r + c * D) + (size_t)(r2 + c2 * D) * V]; };
    const cd I(0, 1);
    if (H) for (int r = 0; r < D; ++r) for (int c = 0; c < D; ++c) for (int m = 0; m < D; ++m) {
        at(r, c, m, c) += -I * H[r + m * D];      // -i H rho
        at(r, c, r, m) += I * H[m + c * D];       // +i rho H
    }
    std::vector<cd> L(D * D), G(D * D);
    for (int j = 0; j < nl; ++j) {
        embed1(k, l_pos[j], L2 + 4 * j, L.data());
        for (int r = 0; r < D; ++r) for (int c = 0; c < D; ++c) { cd acc = 0; for (int m = 0; m < D; ++m) acc += std::conj(L[m + r * D]) * L[m + c * D]; G[r + c * D] = acc; }   // L' L
        for (int r = 0; r < D; ++r) for (int c = 0; c < D; ++c) for (int r2 = 0; r2 < D; ++r2) for (int c2 = 0; c2 < D; ++c2)
            at(r, c, r2, c2) += L[r + r2 * D] * std::conj(L[c + c2 * D]);   // L rho L'
        for (int r = 0; r < D; ++r) for (int c = 0; c < D; ++c) for (int m = 0; m < D; ++m) {
            at(r, c, m, c) += -0.5 * G[r + m * D];
            at(r, c, r, m) += -0.5 * G[m + c * D];
        }
    }
}
// eigen-decomposition of a small Hermitian matrix by cyclic complex Jacobi rotations: A = V diag(w) V'
static void jacobi_herm(int N, std::vector<cd> A, std::vector<double>& w, std::vector<cd>& V) {
    V.assign(N * N, cd(0));
    for (int i = 0; i < N; ++i) V[i + i * N] = 1;
    for (int sweep = 0; sweep < 60; ++sweep) {
        double off = 0;
        for (int p = 0; p < N; ++p) for (int q = p + 1; q < N; ++q) off += std::norm(A[p + q * N]);
        if (off < 1e-32) break;
        for (int p = 0; p < N; ++p) for (int q = p + 1; q < N; ++q) {
            const cd apq = A[p + q * N];
            const double mag = std::abs(apq);
            if (mag < 1e-300) continue;
            const cd ph = apq / mag;                      // A_pq = mag * ph
            const double app = A[p + p * N].real(), aqq = A[q + q * N].real();
            const double theta = 0.5 * atan2(2 * mag, app - aqq);
            const double c = cos(theta), s = sin(theta);
            // rotation J: columns p,q <- (c |p> + s conj(ph) |q>), (-s ph |p> + c |q>)
            for (int i = 0; i < N; ++i) {
                const cd aip = A[i + p * N], aiq = A[i + q * N];
                A[i + p * N] = c * aip + s * std::conj(ph) * aiq;
                A[i + q * N] = -s * ph * aip + c * aiq;
                const cd vip = V[i + p * N], viq = V[i + q * N];
                V[i + p * N] = c * vip + s * std::conj(ph) * viq;
                V[i + q * N] = -s * ph * vip + c * viq;
            }
            for (int i = 0; i < N; ++i) {
                const cd api = A[p + i * N], aqi = A[q + i * N];
                A[p + i * N] = c * api + s * ph * aqi;
                A[q + i * N] = -s * std::conj(ph) * api + c * aqi;
            }
        }
    }
    w.resize(N);
    for (int i = 0; i < N; ++i) w[i] = A[i + i * N].real();
}

int qsbd_evolve(qsb_state* st, const int* idx, int k, const double* H_ri, int nl, const int* l_pos, const double* L_ri, double duration, uint64_t draw_id) {
    qsb_ctx* ctx = st->ctx;
    if (ctx->rec) QSB_FAIL(QSB_ERR_UNSUPPORTED, "qsb_evolve cannot be recorded");
    const int D = 1 << k, V = D * D, n = st->n;
    std::vector<cd> H, L2((size_t)4 * (nl > 0 ? nl : 1));
    if (H_ri) { H.resize(D * D); for (int i = 0; i < D * D; ++i) H[i] = cd(H_ri[2 * i], H_ri[2 * i + 1]); }
    for (int i = 0; i < 4 * nl; ++i) L2[i] = cd(L_ri[2 * i], L_ri[2 * i + 1]);
    if (st->kind == QSB_MCKET && nl == 0) {   // timeevolution.schroedinger: psi <- exp(-i H t) psi
        std::vector<cd> A(D * D, cd(0)), U(D * D);
        if (H_ri) for (int i = 0; i < D * D; ++i) A[i] = cd(0, -duration) * H[i];
        expm(D, A.data(), U.data());
        return qsbd_apply(st, idx, k, reinterpret_cast<const double*>(U.data()), nullptr);
    }
    std::vector<cd> M, S((size_t)V * V);
    liouvillian(k, H_ri ? H.data() : nullptr, nl, l_pos, L2.data(), M);
    for (auto& v : M) v *= duration;
    expm(V, M.data(), S.data());
    if (st->kind == QSB_MCKET) {
        if (k == 2) {
            // Choi matrix of the 2-qubit channel: C[(r,r'),(c,c')] = S[(r,c),(r',c')], indices r + 4 r'; K_m = sqrt(w_m) unvec(v_m)
            std::vector<cd> C(256), Vv; std::vector<double> w;
            for (int r = 0; r < 4; ++r) for (int r2 = 0; r2 < 4; ++r2) for (int c = 0; c < 4; ++c) for (int c2 = 0; c2 < 4; ++c2)
                C[(r + 4 * r2) + 16 * (c + 4 * c2)] = S[(r + 4 * c) + 16 * (r2 + 4 * c2)];
            for (int i = 0; i < 16; ++i) for (int j = i + 1; j < 16; ++j) { cd h = 0.5 * (C[i + 16 * j] + std::conj(C[j + 16 * i])); C[i + 16 * j] = h; C[j + 16 * i] = std::conj(h); }
            jacobi_herm(16, C, w, Vv);
            std::vector<int> order(16);
            for (int i = 0; i < 16; ++i) order[i] = i;
            for (int i = 1; i < 16; ++i) for (int j = i; j > 0 && w[order[j]] > w[order[j - 1]]; --j) { int t2 = order[j]; order[j] = order[j - 1]; order[j - 1] = t2; }
            std::vector<cd> tab(512, cd(0));   // [0,256): E_m = K_m' K_m ; [256,512): K_m
            int nk = 0;
            for (int m = 0; m < 16; ++m) {
                const int e = order[m];
                if (w[e] < 1e-15) continue;
                const double sq = sqrt(w[e]);
                cd* K = tab.data() + 256 + nk * 16; cd* E = tab.data() + nk * 16;
                for (int i = 0; i < 16; ++i) K[i] = sq * Vv[i + 16 * e];                       // i = r + 4 r'
                for (int a = 0; a < 4; ++a) for (int b = 0; b < 4; ++b) { cd acc = 0; for (int r = 0; r < 4; ++r) acc += std::conj(K[r + 4 * a]) * K[r + 4 * b]; E[a + 4 * b] = acc; }
                ++nk;
            }
            void* dtab;
            QSB_TRY(qsb_upload(ctx, tab.data(), tab.size() * sizeof(cd), &dtab));
            const size_t range = (size_t)1 << (n - 2);
            size_t want = ((size_t)ctx->sm_count * 2048 + ctx->npad - 1) / ctx->npad;
            int nsplit = (int)(want < 1 ? 1 : want > range ? range : want > 256 ? 256 : want);
            QSB_TRY(qsb_ensure_scratch(ctx, ((size_t)nsplit * 32 + 2) * ctx->npad * sizeof(double)));
            dim3 grid((unsigned)((ctx->npad + TPB - 1) / TPB), nsplit, 1);
            QSB_PRE(ctx, "k_rdm2_ket");
            k_rdm2_ket<<<grid, TPB, 0, ctx->stream>>>((const double2*)st->data, ctx->npad, idx[0], idx[1], range, nsplit, ctx->scratch);
            QSB_LAUNCH_CHECK(ctx);
            double* scale = ctx->scratch + (size_t)nsplit * 32 * ctx->npad;
            int* sel = (int*)(scale + ctx->npad);
            QSB_PRE(ctx, "k_mc2_finalize");
            k_mc2_finalize<<<(unsigned)((ctx->npad + 127) / 128), 128, 0, ctx->stream>>>(ctx->scratch, nsplit, ctx->npad, (const double2*)dtab, nk, ctx->seed, (uint64_t)ctx->traj_offset, draw_id, scale, sel);
            QSB_LAUNCH_CHECK(ctx);
            unsigned gx2 = (unsigned)((ctx->npad + TPB - 1) / TPB);
            size_t gy2 = ((size_t)ctx->sm_count * 32 + gx2 - 1) / gx2;
            if (gy2 > range) gy2 = range;
            if (gy2 < 1) gy2 = 1;
            if (gy2 > 65535) gy2 = 65535;
            QSB_PRE(ctx, "k_apply_kraus2_sel");
            k_apply_kraus2_sel<<<dim3(gx2, (unsigned)gy2), TPB, 0, ctx->stream>>>((double2*)st->data, ctx->npad, range, idx[0], idx[1], (const double2*)dtab, scale, sel);
            QSB_LAUNCH_CHECK(ctx);
            qsb_account(ctx, "k_apply_kraus2_sel", (uint64_t)32 * ((uint64_t)1 << n) * ctx->ntraj); ctx->ops += ctx->ntraj; ctx->draws++;
            return QSB_OK;
        }
        // Choi matrix C[(r,r'),(c,c')] = S[(r,c),(r',c')] = sum_m vec(K_m) vec(K_m)': K_m = sqrt(w_m) unvec(v_m)
        std::vector<cd> C(16), Vv; std::vector<double> w;
        for (int r = 0; r < 2; ++r) for (int r2 = 0; r2 < 2; ++r2) for (int c = 0; c < 2; ++c) for (int c2 = 0; c2 < 2; ++c2)
            C[(r + 2 * r2) + 4 * (c + 2 * c2)] = S[(r + 2 * c) + 4 * (r2 + 2 * c2)];
        for (int i = 0; i < 4; ++i) for (int j = i + 1; j < 4; ++j) { cd h = 0.5 * (C[i + 4 * j] + std::conj(C[j + 4 * i])); C[i + 4 * j] = h; C[j + 4 * i] = std::conj(h); }
        jacobi_herm(4, C, w, Vv);
        int order[4] = {0, 1, 2, 3};
        for (int i = 1; i < 4; ++i) for (int j = i; j > 0 && w[order[j]] > w[order[j - 1]]; --j) { int s = order[j]; order[j] = order[j - 1]; order[j - 1] = s; }
        KrausArgs ka; memset(&ka, 0, sizeof(ka));
        for (int m = 0; m < 4; ++m) {
            const int e = order[m];
            if (w[e] < 1e-15) continue;
            const double s = sqrt(w[e]);
            for (int i = 0; i < 4; ++i) ka.k[ka.nk][i] = make_double2(s * Vv[i + 4 * e].real(), s * Vv[i + 4 * e].imag());   // i = r + 2 r'
            ka.nk++;
        }
        return qsbd_kraus_sampled(st, idx[0], ka, draw_id);
    }
    if (st->kind == QSB_KET) QSB_TRY(qsbd_promote_dm(st));   // apply_noninstant!(dm(state), ...), noninstant.jl:21-23
    void* dS;
    QSB_TRY(qsb_upload(ctx, S.data(), S.size() * sizeof(cd), &dS));
    SuperArgs a; memset(&a, 0, sizeof(a));
    for (int i = 0; i < k; ++i) { a.pos[i] = idx[i]; a.sorted[i] = idx[i]; a.sorted[k + i] = idx[i] + n; }
    for (int i = 1; i < 2 * k; ++i) for (int j = i; j > 0 && a.sorted[j] < a.sorted[j - 1]; --j) { int s = a.sorted[j]; a.sorted[j] = a.sorted[j - 1]; a.sorted[j - 1] = s; }
    const size_t groups = (size_t)1 << (2 * n - 2 * k);
    unsigned gx = (unsigned)((ctx->npad + TPB - 1) / TPB);
    size_t gy = ((size_t)ctx->sm_count * 32 + gx - 1) / gx;
    if (gy < 1) gy = 1;
    if (gy > groups) gy = groups;
    if (gy > 65535) gy = 65535;
    QSB_PRE(ctx, "k_apply_super");
    if (k == 1) k_apply_super<1><<<dim3(gx, (unsigned)gy), TPB, 0, ctx->stream>>>((double2*)st->data, ctx->npad, groups, n, a, (const double2*)dS);
    else k_apply_super<2><<<dim3(gx, (unsigned)gy), TPB, 0, ctx->stream>>>((double2*)st->data, ctx->npad, groups, n, a, (const double2*)dS);
    QSB_LAUNCH_CHECK(ctx);
    qsb_account(ctx, "k_apply_super", (uint64_t)32 * ((uint64_t)1 << (2 * n)) * ctx->ntraj);
    ctx->ops += ctx->ntraj;
    return QSB_OK;
}
