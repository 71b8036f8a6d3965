// qsb_dexec.cu -- the fused executor of recorded dense programs (see qsb_dexec.cuh) and its host side.
#include <math.h>
#include <algorithm>
#include <map>
#include <vector>

#include "qsb_dexec.cuh"

// ---------------------------------------------------------------------------------------------
// device side
// ---------------------------------------------------------------------------------------------
struct DTile {
    double2* arena;      // [elems][T]
    uint32_t* masks;     // [nmasks]
    double* scr;         // [2][T] measurement scratch: scale, selected branch
    int T, logT, G;      // lanes, log2(lanes), groups = 256 / T
    int lane, grp;
    uint32_t gmask;      // the lanes of this thread's group inside its warp (all 32 when T == 32)
    int gshift;          // position of the group's lane 0 inside the warp
    uint32_t validw;     // lanes of the tile that are trajectories of the job
    __device__ __forceinline__ double2& at(int base, int e) const { return arena[(size_t)(base + e) * T + lane]; }
};

__device__ __forceinline__ int ins0(int g, int pos) { return ((g >> pos) << (pos + 1)) | (g & ((1 << pos) - 1)); }

// rho <- U rho U^dagger on the 2^K x 2^K blocks addressed by row bits pos[] and column bits pos[] + n.
// op.u[0] has bit (r + m*M) set where U[r][m] != 0: the repeater gates (CNOT, CPHASE, Paulis) are permutation-like, so
// most of the M^3 complex multiply-adds of each stage are skipped (uniform branches: U is shared by the ensemble).
// U is read on demand (broadcast loads) instead of living in registers: x alone is 4*M*M registers.
template <int K>
__device__ __forceinline__ void d_apply(const DTile& t, const DExecOp& op, const double* __restrict__ aux) {
    constexpr int M = 1 << K;
    if (op.m[0] >= 0 && !((t.masks[op.m[0]] >> t.lane) & 1u)) return;
    const int n = op.n, base0 = op.s[0];
    const uint32_t nz = op.u[0];
    int pos[K], sorted[2 * K];
#pragma unroll
    for (int i = 0; i < K; ++i) pos[i] = op.q[i];
    if constexpr (K == 1) { sorted[0] = pos[0]; sorted[1] = pos[0] + n; }
    else {
        const int lo = min(pos[0], pos[1]), hi = max(pos[0], pos[1]);
        sorted[0] = lo; sorted[1] = hi; sorted[2] = lo + n; sorted[3] = hi + n;
    }
    const double2* U = reinterpret_cast<const double2*>(aux + op.aux);
    const int groups = 1 << (2 * n - 2 * K);
    for (int g = t.grp; g < groups; g += t.G) {
        int b = g;
#pragma unroll
        for (int i = 0; i < 2 * K; ++i) b = ins0(b, sorted[i]);
        double2 x[M][M];
#pragma unroll
        for (int r = 0; r < M; ++r)
#pragma unroll
            for (int c = 0; c < M; ++c) {
                int e = b;
#pragma unroll
                for (int j = 0; j < K; ++j) e |= (((r >> j) & 1) << pos[j]) | (((c >> j) & 1) << (pos[j] + n));
                x[r][c] = t.at(base0, e);
            }
        // row by row: y[r][:] = (U x)[r][:], then out[r][c] = sum_m y[r][m] conj(U[c][m]); x stays in registers, so the
        // stores may overwrite its elements
#pragma unroll
        for (int r = 0; r < M; ++r) {
            double2 y[M];
#pragma unroll
            for (int c = 0; c < M; ++c) y[c] = make_double2(0, 0);
#pragma unroll
            for (int m = 0; m < M; ++m)
                if ((nz >> (r + m * M)) & 1u) {
                    const double2 w = U[r + m * M];
#pragma unroll
                    for (int c = 0; c < M; ++c) cfma(y[c], w, x[m][c]);
                }
#pragma unroll
            for (int c = 0; c < M; ++c) {
                double2 acc = make_double2(0, 0);
#pragma unroll
                for (int m = 0; m < M; ++m)
                    if ((nz >> (c + m * M)) & 1u) { const double2 w = U[c + m * M]; acc.x += y[m].x * w.x + y[m].y * w.y; acc.y += y[m].y * w.x - y[m].x * w.y; }
                int e = b;
#pragma unroll
                for (int j = 0; j < K; ++j) e |= (((r >> j) & 1) << pos[j]) | (((c >> j) & 1) << (pos[j] + n));
                t.at(base0, e) = acc;
            }
        }
    }
}

// Generalised-permutation gates (one non-zero per row: Paulis, CNOT, CPHASE, ZCY, ...): out[r][c] = w_r conj(w_c) x[p(r)][p(c)].
// op.u[0] packs p (K bits per row), op.u[1] marks the (r,c) whose coefficient differs from 1, aux holds the coefficients.
// No M^3 products and no bit tests: 2 M^2 shared-memory accesses per block and at most M^2 complex multiplies.
template <int K>
__device__ __forceinline__ void d_apply_perm(const DTile& t, const DExecOp& op, const double* __restrict__ aux) {
    constexpr int M = 1 << K;
    if (op.m[0] >= 0 && !((t.masks[op.m[0]] >> t.lane) & 1u)) return;
    const int n = op.n, base0 = op.s[0];
    int pos[K], sorted[2 * K];
#pragma unroll
    for (int i = 0; i < K; ++i) pos[i] = op.q[i];
    if constexpr (K == 1) { sorted[0] = pos[0]; sorted[1] = pos[0] + n; }
    else {
        const int lo = min(pos[0], pos[1]), hi = max(pos[0], pos[1]);
        sorted[0] = lo; sorted[1] = hi; sorted[2] = lo + n; sorted[3] = hi + n;
    }
    int dst[M], src[M];   // row-bit patterns of the block's rows, and of the rows they are fetched from
#pragma unroll
    for (int r = 0; r < M; ++r) {
        int d = 0, sidx = (int)(op.u[0] >> (K * r)) & (M - 1), sbits = 0;
#pragma unroll
        for (int j = 0; j < K; ++j) { d |= ((r >> j) & 1) << pos[j]; sbits |= ((sidx >> j) & 1) << pos[j]; }
        dst[r] = d; src[r] = sbits;
    }
    const uint32_t cm = op.u[1];
    const double2* C = reinterpret_cast<const double2*>(aux + op.aux);
    const int groups = 1 << (2 * n - 2 * K);
    for (int g = t.grp; g < groups; g += t.G) {
        int b = g;
#pragma unroll
        for (int i = 0; i < 2 * K; ++i) b = ins0(b, sorted[i]);
        double2 v[M][M];
#pragma unroll
        for (int r = 0; r < M; ++r)
#pragma unroll
            for (int c = 0; c < M; ++c) v[r][c] = t.at(base0, b | src[r] | (src[c] << n));
#pragma unroll
        for (int r = 0; r < M; ++r)
#pragma unroll
            for (int c = 0; c < M; ++c) {
                double2 o = v[r][c];
                if ((cm >> (r + c * M)) & 1u) o = cmul(C[r + c * M], o);
                t.at(base0, b | dst[r] | (dst[c] << n)) = o;
            }
    }
}

// closed-form single-qubit channel on the 2x2 blocks of qubit q (coefficients a, b, c, dd, f; see qsb_dense.cu)
__device__ __forceinline__ void d_channel(const DTile& t, const DExecOp& op, const double* __restrict__ aux) {
    const double* p = aux + op.aux;
    const double a = p[0], b = p[1], c = p[2], dd = p[3], f = p[4];
    const bool offdiag = p[5] != 0.0;
    const int n = op.n, q = op.q[0], base0 = op.s[0];
    const int rbit = 1 << q, cbit = 1 << (q + n), groups = 1 << (2 * n - 2);
    for (int g = t.grp; g < groups; g += t.G) {
        const int e = ins0(ins0(g, q), q + n);
        double2& p01 = t.at(base0, e | cbit);
        double2& p10 = t.at(base0, e | rbit);
        if (!offdiag) {
            double2& p00 = t.at(base0, e);
            double2& p11 = t.at(base0, e | rbit | cbit);
            const double2 e00 = p00, e11 = p11;
            p00 = make_double2(a * e00.x + b * e11.x, a * e00.y + b * e11.y);
            p11 = make_double2(c * e11.x + dd * e00.x, c * e11.y + dd * e00.y);
        }
        p01 = cscale(p01, f); p10 = cscale(p10, f);
    }
}

// out = a (x) b, a least significant
__device__ __forceinline__ void d_tensor(const DTile& t, const DExecOp& op) {
    const int na = op.n, nb = op.n2, n = na + nb;
    const int Da = 1 << na, D = 1 << n, L = 1 << (2 * n);
    for (int eo = t.grp; eo < L; eo += t.G) {
        const int r = eo & (D - 1), c = eo >> n;
        const int ea = (r & (Da - 1)) + ((c & (Da - 1)) << na), eb = (r >> na) + ((c >> na) << nb);
        t.at(op.s[0], eo) = cmul(t.at(op.s[1], ea), t.at(op.s[2], eb));
    }
    if (threadIdx.x == 0) t.masks[op.al[0]] = t.masks[op.al[1]] & t.masks[op.al[2]];
}

#define DHOLD 8   // output elements a thread keeps in registers across the barrier of an in-place shrink

// project_traceout! phase A (one group): reduced 2x2 matrix of qubit q, branch probabilities, the draw, the branch
__device__ __forceinline__ void d_measure_decide(const DTile& t, const DExecOp& op, const DExecArgs& a, uint64_t traj) {
    const int n = op.n, q = op.q[0], bit = 1 << q, range = 1 << (n - 1), s0 = op.s[0];
    double m00 = 0, m11 = 0, mre = 0, mim = 0;
    for (int i = 0; i < range; ++i) {
        const int b0 = ins0(i, q), b1 = b0 | bit;
        const double2 e00 = t.at(s0, b0 + (b0 << n)), e11 = t.at(s0, b1 + (b1 << n)), e01 = t.at(s0, b0 + (b1 << n));
        m00 += e00.x; m11 += e11.x; mre += e01.x; mim += e01.y;
    }
    const double* kp = a.aux + op.aux;
    double P[2];
#pragma unroll
    for (int j = 0; j < 2; ++j) {
        const double2 a0 = make_double2(kp[4 * j], kp[4 * j + 1]), a1 = make_double2(kp[4 * j + 2], kp[4 * j + 3]);
        const double2 x = cconjmul(a0, a1);
        P[j] = (a0.x * a0.x + a0.y * a0.y) * m00 + (a1.x * a1.x + a1.y * a1.y) * m11 + 2.0 * (x.x * mre - x.y * mim);
    }
    const bool al = (t.masks[op.al[0]] >> t.lane) & 1u;
    const double u = qsb_draw(a.seed, traj, (uint64_t)op.draw + a.draw_base);
    const int j = (P[0] >= u) ? 0 : 1;   // findfirst(>=(rand()), cumsum(P)), quantumoptics.jl:54,61
    t.scr[t.lane] = 1.0 / P[j];
    t.scr[t.T + t.lane] = (double)j;
    if (al && fabs(P[0] + P[1] - 1.0) > 1.5e-8) atomicOr(a.flags, 1);
    const uint32_t w2 = (__ballot_sync(t.gmask, al && j == 1) >> t.gshift);
    if (t.lane == 0) t.masks[op.m[0]] = w2 & (t.T == 32 ? 0xffffffffu : ((1u << t.T) - 1u));
}

// phase B (all groups): rho' = <psi_j| rho |psi_j> / P_j written over the head of the same block
__device__ __forceinline__ void d_measure_project(const DTile& t, const DExecOp& op, const DExecArgs& a) {
    const int n = op.n, q = op.q[0], bit = 1 << q, s0 = op.s[0];
    const int items = 1 << (2 * n - 2), Dn = 1 << (n - 1);
    const double sc = t.scr[t.lane];
    const int j = (int)t.scr[t.T + t.lane];
    const double* kp = a.aux + op.aux + 4 * j;
    const double2 p0 = make_double2(kp[0], kp[1]), p1 = make_double2(kp[2], kp[3]);
    const double2 c0 = make_double2(p0.x, -p0.y), c1 = make_double2(p1.x, -p1.y);
    for (int base = 0; base < items; base += DHOLD * t.G) {
        double2 v[DHOLD];
#pragma unroll
        for (int h = 0; h < DHOLD; ++h) {
            const int g = base + h * t.G + t.grp;
            if (g < items) {
                const int ra = ins0(g & (Dn - 1), q), cb = ins0(g >> (n - 1), q);
                const double2 e00 = t.at(s0, ra + (cb << n)), e10 = t.at(s0, (ra | bit) + (cb << n));
                const double2 e01 = t.at(s0, ra + ((cb | bit) << n)), e11 = t.at(s0, (ra | bit) + ((cb | bit) << n));
                const double2 col0 = cadd(cmul(c0, e00), cmul(c1, e10)), col1 = cadd(cmul(c0, e01), cmul(c1, e11));
                v[h] = cscale(cadd(cmul(col0, p0), cmul(col1, p1)), sc);
            }
        }
        __syncthreads();   // every read of this chunk happens before any of its writes; later chunks read untouched elements
#pragma unroll
        for (int h = 0; h < DHOLD; ++h) {
            const int g = base + h * t.G + t.grp;
            if (g < items) t.at(s0, g) = v[h];
        }
    }
}

// traceout!: rho'[a,b] = sum_s rho[(a,s),(b,s)], in place
__device__ __forceinline__ void d_ptrace(const DTile& t, const DExecOp& op) {
    const int n = op.n, q = op.q[0], bit = 1 << q, s0 = op.s[0];
    const int items = 1 << (2 * n - 2), Dn = 1 << (n - 1);
    for (int base = 0; base < items; base += DHOLD * t.G) {
        double2 v[DHOLD];
#pragma unroll
        for (int h = 0; h < DHOLD; ++h) {
            const int g = base + h * t.G + t.grp;
            if (g < items) {
                const int ra = ins0(g & (Dn - 1), q), cb = ins0(g >> (n - 1), q);
                v[h] = cadd(t.at(s0, ra + (cb << n)), t.at(s0, (ra | bit) + ((cb | bit) << n)));
            }
        }
        __syncthreads();
#pragma unroll
        for (int h = 0; h < DHOLD; ++h) {
            const int g = base + h * t.G + t.grp;
            if (g < items) t.at(s0, g) = v[h];
        }
    }
}

// observable (one group): per-trajectory expectation value, ensemble statistics by warp shuffles + one atomic row update
__device__ __forceinline__ void d_expect(const DTile& t, const DExecOp& op, const DExecArgs& a, int64_t t_loc) {
    const int n = op.n, s0 = op.s[0];
    double re = 0, im = 0;
    if (op.mode == 0) {
        const int D = 1 << n, xm = (int)op.u[0], zm = (int)op.u[1];
        for (int c = 0; c < D; ++c) {
            const double2 v = t.at(s0, c + ((c ^ xm) << n));
            const double sgn = (__popc(c & zm) & 1) ? -1.0 : 1.0;
            re += sgn * v.x; im += sgn * v.y;
        }
    } else if (op.mode == 1) {
        const double* sp = a.aux + op.aux;
        const int nnz = op.nnz;
        for (int i = 0; i < nnz; ++i) {
            const int ri = (int)sp[3 * i];
            const double2 ai = make_double2(sp[3 * i + 1], sp[3 * i + 2]);
            for (int j = 0; j < nnz; ++j) {
                const int rj = (int)sp[3 * j];
                const double2 aj = make_double2(sp[3 * j + 1], sp[3 * j + 2]);
                const double2 v = cmul(cconjmul(ai, t.at(s0, ri + (rj << n))), aj);
                re += v.x; im += v.y;
            }
        }
    } else {
        const int k = op.k, K = 1 << k;
        const double* pp = a.aux + op.aux;
        const double2* O = reinterpret_cast<const double2*>(pp + (k + 1) / 2 * 2);
        int pos[QSB_DEXEC_MAXN], sorted[QSB_DEXEC_MAXN];
        for (int i = 0; i < k; ++i) { pos[i] = (int)pp[i]; sorted[i] = pos[i]; }
        for (int i = 1; i < k; ++i) for (int j = i; j > 0 && sorted[j] < sorted[j - 1]; --j) { const int x = sorted[j]; sorted[j] = sorted[j - 1]; sorted[j - 1] = x; }
        for (int g = 0; g < (1 << (n - k)); ++g) {
            int b = g;
            for (int i = 0; i < k; ++i) b = ins0(b, sorted[i]);
            for (int i = 0; i < K; ++i) {
                int ei = b;
                for (int c = 0; c < k; ++c) ei |= ((i >> c) & 1) << pos[c];
                for (int j = 0; j < K; ++j) {
                    const double2 o = O[i + j * K];
                    if (o.x == 0.0 && o.y == 0.0) continue;
                    int ej = b;
                    for (int c = 0; c < k; ++c) ej |= ((j >> c) & 1) << pos[c];
                    const double2 v = cmul(o, t.at(s0, ej + (ei << n)));   // O_ij rho[j,i]
                    re += v.x; im += v.y;
                }
            }
        }
    }
    switch (op.fpow & 3u) {   // times i^fpow
        case 1: { const double x = re; re = -im; im = x; break; }
        case 2: re = -re; im = -im; break;
        case 3: { const double x = re; re = im; im = -x; break; }
        default: break;
    }
    const bool al = ((t.masks[op.al[0]] & (a.live_word >= 0 ? t.masks[a.live_word] : 0xffffffffu)) >> t.lane) & 1u;   // retired lanes are tallied where they continue
    if (!al) { re = 0; im = 0; }
    if (op.per_traj && ((t.validw >> t.lane) & 1u) && (a.live_word < 0 || ((t.masks[a.live_word] >> t.lane) & 1u))) { op.per_traj[2 * t_loc] = re; op.per_traj[2 * t_loc + 1] = im; }
    if (op.stats) {
        double v[4] = {al ? 1.0 : 0.0, re, re * re, im};
#pragma unroll
        for (int i = 0; i < 4; ++i)
            for (int o = t.T >> 1; o; o >>= 1) v[i] += __shfl_xor_sync(t.gmask, v[i], o);
        if (t.lane == 0 && v[0] != 0.0) { atomicAdd(&op.stats[0], v[0]); atomicAdd(&op.stats[1], v[1]); atomicAdd(&op.stats[2], v[2]); if (v[3] != 0.0) atomicAdd(&op.stats[3], v[3]); }
    }
}

// THREADS x MINB: block size and the blocks per SM the register allocation must allow -- 256x1 (up to 255 registers) for
// programs whose arena fills an SM, 256x2 and 128x4 (128 registers) for smaller arenas where several tiles share an SM, and
// 32xN: ONE WARP per tile of 8 or 4 trajectories (4 or 8 groups inside the warp) -- no idle warps decoding ops they have no
// items for, block barriers degenerate to warp barriers, and 5-10 independent tiles per SM hide each other's latency.
template <int THREADS, int MINB>
__global__ void __launch_bounds__(THREADS, MINB) k_dense_exec(DExecArgs a0) {
    extern __shared__ double2 dsm[];
    constexpr int CHUNK = QSB_DEXEC_CHUNK_OF(THREADS);
    __shared__ DExecOp sops[CHUNK];
    DExecArgs a = a0;
    DTile t;
    t.logT = a.logT; t.T = 1 << a.logT; t.G = THREADS >> a.logT;
    t.lane = threadIdx.x & (t.T - 1); t.grp = threadIdx.x >> a.logT;
    const int pgrp = t.grp;
    t.gshift = (int)(threadIdx.x & 31u) & ~(t.T - 1);
    t.gmask = t.T == 32 ? 0xffffffffu : (((1u << t.T) - 1u) << t.gshift);
    t.arena = dsm;
    t.scr = reinterpret_cast<double*>(dsm + (size_t)a.elems * t.T);
    t.masks = reinterpret_cast<uint32_t*>(t.scr + 2 * t.T);
    if (a.aux_smem) {   // the constants of every op (gate matrices, channel coefficients, kets) next to the arena
        double* saux = reinterpret_cast<double*>(dsm) + ((size_t)a.elems * t.T * 2 + 2 * t.T + (a.nmasks + 3) / 4 * 2);   // 16-byte aligned
        for (int k = threadIdx.x; k < a.aux_smem; k += THREADS) saux[k] = __ldg(a0.aux + k);
        a.aux = saux;   // published by the barrier of the first op-chunk refill
    }
    const uint32_t tmask = t.T == 32 ? 0xffffffffu : ((1u << t.T) - 1u);
    for (int64_t tile = blockIdx.x; tile < a.ntiles; tile += gridDim.x) {
        const int64_t t0 = tile << a.logT, pos = t0 + t.lane;                 // position inside the job
        // lane subset (cohort re-runs): position p is the trajectory whose slot in the context is lanes[p]; draws, per-trajectory rows
        // and retire codes are indexed by that slot, so every program of a tree writes into the same buffers
        const int64_t t_loc = a.lanes ? (pos < a.count ? (int64_t)__ldg(a.lanes + pos) : 0) : pos;
        const uint64_t traj = a.traj0 + (uint64_t)t_loc;
        uint32_t validw;
        if (a.lanes) { const int64_t left = a.count - t0; validw = left >= t.T ? tmask : left > 0 ? ((1u << left) - 1u) : 0u; }
        else validw = (a.valid[t0 >> 5] >> (t0 & 31)) & tmask;
        t.validw = validw;
        if (a.live_word >= 0 && threadIdx.x == 0) t.masks[a.live_word] = validw;   // published by the barrier of the first op-chunk refill
        for (int i = 0; i < a.nops; ++i) {
            if ((i % CHUNK) == 0) {   // refill (also the barrier between tiles)
                __syncthreads();
                const int nchunk = min(CHUNK, a.nops - i);
                const uint4* src = reinterpret_cast<const uint4*>(a.ops + i);
                uint4* dst = reinterpret_cast<uint4*>(sops);
                for (int k = threadIdx.x; k < nchunk * (int)(sizeof(DExecOp) / 16); k += THREADS) dst[k] = __ldg(src + k);
                __syncthreads();
            }
            const DExecOp& op = sops[i % CHUNK];
            if (op.flags & 1) __syncthreads();
            t.grp = (pgrp - (int)op.g0) & (t.G - 1);   // the op deals its items starting at group g0 (level scheduling)
            switch (op.type) {
            case DOP_CONST: {
                const int L = 1 << (2 * op.n);
                const double2* src = reinterpret_cast<const double2*>(a.aux + op.aux);
                for (int e = t.grp; e < L; e += t.G) t.at(op.s[0], e) = src[e];
                if (threadIdx.x == 0) t.masks[op.al[0]] = validw;
                break;
            }
            case DOP_COPY: {
                const int L = 1 << (2 * op.n);
                for (int e = t.grp; e < L; e += t.G) t.at(op.s[0], e) = t.at(op.s[1], e);
                if (threadIdx.x == 0) t.masks[op.al[0]] = t.masks[op.al[1]];
                break;
            }
            case DOP_TENSOR: d_tensor(t, op); break;
            case DOP_APPLY:
                if (op.mode == 1) { if (op.k == 1) d_apply_perm<1>(t, op, a.aux); else d_apply_perm<2>(t, op, a.aux); }
                else if (op.k == 1) d_apply<1>(t, op, a.aux); else d_apply<2>(t, op, a.aux);
                break;
            case DOP_CHANNEL: d_channel(t, op, a.aux); break;
            case DOP_MEASURE:
                if (t.grp == op.grp) d_measure_decide(t, op, a, traj);
                __syncthreads();
                if (op.n > 1) d_measure_project(t, op, a);
                break;
            case DOP_PTRACE: d_ptrace(t, op); break;
            case DOP_EXPECT: if (t.grp == op.grp) d_expect(t, op, a, t_loc); break;
            case TOP_MASK_FILL: if (threadIdx.x == 0) t.masks[op.m[0]] = op.mode ? validw : 0u; break;
            case TOP_MASK_LOGIC:
                if (threadIdx.x == 0) {
                    const uint32_t x = t.masks[op.m[1]], y = op.m[2] >= 0 ? t.masks[op.m[2]] : 0u;
                    uint32_t r;
                    switch (op.mode) {
                        case QSB_MASK_AND: r = x & y; break;
                        case QSB_MASK_OR: r = x | y; break;
                        case QSB_MASK_XOR: r = x ^ y; break;
                        case QSB_MASK_ANDNOT: r = x & ~y; break;
                        case QSB_MASK_NOT: r = ~x; break;
                        default: r = ~(x ^ y); break;
                    }
                    t.masks[op.m[0]] = r & validw;
                }
                break;
            case TOP_MASK_RANDOM:
                if (t.grp == op.grp) {
                    const double p = a.aux[op.aux];
                    const bool b = qsb_draw(a.seed, traj, (uint64_t)op.draw + a.draw_base) < p;
                    const uint32_t w = (__ballot_sync(t.gmask, b) >> t.gshift) & tmask;
                    if (t.lane == 0) t.masks[op.m[0]] = w & validw;
                }
                break;
            case TOP_KILL: if (threadIdx.x == 0) t.masks[op.al[0]] &= ~t.masks[op.m[0]]; break;
            case TOP_RETIRE:   // cohort forking: these lanes leave the program (no later statistic counts them) and are tagged for the tree driver
                if (t.grp == op.grp) {
                    const uint32_t r = t.masks[op.m[0]] & t.masks[a.live_word];
                    int32_t* const codes = reinterpret_cast<int32_t*>(op.per_traj);
                    if (codes && ((r >> t.lane) & 1u)) codes[t_loc] = (int32_t)op.u[0];
                    __syncwarp(t.gmask);
                    if (t.lane == 0) t.masks[a.live_word] &= ~r;
                }
                break;
            case TOP_ALIVE_COPY: if (threadIdx.x == 0) t.masks[op.m[0]] = t.masks[op.al[0]]; break;
            default: break;
            }
        }
    }
}

// ---------------------------------------------------------------------------------------------
// host side: recorded ops -> DExecOp list, barrier placement, launch shape
// ---------------------------------------------------------------------------------------------
struct qsb_dprogram {
    qsb_ctx* ctx;
    DExecOp* d_ops; DExecOp* h_ops; int nops;
    double* d_aux; int naux;
    int elems, nmasks, logT, blocks_per_sm, smem_bytes, barriers, threads, variant, aux_smem, regs, levels;
    int live_word;                   // mask word of the cohort's live lanes (-1: no TOP_RETIRE in the program)
    uint64_t per_traj_rows, nstats;
};

namespace {
struct Span { int lo, hi; };                     // [lo, hi) in the element arena, or a mask word as [w, w+1) shifted by MASK0
const int MASK0 = 1 << 28, SCR = (1 << 28) - 1;  // address-space tags: mask words live above MASK0, the measurement scratch is SCR
struct Access { std::vector<Span> r, w; };
bool overlap(const std::vector<Span>& a, const std::vector<Span>& b) {
    for (auto& x : a) for (auto& y : b) if (x.lo < y.hi && y.lo < x.hi) return true;
    return false;
}
Access access_of(const TabOp& o, int live = -1) {
    Access A;
    auto E = [](int base, int n) { return Span{base, base + (1 << (2 * n))}; };
    auto M = [](int w) { return Span{MASK0 + w, MASK0 + w + 1}; };
    switch (o.type) {
    case DOP_CONST: A.w = {E(o.s[0], o.n), M(o.al[0])}; break;
    case DOP_COPY: A.r = {E(o.s[1], o.n), M(o.al[1])}; A.w = {E(o.s[0], o.n), M(o.al[0])}; break;
    case DOP_TENSOR: A.r = {E(o.s[1], o.n), E(o.s[2], o.n2), M(o.al[1]), M(o.al[2])}; A.w = {E(o.s[0], o.n + o.n2), M(o.al[0])}; break;
    case DOP_APPLY: A.r = {E(o.s[0], o.n)}; if (o.m[0] >= 0) A.r.push_back(M(o.m[0])); A.w = {E(o.s[0], o.n)}; break;
    case DOP_CHANNEL: case DOP_PTRACE: A.r = {E(o.s[0], o.n)}; A.w = {E(o.s[0], o.n)}; break;
    case DOP_MEASURE: A.r = {E(o.s[0], o.n), M(o.al[0]), Span{SCR, SCR + 1}}; A.w = {E(o.s[0], o.n), M(o.m[0]), Span{SCR, SCR + 1}}; break;
    case DOP_EXPECT: A.r = {E(o.s[0], o.n), M(o.al[0])}; if (live >= 0) A.r.push_back(M(live)); break;
    case TOP_RETIRE: A.r = {M(o.m[0]), M(live)}; A.w = {M(live)}; break;
    case TOP_MASK_FILL: case TOP_MASK_RANDOM: A.w = {M(o.m[0])}; break;
    case TOP_MASK_LOGIC: A.r = {M(o.m[1])}; if (o.m[2] >= 0) A.r.push_back(M(o.m[2])); A.w = {M(o.m[0])}; break;
    case TOP_KILL: A.r = {M(o.al[0]), M(o.m[0])}; A.w = {M(o.al[0])}; break;
    case TOP_ALIVE_COPY: A.r = {M(o.al[0])}; A.w = {M(o.m[0])}; break;
    default: break;
    }
    return A;
}
}  // namespace

int qsb_dexec_build(qsb_ctx* ctx, qsb_recorder* R, int elems, int nmasks, qsb_dprogram** out) {
    qsb_dprogram* P = new qsb_dprogram();
    memset(P, 0, sizeof(*P));
    P->ctx = ctx; P->elems = elems > 0 ? elems : 1; P->nmasks = nmasks > 0 ? nmasks : 1;
    P->live_word = -1;
    for (auto& o : R->ops) if (o.type == TOP_RETIRE) { P->live_word = P->nmasks; P->nmasks += 1; break; }   // one more mask word: the cohort's live lanes
    // ops, barrier flags, group assignment
    std::vector<DExecOp> xops;
    std::vector<double> aux = R->daux;
    std::vector<int> src;   // xops[j] came from R->ops[src[j]]
    std::map<int, int> coef_at;
    for (size_t i = 0; i < R->ops.size(); ++i) {
        const TabOp& o = R->ops[i];
        if (o.type == TOP_NOP) continue;
        DExecOp x; memset(&x, 0, sizeof(x));
        x.type = (uint8_t)o.type; x.n = (uint8_t)o.n; x.n2 = (uint8_t)o.n2; x.k = (uint8_t)o.k;
        x.q[0] = (uint8_t)o.q[0]; x.q[1] = (uint8_t)o.q[1];
        for (int j = 0; j < 3; ++j) { x.al[j] = (int16_t)o.al[j]; x.m[j] = (int16_t)o.m[j]; x.s[j] = o.s[j]; }
        x.aux = o.obs; x.u[0] = o.u[0]; x.u[1] = o.u[1]; x.draw = (uint32_t)o.draw; x.stats = o.stats; x.per_traj = o.per_traj;
        if (o.type == DOP_EXPECT) {
            x.mode = (uint8_t)o.iv[0]; x.nnz = (uint16_t)o.iv[1];
            // f = (+-1) * i^ny as a power of i
            x.fpow = o.f[0] > 0.5 ? 0u : o.f[0] < -0.5 ? 2u : o.f[1] > 0.5 ? 1u : 3u;
            if (o.per_traj) P->per_traj_rows++;
        } else if (o.type == DOP_APPLY) {
            const int M = 1 << o.k, MM = M * M;
            for (int e = 0; e < MM; ++e) if (aux[o.obs + 2 * e] != 0.0 || aux[o.obs + 2 * e + 1] != 0.0) x.u[0] |= 1u << e;
            // one non-zero per row and per column: a generalised permutation, applied by moving elements (d_apply_perm)
            int p[4] = {0, 0, 0, 0}, colseen = 0; bool gperm = true;
            for (int r = 0; r < M && gperm; ++r) {
                int cnt = 0;
                for (int m = 0; m < M; ++m) if ((x.u[0] >> (r + m * M)) & 1u) { p[r] = m; ++cnt; }
                if (cnt != 1 || ((colseen >> p[r]) & 1)) gperm = false;
                colseen |= 1 << p[r];
            }
            if (gperm) {
                std::vector<double> coef(2 * MM);
                uint32_t perm = 0, cmask = 0;
                for (int r = 0; r < M; ++r) perm |= (uint32_t)p[r] << (o.k * r);
                for (int c = 0; c < M; ++c) for (int r = 0; r < M; ++r) {   // coefficient w_r conj(w_c), w_r = U[r][p(r)]
                    const double ar = aux[o.obs + 2 * (r + p[r] * M)], ai = aux[o.obs + 2 * (r + p[r] * M) + 1];
                    const double br = aux[o.obs + 2 * (c + p[c] * M)], bi = aux[o.obs + 2 * (c + p[c] * M) + 1];
                    const double re = ar * br + ai * bi, im = ai * br - ar * bi;
                    coef[2 * (r + c * M)] = re; coef[2 * (r + c * M) + 1] = im;
                    if (re != 1.0 || im != 0.0) cmask |= 1u << (r + c * M);
                }
                auto hit = coef_at.find(o.obs);   // the same gate matrix (deduplicated payload) shares one coefficient table
                if (hit == coef_at.end()) {
                    if (aux.size() & 1) aux.push_back(0.0);
                    hit = coef_at.insert({o.obs, (int)aux.size()}).first;
                    aux.insert(aux.end(), coef.begin(), coef.end());
                }
                x.aux = hit->second; x.mode = 1; x.u[0] = perm; x.u[1] = cmask;
            }
        } else if (o.type == TOP_MASK_FILL || o.type == TOP_MASK_LOGIC) x.mode = (uint8_t)o.iv[0];
        else if (o.type == TOP_MASK_RANDOM) { if (aux.size() & 1) aux.push_back(0.0); x.aux = (int32_t)aux.size(); aux.push_back(o.f[0]); }
        if (o.draw > 0xffffffffull || P->nmasks > 32767) { delete P; QSB_FAIL(QSB_ERR_UNSUPPORTED, "program too large for the dense executor's op encoding"); }
        src.push_back((int)i);
        xops.push_back(x);
    }
    // Level scheduling: the recorded order is the host's call order, but ops on different pairs (background noise of every
    // idle pair at a time step, observables, the two halves of a swap) are independent.  Ops are re-ordered by their depth
    // in the hazard graph (RAW / WAR / WAW on arena spans, mask words and the measurement scratch), so that the ops of one
    // level run without barriers between them and spread over the groups of a block.  Draw ids and statistics rows are
    // carried by the ops themselves, so the result does not depend on the order.  QSB_DEXEC_NOSCHED=1 keeps call order.
    std::vector<int> level(xops.size(), 0);
    if (!getenv("QSB_DEXEC_NOSCHED") && xops.size() > 1) {
        std::vector<Access> accs(xops.size());
        for (size_t j = 0; j < xops.size(); ++j) accs[j] = access_of(R->ops[src[j]], P->live_word);
        int nlev = 0;
        for (size_t j = 0; j < xops.size(); ++j) {
            int lv = 0;
            for (size_t i = j; i-- > 0;) {
                if (level[i] < lv) continue;   // cannot raise lv
                if (overlap(accs[j].w, accs[i].r) || overlap(accs[j].w, accs[i].w) || overlap(accs[j].r, accs[i].w)) lv = level[i] + 1;
            }
            level[j] = lv; nlev = std::max(nlev, lv + 1);
        }
        if (getenv("QSB_REC_TRACE")) {   // how long would the chain be without storage reuse?  (tuning aid)
            auto depth = [&](bool raw_only, bool skip_small) {
                auto filt = [&](const std::vector<Span>& v) { std::vector<Span> o; for (auto& x : v) if (!skip_small || x.lo < SCR) o.push_back(x); return o; };
                std::vector<int> lv(xops.size(), 0); int d = 0;
                std::vector<Access> fa(xops.size());
                for (size_t j = 0; j < xops.size(); ++j) { fa[j].r = filt(accs[j].r); fa[j].w = filt(accs[j].w); }
                for (size_t j = 0; j < xops.size(); ++j) {
                    int l = 0;
                    for (size_t i = j; i-- > 0;) {
                        if (lv[i] < l) continue;
                        bool dep = overlap(fa[j].r, fa[i].w) || (!raw_only && (overlap(fa[j].w, fa[i].r) || overlap(fa[j].w, fa[i].w)));
                        if (dep) l = lv[i] + 1;
                    }
                    lv[j] = l; d = std::max(d, l + 1);
                }
                return d;
            };
            fprintf(stderr, "[rec] dense hazard-graph depth: %d (all hazards), %d (ignoring mask words and scratch), %d (true dependencies only), %d (true, ignoring masks/scratch)\n",
                    nlev, depth(false, true), depth(true, false), depth(true, true));
        }
        std::vector<int> order(xops.size());
        for (size_t j = 0; j < order.size(); ++j) order[j] = (int)j;
        std::stable_sort(order.begin(), order.end(), [&](int x, int y) { return level[x] < level[y]; });
        std::vector<DExecOp> xo(xops.size()); std::vector<int> so(xops.size()), lo(xops.size());
        for (size_t j = 0; j < order.size(); ++j) { xo[j] = xops[order[j]]; so[j] = src[order[j]]; lo[j] = level[order[j]]; }
        xops.swap(xo); src.swap(so); level.swap(lo);
        P->levels = nlev;
    }
    P->nops = (int)xops.size(); P->naux = (int)aux.size(); P->nstats = R->stats_ptrs.size();
    // launch shape: T lanes per tile (32, 16 or 8) x block variant.  The executor is latency bound (a tile walks its op list
    // serially), so throughput follows the number of trajectories resident per SM: rank by resident trajectories, then by tile width.  QSB_DEXEC_T / QSB_DEXEC_THREADS pin the shape (tuning aid).
    int dev_smem = 0, sm_smem = 0;
    cudaDeviceGetAttribute(&dev_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, ctx->device);
    cudaDeviceGetAttribute(&sm_smem, cudaDevAttrMaxSharedMemoryPerMultiprocessor, ctx->device);
    const int NV = 7;
    cudaFuncAttributes fa[NV];
    const int vthreads[NV] = {256, 256, 128, 32, 32, 512, 1024};
    const void* vfn[NV] = {(const void*)k_dense_exec<256, 1>, (const void*)k_dense_exec<256, 2>, (const void*)k_dense_exec<128, 4>,
                           (const void*)k_dense_exec<32, 8>, (const void*)k_dense_exec<32, 16>, (const void*)k_dense_exec<512, 1>,
                           (const void*)k_dense_exec<1024, 1>};
    for (int v = 0; v < NV; ++v)
        if (cudaFuncGetAttributes(&fa[v], vfn[v]) != cudaSuccess) { cudaGetLastError(); delete P; QSB_FAIL(QSB_ERR_CUDA, "k_dense_exec is not loadable on this device"); }
    const char *force_t = getenv("QSB_DEXEC_T"), *force_v = getenv("QSB_DEXEC_VARIANT");
    const int aux_smem = P->naux * 8 <= 16384 ? (P->naux + 1) / 2 * 2 : 0;
    long best = -1;
    for (int v = 0; v < NV; ++v) for (int logT = 5; logT >= 2; --logT) {
        const int T = 1 << logT;
        if (T > vthreads[v] || (vthreads[v] == 32 && T > 8)) continue;
        if ((force_t && atoi(force_t) != T) || (force_v && atoi(force_v) != v)) continue;
        if (v >= 5 && !force_v) continue;   // the wide blocks (more warps per tile) are opt-in until measured on a workload
        size_t smem = (size_t)P->elems * T * sizeof(double2) + 2 * T * sizeof(double) + (size_t)(P->nmasks + 3) / 4 * 16 + (size_t)aux_smem * 8;
        smem = (smem + 15) / 16 * 16;
        if (smem + fa[v].sharedSizeBytes > (size_t)dev_smem) continue;
        int bps = (int)((size_t)sm_smem / (smem + fa[v].sharedSizeBytes + 1024));
        const int by_regs = 65536 / (fa[v].numRegs * vthreads[v]);
        if (bps > by_regs) bps = by_regs;
        if (bps > 24) bps = 24;
        if (bps < 1) continue;
        if (v == 0 && bps > 1) bps = 1;   // the 255-register variant is for arenas that fill an SM
        // measured on B200 (tools/sweep_dexec.sh, firstgenrepeater dm): at equal resident trajectories the widest tile wins
        // (32 x 1 block: 349 ms, 16 x 2: 398 ms, 8 x 4: 456 ms per 10^6 trajectories)
        const long rank = (long)T * bps * 10000 + (long)T * 10 + (NV - v);
        if (rank > best) { best = rank; P->logT = logT; P->blocks_per_sm = bps; P->smem_bytes = (int)smem; P->variant = v; P->threads = vthreads[v]; P->regs = fa[v].numRegs; }
    }
    if (best < 0) { delete P; QSB_FAIL(QSB_ERR_UNSUPPORTED, "the recorded dense program keeps %d matrix elements per trajectory alive: it does not fit in shared memory even 8 trajectories at a time", elems); }
    P->aux_smem = aux_smem;
    {   // the single-group parts of ops (measurement decision, observables, random masks) rotate over the groups
        const int G = P->threads >> P->logT;
        // within a level the groups are dealt out round-robin: an op with fewer items than groups starts at the first group
        // the previous op of the level left idle (g0), single-group parts (measurement decision, observables, random masks) take one group
        int rr = 0, cur = -1;
        for (size_t j = 0; j < xops.size(); ++j) {
            DExecOp& x = xops[j];
            if (level[j] != cur) { cur = level[j]; if (P->levels) rr = 0; }
            int items = 0;
            switch (x.type) {
                case DOP_MEASURE: case DOP_EXPECT: case TOP_MASK_RANDOM: case TOP_RETIRE: x.grp = (uint16_t)(rr % G); items = 1; break;
                case DOP_CONST: case DOP_COPY: items = 1 << (2 * x.n); break;
                case DOP_TENSOR: items = 1 << (2 * (x.n + x.n2)); break;
                case DOP_APPLY: items = 1 << (2 * x.n - 2 * x.k); break;
                case DOP_CHANNEL: case DOP_PTRACE: items = 1 << (2 * x.n - 2); break;
                default: break;
            }
            if (items > 1 && P->levels) x.g0 = (uint8_t)(rr % G);
            rr += std::min(items, G);
        }
        // barrier before an op when it touches storage an op since the last barrier wrote (or writes what one read); the
        // op-chunk refill of the chosen block variant is a barrier already
        const int chunk = QSB_DEXEC_CHUNK_OF(P->threads);
        Access acc;
        for (size_t j = 0; j < xops.size(); ++j) {
            const TabOp& o = R->ops[src[j]];
            Access A = access_of(o, P->live_word);
            bool need = overlap(A.w, acc.r) || overlap(A.w, acc.w) || overlap(A.r, acc.w);
            if ((j % chunk) == 0) { acc = Access(); need = false; }
            if (need) { xops[j].flags |= 1; P->barriers++; acc = Access(); }
            if (o.type == DOP_MEASURE || o.type == DOP_PTRACE) { acc = Access(); acc.w = {Span{o.s[0], o.s[0] + (1 << (2 * o.n))}}; if (o.type == DOP_MEASURE) acc.w.push_back(Span{MASK0 + o.m[0], MASK0 + o.m[0] + 1}); }
            else { acc.r.insert(acc.r.end(), A.r.begin(), A.r.end()); acc.w.insert(acc.w.end(), A.w.begin(), A.w.end()); }
        }
    }
    int rc = QSB_OK;
    if (!xops.empty()) {
        if (cudaMallocHost((void**)&P->h_ops, xops.size() * sizeof(DExecOp)) != cudaSuccess) { cudaGetLastError(); qsb_set_error("pinned allocation of the op list failed"); rc = QSB_ERR_OOM; }
        else memcpy(P->h_ops, xops.data(), xops.size() * sizeof(DExecOp));
    }
    if (rc == QSB_OK) rc = qsb_dev_alloc(ctx, (void**)&P->d_ops, xops.empty() ? 16 : xops.size() * sizeof(DExecOp));
    if (rc == QSB_OK) rc = qsb_dev_alloc(ctx, (void**)&P->d_aux, aux.empty() ? 16 : aux.size() * sizeof(double));
    if (rc == QSB_OK && !xops.empty() && cudaMemcpyAsync(P->d_ops, P->h_ops, xops.size() * sizeof(DExecOp), cudaMemcpyHostToDevice, ctx->stream) != cudaSuccess) rc = QSB_ERR_CUDA;
    if (rc == QSB_OK && !aux.empty() && cudaMemcpyAsync(P->d_aux, aux.data(), aux.size() * sizeof(double), cudaMemcpyHostToDevice, ctx->stream) != cudaSuccess) rc = QSB_ERR_CUDA;
    if (rc == QSB_OK && cudaStreamSynchronize(ctx->stream) != cudaSuccess) rc = QSB_ERR_CUDA;   // `aux` dies below
    for (int v = 0; v < NV && rc == QSB_OK; ++v)
        if (cudaFuncSetAttribute(vfn[v], cudaFuncAttributeMaxDynamicSharedMemorySize, dev_smem - (int)fa[v].sharedSizeBytes) != cudaSuccess) rc = QSB_ERR_CUDA;
    if (rc != QSB_OK) { if (rc == QSB_ERR_CUDA) { qsb_set_error("dense program upload failed: %s", cudaGetErrorString(cudaGetLastError())); } qsb_dexec_free(P); return rc; }
    if (getenv("QSB_REC_TRACE")) fprintf(stderr, "[rec] dense program: %d ops, %d levels, %d barriers, %d elements, %d mask words, %d aux doubles (%s), T=%d, %d threads, %d block(s)/SM, %d B smem, %d regs\n",
                                         P->nops, P->levels, P->barriers, P->elems, P->nmasks, P->naux, P->aux_smem ? "smem" : "global", 1 << P->logT, P->threads, P->blocks_per_sm, P->smem_bytes, P->regs);
    *out = P;
    return QSB_OK;
}

int qsb_dexec_run(qsb_dprogram* P, const uint32_t* lanes_dev, int64_t count, uint64_t traj_shift, uint64_t draw_base, cudaStream_t stream) {
    qsb_ctx* ctx = P->ctx;
    if (!stream) stream = ctx->stream;
    DExecArgs a;
    a.ops = P->d_ops; a.nops = P->nops; a.aux = P->d_aux; a.elems = P->elems; a.nmasks = P->nmasks; a.logT = P->logT; a.aux_smem = P->aux_smem;
    a.valid = ctx->valid; a.ntiles = lanes_dev ? ((count + (1 << P->logT) - 1) >> P->logT) : (ctx->npad >> P->logT);
    a.lanes = lanes_dev; a.count = lanes_dev ? count : ctx->ntraj; a.live_word = P->live_word;
    if (a.ntiles == 0) return QSB_OK;
    a.seed = ctx->seed; a.traj0 = (uint64_t)ctx->traj_offset + traj_shift; a.draw_base = draw_base; a.flags = ctx->dev_flags;
    int64_t blocks = (int64_t)ctx->sm_count * P->blocks_per_sm;
    if (blocks > a.ntiles) blocks = a.ntiles;
    QSB_PRE(ctx, "k_dense_exec");
    if (P->variant == 0) k_dense_exec<256, 1><<<(unsigned)blocks, 256, P->smem_bytes, stream>>>(a);
    else if (P->variant == 1) k_dense_exec<256, 2><<<(unsigned)blocks, 256, P->smem_bytes, stream>>>(a);
    else if (P->variant == 2) k_dense_exec<128, 4><<<(unsigned)blocks, 128, P->smem_bytes, stream>>>(a);
    else if (P->variant == 3) k_dense_exec<32, 8><<<(unsigned)blocks, 32, P->smem_bytes, stream>>>(a);
    else if (P->variant == 5) k_dense_exec<512, 1><<<(unsigned)blocks, 512, P->smem_bytes, stream>>>(a);
    else if (P->variant == 6) k_dense_exec<1024, 1><<<(unsigned)blocks, 1024, P->smem_bytes, stream>>>(a);
    else k_dense_exec<32, 16><<<(unsigned)blocks, 32, P->smem_bytes, stream>>>(a);
    QSB_LAUNCH_CHECK(ctx);
    // compulsory HBM traffic of the fused launch: op list + constants once, statistics and per-trajectory rows written once
    qsb_account(ctx, "k_dense_exec", (uint64_t)P->nops * sizeof(DExecOp) + (uint64_t)P->naux * 8 + P->nstats * 32 + P->per_traj_rows * (uint64_t)ctx->npad * 16);
    return QSB_OK;
}

int qsb_dexec_upload(qsb_dprogram* P, int64_t* bytes) {
    size_t b = (size_t)P->nops * sizeof(DExecOp);
    if (b) QSB_CUDA(cudaMemcpyAsync(P->d_ops, P->h_ops, b, cudaMemcpyHostToDevice, P->ctx->stream));
    if (bytes) *bytes = (int64_t)b;
    return QSB_OK;
}

void qsb_dexec_info(const qsb_dprogram* P, int64_t info[8]) {
    info[0] = P->nops; info[1] = P->elems; info[3] = P->threads; info[4] = P->smem_bytes; info[5] = 0;
}
int qsb_dexec_tile(const qsb_dprogram* P) { return 1 << P->logT; }
int qsb_dexec_barriers(const qsb_dprogram* P) { return P->barriers; }

void qsb_dexec_free(qsb_dprogram* P) {
    if (!P) return;
    qsb_dev_free(P->ctx, P->d_ops); qsb_dev_free(P->ctx, P->d_aux);
    if (P->h_ops) cudaFreeHost(P->h_ops);
    delete P;
}
