// qsb_tab_core.cuh -- the bit-sliced tableau algorithms as host+device inline functions.
// The CUDA kernels in qsb_tab.cu are thin per-word wrappers around these; tests/cpu_emul compiles the
// same functions with g++ to exercise the lane-masked pivot logic on 32-lane words without a GPU
// (a debugging aid for the algorithm only -- never a product path, never timed).
#pragma once
#include <stdint.h>
#include <stddef.h>
#if defined(__CUDACC__)
#define QSB_HD __host__ __device__ __forceinline__
#else
#define QSB_HD inline
#endif
#ifndef QSB_TAB_MAXN
#define QSB_TAB_MAXN 32
#endif
#define MAXN QSB_TAB_MAXN
// Every loop over rows / columns is marked for unrolling: the fused executor inlines these functions with a LITERAL qubit count,
// and only fully unrolled loops keep the per-row scratch arrays (Analysis, pivots, products) in registers -- a rolled loop indexes
// them dynamically, which puts them in local memory, and with the shared-memory carve-out at its maximum an LDL is an L2 round trip.
#if defined(__CUDACC__)
#define QSB_UNROLL _Pragma("unroll")
#else
#define QSB_UNROLL
#endif
// Storage type of one bit plane element: the number of trajectories ("lanes") one thread advances.  The per-op
// kernels use 32-lane words; the fused executor (qsb_exec.cu) can compile this header with narrower planes (more threads
// out of the same ensemble; see qsb_exec.cu).  Arithmetic is always done in 32-bit registers; QSB_FULL masks the lanes.
#ifndef QSB_PLANE_T
#define QSB_PLANE_T uint32_t
#define QSB_CORE_NS tabcore32
#endif
namespace QSB_CORE_NS {
typedef QSB_PLANE_T plane_t;
#define QSB_LANES ((int)(8 * sizeof(plane_t)))
#define QSB_FULL ((uint32_t)(plane_t)~(plane_t)0)

QSB_HD uint32_t qsb_mulhi32(uint32_t a, uint32_t b) {
#if defined(__CUDA_ARCH__)
    return __umulhi(a, b);
#else
    return (uint32_t)(((uint64_t)a * b) >> 32);
#endif
}
// Philox4x32-10; identical to the oracle's qso_draw (the "injected draw").
// One Philox block = 128 bits = TWO 53-bit uniforms: draw 2b uses words (0,1), draw 2b+1 words (2,3) of the block with
// counter (trajectory, b).  Consecutive draw ids of a trajectory therefore share one block (the fused draw pass
// keeps it in registers); a lone draw costs the same ten rounds as before.
QSB_HD void qsb_philox_block(uint64_t seed, uint64_t traj, uint64_t block, uint32_t* out) {
    uint32_t c0 = (uint32_t)traj, c1 = (uint32_t)(traj >> 32), c2 = (uint32_t)block, c3 = (uint32_t)(block >> 32);
    uint32_t k0 = (uint32_t)seed, k1 = (uint32_t)(seed >> 32);
    for (int r = 0; r < 10; ++r) {
        uint32_t h0 = qsb_mulhi32(0xD2511F53u, c0), l0 = 0xD2511F53u * c0;
        uint32_t h1 = qsb_mulhi32(0xCD9E8D57u, c2), l1 = 0xCD9E8D57u * c2;
        uint32_t n0 = h1 ^ c1 ^ k0, n2 = h0 ^ c3 ^ k1;
        c0 = n0; c1 = l1; c2 = n2; c3 = l0;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
    out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}
QSB_HD double qsb_uniform53(uint32_t lo, uint32_t hi) {
    uint64_t bits = ((uint64_t)hi << 32) | lo;
    return (double)(bits >> 11) * (1.0 / 9007199254740992.0);
}
QSB_HD double qsb_draw(uint64_t seed, uint64_t traj, uint64_t draw_id) {
    uint32_t o[4];
    qsb_philox_block(seed, traj, draw_id >> 1, o);
    return (draw_id & 1) ? qsb_uniform53(o[2], o[3]) : qsb_uniform53(o[0], o[1]);
}

struct TabView {
    plane_t* p;   // this thread's column: plane k of the state is p[k * W]
    size_t W; int n; int S;
    QSB_HD uint32_t ld(int plane) const { return (uint32_t)p[(size_t)(unsigned)plane * W]; }
    QSB_HD void st(int plane, uint32_t v) const { p[(size_t)(unsigned)plane * W] = (plane_t)v; }
    QSB_HD int dx(int i, int c) const { return i * S + c; }
    QSB_HD int dz(int i, int c) const { return i * S + n + c; }
    QSB_HD int sx(int i, int c) const { return i * S + 2 * n + c; }
    QSB_HD int sz(int i, int c) const { return i * S + 3 * n + c; }
    QSB_HD int sg(int i) const { return i * S + 4 * n; }
    QSB_HD int ac(int i) const { return i * S + 4 * n + 1; }
};
QSB_HD TabView make_view(plane_t* p, size_t W, size_t w, int n) { TabView v; v.p = p + w; v.W = W; v.n = n; v.S = 4 * n + 2; return v; }


// dst row (planes x0.., z0.., optional sign plane) <- src * dst on lanes m, src given in registers/local
// arrays.  The exponent of i is tracked in a bit-sliced 2-bit counter (lo,hi); rows always commute here,
// so lo ends at 0 and the sign flip is hi.
QSB_HD void row_mul(const TabView& v, int x0, int z0, int sgplane, uint32_t m,
                                        const uint32_t* sx, const uint32_t* sz, uint32_t ssign) {
    if (m == 0) return;
    uint32_t lo = 0, hi = 0;
    QSB_UNROLL
    for (int c = 0; c < v.n; ++c) {
        uint32_t x1 = sx[c], z1 = sz[c];
        if ((x1 | z1) == 0) continue;
        uint32_t x2 = v.ld(x0 + c), z2 = v.ld(z0 + c);
        uint32_t anti = (x1 & z2) ^ (z1 & x2);
        // g = -1 for (X,Z), (Y,X), (Z,Y); +1 for the cyclic pairs
        uint32_t neg = anti & ((x1 & ~z1 & ~x2) | (x1 & z1 & ~z2) | (~x1 & x2 & z2));
        hi ^= (lo & anti) ^ neg;
        lo ^= anti;
        v.st(x0 + c, x2 ^ (x1 & m));
        v.st(z0 + c, z2 ^ (z1 & m));
    }
    if (sgplane >= 0) v.st(sgplane, v.ld(sgplane) ^ ((ssign ^ hi) & m));
}

struct Analysis {
    uint32_t acS[MAXN], acD[MAXN], act[MAXN], piv[MAXN];
    uint32_t found;
};

// anticommutation words of every row with the Pauli (px,pz) and lane-wise pivot choice:
// first active stabilizer that anticommutes, else an inactive (logical) row that anticommutes.
QSB_HD void analyze(const TabView& v, uint32_t px, uint32_t pz, Analysis& A, uint32_t* pivD) {
    int n = v.n;
    QSB_UNROLL
    for (int i = 0; i < n; ++i) {
        uint32_t s = 0, d = 0;
        QSB_UNROLL
        for (int c = 0; c < n; ++c) {
            if ((pz >> c) & 1) { s ^= v.ld(v.sx(i, c)); d ^= v.ld(v.dx(i, c)); }
            if ((px >> c) & 1) { s ^= v.ld(v.sz(i, c)); d ^= v.ld(v.dz(i, c)); }
        }
        A.acS[i] = s; A.acD[i] = d; A.act[i] = v.ld(v.ac(i));
    }
    uint32_t found = 0;
    QSB_UNROLL
    for (int i = 0; i < n; ++i) { A.piv[i] = A.acS[i] & A.act[i] & ~found; found |= A.piv[i]; }
    QSB_UNROLL
    for (int i = 0; i < n; ++i) { uint32_t t = A.acS[i] & ~A.act[i] & ~found; A.piv[i] |= t; found |= t; }
    QSB_UNROLL
    for (int i = 0; i < n; ++i) { uint32_t t = A.acD[i] & ~A.act[i] & ~found; pivD[i] = t; found |= t; }
    A.found = found;
}

// sign (1 = minus) of the group element equal to the Pauli bits on the lanes where it is determined:
// product of the active S_i whose destabilizer anticommutes with P.
QSB_HD uint32_t determined_sign(const TabView& v, const Analysis& A, uint32_t det) {
    uint32_t accx[MAXN], accz[MAXN];
    int n = v.n;
    QSB_UNROLL
    for (int c = 0; c < n; ++c) { accx[c] = 0; accz[c] = 0; }
    uint32_t lo = 0, hi = 0, sg = 0;
    QSB_UNROLL
    for (int i = 0; i < n; ++i) {
        uint32_t m = A.acD[i] & A.act[i] & det;
        if (m == 0) continue;
        QSB_UNROLL
        for (int c = 0; c < n; ++c) {
            uint32_t x1 = v.ld(v.sx(i, c)) & m, z1 = v.ld(v.sz(i, c)) & m;
            uint32_t x2 = accx[c], z2 = accz[c];
            uint32_t anti = (x1 & z2) ^ (z1 & x2);
            uint32_t neg = anti & ((x1 & ~z1 & ~x2) | (x1 & z1 & ~z2) | (~x1 & x2 & z2));
            hi ^= (lo & anti) ^ neg;
            lo ^= anti;
            accx[c] = x2 ^ x1; accz[c] = z2 ^ z1;
        }
        sg ^= v.ld(v.sg(i)) & m;
    }
    return (sg ^ hi) & det;
}

// Full projection update.  Returns in A.piv the (lane-wise one-hot) pair whose S row now holds +/-P.
// random lanes (A.found): S_piv <- (-1)^rbit P, D_piv <- old S_piv, other anticommuting rows *= old S_piv.
// determined lanes: outcome from determined_sign; when `settle` the pair of the first contributing
// destabilizer is rewritten so that S_piv = +/-P as well (needed before a qubit is removed).
QSB_HD uint32_t project_update(const TabView& v, uint32_t px, uint32_t pz, uint32_t rbits, bool settle, Analysis& A) {
    int n = v.n;
    uint32_t pivD[MAXN];
    analyze(v, px, pz, A, pivD);
    uint32_t found = A.found, det = ~found & QSB_FULL;
    uint32_t dsign = determined_sign(v, A, det);
    // logical-row pivots found on a D row: exchange D_i and S_i there (both are logical operators)
    QSB_UNROLL
    for (int i = 0; i < n; ++i) {
        uint32_t m = pivD[i];
        if (m == 0) continue;
        QSB_UNROLL
        for (int c = 0; c < n; ++c) {
            uint32_t a = v.ld(v.dx(i, c)), b = v.ld(v.sx(i, c)), t = (a ^ b) & m;
            v.st(v.dx(i, c), a ^ t); v.st(v.sx(i, c), b ^ t);
            a = v.ld(v.dz(i, c)); b = v.ld(v.sz(i, c)); t = (a ^ b) & m;
            v.st(v.dz(i, c), a ^ t); v.st(v.sz(i, c), b ^ t);
        }
        v.st(v.sg(i), v.ld(v.sg(i)) & ~m);
        uint32_t t = (A.acS[i] ^ A.acD[i]) & m;
        A.acS[i] ^= t; A.acD[i] ^= t;
        A.piv[i] |= m;
    }
    if (found) {
        uint32_t pvx[MAXN], pvz[MAXN], pvs = 0;
        QSB_UNROLL
        for (int c = 0; c < n; ++c) { pvx[c] = 0; pvz[c] = 0; }
        QSB_UNROLL
        for (int i = 0; i < n; ++i) {
            uint32_t m = A.piv[i];
            if (m == 0) continue;
            QSB_UNROLL
            for (int c = 0; c < n; ++c) { pvx[c] |= v.ld(v.sx(i, c)) & m; pvz[c] |= v.ld(v.sz(i, c)) & m; }
            pvs |= v.ld(v.sg(i)) & m;
        }
        QSB_UNROLL
        for (int i = 0; i < n; ++i) {
            row_mul(v, v.sx(i, 0), v.sz(i, 0), v.sg(i), A.acS[i] & found & ~A.piv[i], pvx, pvz, pvs);
            row_mul(v, v.dx(i, 0), v.dz(i, 0), -1, A.acD[i] & found & ~A.piv[i], pvx, pvz, pvs);
        }
        QSB_UNROLL
        for (int i = 0; i < n; ++i) {
            uint32_t m = A.piv[i];
            if (m == 0) continue;
            QSB_UNROLL
            for (int c = 0; c < n; ++c) {
                uint32_t pxc = ((px >> c) & 1) ? QSB_FULL : 0u, pzc = ((pz >> c) & 1) ? QSB_FULL : 0u;
                v.st(v.dx(i, c), (v.ld(v.dx(i, c)) & ~m) | (pvx[c] & m));
                v.st(v.dz(i, c), (v.ld(v.dz(i, c)) & ~m) | (pvz[c] & m));
                v.st(v.sx(i, c), (v.ld(v.sx(i, c)) & ~m) | (pxc & m));
                v.st(v.sz(i, c), (v.ld(v.sz(i, c)) & ~m) | (pzc & m));
            }
            v.st(v.sg(i), (v.ld(v.sg(i)) & ~m) | (rbits & m));
            A.act[i] |= m;
            v.st(v.ac(i), A.act[i]);
        }
    }
    if (settle && det) {
        // choose p = first active pair whose D anticommutes with P; D_i <- D_i D_p for the others; S_p <- +/-P
        uint32_t pd[MAXN], fd = 0;
        QSB_UNROLL
        for (int i = 0; i < n; ++i) { pd[i] = A.acD[i] & A.act[i] & det & ~fd; fd |= pd[i]; }
        uint32_t dpx[MAXN], dpz[MAXN];
        QSB_UNROLL
        for (int c = 0; c < n; ++c) { dpx[c] = 0; dpz[c] = 0; }
        QSB_UNROLL
        for (int i = 0; i < n; ++i) {
            if (pd[i] == 0) continue;
            QSB_UNROLL
            for (int c = 0; c < n; ++c) { dpx[c] |= v.ld(v.dx(i, c)) & pd[i]; dpz[c] |= v.ld(v.dz(i, c)) & pd[i]; }
        }
        QSB_UNROLL
        for (int i = 0; i < n; ++i) {
            uint32_t m = A.acD[i] & A.act[i] & det & ~pd[i];
            if (m) {
                QSB_UNROLL
                for (int c = 0; c < n; ++c) { v.st(v.dx(i, c), v.ld(v.dx(i, c)) ^ (dpx[c] & m)); v.st(v.dz(i, c), v.ld(v.dz(i, c)) ^ (dpz[c] & m)); }
            }
        }
        QSB_UNROLL
        for (int i = 0; i < n; ++i) {
            uint32_t m = pd[i];
            if (m == 0) continue;
            QSB_UNROLL
            for (int c = 0; c < n; ++c) {
                uint32_t pxc = ((px >> c) & 1) ? QSB_FULL : 0u, pzc = ((pz >> c) & 1) ? QSB_FULL : 0u;
                v.st(v.sx(i, c), (v.ld(v.sx(i, c)) & ~m) | (pxc & m));
                v.st(v.sz(i, c), (v.ld(v.sz(i, c)) & ~m) | (pzc & m));
            }
            v.st(v.sg(i), (v.ld(v.sg(i)) & ~m) | (dsign & m));
            A.piv[i] |= m;
        }
    }
    return (rbits & found) | dsign;  // outcome bit per lane (1 = eigenvalue -1)
}

// drop qubit q and the pivot pair: every other row has I or the measured Pauli on q, multiplying by
// S_piv (= +/-P_q) clears it and only touches the sign; pair n-1 moves into the pivot's slot.
QSB_HD void remove_qubit(const TabView& v, const TabView& o, int q, const Analysis& A, uint32_t outcome, uint32_t skip_sign_fix) {
    int n = v.n;
    QSB_UNROLL
    for (int i = 0; i < n; ++i) {
        uint32_t nz = (v.ld(v.sx(i, q)) | v.ld(v.sz(i, q))) & ~A.piv[i] & ~skip_sign_fix;
        if (nz & outcome) v.st(v.sg(i), v.ld(v.sg(i)) ^ (nz & outcome));
    }
    int last = n - 1;
    QSB_UNROLL
    for (int i = 0; i < n - 1; ++i) {
        uint32_t m = A.piv[i];  // lanes whose pivot sits here take pair `last`
        // plane order (pair, kind, column) ascending on both sides: every write lands at or below the plane
        // just read, so `o` may alias `v` (in-place shrink inside a recorded program's arena)
        QSB_UNROLL
        for (int kind = 0; kind < 4; ++kind) {
            int cc = 0;
            QSB_UNROLL
            for (int c = 0; c < n; ++c) {
                if (c == q) continue;
                o.st(i * o.S + kind * (n - 1) + cc, (v.ld(i * v.S + kind * n + c) & ~m) | (v.ld(last * v.S + kind * n + c) & m));
                ++cc;
            }
        }
        o.st(o.sg(i), (v.ld(v.sg(i)) & ~m) | (v.ld(v.sg(last)) & m));
        o.st(o.ac(i), (v.ld(v.ac(i)) & ~m) | (v.ld(v.ac(last)) & m));
    }
}


QSB_HD uint32_t lane_bits_ge_half(uint64_t seed, uint64_t traj0, size_t w, uint64_t draw_id) {
    uint32_t r = 0;
    for (int l = 0; l < QSB_LANES; ++l) if (qsb_draw(seed, traj0 + (uint64_t)(w * QSB_LANES + l), draw_id) >= 0.5) r |= 1u << l;
    return r;
}


// ---------------------------------------------------------------------------------------------
// per-word bodies of the kernels in qsb_tab.cu
// ---------------------------------------------------------------------------------------------
QSB_HD void tab_init_word(plane_t* p, size_t W, size_t w, int n, int active) {
    TabView v = make_view(p, W, w, n);
    QSB_UNROLL
    for (int i = 0; i < n; ++i) {
        QSB_UNROLL
        for (int c = 0; c < n; ++c) {
            v.st(v.dx(i, c), c == i ? QSB_FULL : 0u); v.st(v.dz(i, c), 0u);
            v.st(v.sx(i, c), 0u); v.st(v.sz(i, c), c == i ? QSB_FULL : 0u);
        }
        v.st(v.sg(i), 0u);
        v.st(v.ac(i), active ? QSB_FULL : 0u);
    }
}

QSB_HD void tab_tensor_word(plane_t* out, const plane_t* a, const plane_t* b, size_t W, size_t w, int na, int nb) {
    int n = na + nb;
    TabView o = make_view(out, W, w, n), va = make_view((plane_t*)a, W, w, na), vb = make_view((plane_t*)b, W, w, nb);
    QSB_UNROLL
    for (int i = 0; i < n; ++i) {
        bool fa = i < na;
        int j = fa ? i : i - na;
        QSB_UNROLL
        for (int c = 0; c < n; ++c) {
            bool in_a = c < na;
            uint32_t dxv = 0, dzv = 0, sxv = 0, szv = 0;
            if (fa && in_a) { dxv = va.ld(va.dx(j, c)); dzv = va.ld(va.dz(j, c)); sxv = va.ld(va.sx(j, c)); szv = va.ld(va.sz(j, c)); }
            if (!fa && !in_a) { int cc = c - na; dxv = vb.ld(vb.dx(j, cc)); dzv = vb.ld(vb.dz(j, cc)); sxv = vb.ld(vb.sx(j, cc)); szv = vb.ld(vb.sz(j, cc)); }
            o.st(o.dx(i, c), dxv); o.st(o.dz(i, c), dzv); o.st(o.sx(i, c), sxv); o.st(o.sz(i, c), szv);
        }
        o.st(o.sg(i), fa ? va.ld(va.sg(j)) : vb.ld(vb.sg(j)));
        o.st(o.ac(i), fa ? va.ld(va.ac(j)) : vb.ld(vb.ac(j)));
    }
}

struct TabGateArgs { int n, k, q[2]; uint8_t outmask[4]; uint32_t lut, anf; /* anf = tab_lut_to_anf(lut, 2k), filled by the host */ };
// Sign truth table (bit t of lut: the image of the Pauli with input bits t picks up a minus) -> algebraic normal
// form: bit t of the result = monomial AND_{j in t} in[j] is present.  Standard gates have 0-3 monomials
// (Paulis: linear; CNOT: x_c z_t x_t + x_c z_t z_c + x_c z_t), so the kernels XOR a few ANDs instead of
// OR-ing up to 16 minterms.
QSB_HD uint32_t tab_lut_to_anf(uint32_t lut, int nb) {
    uint32_t f = lut;
    for (int j = 0; j < nb; ++j)
        for (int t = 0; t < (1 << nb); ++t)
            if ((t >> j) & 1) f ^= ((f >> (t ^ (1 << j))) & 1u) << t;
    return f;
}
QSB_HD int qsb_ctz32(uint32_t x) {
#if defined(__CUDA_ARCH__)
    return __ffs((int)x) - 1;
#else
    return __builtin_ctz(x);
#endif
}
// conjugation of every row by a K-qubit Clifford gate on the lanes m: the (x,z) bits of the target qubits go
// through the GF(2)-linear map `outmask` (only outputs that differ from their input are recomputed and stored),
// stabilizer signs flip where the ANF of the sign table evaluates to 1.
template <int K>
QSB_HD void tab_gate_rows(const TabView& v, int q0, int q1, const uint8_t* om, uint32_t anf, uint32_t m) {
    constexpr int NB = 2 * K;
    const int n = v.n;
    uint32_t chg = 0;
    QSB_UNROLL
    for (int b = 0; b < NB; ++b) if (om[b] != (1u << b)) chg |= 1u << b;
    QSB_UNROLL
    for (int i = 0; i < n; ++i) {
        QSB_UNROLL
        for (int half = 0; half < 2; ++half) {
            if (!half && !chg) continue;                      // Pauli gates touch signs only
            const int xb = i * v.S + half * 2 * n, zb = xb + n;
            uint32_t in[NB];
            in[0] = v.ld(xb + q0); in[1] = v.ld(zb + q0);
            if (K == 2) { in[NB - 2] = v.ld(xb + q1); in[NB - 1] = v.ld(zb + q1); }
            QSB_UNROLL
            for (int b = 0; b < NB; ++b) {
                if (!((chg >> b) & 1)) continue;
                uint32_t o = 0;
                QSB_UNROLL
                for (int j = 0; j < NB; ++j) o ^= ((om[b] >> j) & 1) ? in[j] : 0u;
                v.st(((b & 1) ? zb : xb) + ((b >> 1) ? q1 : q0), in[b] ^ ((in[b] ^ o) & m));
            }
            if (half && anf) {
                uint32_t flip = 0;
                for (uint32_t t = anf; t; t &= t - 1) {
                    const int k = qsb_ctz32(t);
                    uint32_t term = QSB_FULL;
                    QSB_UNROLL
                    for (int j = 0; j < NB; ++j) term &= ((k >> j) & 1) ? in[j] : QSB_FULL;
                    flip ^= term;
                }
                flip &= m;
                if (flip) v.st(v.sg(i), v.ld(v.sg(i)) ^ flip);
            }
        }
    }
}
QSB_HD void tab_gate_anf(plane_t* p, size_t W, size_t w, int n, int k, int q0, int q1, const uint8_t* om, uint32_t anf, uint32_t m) {
    if (m == 0) return;
    TabView v = make_view(p, W, w, n);
    if (k == 1) tab_gate_rows<1>(v, q0, q0, om, anf, m);
    else tab_gate_rows<2>(v, q0, q1, om, anf, m);
}
QSB_HD void tab_gate_word(plane_t* p, size_t W, size_t w, const TabGateArgs& a, uint32_t m) {
    tab_gate_anf(p, W, w, a.n, a.k, a.q[0], a.q[1], a.outmask, a.anf, m);
}

struct TabNoiseArgs { int n, q, np; double cum[3]; int op[3]; /* 1 X, 2 Y, 3 Z */ };
// which Pauli (bit 0: X component, bit 1: Z component) the draw r selects -- clifford/uptotime.jl:1-12
QSB_HD uint32_t tab_noise_decide(const TabNoiseArgs& a, double r) {
    for (int i = 0; i < a.np; ++i) {
        if (r > a.cum[i]) {
            int P = a.op[i];
            return ((P == 1 || P == 2) ? 1u : 0u) | ((P == 3 || P == 2) ? 2u : 0u);
        }
    }
    return 0u;
}
// fx / fz: lanes receiving a Pauli with X / Z component
QSB_HD void tab_noise_apply(plane_t* p, size_t W, size_t w, int n, int q, uint32_t fx, uint32_t fz) {
    if ((fx | fz) == 0) return;
    TabView v = make_view(p, W, w, n);
    QSB_UNROLL
    for (int i = 0; i < n; ++i) {  // sign of S_i flips where it anticommutes with the applied Pauli
        uint32_t flip = (v.ld(v.sx(i, q)) & fz) ^ (v.ld(v.sz(i, q)) & fx);
        if (flip) v.st(v.sg(i), v.ld(v.sg(i)) ^ flip);
    }
}
QSB_HD void tab_noise_word(plane_t* p, size_t W, size_t w, size_t wg, const TabNoiseArgs& a, uint64_t seed, uint64_t traj0, uint64_t draw_id) {
    uint32_t fx = 0, fz = 0;
    for (int l = 0; l < QSB_LANES; ++l) {
        uint32_t d = tab_noise_decide(a, qsb_draw(seed, traj0 + (uint64_t)(wg * QSB_LANES + l), draw_id));
        fx |= (d & 1u) << l; fz |= (d >> 1) << l;
    }
    tab_noise_apply(p, W, w, a.n, a.q, fx, fz);
}

// project_traceout! on a tableau (a9): projectremoverand!(state, projectX!/Y!/Z!, idx).
// returns the outcome word (bit = 1 for outcome 2); *rnd = lanes whose outcome consumed the draw.
// rbits: per-lane random eigenvalue bit (draw >= 0.5), used only on the lanes where the outcome is random
QSB_HD uint32_t tab_measure_bits(plane_t* in, plane_t* out, size_t W, size_t w, int n, int q, int tx, int tz, uint32_t rbits, uint32_t* rnd) {
    TabView v = make_view(in, W, w, n), o = make_view(out, W, w, n - 1);
    Analysis A;
    uint32_t outcome = project_update(v, (uint32_t)tx << q, (uint32_t)tz << q, rbits, true, A);
    *rnd = A.found;
    if (n > 1) remove_qubit(v, o, q, A, outcome, 0u);
    return outcome;
}
QSB_HD uint32_t tab_measure_word(plane_t* in, plane_t* out, size_t W, size_t w, size_t wg, int n, int q, int tx, int tz,
                                 uint64_t seed, uint64_t traj0, uint64_t draw_id, uint32_t* rnd) {
    TabView v = make_view(in, W, w, n), o = make_view(out, W, w, n - 1);
    Analysis A;
    uint32_t rbits = lane_bits_ge_half(seed, traj0, wg, draw_id);
    uint32_t outcome = project_update(v, (uint32_t)tx << q, (uint32_t)tz << q, rbits, true, A);
    *rnd = A.found;
    if (n > 1) remove_qubit(v, o, q, A, outcome, 0u);
    return outcome;
}

// traceout! on a tableau (a10): the group of ptrace_q(rho) is { g in G : g|_q = I }.
QSB_HD void tab_ptrace_word(plane_t* in, plane_t* out, size_t W, size_t w, int n, int q) {
    TabView v = make_view(in, W, w, n), o = make_view(out, W, w, n - 1);
    Analysis A;
    uint32_t outcome = project_update(v, 0u, 1u << q, 0u, true, A);  // Z_q; the sign is forgotten on random lanes
    uint32_t rnd = A.found;
    if (rnd) {
        // the other stabilizers with Z on q only survive in pairs: keep products with the first one, which is lost
        uint32_t pj[MAXN], fj = 0;
        QSB_UNROLL
        for (int i = 0; i < n; ++i) { pj[i] = A.act[i] & v.ld(v.sz(i, q)) & ~A.piv[i] & rnd & ~fj; fj |= pj[i]; }
        if (fj) {
            uint32_t jx[MAXN], jz[MAXN], js = 0, dax[MAXN], daz[MAXN];
            QSB_UNROLL
            for (int c = 0; c < n; ++c) { jx[c] = 0; jz[c] = 0; dax[c] = 0; daz[c] = 0; }
            QSB_UNROLL
            for (int i = 0; i < n; ++i) {
                if (pj[i] == 0) continue;
                QSB_UNROLL
                for (int c = 0; c < n; ++c) { jx[c] |= v.ld(v.sx(i, c)) & pj[i]; jz[c] |= v.ld(v.sz(i, c)) & pj[i]; }
                js |= v.ld(v.sg(i)) & pj[i];
            }
            QSB_UNROLL
            for (int k = 0; k < n; ++k) {
                uint32_t mk = A.act[k] & v.ld(v.sz(k, q)) & ~A.piv[k] & ~pj[k] & fj;
                if (mk == 0) continue;
                QSB_UNROLL
                for (int c = 0; c < n; ++c) { dax[c] ^= v.ld(v.dx(k, c)) & mk; daz[c] ^= v.ld(v.dz(k, c)) & mk; }
                row_mul(v, v.sx(k, 0), v.sz(k, 0), v.sg(k), mk, jx, jz, js);
            }
            QSB_UNROLL
            for (int i = 0; i < n; ++i) {
                if (pj[i] == 0) continue;
                QSB_UNROLL
                for (int c = 0; c < n; ++c) { v.st(v.dx(i, c), v.ld(v.dx(i, c)) ^ (dax[c] & pj[i])); v.st(v.dz(i, c), v.ld(v.dz(i, c)) ^ (daz[c] & pj[i])); }
                v.st(v.ac(i), v.ld(v.ac(i)) & ~pj[i]);
            }
        }
    }
    remove_qubit(v, o, q, A, outcome, rnd);
}

// force a (multi-qubit) Pauli with a given sign into the group: used to build a state from generators
QSB_HD void tab_force_word(plane_t* p, size_t W, size_t w, int n, uint32_t px, uint32_t pz, int sign) {
    TabView v = make_view(p, W, w, n);
    Analysis A;
    project_update(v, px, pz, sign ? QSB_FULL : 0u, false, A);
}

struct TabObsArgs { int n, k, nelem; uint32_t px[64], pz[64]; uint8_t sign[64]; };
// Each element h of the observable's Pauli group contributes <h> in {-1,0,+1}; a single Pauli is the
// 1-element case.  value = sum_h <h> / nelem (for a projector the signs are a character of the subgroup
// with non-zero expectation, so the sum is |H| or 0).  Bit-sliced tallies of the +1 and -1 terms.
QSB_HD void tab_expect_word_n(plane_t* p, size_t W, size_t w, int n, const TabObsArgs& a, uint32_t* cnt, uint32_t* cntm);
QSB_HD void tab_expect_word(plane_t* p, size_t W, size_t w, const TabObsArgs& a, uint32_t* cnt, uint32_t* cntm) {
    tab_expect_word_n(p, W, w, a.n, a, cnt, cntm);
}
QSB_HD void tab_expect_word_n(plane_t* p, size_t W, size_t w, int n, const TabObsArgs& a, uint32_t* cnt, uint32_t* cntm) {
    TabView v = make_view(p, W, w, n);
    for (int b = 0; b < 8; ++b) { cnt[b] = 0; cntm[b] = 0; }
    Analysis A;
    uint32_t pivD[MAXN];
    for (int e = 0; e < a.nelem; ++e) {
        uint32_t nonzero, neg;
        if ((a.px[e] | a.pz[e]) == 0) { nonzero = QSB_FULL; neg = a.sign[e] ? QSB_FULL : 0u; }
        else {
            analyze(v, a.px[e], a.pz[e], A, pivD);
            nonzero = ~A.found & QSB_FULL;
            neg = determined_sign(v, A, nonzero) ^ (a.sign[e] ? nonzero : 0u);
        }
        uint32_t carry = nonzero & ~neg;
        for (int b = 0; b < 8 && carry; ++b) { uint32_t t = cnt[b] & carry; cnt[b] ^= carry; carry = t; }
        carry = nonzero & neg;
        for (int b = 0; b < 8 && carry; ++b) { uint32_t t = cntm[b] & carry; cntm[b] ^= carry; carry = t; }
    }
}
// a single Pauli (the common observable): lanes with <P> = +1 / -1; everything the generic routine reads from its argument
// table comes in registers here (the fused executor embeds px, pz and the sign in the op record)
QSB_HD void tab_expect_pauli(plane_t* p, size_t W, size_t w, int n, uint32_t px, uint32_t pz, int sign, uint32_t* plus, uint32_t* minus) {
    TabView v = make_view(p, W, w, n);
    uint32_t nonzero, neg;
    if ((px | pz) == 0) { nonzero = QSB_FULL; neg = sign ? QSB_FULL : 0u; }
    else {
        Analysis A;
        uint32_t pivD[MAXN];
        analyze(v, px, pz, A, pivD);
        nonzero = ~A.found & QSB_FULL;
        neg = determined_sign(v, A, nonzero) ^ (sign ? nonzero : 0u);
    }
    *plus = nonzero & ~neg; *minus = nonzero & neg;
}
}  // namespace QSB_CORE_NS
using namespace QSB_CORE_NS;
