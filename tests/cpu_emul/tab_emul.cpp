// tab_emul.cpp -- host build of the bit-sliced tableau core (csrc/qsb_tab_core.cuh) for CPU-only tests.
// A debugging aid for the lane-masked pivot algorithm: it runs the very same inline functions the CUDA
// kernels wrap, one 32-lane word at a time.  Never used by the product, bench.py or any timing.
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <vector>
#include "../../quantumsavory.jl_b200/csrc/qsb_tab_core.cuh"

struct ETab { int n; size_t W; std::vector<uint32_t> p; };
static int planes(int n) { return n * (4 * n + 2); }

extern "C" {
ETab* et_new(int n, size_t W, int active) {
    ETab* t = new ETab(); t->n = n; t->W = W; t->p.assign((size_t)planes(n) * W + 4, 0);
    for (size_t w = 0; w < W; ++w) tab_init_word(t->p.data(), W, w, n, active);
    return t;
}
void et_free(ETab* t) { delete t; }
int et_n(ETab* t) { return t->n; }
ETab* et_copy(ETab* t) { return new ETab(*t); }
void et_force(ETab* t, uint32_t px, uint32_t pz, int sign) { for (size_t w = 0; w < t->W; ++w) tab_force_word(t->p.data(), t->W, w, t->n, px, pz, sign); }
ETab* et_tensor(ETab* a, ETab* b) {
    ETab* t = new ETab(); t->n = a->n + b->n; t->W = a->W; t->p.assign((size_t)planes(t->n) * t->W + 4, 0);
    for (size_t w = 0; w < t->W; ++w) tab_tensor_word(t->p.data(), a->p.data(), b->p.data(), t->W, w, a->n, b->n);
    return t;
}
int et_gate(ETab* t, const int* idx, int k, const uint8_t* table, const uint32_t* mask) {
    TabGateArgs a; memset(&a, 0, sizeof(a));
    a.n = t->n; a.k = k;
    for (int j = 0; j < k; ++j) a.q[j] = idx[j];
    int nb = 2 * k, full = (1 << nb) - 1, cols[4];
    for (int j = 0; j < nb; ++j) cols[j] = table[1 << j] & full;
    for (int x = 0; x < (1 << nb); ++x) {
        int img = 0;
        for (int j = 0; j < nb; ++j) if ((x >> j) & 1) img ^= cols[j];
        if (img != (table[x] & full)) return 5;
        if ((table[x] >> nb) & 1) a.lut |= 1u << x;
    }
    for (int b = 0; b < nb; ++b) for (int j = 0; j < nb; ++j) if ((cols[j] >> b) & 1) a.outmask[b] |= 1 << j;
    a.anf = tab_lut_to_anf(a.lut, nb);
    for (size_t w = 0; w < t->W; ++w) tab_gate_word(t->p.data(), t->W, w, a, mask ? mask[w] : 0xFFFFFFFFu);
    return 0;
}
void et_noise(ETab* t, int q, int np, const double* cum, const int* op, uint64_t seed, uint64_t traj0, uint64_t draw_id) {
    TabNoiseArgs a; memset(&a, 0, sizeof(a));
    a.n = t->n; a.q = q; a.np = np;
    for (int i = 0; i < np; ++i) { a.cum[i] = cum[i]; a.op[i] = op[i]; }
    for (size_t w = 0; w < t->W; ++w) tab_noise_word(t->p.data(), t->W, w, w, a, seed, traj0, draw_id);
}
void et_measure(ETab* t, int q, int tx, int tz, uint64_t seed, uint64_t traj0, uint64_t draw_id, uint32_t* outcome2, uint32_t* rnd) {
    int n = t->n;
    std::vector<uint32_t> o((size_t)planes(n - 1) * t->W + 4, 0);
    for (size_t w = 0; w < t->W; ++w) outcome2[w] = tab_measure_word(t->p.data(), o.data(), t->W, w, w, n, q, tx, tz, seed, traj0, draw_id, &rnd[w]);
    t->p.swap(o); t->n = n - 1;
}
void et_ptrace(ETab* t, int q) {
    int n = t->n;
    std::vector<uint32_t> o((size_t)planes(n - 1) * t->W + 4, 0);
    for (size_t w = 0; w < t->W; ++w) tab_ptrace_word(t->p.data(), o.data(), t->W, w, n, q);
    t->p.swap(o); t->n = n - 1;
}
// values[lane] = (plus - minus) for each of the 32*W lanes; caller divides by nelem
void et_expect(ETab* t, int nelem, const uint32_t* px, const uint32_t* pz, const uint8_t* sign, int* values) {
    TabObsArgs a; memset(&a, 0, sizeof(a));
    a.n = t->n; a.nelem = nelem;
    for (int e = 0; e < nelem; ++e) { a.px[e] = px[e]; a.pz[e] = pz[e]; a.sign[e] = sign[e]; }
    for (size_t w = 0; w < t->W; ++w) {
        uint32_t cnt[8], cntm[8];
        tab_expect_word(t->p.data(), t->W, w, a, cnt, cntm);
        for (int l = 0; l < 32; ++l) {
            int plus = 0, minus = 0;
            for (int b = 0; b < 8; ++b) { plus |= ((cnt[b] >> l) & 1) << b; minus |= ((cntm[b] >> l) & 1) << b; }
            values[w * 32 + l] = plus - minus;
        }
    }
}
// rows: n D rows then n S rows
void et_lane(ETab* t, size_t lane, uint8_t* xs, uint8_t* zs, uint8_t* signs, uint8_t* active) {
    int n = t->n, S = 4 * n + 2; size_t w = lane / 32; int l = lane % 32;
    auto bit = [&](int plane) { return (uint8_t)((t->p[(size_t)plane * t->W + w] >> l) & 1u); };
    for (int i = 0; i < n; ++i) {
        for (int c = 0; c < n; ++c) {
            xs[i * n + c] = bit(i * S + c); zs[i * n + c] = bit(i * S + n + c);
            xs[(n + i) * n + c] = bit(i * S + 2 * n + c); zs[(n + i) * n + c] = bit(i * S + 3 * n + c);
        }
        signs[i] = 0; signs[n + i] = bit(i * S + 4 * n);
        active[i] = bit(i * S + 4 * n + 1);
    }
}
double et_draw(uint64_t seed, uint64_t traj, uint64_t draw_id) { return qsb_draw(seed, traj, draw_id); }
}
