// qsb_bitsliced.cpp -- the OPTIMISED CPU baseline of SURVEY 8d / BASELINE.md section 2 item 3 (TEST / BENCH INFRASTRUCTURE ONLY).
//
// Where oracle/qsb_replay.c restates the reference's per-trajectory algorithm, this file is the fair "what a good CPU
// implementation of the SAME batch-major design achieves" number: the recorded op list of a tableau program
// (qsb_program_export: the very records the GPU executor reads) is replayed on the host with the bit-sliced tableau core
// (csrc/qsb_tab_core.cuh compiled by g++, 32 trajectories per word) -- one word column per task, `#pragma omp parallel for`
// over words, Philox draws evaluated per lane exactly as the per-op kernels compare them.  Statistics rows are integer
// tallies, so they must equal the GPU's bit for bit: bench.py and tests/test_gpu_bitsliced_cpu.py check that.
// Nothing under quantumsavory.jl_b200/ links or loads this file.
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <vector>
#ifdef _OPENMP
#include <omp.h>
#endif
#include "../quantumsavory.jl_b200/csrc/qsb_exec_kernel.cuh"   // ExecOp, DrawSpec, TabObsArgs, TOP_* (host-only parts)

namespace {
struct Meta { int32_t planes, nslots, nops, nspecs, nobs, naux, nstats, first_dslot; };

// decision words of one stochastic op for the 32 lanes of word `wg` (the draw pass of the GPU, restated per lane)
inline void decide(const DrawSpec& s, uint64_t seed, uint64_t traj0, size_t wg, uint64_t draw_base, uint32_t out[3]) {
    out[0] = out[1] = out[2] = 0;
    const uint64_t id = s.draw + draw_base;
    for (int l = 0; l < 32; ++l) {
        uint32_t blk[4];
        qsb_philox_block(seed, traj0 + (uint64_t)(wg * 32 + l), id >> 1, blk);
        const uint64_t k = ((id & 1) ? (((uint64_t)blk[3] << 32) | blk[2]) : (((uint64_t)blk[1] << 32) | blk[0])) >> 11;   // u = k / 2^53
        if (s.type == TOP_NOISE) {
            uint32_t d = 0;
            for (int j = s.np - 1; j >= 0; --j) if (k > s.thr[j]) d = s.code[j];
            out[0] |= (d & 1u) << l; out[1] |= ((d >> 1) & 1u) << l;
        } else if (s.type == TOP_SELECT) {
            uint32_t d = (uint32_t)(s.np - 1);
            for (int j = s.np - 2; j >= 0; --j) if (k < s.thr[j]) d = (uint32_t)j;
            out[0] |= (d & 1u) << l; out[1] |= ((d >> 1) & 1u) << l; out[2] |= ((d >> 2) & 1u) << l;
        } else {
            const bool below = k < s.thr[0];
            out[0] |= (uint32_t)(s.type == TOP_MEASURE ? !below : below) << l;
        }
    }
}
}  // namespace

// stats_out: nstats rows of {n_alive, sum, sum of squares, 0} (row order = the program's statistics row order)
extern "C" int bs_run(const void* ops_v, const void* specs_v, const void* obs_v, const uint8_t* aux, const uint64_t* stats_ptrs, const int32_t* meta_v,
                      uint64_t seed, uint64_t traj0, uint64_t draw_base, int64_t ntraj, double* stats_out, int* threads_out) {
    const ExecOp* ops = (const ExecOp*)ops_v;
    const DrawSpec* specs = (const DrawSpec*)specs_v;
    const TabObsArgs* obs = (const TabObsArgs*)obs_v;
    const Meta& M = *(const Meta*)meta_v;
    const int64_t words = (ntraj + 31) / 32;
    // spec of every stochastic op, by its first decision slot
    std::vector<int> spec_of_slot((size_t)M.nslots + 4, -1);
    for (int i = 0; i < M.nspecs; ++i) spec_of_slot[specs[i].slot] = i;
    // statistics row of an op = index of its pointer in the program's row list
    std::vector<int> row_of_op(M.nops, -1);
    for (int i = 0; i < M.nops; ++i) if (ops[i].type == TOP_EXPECT && ops[i].stats)
        for (int r = 0; r < M.nstats; ++r) if (stats_ptrs[r] == (uint64_t)(uintptr_t)ops[i].stats) { row_of_op[i] = r; break; }
    std::vector<long long> tally((size_t)M.nstats * 3, 0);
    std::vector<int> nelem_row(M.nstats, 1);
    int nthreads = 1;
#pragma omp parallel
    {
#ifdef _OPENMP
#pragma omp single
        nthreads = omp_get_num_threads();
#endif
        std::vector<uint32_t> col((size_t)M.planes + 8);
        std::vector<long long> mine((size_t)M.nstats * 3, 0);
#pragma omp for schedule(dynamic, 16)
        for (int64_t wg = 0; wg < words; ++wg) {
            uint32_t* const c = col.data();
            const int64_t left = ntraj - wg * 32;
            const uint32_t validw = left >= 32 ? 0xffffffffu : ((1u << left) - 1u);
            uint32_t live = validw;
#define PL(idx) (c + (idx))
            for (int i = 0; i < M.nops; ++i) {
                const ExecOp& op = ops[i];
                uint32_t d[3] = {0, 0, 0};
                if (op.type == TOP_SELECT || op.type == TOP_SELECT_CONST || op.type == TOP_NOISE || op.type == TOP_MEASURE || op.type == TOP_MASK_RANDOM)
                    decide(specs[spec_of_slot[op.dslot]], seed, traj0, (size_t)wg, draw_base, d);
                switch (op.type) {
                case TOP_INIT: tab_init_word(PL(op.s[0]), 1, 0, op.n, op.iv[0]); *PL(op.al[0]) = validw; break;
                case TOP_TENSOR: tab_tensor_word(PL(op.s[0]), PL(op.s[1]), PL(op.s[2]), 1, 0, op.n, op.n2); *PL(op.al[0]) = *PL(op.al[1]) & *PL(op.al[2]); break;
                case TOP_COPY: { const int planes = op.n * (4 * op.n + 2); for (int p = 0; p < planes; ++p) PL(op.s[0] + p)[0] = PL(op.s[1] + p)[0]; *PL(op.al[0]) = *PL(op.al[1]); break; }
                case TOP_SELECT: case TOP_SELECT_CONST: {
                    uint32_t m[8];
                    for (int j = 0; j < 8; ++j) m[j] = ((j & 1) ? d[0] : ~d[0]) & ((j & 2) ? d[1] : ~d[1]) & ((j & 4) ? d[2] : ~d[2]);
                    const int planes = op.n * (4 * op.n + 2);
                    if (op.type == TOP_SELECT) {
                        for (int p = 0; p < planes; ++p) { uint32_t v = 0; for (int j = 0; j < op.n2; ++j) v |= m[j] & PL(op.s[1 + j] + p)[0]; PL(op.s[0] + p)[0] = v; }
                        uint32_t al = 0; for (int j = 0; j < op.n2; ++j) al |= m[j] & *PL(op.al[1 + j]);
                        *PL(op.al[0]) = al;
                    } else {
                        const uint8_t* tab = aux + op.obs;
                        for (int p = 0; p < planes; ++p) { uint32_t v = 0; for (int j = 0; j < 5; ++j) v |= ((tab[p] >> j) & 1) ? m[j] : 0u; PL(op.s[0] + p)[0] = v; }
                        *PL(op.al[0]) = validw;
                    }
                    break;
                }
                case TOP_GATE: { const uint32_t gm = op.m[0] >= 0 ? *PL(op.m[0]) : 0xffffffffu; tab_gate_anf(PL(op.s[0]), 1, 0, op.n, op.k, op.q[0], op.q[1], op.outmask, op.u[0], gm); break; }
                case TOP_NOISE: tab_noise_apply(PL(op.s[0]), 1, 0, op.n, op.q[0], d[0], d[1]); break;
                case TOP_MEASURE: {
                    uint32_t rnd;
                    const uint32_t outcome = tab_measure_bits(PL(op.s[1]), PL(op.s[0]), 1, 0, op.n, op.q[0], op.iv[0], op.iv[1], d[0], &rnd);
                    const uint32_t al = *PL(op.al[1]);
                    *PL(op.m[0]) = outcome & al;
                    if (op.m[1] >= 0) *PL(op.m[1]) = rnd & al;
                    break;
                }
                case TOP_PTRACE: tab_ptrace_word(PL(op.s[1]), PL(op.s[0]), 1, 0, op.n, op.q[0]); break;
                case TOP_FORCE: tab_force_word(PL(op.s[0]), 1, 0, op.n, op.u[0], op.u[1], op.iv[0]); break;
                case TOP_EXPECT: {
                    uint32_t cnt[8], cntm[8];
                    int nelem = 1;
                    if (op.k == 1) { for (int b = 1; b < 8; ++b) { cnt[b] = 0; cntm[b] = 0; } tab_expect_pauli(PL(op.s[0]), 1, 0, op.n, op.u[0], op.u[1], op.iv[0], &cnt[0], &cntm[0]); }
                    else { const TabObsArgs& o = obs[op.obs]; nelem = o.nelem; tab_expect_word_n(PL(op.s[0]), 1, 0, op.n, o, cnt, cntm); }
                    const uint32_t al = *PL(op.al[0]) & live;
                    long long num = 0, num2 = 0;
                    for (int l = 0; l < 32; ++l) {
                        if (!((al >> l) & 1u)) continue;
                        int plus = 0, minus = 0;
                        for (int b = 0; b < 8; ++b) { plus |= ((cnt[b] >> l) & 1) << b; minus |= ((cntm[b] >> l) & 1) << b; }
                        const int dd = plus - minus;
                        num += dd; num2 += (long long)dd * dd;
                    }
                    const int r = row_of_op[i];
                    if (r >= 0) { mine[3 * r] += __builtin_popcount(al); mine[3 * r + 1] += num; mine[3 * r + 2] += num2; nelem_row[r] = nelem; }
                    break;
                }
                case TOP_MASK_FILL: *PL(op.m[0]) = op.iv[0] ? validw : 0u; break;
                case TOP_MASK_LOGIC: {
                    const uint32_t x = *PL(op.m[1]), y = op.m[2] >= 0 ? *PL(op.m[2]) : 0u;
                    uint32_t r;
                    switch (op.iv[0]) {
                        case QSB_MASK_AND: r = x & y; break; case QSB_MASK_OR: r = x | y; break; case QSB_MASK_XOR: r = x ^ y; break;
                        case QSB_MASK_ANDNOT: r = x & ~y; break; case QSB_MASK_NOT: r = ~x; break; default: r = ~(x ^ y); break;
                    }
                    *PL(op.m[0]) = r & validw;
                    break;
                }
                case TOP_MASK_RANDOM: *PL(op.m[0]) = d[0] & validw; break;
                case TOP_KILL: *PL(op.al[0]) &= ~*PL(op.m[0]); break;
                case TOP_ALIVE_COPY: *PL(op.m[0]) = *PL(op.al[0]); break;
                case TOP_RETIRE: live &= ~(*PL(op.m[0]) & live); break;
                default: break;
                }
            }
#undef PL
        }
#pragma omp critical
        for (size_t q = 0; q < mine.size(); ++q) tally[q] += mine[q];
    }
    for (int r = 0; r < M.nstats; ++r) {
        const double inv = 1.0 / (double)nelem_row[r];
        stats_out[4 * r] = (double)tally[3 * r]; stats_out[4 * r + 1] = (double)tally[3 * r + 1] * inv;
        stats_out[4 * r + 2] = (double)tally[3 * r + 2] * inv * inv; stats_out[4 * r + 3] = 0.0;
    }
    if (threads_out) *threads_out = nthreads;
    return 0;
}
