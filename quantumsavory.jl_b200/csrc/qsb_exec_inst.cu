// qsb_exec_inst.cu -- one instantiation of the fused executor kernel per compilation (-DQSB_EXEC_T=0|1).
#include "qsb_exec_kernel.cuh"
#include <cuda_pipeline.h>
#ifndef QSB_EXEC_T
#error "compile with -DQSB_EXEC_T=1 (arena tile in shared memory) or 0 (arena in global memory)"
#endif
#ifndef QSB_EXEC_G_MINB
#define QSB_EXEC_G_MINB 6   // resident 128-thread blocks per SM the global-arena executor is compiled for (register cap 80)
#endif

// The tableau functions are force-inlined with a LITERAL qubit count so that their loops unroll and their
// per-row scratch arrays live in registers (a run-time n keeps them in local memory: 2.2 KB of stack per thread).
// States of more than 4 qubits take a NOINLINE fallback with a run-time n (its arrays live in local memory); keeping it out of
// line lets the compiler allocate registers for the literal cases alone (inline it cost 2.2 KB of stack frame and capped the
// kernel at 64 registers with spills).
#define N_CASE(K, ...) case K: { constexpr int N = K; __VA_ARGS__; } break;
// Literal cases stop at 4 qubits (the largest state of the BASELINE workloads): the kernel's code size grows with the square of the
// literal n, and with 1-2 warps per scheduler instruction-fetch stalls are not hidden (ncu: 4.8 no-instruction stalls per issue on the
// one-warp blocks of tree level 3, 13 on the global-arena executor, with a fifth literal case that no workload used).
// (The global-arena instantiation keeps its fifth case: without it ptxas spills at the 80-register cap and the grid ran 13 % slower.)
#if QSB_EXEC_T
#define N_CASE5(...)
#else
#define N_CASE5(...) N_CASE(5, __VA_ARGS__)
#endif
#define N_SWITCH(NV, FALLBACK, ...) switch (NV) { N_CASE(1, __VA_ARGS__) N_CASE(2, __VA_ARGS__) N_CASE(3, __VA_ARGS__) N_CASE(4, __VA_ARGS__) \
                                        N_CASE5(__VA_ARGS__) default: { FALLBACK; } break; }
static __device__ __noinline__ void big_init(plane_t* p, size_t W, int n, int active) { tab_init_word(p, W, 0, n, active); }
static __device__ __noinline__ void big_gate(plane_t* p, size_t W, int n, int k, int q0, int q1, const uint8_t* om, uint32_t anf, uint32_t m) { tab_gate_anf(p, W, 0, n, k, q0, q1, om, anf, m); }
static __device__ __noinline__ void big_noise(plane_t* p, size_t W, int n, int q, uint32_t fx, uint32_t fz) { tab_noise_apply(p, W, 0, n, q, fx, fz); }
static __device__ __noinline__ uint32_t big_measure(plane_t* in, plane_t* out, size_t W, int n, int q, int tx, int tz, uint32_t rb, uint32_t* rnd) { return tab_measure_bits(in, out, W, 0, n, q, tx, tz, rb, rnd); }
static __device__ __noinline__ void big_ptrace(plane_t* in, plane_t* out, size_t W, int n, int q) { tab_ptrace_word(in, out, W, 0, n, q); }
static __device__ __noinline__ void big_force(plane_t* p, size_t W, int n, uint32_t px, uint32_t pz, int sign) { tab_force_word(p, W, 0, n, px, pz, sign); }
static __device__ __noinline__ void big_expect1(plane_t* p, size_t W, int n, uint32_t px, uint32_t pz, int sign, uint32_t* plus, uint32_t* minus) { tab_expect_pauli(p, W, 0, n, px, pz, sign, plus, minus); }
static __device__ __noinline__ void big_expect(plane_t* p, size_t W, int n, const TabObsArgs& a, uint32_t* cnt, uint32_t* cntm) { tab_expect_word_n(p, W, 0, n, a, cnt, cntm); }

// T == 1: the arena is a [blockDim.x][tile_stride] tile of shared memory.  The block size is a RUN-TIME value (any thread count
//         up to 256, partial last warp allowed) so that the host can size the blocks to cover the ensemble in whole waves
//         (qsb_exec.cu: a 9 % tail wave used to double the run time).
// T == 0: the arena is in global memory with stride `words` (programs whose tile exceeds the shared memory of an SM).
// One thread owns one 32-trajectory word column of every plane of its job; nothing is exchanged between threads, so
// consecutive ops need no barrier.  `live` holds the lanes that still belong to the job's cohort (TOP_RETIRE).
template <int T>
__global__ void __launch_bounds__(T == 1 ? 256 : 128, T == 1 ? 1 : QSB_EXEC_G_MINB) k_tab_exec(WaveArgs a) {
    extern __shared__ plane_t tile[];
    __shared__ ExecJob sjob;
    __shared__ __align__(16) ExecOp sops[2][QSB_EXEC_CHUNK];   // double buffered: chunk c + 1 arrives (cp.async) while chunk c executes
    if (threadIdx.x == 0) {   // the job of this block: largest block0_exec <= blockIdx.x
        int lo = 0, hi = a.njobs - 1;
        while (lo < hi) { const int mid = (lo + hi + 1) >> 1; if (a.jobs[mid].block0_exec <= blockIdx.x) lo = mid; else hi = mid - 1; }
        sjob = a.jobs[lo];
    }
    __syncthreads();
    const ExecOp* const ops = sjob.ops;
    const int nops = sjob.nops;
    const uint32_t* const lanes = sjob.lanes;
    const size_t wg = (blockIdx.x - sjob.block0_exec) * (size_t)blockDim.x + threadIdx.x;   // word of the job (trajectories 32*wg ..)
    const bool on = wg < sjob.words;
    // T == 1: the tile is [thread][tile_stride] -- a thread's planes are CONTIGUOUS (plane k at col[k]), so plane offsets fold
    // into the immediate field of LDS/STS, distinct planes are provably distinct addresses (loads may be hoisted above stores:
    // instruction-level parallelism inside an op body) and no address multiply is needed; tile_stride is odd, so the 32
    // threads of a warp hit 32 different banks.  T == 0: plane-major global arena, plane k at col[k * words] (a warp reads one
    // 128-byte line per plane; the [word][plane] layout was measured 2.8x slower there: 32 lines per request thrash the L1).
    plane_t* const col = T ? tile + (size_t)threadIdx.x * (size_t)a.tile_stride : sjob.arena + (on ? wg : 0);
    const size_t W = T ? (size_t)1 : (size_t)sjob.words;
    const size_t w = 0;
    const uint32_t left = on ? sjob.count - (uint32_t)wg * 32u : 0u;
    const uint32_t validw = left >= 32u ? 0xffffffffu : ((1u << left) - 1u);
    uint32_t live = validw;
#define PL(idx) (col + (size_t)(idx) * W)
#define LANE_ID(l) (lanes ? __ldg(lanes + wg * 32 + (l)) : (uint32_t)(wg * 32 + (l)))
    // decision words: bits[slot][wstride], this thread's column; a warp reads one 128-byte line per slot.  The three words of
    // the NEXT stochastic op are fetched while the current one executes (ExecOp::next_dslot), so their latency is off the
    // critical path; DBITS(j) is word j of the current stochastic op.
    const uint32_t* const bcol = sjob.bits + (on ? wg : 0);
    const size_t bstride = sjob.wstride;
    uint32_t pre0 = 0, pre1 = 0, pre2 = 0, cur0 = 0, cur1 = 0, cur2 = 0;
#define FETCH(slot) do { const uint32_t* q_ = bcol + (size_t)(slot) * bstride; pre0 = __ldg(q_); pre1 = __ldg(q_ + bstride); pre2 = __ldg(q_ + 2 * bstride); } while (0)
#define ADVANCE(op) do { cur0 = pre0; cur1 = pre1; cur2 = pre2; if ((op).next_dslot >= 0) FETCH((op).next_dslot); } while (0)
    if (sjob.first_dslot >= 0) FETCH(sjob.first_dslot);
    for (int i = 0; i < nops; ++i) {
        if ((i % QSB_EXEC_CHUNK) == 0) {   // chunk boundary: the chunk was requested one chunk ago (the first one just below); request the next
            const int c = i / QSB_EXEC_CHUNK;
            auto request = [&](int first, int buf) {
                const int nchunk = min(QSB_EXEC_CHUNK, nops - first);
                const uint4* src = reinterpret_cast<const uint4*>(ops + first);
                uint4* dst = reinterpret_cast<uint4*>(sops[buf]);
                for (int k = threadIdx.x; k < nchunk * (int)(sizeof(ExecOp) / 16); k += blockDim.x) __pipeline_memcpy_async(dst + k, src + k, 16);
                __pipeline_commit();
            };
            if (i == 0) request(0, 0);
            __pipeline_wait_prior(0);
            __syncthreads();               // chunk c is complete for every thread, and nobody still reads the buffer chunk c + 1 goes to
            if (i + QSB_EXEC_CHUNK < nops) request(i + QSB_EXEC_CHUNK, (c + 1) & 1);
        }
        const ExecOp& op = sops[(i / QSB_EXEC_CHUNK) & 1][i % QSB_EXEC_CHUNK];
        switch (op.type) {
        case TOP_INIT:
            if (on) { N_SWITCH(op.n, big_init(PL(op.s[0]), W, op.n, op.iv[0]), tab_init_word(PL(op.s[0]), W, w, N, op.iv[0])) PL(op.al[0])[w] = validw; }
            break;
        case TOP_TENSOR:
            if (on) {
                if (op.n == 2 && op.n2 == 2) tab_tensor_word(PL(op.s[0]), PL(op.s[1]), PL(op.s[2]), W, w, 2, 2);
                else tab_tensor_word(PL(op.s[0]), PL(op.s[1]), PL(op.s[2]), W, w, op.n, op.n2);
                PL(op.al[0])[w] = PL(op.al[1])[w] & PL(op.al[2])[w]; }
            break;
        case TOP_COPY:
            if (on) {
                int planes = op.n * (4 * op.n + 2);
                for (int p = 0; p < planes; ++p) PL(op.s[0] + p)[w] = PL(op.s[1] + p)[w];
                PL(op.al[0])[w] = PL(op.al[1])[w];
            }
            break;
        case TOP_SELECT:
            if (on) {
                uint32_t m[5];
                ADVANCE(op);
                const uint32_t b0 = cur0, b1 = cur1, b2 = cur2;   // bit planes of the selected index
                for (int j = 0; j < 5; ++j) m[j] = ((j & 1) ? b0 : ~b0) & ((j & 2) ? b1 : ~b1) & ((j & 4) ? b2 : ~b2);
                int planes = op.n * (4 * op.n + 2);
                for (int p = 0; p < planes; ++p) {
                    uint32_t v = 0;
                    for (int j = 0; j < op.n2; ++j) v |= m[j] & PL(op.s[1 + j] + p)[w];
                    PL(op.s[0] + p)[w] = v;
                }
                uint32_t al = 0;
                for (int j = 0; j < op.n2; ++j) al |= m[j] & PL(op.al[1 + j])[w];
                PL(op.al[0])[w] = al;
            }
            break;
        case TOP_SELECT_CONST:
            if (on) {
                uint32_t m[5];
                ADVANCE(op);
                const uint32_t b0 = cur0, b1 = cur1, b2 = cur2;
                for (int j = 0; j < 5; ++j) m[j] = ((j & 1) ? b0 : ~b0) & ((j & 2) ? b1 : ~b1) & ((j & 4) ? b2 : ~b2);
                const uint8_t* tab = sjob.aux + op.obs;
                const int planes = op.n * (4 * op.n + 2);
                for (int p = 0; p < planes; ++p) {
                    const uint32_t c = tab[p];
                    uint32_t v = 0;
                    for (int j = 0; j < 5; ++j) v |= ((c >> j) & 1) ? m[j] : 0u;
                    PL(op.s[0] + p)[w] = (plane_t)v;
                }
                PL(op.al[0])[w] = (plane_t)validw;
            }
            break;
        case TOP_GATE:
            if (on) {
                const uint32_t gm = op.m[0] >= 0 ? (uint32_t)PL(op.m[0])[w] : QSB_FULL;
                N_SWITCH(op.n, big_gate(PL(op.s[0]), W, op.n, op.k, op.q[0], op.q[1], op.outmask, op.u[0], gm), tab_gate_anf(PL(op.s[0]), W, w, N, op.k, op.q[0], op.q[1], op.outmask, op.u[0], gm))
            }
            break;
        case TOP_NOISE:
            if (on) {
                ADVANCE(op);
                const uint32_t fx = cur0, fz = cur1;
                N_SWITCH(op.n, big_noise(PL(op.s[0]), W, op.n, op.q[0], fx, fz), tab_noise_apply(PL(op.s[0]), W, w, N, op.q[0], fx, fz))
            }
            break;
        case TOP_MEASURE:
            if (on) {
                uint32_t rnd;
                ADVANCE(op);
                uint32_t outcome = 0; const uint32_t rb = cur0;
                N_SWITCH(op.n, outcome = big_measure(PL(op.s[1]), PL(op.s[0]), W, op.n, op.q[0], op.iv[0], op.iv[1], rb, &rnd), outcome = tab_measure_bits(PL(op.s[1]), PL(op.s[0]), W, w, N, op.q[0], op.iv[0], op.iv[1], rb, &rnd))
                uint32_t al = PL(op.al[1])[w];
                PL(op.m[0])[w] = outcome & al;
                if (op.m[1] >= 0) PL(op.m[1])[w] = rnd & al;
            }
            break;
        case TOP_PTRACE:
            if (on) { N_SWITCH(op.n, big_ptrace(PL(op.s[1]), PL(op.s[0]), W, op.n, op.q[0]), tab_ptrace_word(PL(op.s[1]), PL(op.s[0]), W, w, N, op.q[0])) }
            break;
        case TOP_FORCE:
            if (on) { N_SWITCH(op.n, big_force(PL(op.s[0]), W, op.n, op.u[0], op.u[1], op.iv[0]), tab_force_word(PL(op.s[0]), W, w, N, op.u[0], op.u[1], op.iv[0])) }
            break;
        case TOP_EXPECT: {
            // integer tallies (exact, order independent): alive lanes, sum of (plus - minus), sum of its square
            int n_alive = 0, num = 0, num2 = 0;
            int nelem = 1;
            if (on) {
                uint32_t cnt[8], cntm[8];
                if (op.k == 1) {     // a single Pauli: px, pz and the sign travel in the op record (no loads from the argument table)
#pragma unroll
                    for (int b = 1; b < 8; ++b) { cnt[b] = 0; cntm[b] = 0; }
                    N_SWITCH(op.n, big_expect1(PL(op.s[0]), W, op.n, op.u[0], op.u[1], op.iv[0], &cnt[0], &cntm[0]),
                             tab_expect_pauli(PL(op.s[0]), W, w, N, op.u[0], op.u[1], op.iv[0], &cnt[0], &cntm[0]))
                } else {
                    const TabObsArgs& o = sjob.obs[op.obs];
                    nelem = o.nelem;
                    N_SWITCH(op.n, big_expect(PL(op.s[0]), W, op.n, o, cnt, cntm), tab_expect_word_n(PL(op.s[0]), W, w, N, o, cnt, cntm))
                }
                const uint32_t al = PL(op.al[0])[w] & live;   // retired lanes left this cohort: they are tallied where they continue
                n_alive = __popc(al);
                if (nelem == 1 && !op.per_traj) {   // a single Pauli: each lane holds -1, 0 or +1
                    num = __popc(cnt[0] & al) - __popc(cntm[0] & al);
                    num2 = __popc((cnt[0] | cntm[0]) & al);
                } else {
                    const double inv = 1.0 / (double)nelem;
                    for (int l = 0; l < 32; ++l) {
                        int plus = 0, minus = 0;
                        for (int b = 0; b < 8; ++b) { plus |= ((cnt[b] >> l) & 1) << b; minus |= ((cntm[b] >> l) & 1) << b; }
                        const bool lane = (al >> l) & 1u;
                        const int d = lane ? plus - minus : 0;
                        if (op.per_traj && ((live >> l) & 1u)) { const size_t t = LANE_ID(l); op.per_traj[2 * t] = (double)d * inv; op.per_traj[2 * t + 1] = 0.0; }
                        num += d; num2 += d * d;
                    }
                }
            }
            if (op.stats) {  // whole warps reach this point together (the op list is uniform)
                // the last warp of a run-time sized block may be partial: the member mask follows from the block geometry
                // (__activemask() would only report who happens to be converged here, and a split warp would drop tallies)
                const unsigned wbase = threadIdx.x & ~31u, wthreads = blockDim.x - wbase;
                const unsigned am = wthreads >= 32u ? 0xffffffffu : ((1u << wthreads) - 1u);
                n_alive = __reduce_add_sync(am, n_alive);
                num = __reduce_add_sync(am, num);
                num2 = __reduce_add_sync(am, num2);
                nelem = __reduce_max_sync(am, nelem);
                if ((threadIdx.x & 31) == 0 && n_alive) {
                    const double inv = 1.0 / (double)nelem;   // nelem = 2^k: the scaled tallies are exact
                    atomicAdd(&op.stats[0], (double)n_alive); atomicAdd(&op.stats[1], (double)num * inv); atomicAdd(&op.stats[2], (double)num2 * inv * inv);
                }
            }
            break;
        }
        case TOP_MASK_FILL:
            if (on) PL(op.m[0])[w] = op.iv[0] ? validw : 0u;
            break;
        case TOP_MASK_LOGIC:
            if (on) {
                uint32_t x = PL(op.m[1])[w], y = op.m[2] >= 0 ? PL(op.m[2])[w] : 0u, r;
                switch (op.iv[0]) {
                    case QSB_MASK_AND: r = x & y; break;
                    case QSB_MASK_OR: r = x | y; break;
                    case QSB_MASK_XOR: r = x ^ y; break;
                    case QSB_MASK_ANDNOT: r = x & ~y; break;
                    case QSB_MASK_NOT: r = ~x; break;
                    default: r = ~(x ^ y); break;
                }
                PL(op.m[0])[w] = (plane_t)(r & validw);
            }
            break;
        case TOP_MASK_RANDOM:
            if (on) {
                ADVANCE(op);
                PL(op.m[0])[w] = (plane_t)(cur0 & validw);
            }
            break;
        case TOP_KILL:
            if (on) PL(op.al[0])[w] = (plane_t)(PL(op.al[0])[w] & ~PL(op.m[0])[w]);
            break;
        case TOP_ALIVE_COPY:
            if (on) PL(op.m[0])[w] = PL(op.al[0])[w];
            break;
        case TOP_RETIRE:
            // cohort forking: the lanes of the mask leave this cohort (their classical history differs from here on); each
            // is marked with the op's code so that the host can route it to the program of its own history.  A retired lane
            // keeps computing (its column is never read by anyone else) but no longer counts in any statistic.
            if (on) {
                const uint32_t r = (uint32_t)PL(op.m[0])[w] & live;
                live &= ~r;
                int32_t* const codes = reinterpret_cast<int32_t*>(op.per_traj);
                if (codes) for (uint32_t t = r; t; t &= t - 1) codes[LANE_ID(__ffs((int)t) - 1)] = (int32_t)op.u[0];
            }
            break;
        }
    }
#undef PL
#undef FETCH
#undef ADVANCE
#undef LANE_ID
}

#define QSB_CAT2(a, b) a##b
#define QSB_CAT(a, b) QSB_CAT2(a, b)
cudaError_t QSB_CAT(qsb_exec_launch_, QSB_EXEC_T)(const WaveArgs& a, unsigned blocks, int threads, int smem_bytes, cudaStream_t stream) {
    // the opt-in is per device (and per function): remember it for the device that is current for this launch
    static int configured[64];
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev < 0 || dev >= 64 || smem_bytes > configured[dev]) {
        cudaError_t e = cudaFuncSetAttribute(k_tab_exec<QSB_EXEC_T>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes);
        if (e != cudaSuccess) return e;
        if (dev >= 0 && dev < 64) configured[dev] = smem_bytes;
    }
    k_tab_exec<QSB_EXEC_T><<<blocks, threads, smem_bytes, stream>>>(a);
    return cudaGetLastError();
}
