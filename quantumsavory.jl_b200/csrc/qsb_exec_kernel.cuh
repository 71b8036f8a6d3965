// qsb_exec_kernel.cuh -- the fused executor kernel of recorded tableau programs (see qsb_exec.cu).
// Included by qsb_exec.cu (host side: ExecJob / WaveArgs) and by qsb_exec_inst.cu, which is compiled once per
// arena placement (-DQSB_EXEC_T=1 shared memory, 0 global memory) so that the instantiations build in parallel.
#pragma once
#define QSB_PLANE_T uint32_t    // measured on B200 (round 1): 32-lane words 5.21 ms, 16-lane 5.22 ms, 8-lane bytes 5.55 ms per 10^6 trajectories
#define QSB_CORE_NS tabcorex
#include "qsb_internal.cuh"
#include "qsb_tab_core.cuh"

// The op as the kernel reads it: TabOp without the host-only fields (mixture weights, draw ids), 80 bytes.
// Ops are staged through shared memory in chunks, so a thread decodes them with broadcast LDS instead of a
// chain of dependent global loads per op.
struct ExecOp {
    uint8_t type, n, n2, k, q[2], iv[2];
    uint8_t outmask[4];
    uint32_t u[2];
    int16_t s[6], al[6], m[3];
    int16_t pad;
    int32_t obs, dslot;
    int32_t next_dslot;       // first decision slot of the NEXT stochastic op (-1: none): its words are fetched one stochastic op ahead
    double* stats;
    double* per_traj;         // TOP_EXPECT: per-trajectory row; TOP_RETIRE: int32 code row (one entry per lane of the context)
};
#define QSB_EXEC_CHUNK 32   // ops per staged chunk, two buffers (64-op chunks save barriers at level 0 but their 10 KB per block cost a wave at the small-block tree levels)

// One (program, lane subset) pair of a launch.  A launch ("wave") executes any number of jobs: block b belongs to the job
// with the largest block0 <= b.  The full-ensemble replay is the one-job case with lanes == nullptr.
struct ExecJob {
    const ExecOp* ops;
    const TabObsArgs* obs;
    const uint8_t* aux;
    const void* specs;        // DrawSpec[nspecs] (qsb_exec.cu)
    const void* items;        // DrawItem[nitems]: the specs grouped by Philox block (two draws per block), the draw pass's work list
    int nitems, pad_items;
    const uint32_t* lanes;    // lane index (trajectory slot of the context) of each of the job's `count` trajectories; nullptr: identity
    uint32_t* bits;           // decision words of every stochastic op, [slot][wstride] (a word = 32 trajectories of the job): slot-major,
                              // so the 32 threads of an executor warp read one 128-byte line per slot
    uint32_t* arena;          // global-memory arena [planes][words] (programs whose tile does not fit in shared memory), else nullptr
    int nops, nspecs, nslots, planes;
    uint32_t wstride;         // words per slot row (job words rounded up to 8: rows start on a 32-byte sector)
    int32_t first_dslot;      // decision slot of the first stochastic op (-1: none)
    uint32_t count, words;    // trajectories of the job, ceil(count / 32)
    uint32_t block0_exec, block0_draw;   // first block of the job in the executor grid / in the draw grid
};
struct WaveArgs {
    const ExecJob* jobs; int njobs;
    uint64_t seed, traj0, draw_base;   // global trajectory id of lane index 0, offset of every draw id
    uint32_t rk[20];                   // Philox round keys (k0, k1 of rounds 0..9) of `seed`: constant-bank operands of the draw pass
    int tile_stride;                   // shared-memory executor: words per thread of the [thread][plane] tile (odd: conflict-free banks)
};

// one stochastic op of the program as the draw pass sees it
struct DrawSpec {
    uint64_t draw;
    int type, np, slot;        // type: TOP_NOISE / TOP_SELECT / TOP_MEASURE / TOP_MASK_RANDOM
    uint32_t thr_hi_any;       // noise: high word of ((thr_any + 1) << 11) -- hi32(draw bits) < thr_hi_any on every lane proves "no error in this word"
    // the comparisons of the per-op kernels (u = k / 2^53 against a double c) restated EXACTLY on the 53-bit integer k:
    //   u > c  <=>  k > floor(c * 2^53)        u < c  <=>  k < ceil(c * 2^53)        (c * 2^53 is exact in a double)
    uint64_t thr[5];
    uint64_t thr_any;          // noise: the lowest threshold -- k <= thr_any on every lane means "no error anywhere in this word"
    uint8_t code[4];           // noise: 2-bit decision (bit 0: X component, bit 1: Z component) of branch j
    int pad2;
};

// launchers: [0] arena in global memory (one job per launch), [1] shared-memory tile with a run-time block size
typedef cudaError_t (*qsb_exec_launch_fn)(const WaveArgs&, unsigned blocks, int threads, int smem_bytes, cudaStream_t stream);
qsb_exec_launch_fn qsb_exec_launcher(int variant);
static inline int qsb_exec_stride(int planes) { return planes | 1; }
