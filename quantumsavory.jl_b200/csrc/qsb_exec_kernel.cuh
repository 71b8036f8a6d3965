// qsb_exec_kernel.cuh -- the fused executor kernel of recorded tableau programs (see qsb_exec.cu).
// Included by qsb_exec.cu (host side: ExecArgs) and by qsb_exec_inst.cu, which is compiled once per block
// size T (-DQSB_EXEC_T=...) so that the instantiations build in parallel.
#pragma once
#ifndef QSB_EXEC_PLANE_T
#define QSB_EXEC_PLANE_T uint32_t   // measured on B200: 32-lane words 5.21 ms, 16-lane 5.22 ms, 8-lane bytes 5.55 ms per 10^6 firstgenrepeater trajectories
#endif
#define QSB_PLANE_T QSB_EXEC_PLANE_T
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
    double* stats;
    double* per_traj;
};
#define QSB_EXEC_CHUNK 32   // ops staged per shared-memory refill

struct ExecArgs {
    const ExecOp* ops; int nops;
    const TabObsArgs* obs;
    const uint8_t* aux;
    plane_t* arena;           // nullptr -> shared-memory tile
    const uint32_t* bits;     // decision words of every stochastic op, [word][slot_stride] (a word = 32 trajectories)
    int slot_stride;
    const plane_t* valid;
    size_t W;                 // lane groups (threads) in the ensemble = trajectories / QSB_LANES
    uint64_t seed, traj0, draw_base;
    int dyn_threads;          // block size of the run-time sized variant (k_tab_exec<1>)
};

// launchers, one per compiled block size (0 = arena in global memory)
typedef cudaError_t (*qsb_exec_launch_fn)(const ExecArgs&, unsigned blocks, int smem_bytes, cudaStream_t stream);
qsb_exec_launch_fn qsb_exec_launcher(int threads);   // nullptr when that block size is not compiled in
extern const int qsb_exec_block_sizes[];             // 0-terminated list of the shared-memory block sizes
