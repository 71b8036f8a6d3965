// qsb_dexec.cuh -- recorded DENSE programs: the op list of a lockstep schedule on small density matrices, executed
// in ONE launch with every live register state of a tile of trajectories resident in shared memory.
//
// Why: in the repeater workloads every live state is a 2..4-qubit density matrix (256 B .. 4 KiB per trajectory).
// The per-op kernels stream each op through HBM at ~0.95 of the measured bandwidth, but an op moves the whole
// state in and out for a handful of flops; keeping the states of T trajectories on chip for the WHOLE schedule
// removes that traffic (north_star: "small registers stay resident in shared memory").
//
// Layout: a tile is T consecutive trajectories (T = 32, 16 or 8, the largest whose arena fits).  The element
// arena is double2 arena[elem * T + lane] -- trajectory fastest, so a warp's 16-byte accesses to one element
// row are contiguous and bank-conflict free.  A block is 256 threads = T lanes x G groups; an op's independent
// work items (2x2 / 4x4 blocks, output elements) are dealt to the groups, reductions and draws of an op run in
// one group.  Masks and alive bits are one 32-bit word per tile.  __syncthreads() only where consecutive ops
// touch the same storage (flag computed at qsb_record_end).
#pragma once
#include <map>
#include <vector>

#include "qsb_internal.cuh"

struct DExecOp {              // 80 bytes (five 16-byte words: staged through shared memory with uint4 copies)
    uint8_t type, n, n2, k;
    uint8_t q[2], mode, flags;       // flags bit 0: barrier before this op
    int16_t al[3], m[3];             // alive / mask words (-1: none)
    uint16_t grp, nnz;               // group that runs the single-group part of the op; mode-1 projector entries
    int32_t s[3];                    // element offsets of the state blocks (s[0] = destination)
    int32_t aux;                     // offset (in doubles) of the op's constants in the aux table
    uint32_t u[2];                   // Pauli x / z masks
    uint32_t draw;
    uint32_t fpow;                   // DOP_EXPECT: the value is multiplied by i^fpow (sign and the i^ny of a Pauli string)
    double* stats;                   // absolute device pointers captured at record time
    double* per_traj;
    uint8_t g0, reserved0[7];        // level scheduling: the op deals its items to the groups starting at group g0
};
static_assert(sizeof(DExecOp) == 80, "DExecOp layout");
#define QSB_DEXEC_CHUNK_OF(threads) ((threads) >= 128 ? 32 : 8)   // ops staged per shared-memory refill (80 B each, per block)
#define QSB_DEXEC_THREADS 256

struct DExecArgs {
    const DExecOp* ops; int nops;
    const double* aux;
    int elems, nmasks, logT;         // arena elements per trajectory, mask words, log2(T)
    int aux_smem;                    // > 0: the aux table (this many doubles) is copied into shared memory at kernel start
    const uint32_t* valid;
    int64_t ntiles;
    uint64_t seed, traj0, draw_base;
    int* flags;                      // sticky device-side error flags of the context
    // cohort forking (qsb_ctx_retire on a dense recording): the launch may cover a lane subset of the context
    const uint32_t* lanes;           // lane index (trajectory slot of the context) of job position p; nullptr: identity
    int64_t count;                   // trajectories of the job (lanes != nullptr)
    int live_word;                   // mask word that holds the lanes still in this cohort (-1: the program never retires lanes)
};

// ---- the recorder (shared by qsb_exec.cu and qsb_dexec.cu) ------------------------------------------------
// one address space of a recording: bit planes (tableau states, masks, alive bits) or dense-matrix elements
struct RecArena {
    int next_plane = 0, high = 0;
    std::multimap<int, int> free_blocks;  // size -> base
    // lifetime log of every block, for the offline re-packing at qsb_record_end:
    // live for ops [t0, t1); sizes.back().second units from op sizes.back().first on (in-place shrinks)
    struct Blk { int base, t0, t1; std::vector<std::pair<int, int>> sizes; int nbase; bool used; };
    std::vector<Blk> blks;
    std::map<int, int> live;              // online base -> index into blks
};
struct qsb_recorder {
    std::vector<TabOp> ops;
    std::vector<TabObsArgs> obs;
    std::vector<double*> stats_ptrs;
    RecArena A[2];                        // [0] planes (tableaux, masks, alive), [1] elements of dense programs
    bool dense = false, tab = false;      // a program is either all-tableau or all-dense
    std::vector<double> daux;             // constant payloads of dense ops (gate matrices, channel coefficients, ...), deduplicated
    std::map<std::vector<double>, int> daux_index;
    uint64_t ops0 = 0, draws0 = 0, bytes0 = 0;   // context counters at qsb_record_begin (recording itself executes nothing)
    int now() const { return (int)ops.size(); }
};

struct qsb_dprogram;                 // built by qsb_dexec_build from a dense recording
// elems / nmasks: arena sizes after re-packing; the recorder's ops already carry the re-packed references
int qsb_dexec_build(qsb_ctx* ctx, qsb_recorder* R, int elems, int nmasks, qsb_dprogram** out);
int qsb_dexec_run(qsb_dprogram* P, const uint32_t* lanes_dev, int64_t count, uint64_t traj_shift, uint64_t draw_base, cudaStream_t stream = nullptr);   // stream: nullptr = the context's
int qsb_dexec_upload(qsb_dprogram* P, int64_t* bytes);
void qsb_dexec_info(const qsb_dprogram* P, int64_t info[8]);
void qsb_dexec_free(qsb_dprogram* P);
