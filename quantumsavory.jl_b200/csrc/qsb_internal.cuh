// qsb_internal.cuh -- shared declarations of libqsb200 (not part of the public ABI).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include <set>
#include <string>
#include <vector>

#include "../../include/qsb200.h"

// ---------------------------------------------------------------------------------------------
// Memory layout (DESIGN.md "Data layout in HBM")
//   dense kinds : double2 d[elem * Npad + traj]      -- batch-major, trajectory fastest, so that a
//                 warp reads the same matrix element of 32 consecutive trajectories as one 512-byte
//                 fully coalesced request of 128-bit lanes.  ket: elem in [0,2^n); dm: elem = r + c*2^n.
//   tableau     : uint32 plane[p * W + w]            -- bit-sliced ACROSS trajectories: bit l of word w
//                 of plane p is tableau bit p of trajectory 32*w+l.  W = Npad/32.
//                 pair i (i<n) owns planes [i*(4n+2), (i+1)*(4n+2)):
//                    Dx[0..n) Dz[0..n) Sx[0..n) Sz[0..n) Ssign active
//                 (D_i,S_i) is a symplectic pair; S_i is a real stabilizer where `active` is set,
//                 otherwise (D_i,S_i) are the logical X/Z of a maximally mixed degree of freedom.
//   masks       : uint32 m[w]                        -- same lane mapping as the tableau words.
//   Npad        : ntraj rounded up to a multiple of 32; pad lanes are permanently dead.
// ---------------------------------------------------------------------------------------------

#define QSB_TAB_MAXN 32
#define QSB_DENSE_MAXN 14 /* dm: 4^14 elements per trajectory is already 4 GiB */

struct qsb_ctx {
    int device;
    int64_t ntraj, npad, nwords, traj_offset;
    uint64_t seed;
    cudaStream_t stream;
    bool own_stream;
    uint32_t* valid;     // [nwords] 1 for real trajectories
    int* dev_flags;      // [4] sticky device-side error flags
    double* scratch;     // reduction partials
    size_t scratch_bytes;
    double* stats_tmp;   // [8]
    long long* counts_tmp;  // [8]
    void* obs_buf;       // uploaded observable data (kets, operators, generator tables)
    size_t obs_bytes;
    void* host_stage;    // pinned staging for uploads
    size_t host_stage_bytes;
    uint64_t launches, algo_bytes, ops, draws;
    int sm_count;
    struct qsb_prof* prof;  // per-kernel CUDA-event timing, off unless qsb_ctx_profile(ctx, 1)
    std::set<struct qsb_state*>* live_states;  // handles not yet freed by the caller (released at destroy)
    std::set<struct qsb_mask*>* live_masks;
    struct qsb_recorder* rec;  // non-null between qsb_record_begin and qsb_record_end
    int rec_gen;
    // scratch of the fused launches (qsb_exec.cu), shared by every program of the context and grown on demand:
    uint32_t* wave_bits; size_t wave_bits_bytes;   // decision words of the launch in flight (draw pass -> executor)
    void* wave_jobs; size_t wave_jobs_bytes;       // job table of the launch in flight
    int wave_threads, wave_blocks;                 // shape of the last executor launch (reported by tools)
    int32_t* route_bins; size_t route_bins_bytes;  // cohort routing scratch: counts / cursors, offsets, child programs per key
};

struct qsb_state {
    qsb_ctx* ctx;
    int kind;
    int n;
    void* data;       // double2* or uint32_t*
    size_t bytes;
    uint32_t* alive;  // [nwords]
    int vbase, valive; // record mode: plane indices inside the program arena (data == alive == nullptr)
    int vcap;          // record mode, dense: elements the state's arena block was allocated with
    int gen;           // recording generation the virtual planes belong to
};

struct qsb_mask {
    qsb_ctx* ctx;
    uint32_t* w;
    int vplane;        // record mode: plane index inside the program arena (w == nullptr)
    int gen;
};

// Several contexts on different GPUs may live in one process (the Julia glue drives 8 shards from one host loop):
// every path that allocates, launches or copies first makes the context's device current for the calling thread.
static inline void qsb_use_device(const qsb_ctx* ctx) { int cur = -1; if (cudaGetDevice(&cur) != cudaSuccess || cur != ctx->device) cudaSetDevice(ctx->device); }

// ---- error plumbing ---------------------------------------------------------------------------
void qsb_set_error(const char* fmt, ...);
#define QSB_FAIL(code, ...) do { qsb_set_error(__VA_ARGS__); return (code); } while (0)
#define QSB_CUDA(expr) do { cudaError_t e_ = (expr); if (e_ != cudaSuccess) { \
    qsb_set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(e_), __FILE__, __LINE__); \
    return e_ == cudaErrorMemoryAllocation ? QSB_ERR_OOM : QSB_ERR_CUDA; } } while (0)
#define QSB_TRY(expr) do { int rc_ = (expr); if (rc_ != QSB_OK) return rc_; } while (0)
#define QSB_LAUNCH_CHECK(ctx) do { (ctx)->launches++; if ((ctx)->prof) qsb_prof_post(ctx); cudaError_t e_ = cudaGetLastError(); if (e_ != cudaSuccess) { \
    qsb_set_error("kernel launch failed: %s (%s:%d)", cudaGetErrorString(e_), __FILE__, __LINE__); return QSB_ERR_CUDA; } } while (0)
// brackets the next launch with CUDA events on the context's stream when profiling is on
#define QSB_PRE(ctx, name) do { qsb_use_device(ctx); if ((ctx)->prof) qsb_prof_pre(ctx, name); } while (0)
void qsb_prof_pre(qsb_ctx* ctx, const char* kernel);
void qsb_prof_post(qsb_ctx* ctx);
// adds an op's algorithmic bytes (SURVEY 8d table x trajectories) to the context total and to `kernel`'s row
void qsb_account(qsb_ctx* ctx, const char* kernel, uint64_t bytes);

// ---- allocation -------------------------------------------------------------------------------
int qsb_dev_alloc(qsb_ctx* ctx, void** p, size_t bytes);
int qsb_dev_free(qsb_ctx* ctx, void* p);
int qsb_ensure_scratch(qsb_ctx* ctx, size_t bytes);
int qsb_upload(qsb_ctx* ctx, const void* host, size_t bytes, void** dev_out);  // into ctx->obs_buf (stream ordered)
int qsb_state_alloc(qsb_ctx* ctx, int kind, int n, qsb_state** out);
size_t qsb_state_bytes(const qsb_ctx* ctx, int kind, int n);
int qsb_mask_alloc(qsb_ctx* ctx, qsb_mask** out);

static inline int qsb_tab_planes(int n) { return n * (4 * n + 2); }
static inline bool qsb_is_dense(int kind) { return kind == QSB_KET || kind == QSB_DM || kind == QSB_MCKET; }
static inline bool qsb_is_ketlike(int kind) { return kind == QSB_KET || kind == QSB_MCKET; }

// ---- dense implementation (qsb_dense.cu) ---------------------------------------------------------
int qsbd_new(qsb_ctx* ctx, int kind, int n, qsb_state** out);
int qsbd_from_host(qsb_ctx* ctx, int kind, int n, const double* data, qsb_state** out);
int qsbd_select(qsb_ctx* ctx, const qsb_state* const* states, const double* w, int m, uint64_t draw_id, qsb_state** out);
int qsbd_promote_dm(qsb_state* st);
int qsbd_to_host(const qsb_state* st, int64_t traj, double* out);
int qsbd_tensor2(qsb_ctx* ctx, const qsb_state* a, const qsb_state* b, qsb_state** out);
int qsbd_apply(qsb_state* st, const int* idx, int k, const double* U, const qsb_mask* mask);
int qsbd_channel(qsb_state* st, int idx, int bg, const double* prm, double dt, uint64_t draw_id);
struct KrausArgs { double2 k[4][4]; int nk; };   // up to four 2x2 Kraus operators, column-major k[m][r + 2c]
int qsbd_kraus_sampled(qsb_state* st, int idx, const KrausArgs& ka, uint64_t draw_id);
int qsbd_evolve(qsb_state* st, const int* idx, int k, const double* H_ri, int nl, const int* l_pos, const double* L_ri, double duration, uint64_t draw_id);
int qsbd_measure(qsb_state* st, int idx, const double* kets, uint64_t draw_id, qsb_mask* outcome2, int64_t* counts_host);
int qsbd_ptrace(qsb_state* st, int idx, uint64_t draw_id);
int qsbd_expect(qsb_state* st, const int* idx, int k, int obs_kind, const void* obs, int obs_sign,
                double* per_traj, double* stats_dev, double* stats_host);

// ---- tableau implementation (qsb_tab.cu) ---------------------------------------------------------
int qsbt_new(qsb_ctx* ctx, int n, int active, qsb_state** out);
int qsbt_from_host(qsb_ctx* ctx, int n, int ngen, const uint8_t* xs, const uint8_t* zs, const uint8_t* signs, qsb_state** out);
int qsbt_select(qsb_ctx* ctx, const qsb_state* const* states, const double* w, int m, uint64_t draw_id, qsb_state** out);
int qsbt_to_host(const qsb_state* st, int64_t traj, uint8_t* xs, uint8_t* zs, uint8_t* signs, uint8_t* active);
int qsbt_tensor2(qsb_ctx* ctx, const qsb_state* a, const qsb_state* b, qsb_state** out);
int qsbt_apply(qsb_state* st, const int* idx, int k, const uint8_t* table, const qsb_mask* mask);
int qsbt_channel(qsb_state* st, int idx, int bg, const double* prm, double dt, uint64_t draw_id);
int qsbt_measure(qsb_state* st, int idx, int basis, uint64_t draw_id, qsb_mask* outcome2, qsb_mask* was_random, int64_t* counts_host);
int qsbt_ptrace(qsb_state* st, int idx);
int qsbt_expect(qsb_state* st, const int* idx, int k, int obs_kind, const void* obs, int obs_sign,
                double* per_traj, double* stats_dev, double* stats_host);

// ---- recorded programs (qsb_exec.cu): tableau op lists executed by ONE fused launch ----------------
enum { TOP_INIT, TOP_TENSOR, TOP_SELECT, TOP_COPY, TOP_GATE, TOP_NOISE, TOP_MEASURE, TOP_PTRACE, TOP_FORCE, TOP_EXPECT,
       TOP_MASK_FILL, TOP_MASK_LOGIC, TOP_MASK_RANDOM, TOP_KILL, TOP_ALIVE_COPY,
       TOP_NOP, TOP_SELECT_CONST /* produced by constant folding at qsb_record_end */, TOP_RETIRE /* cohort forking */,
       // dense programs (qsb_dexec.cu): every state is held as a density matrix of 4^n elements in the element arena
       DOP_CONST, DOP_TENSOR, DOP_COPY, DOP_APPLY, DOP_CHANNEL, DOP_MEASURE, DOP_PTRACE, DOP_EXPECT };
#define QSB_DEXEC_MAXN 5   /* largest state of a dense program: 4^5 elements per trajectory */
struct TabOp {
    int type, n, n2, k;
    int s[6];       // state base planes (s[0] is the destination where there is one)
    int al[6];      // alive planes matching s[]
    int m[3];       // mask planes (-1: none)
    int q[2];
    int iv[4];
    uint32_t u[2], lut;
    uint8_t outmask[4];
    double f[8];
    uint64_t draw;
    double* stats;      // absolute device pointers captured at record time
    double* per_traj;
    int obs;            // index into the program's TabObsArgs table
    int dslot;          // first decision-word slot of a stochastic op (filled at qsb_record_end)
};
// arena 0: bit planes (tableaux, masks, alive bits); arena 1: matrix elements of dense programs
int qsb_rec_alloc(qsb_ctx* ctx, int arena, int units);
void qsb_rec_free(qsb_ctx* ctx, int arena, int base);
void qsb_rec_shrink_to(qsb_ctx* ctx, int arena, int base, int new_units);   // keep the head of a block, recycle its tail
int qsb_rec_set_family(qsb_ctx* ctx, bool dense);                            // a program is all-tableau or all-dense
int qsb_rec_daux(qsb_ctx* ctx, const double* v, int n);                      // constant payload -> offset into the aux table
int qsb_rec_alloc_planes(qsb_ctx* ctx, int nplanes);
void qsb_rec_free_planes(qsb_ctx* ctx, int base, int nplanes);
void qsb_rec_shrink(qsb_ctx* ctx, int base, int old_planes, int new_planes);  // keep the head of a block, recycle its tail
TabOp* qsb_rec_push(qsb_ctx* ctx, int type);           // zero-initialised op with all plane refs = -1
int qsb_rec_push_obs(qsb_ctx* ctx, const void* tab_obs_args, size_t bytes);
void qsb_rec_note_stats(qsb_ctx* ctx, double* p);
// a handle is usable when it is real outside a recording, or virtual and owned by the active recording
int qsb_handle_ok(const qsb_ctx* ctx, bool is_virtual, int gen);

// ---- device helpers ---------------------------------------------------------------------------
#ifdef __CUDACC__
#include "qsb_tab_core.cuh"  // qsb_draw (Philox) shared by all kernels
__device__ __forceinline__ size_t qsb_insert0(size_t g, int pos) {
    size_t low = g & (((size_t)1 << pos) - 1);
    return ((g >> pos) << (pos + 1)) | low;
}
__device__ __forceinline__ double2 cmul(double2 a, double2 b) { return make_double2(a.x * b.x - a.y * b.y, a.x * b.y + a.y * b.x); }
__device__ __forceinline__ double2 cmulc(double2 a, double2 b) { /* a * conj(b) */ return make_double2(a.x * b.x + a.y * b.y, a.y * b.x - a.x * b.y); }
__device__ __forceinline__ double2 cconjmul(double2 a, double2 b) { /* conj(a) * b */ return make_double2(a.x * b.x + a.y * b.y, a.x * b.y - a.y * b.x); }
__device__ __forceinline__ void cfma(double2& acc, double2 a, double2 b) { acc.x += a.x * b.x - a.y * b.y; acc.y += a.x * b.y + a.y * b.x; }
__device__ __forceinline__ double2 cadd(double2 a, double2 b) { return make_double2(a.x + b.x, a.y + b.y); }
__device__ __forceinline__ double2 cscale(double2 a, double s) { return make_double2(a.x * s, a.y * s); }
#endif
