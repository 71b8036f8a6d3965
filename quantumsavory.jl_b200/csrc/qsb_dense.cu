// qsb_dense.cu -- fp64-complex ket / density-matrix kernels (QSB_KET, QSB_DM, QSB_MCKET), sm_100a.
//
// All kernels use the batch-major layout d[elem*Npad + traj] (qsb_internal.cuh).  A thread block is
// 128 consecutive trajectories (blockIdx.x) and walks element groups with stride gridDim.y, so every
// load/store instruction of a warp is one contiguous 512-byte request of 16-byte lanes and no
// integer division is needed.  Every op is a single streaming pass: the roofline that bounds each
// kernel is HBM bandwidth and its algorithmic bytes per trajectory are listed in DESIGN.md.
#include <math.h>
#include <stdlib.h>

#include <vector>

#include "qsb_internal.cuh"

#define TPB 128

// ---- record mode (dense programs, qsb_dexec.cu) --------------------------------------------------
// Inside a recording a dense state is a block of 4^n elements in the program's element arena and is ALWAYS kept as
// a density matrix (a ket is recorded as |psi><psi|); st->kind keeps the logical kind for the promotion rules.
static inline uint64_t dm_elems(int n) { return (uint64_t)1 << (2 * n); }
static int rec_const(qsb_ctx* ctx, qsb_state* st, const std::vector<double>& rho) {
    TabOp* o = qsb_rec_push(ctx, DOP_CONST);
    o->n = st->n; o->s[0] = st->vbase; o->al[0] = st->valive; o->obs = qsb_rec_daux(ctx, rho.data(), (int)rho.size());
    return QSB_OK;
}
// After an in-place shrink a small state would keep sitting at the base of the big block it was projected out of (the
// pair left by a swap at the base of the 256-element composite), fragmenting the arena for as long as it lives.  Move it
// to a block of its own size: a 16-element copy in shared memory buys a much smaller arena (more trajectories per SM).
static void rec_relocate_small(qsb_ctx* ctx, qsb_state* st) {
    const int need = st->n >= 1 ? 1 << (2 * st->n) : 1;
    if (st->n < 1 || st->n > 2 || st->vcap <= need) return;
    const int nb = qsb_rec_alloc(ctx, 1, need);
    TabOp* o = qsb_rec_push(ctx, DOP_COPY); o->n = st->n; o->s[0] = nb; o->s[1] = st->vbase; o->al[0] = st->valive; o->al[1] = st->valive;
    qsb_rec_free(ctx, 1, st->vbase);
    st->vbase = nb; st->vcap = need;
}
static std::vector<double> outer_ri(const double* psi, size_t D) {   // column-major rho[r + c D] = psi[r] conj(psi[c])
    std::vector<double> rho(2 * D * D);
    for (size_t c = 0; c < D; ++c) for (size_t r = 0; r < D; ++r) {
        rho[2 * (r + c * D)] = psi[2 * r] * psi[2 * c] + psi[2 * r + 1] * psi[2 * c + 1];
        rho[2 * (r + c * D) + 1] = psi[2 * r + 1] * psi[2 * c] - psi[2 * r] * psi[2 * c + 1];
    }
    return rho;
}

struct Grid2 { dim3 grid; };
static Grid2 plan_grid(const qsb_ctx* ctx, size_t groups) {
    unsigned gx = (unsigned)((ctx->npad + TPB - 1) / TPB);
    size_t target = (size_t)ctx->sm_count * 32;  // blocks in flight we aim for
    size_t gy = (target + gx - 1) / gx;
    if (gy < 1) gy = 1;
    if (gy > groups) gy = groups;
    if (gy > 65535) gy = 65535;
    Grid2 g; g.grid = dim3(gx, (unsigned)gy, 1);
    return g;
}
static int plan_split(const qsb_ctx* ctx, size_t range) {
    size_t target_threads = (size_t)ctx->sm_count * 2048;
    size_t s = (target_threads + ctx->npad - 1) / ctx->npad;
    if (s < 1) s = 1;
    if (s > range) s = range;
    if (s > 4096) s = 4096;
    return (int)s;
}

// =============================================================================================
// construction
// =============================================================================================
__global__ void k_alive_init(uint32_t* alive, const uint32_t* valid, int64_t nwords) {
    int64_t w = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (w < nwords) alive[w] = valid[w];
}
__global__ void k_alive_and(uint32_t* out, const uint32_t* a, const uint32_t* b, int64_t nwords) {
    int64_t w = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (w < nwords) out[w] = a[w] & b[w];
}
static int alive_init(qsb_ctx* ctx, qsb_state* st) {
    QSB_PRE(ctx, "k_alive_init");
    k_alive_init<<<(unsigned)((ctx->nwords + 255) / 256), 256, 0, ctx->stream>>>(st->alive, ctx->valid, ctx->nwords);
    QSB_LAUNCH_CHECK(ctx);
    return QSB_OK;
}
static int alive_copy(qsb_ctx* ctx, qsb_state* dst, const qsb_state* src) {
    QSB_CUDA(cudaMemcpyAsync(dst->alive, src->alive, ctx->nwords * sizeof(uint32_t), cudaMemcpyDeviceToDevice, ctx->stream));
    return QSB_OK;
}

__global__ void k_fill0(double2* d, int64_t npad, size_t elems) {
    int64_t t = blockIdx.x * (int64_t)TPB + threadIdx.x;
    if (t >= npad) return;
    for (size_t e = blockIdx.y; e < elems; e += gridDim.y) d[e * npad + t] = make_double2(e == 0 ? 1.0 : 0.0, 0.0);
}
__global__ void k_broadcast(double2* d, const double2* __restrict__ src, int64_t npad, size_t elems) {
    int64_t t = blockIdx.x * (int64_t)TPB + threadIdx.x;
    if (t >= npad) return;
    for (size_t e = blockIdx.y; e < elems; e += gridDim.y) d[e * npad + t] = src[e];
}
int qsbd_new(qsb_ctx* ctx, int kind, int n, qsb_state** out) {
    qsb_state* st;
    QSB_TRY(qsb_state_alloc(ctx, kind, n, &st));
    if (ctx->rec) {
        std::vector<double> rho(2 * dm_elems(n), 0.0);
        rho[0] = 1.0;
        *out = st;
        return rec_const(ctx, st, rho);
    }
    size_t elems = st->bytes / (ctx->npad * sizeof(double2));
    Grid2 g = plan_grid(ctx, elems);
    QSB_PRE(ctx, "k_fill0");
    k_fill0<<<g.grid, TPB, 0, ctx->stream>>>((double2*)st->data, ctx->npad, elems);
    QSB_LAUNCH_CHECK(ctx);
    QSB_TRY(alive_init(ctx, st));
    *out = st;
    return QSB_OK;
}
int qsbd_from_host(qsb_ctx* ctx, int kind, int n, const double* data, qsb_state** out) {
    qsb_state* st;
    QSB_TRY(qsb_state_alloc(ctx, kind, n, &st));
    size_t elems = st->bytes / (ctx->npad * sizeof(double2));
    if (ctx->rec) {
        std::vector<double> rho = kind == QSB_DM ? std::vector<double>(data, data + 2 * dm_elems(n)) : outer_ri(data, (size_t)1 << n);
        qsb_account(ctx, "k_broadcast", (uint64_t)elems * 16 * ctx->ntraj); ctx->ops += ctx->ntraj;
        *out = st;
        return rec_const(ctx, st, rho);
    }
    void* src;
    QSB_TRY(qsb_upload(ctx, data, elems * sizeof(double2), &src));
    Grid2 g = plan_grid(ctx, elems);
    QSB_PRE(ctx, "k_broadcast");
    k_broadcast<<<g.grid, TPB, 0, ctx->stream>>>((double2*)st->data, (const double2*)src, ctx->npad, elems);
    QSB_LAUNCH_CHECK(ctx);
    QSB_TRY(alive_init(ctx, st));
    qsb_account(ctx, "k_broadcast", (uint64_t)elems * 16 * ctx->ntraj); ctx->ops += ctx->ntraj;
    *out = st;
    return QSB_OK;
}

struct SelArgs { const double2* src[8]; const uint32_t* alive[8]; double cum[8]; int m; };
__global__ void k_select(double2* d, uint32_t* alive_out, SelArgs a, int64_t npad, size_t elems, uint64_t seed, uint64_t traj0, uint64_t draw_id) {
    int64_t t = blockIdx.x * (int64_t)TPB + threadIdx.x;
    if (t >= npad) return;
    double u = qsb_draw(seed, traj0 + (uint64_t)t, draw_id);
    int sel = a.m - 1;
    for (int i = 0; i < a.m; ++i) if (a.cum[i] > u) { sel = i; break; }
    if (blockIdx.y == 0) {
        bool al = (a.alive[sel][t >> 5] >> (t & 31)) & 1u;
        uint32_t word = __ballot_sync(0xffffffffu, al);
        if ((threadIdx.x & 31) == 0) alive_out[t >> 5] = word;
    }
    const double2* s = a.src[sel];
    for (size_t e = blockIdx.y; e < elems; e += gridDim.y) d[e * npad + t] = s[e * npad + t];
}
int qsbd_select(qsb_ctx* ctx, const qsb_state* const* states, const double* w, int m, uint64_t draw_id, qsb_state** out) {
    if (m > 8) QSB_FAIL(QSB_ERR_UNSUPPORTED, "mixtures of more than 8 states are not implemented");
    if (ctx->rec) QSB_FAIL(QSB_ERR_UNSUPPORTED, "sampled mixtures of dense states cannot be recorded (express the mixture as a density matrix)");
    qsb_state* st;
    QSB_TRY(qsb_state_alloc(ctx, states[0]->kind, states[0]->n, &st));
    size_t elems = st->bytes / (ctx->npad * sizeof(double2));
    SelArgs a; a.m = m;
    double total = 0, cum = 0;
    for (int i = 0; i < m; ++i) total += w[i];
    for (int i = 0; i < m; ++i) { cum += w[i] / total; a.cum[i] = cum; a.src[i] = (const double2*)states[i]->data; a.alive[i] = states[i]->alive; }
    Grid2 g = plan_grid(ctx, elems);
    QSB_PRE(ctx, "k_select");
    k_select<<<g.grid, TPB, 0, ctx->stream>>>((double2*)st->data, st->alive, a, ctx->npad, elems, ctx->seed, (uint64_t)ctx->traj_offset, draw_id);
    QSB_LAUNCH_CHECK(ctx);
    qsb_account(ctx, "k_select", (uint64_t)elems * 32 * ctx->ntraj); ctx->ops += ctx->ntraj;
    *out = st;
    return QSB_OK;
}

// dm(ket): rho[r + c D] = psi[r] conj(psi[c])   (uptotime.jl:2, quantumoptics.jl:5-6)
__global__ void k_outer(double2* out, const double2* __restrict__ psi, int64_t npad, int n) {
    int64_t t = blockIdx.x * (int64_t)TPB + threadIdx.x;
    if (t >= npad) return;
    size_t D = (size_t)1 << n, elems = D * D;
    for (size_t e = blockIdx.y; e < elems; e += gridDim.y) {
        size_t r = e & (D - 1), c = e >> n;
        out[e * npad + t] = cmulc(psi[r * npad + t], psi[c * npad + t]);
    }
}
int qsbd_promote_dm(qsb_state* st) {
    qsb_ctx* ctx = st->ctx;
    if (ctx->rec) { st->kind = QSB_DM; st->bytes = qsb_state_bytes(ctx, QSB_DM, st->n); return QSB_OK; }   // already held as a dm
    size_t nbytes = qsb_state_bytes(ctx, QSB_DM, st->n);
    void* nd;
    QSB_TRY(qsb_dev_alloc(ctx, &nd, nbytes));
    Grid2 g = plan_grid(ctx, (size_t)1 << (2 * st->n));
    QSB_PRE(ctx, "k_outer");
    k_outer<<<g.grid, TPB, 0, ctx->stream>>>((double2*)nd, (const double2*)st->data, ctx->npad, st->n);
    QSB_LAUNCH_CHECK(ctx);
    QSB_TRY(qsb_dev_free(ctx, st->data));
    st->data = nd; st->bytes = nbytes; st->kind = QSB_DM;
    return QSB_OK;
}
int qsbd_to_host(const qsb_state* st, int64_t traj, double* out) {
    qsb_ctx* ctx = st->ctx;
    size_t elems = st->bytes / (ctx->npad * sizeof(double2));
    QSB_CUDA(cudaMemcpy2DAsync(out, sizeof(double2), (const double2*)st->data + traj, ctx->npad * sizeof(double2),
                               sizeof(double2), elems, cudaMemcpyDeviceToHost, ctx->stream));
    QSB_CUDA(cudaStreamSynchronize(ctx->stream));
    return QSB_OK;
}

// =============================================================================================
// subsystemcompose: out = a (x) b with a least significant (a3)
// =============================================================================================
// Each thread keeps a tile of up to TB elements of b in registers and streams a through it, so a and b are
// read (about) once and the pass is bounded by the output write: 16*(4^na + 4^nb) read + 16*4^(na+nb) written.
#define TENSOR_TB 16
template <bool DM>
__global__ void __launch_bounds__(TPB) k_tensor(double2* __restrict__ out, const double2* __restrict__ a, const double2* __restrict__ b, int64_t npad, int na, int nb) {
    int64_t t = blockIdx.x * (int64_t)TPB + threadIdx.x;
    if (t >= npad) return;
    size_t Da = (size_t)1 << na, Db = (size_t)1 << nb;
    size_t La = DM ? Da * Da : Da, Lb = DM ? Db * Db : Db;
    int n = na + nb;
    size_t ntiles = (Lb + TENSOR_TB - 1) / TENSOR_TB;
    // blockIdx.y walks (b tile, a chunk) pairs
    size_t chunks = (gridDim.y + ntiles - 1) / ntiles;  // a-chunks per tile
    if (chunks < 1) chunks = 1;
    for (size_t job = blockIdx.y; job < ntiles * chunks; job += gridDim.y) {
        size_t tile = job / chunks, chunk = job % chunks;
        size_t eb0 = tile * TENSOR_TB;
        double2 bv[TENSOR_TB];
#pragma unroll
        for (int j = 0; j < TENSOR_TB; ++j) bv[j] = (eb0 + j < Lb) ? b[(eb0 + j) * npad + t] : make_double2(0, 0);
        size_t ea_lo = La * chunk / chunks, ea_hi = La * (chunk + 1) / chunks;
        for (size_t ea = ea_lo; ea < ea_hi; ++ea) {
            double2 av = a[ea * npad + t];
            size_t ra = DM ? (ea & (Da - 1)) : ea, ca = DM ? (ea >> na) : 0;
#pragma unroll
            for (int j = 0; j < TENSOR_TB; ++j) {
                size_t eb = eb0 + j;
                if (eb < Lb) {
                    size_t rb = DM ? (eb & (Db - 1)) : eb, cb = DM ? (eb >> nb) : 0;
                    size_t eo = DM ? ((ra | (rb << na)) + ((ca | (cb << na)) << n)) : (ra | (rb << na));
                    out[eo * npad + t] = cmul(av, bv[j]);
                }
            }
        }
    }
}
int qsbd_tensor2(qsb_ctx* ctx, const qsb_state* a, const qsb_state* b, qsb_state** out) {
    // kind promotion rules: quantumoptics.jl:3-6, mcket.jl:35-44
    int kind;
    if (a->kind == QSB_DM || b->kind == QSB_DM) kind = QSB_DM;
    else if (a->kind == QSB_MCKET && b->kind == QSB_MCKET) kind = QSB_MCKET;
    else kind = QSB_KET;
    if (a->n + b->n > QSB_DENSE_MAXN) QSB_FAIL(QSB_ERR_UNSUPPORTED, "composite state would exceed %d qubits", QSB_DENSE_MAXN);
    if (ctx->rec) {
        qsb_state* st;
        QSB_TRY(qsb_state_alloc(ctx, kind, a->n + b->n, &st));
        TabOp* o = qsb_rec_push(ctx, DOP_TENSOR); o->n = a->n; o->n2 = b->n;
        o->s[0] = st->vbase; o->s[1] = a->vbase; o->s[2] = b->vbase; o->al[0] = st->valive; o->al[1] = a->valive; o->al[2] = b->valive;
        qsb_account(ctx, "k_tensor", (uint64_t)(qsb_state_bytes(ctx, kind, a->n) + qsb_state_bytes(ctx, kind, b->n) + st->bytes) / ctx->npad * ctx->ntraj); ctx->ops += ctx->ntraj;
        *out = st;
        return QSB_OK;
    }
    qsb_state *pa = nullptr, *pb = nullptr;
    const qsb_state *xa = a, *xb = b;
    if (kind == QSB_DM && a->kind != QSB_DM) { QSB_TRY(qsb_state_copy(a, &pa)); QSB_TRY(qsbd_promote_dm(pa)); xa = pa; }
    if (kind == QSB_DM && b->kind != QSB_DM) { QSB_TRY(qsb_state_copy(b, &pb)); QSB_TRY(qsbd_promote_dm(pb)); xb = pb; }
    qsb_state* st;
    QSB_TRY(qsb_state_alloc(ctx, kind, a->n + b->n, &st));
    size_t La = kind == QSB_DM ? (size_t)1 << (2 * a->n) : (size_t)1 << a->n;
    size_t Lb = kind == QSB_DM ? (size_t)1 << (2 * b->n) : (size_t)1 << b->n;
    Grid2 g = plan_grid(ctx, ((Lb + TENSOR_TB - 1) / TENSOR_TB) * La);
    QSB_PRE(ctx, "k_tensor");
    if (kind == QSB_DM) k_tensor<true><<<g.grid, TPB, 0, ctx->stream>>>((double2*)st->data, (const double2*)xa->data, (const double2*)xb->data, ctx->npad, a->n, b->n);
    else k_tensor<false><<<g.grid, TPB, 0, ctx->stream>>>((double2*)st->data, (const double2*)xa->data, (const double2*)xb->data, ctx->npad, a->n, b->n);
    QSB_LAUNCH_CHECK(ctx);
    QSB_PRE(ctx, "k_alive_and");
    k_alive_and<<<(unsigned)((ctx->nwords + 255) / 256), 256, 0, ctx->stream>>>(st->alive, a->alive, b->alive, ctx->nwords);
    QSB_LAUNCH_CHECK(ctx);
    if (pa) qsb_state_free(pa);
    if (pb) qsb_state_free(pb);
    size_t ea = xa->bytes / ctx->npad, eb = xb->bytes / ctx->npad, eo = st->bytes / ctx->npad;
    qsb_account(ctx, "k_tensor", (uint64_t)(ea + eb + eo) * ctx->ntraj); ctx->ops += ctx->ntraj;
    *out = st;
    return QSB_OK;
}

// =============================================================================================
// apply!: strided 2x2 / 4x4 gate application (a4)
// =============================================================================================
struct GateArgs {
    double2 u[16];   // column-major 2^K x 2^K:  u[r + c*2^K]
    int pos[2];      // target bit positions, operator factor order
    int sorted[4];   // all affected bit positions ascending (ket: K; dm: 2K)
};

__device__ __forceinline__ bool lane_on(const uint32_t* mask, int64_t t) { return mask == nullptr || ((mask[t >> 5] >> (t & 31)) & 1u); }

template <int K>
__global__ void k_apply_ket(double2* d, int64_t npad, size_t groups, GateArgs a, const uint32_t* __restrict__ mask) {
    int64_t t = blockIdx.x * (int64_t)TPB + threadIdx.x;
    if (t >= npad || !lane_on(mask, t)) return;
    constexpr int M = 1 << K;
    for (size_t g = blockIdx.y; g < groups; g += gridDim.y) {
        size_t base = g;
#pragma unroll
        for (int i = 0; i < K; ++i) base = qsb_insert0(base, a.sorted[i]);
        double2 x[M], y[M];
#pragma unroll
        for (int i = 0; i < M; ++i) {
            size_t e = base;
#pragma unroll
            for (int b = 0; b < K; ++b) e |= (size_t)((i >> b) & 1) << a.pos[b];
            x[i] = d[e * npad + t];
        }
#pragma unroll
        for (int r = 0; r < M; ++r) {
            double2 acc = make_double2(0, 0);
#pragma unroll
            for (int c = 0; c < M; ++c) cfma(acc, a.u[r + c * M], x[c]);
            y[r] = acc;
        }
#pragma unroll
        for (int i = 0; i < M; ++i) {
            size_t e = base;
#pragma unroll
            for (int b = 0; b < K; ++b) e |= (size_t)((i >> b) & 1) << a.pos[b];
            d[e * npad + t] = y[i];
        }
    }
}

// rho <- U rho U^dagger on the 2^K x 2^K block addressed by row bits pos[] and column bits pos[]+n
template <int K>
__global__ void __launch_bounds__(TPB) k_apply_dm(double2* d, int64_t npad, size_t groups, int n, GateArgs a, const uint32_t* __restrict__ mask) {
    int64_t t = blockIdx.x * (int64_t)TPB + threadIdx.x;
    if (t >= npad || !lane_on(mask, t)) return;
    constexpr int M = 1 << K;
    for (size_t g = blockIdx.y; g < groups; g += gridDim.y) {
        size_t base = g;
#pragma unroll
        for (int i = 0; i < 2 * K; ++i) base = qsb_insert0(base, a.sorted[i]);
        double2 x[M][M];  // x[row][col]
#pragma unroll
        for (int r = 0; r < M; ++r)
#pragma unroll
            for (int c = 0; c < M; ++c) {
                size_t e = base;
#pragma unroll
                for (int b = 0; b < K; ++b) e |= ((size_t)((r >> b) & 1) << a.pos[b]) | ((size_t)((c >> b) & 1) << (a.pos[b] + n));
                x[r][c] = d[e * npad + t];
            }
        double2 y[M][M];
#pragma unroll
        for (int c = 0; c < M; ++c)  // y = U x
#pragma unroll
            for (int r = 0; r < M; ++r) {
                double2 acc = make_double2(0, 0);
#pragma unroll
                for (int m = 0; m < M; ++m) cfma(acc, a.u[r + m * M], x[m][c]);
                y[r][c] = acc;
            }
#pragma unroll
        for (int r = 0; r < M; ++r)  // out = y U^dagger : out[r][c] = sum_m y[r][m] conj(U[c][m])
#pragma unroll
            for (int c = 0; c < M; ++c) {
                double2 acc = make_double2(0, 0);
#pragma unroll
                for (int m = 0; m < M; ++m) { double2 u = a.u[c + m * M]; acc.x += y[r][m].x * u.x + y[r][m].y * u.y; acc.y += y[r][m].y * u.x - y[r][m].x * u.y; }
                size_t e = base;
#pragma unroll
                for (int b = 0; b < K; ++b) e |= ((size_t)((r >> b) & 1) << a.pos[b]) | ((size_t)((c >> b) & 1) << (a.pos[b] + n));
                d[e * npad + t] = acc;
            }
    }
}

// ---- the same 4x4 conjugation with the 16 rows of a group staged through shared memory by the bulk-copy engine --------
// (north_star: "large registers stream through TMA-staged tiles").  A block owns 128 consecutive trajectories, so element e
// of a group is ONE contiguous 2 KB row d[e * npad + t0 .. t0 + 128): a stage is 16 rows = 32 KB moved by 16
// cp.async.bulk copies that complete on an mbarrier; two stages, so the rows of group g + stride are in flight while group g
// is computed; the results go back with cp.async.bulk shared -> global.  A thread only touches its own 16-byte column of
// every row, so the update is in place in the stage and needs no barrier between load and compute.
// Measured against the register path in profiles/ (tools/tma_experiment.py); selected with QSB_APPLY_TMA=1.
#define TMA_ROWS 16
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, int count) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count)); }
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) { asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory"); }
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    asm volatile("{\n.reg .pred p;\nWAIT_%=:\nmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n@p bra DONE_%=;\nbra WAIT_%=;\nDONE_%=:\n}" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* dst_smem, const void* src_gmem, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst_smem)), "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void bulk_s2g(void* dst_gmem, const void* src_smem, uint32_t bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst_gmem), "r"(smem_u32(src_smem)), "r"(bytes) : "memory");
}
__global__ void __launch_bounds__(TPB) k_apply_dm2_tma(double2* d, int64_t npad, size_t groups, int n, GateArgs a) {
    extern __shared__ __align__(128) unsigned char tma_smem[];
    double2* const stage0 = reinterpret_cast<double2*>(tma_smem);
#define STAGE(s) (stage0 + (s) * (TMA_ROWS * TPB))
    __shared__ __align__(8) uint64_t full[2];
    const size_t xtiles = (size_t)((npad + TPB - 1) / TPB);
    // lanes 0..15 of warp 0 each own ONE row of the stage: lane i computes the address of row i = r + 4 c and issues its copy,
    // so the 32 bulk copies of an item are not serialised behind one thread's address arithmetic
    const int row = threadIdx.x & 15;
    const bool issuer = threadIdx.x < TMA_ROWS;
    size_t row_bits = 0;
    {
        const int r = row & 3, c = row >> 2;
#pragma unroll
        for (int b = 0; b < 2; ++b) row_bits |= ((size_t)((r >> b) & 1) << a.pos[b]) | ((size_t)((c >> b) & 1) << (a.pos[b] + n));
    }
    auto row_ptr = [&](size_t g, int64_t t0) {
        size_t base = g;
#pragma unroll
        for (int j = 0; j < 4; ++j) base = qsb_insert0(base, a.sorted[j]);
        return d + (base | row_bits) * npad + t0;
    };
    auto bytes_of = [&](int64_t t0) { const int64_t left = npad - t0; return (uint32_t)((left < TPB ? left : TPB) * sizeof(double2)); };   // npad % 32 == 0: 16-byte granularity holds
    if (threadIdx.x == 0) { mbar_init(&full[0], 1); mbar_init(&full[1], 1); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
    __syncthreads();
    auto issue_load = [&](size_t g, int64_t t0, int s) {   // called by warp 0 (converged)
        const uint32_t rb = bytes_of(t0);
        if (threadIdx.x == 0) mbar_expect_tx(&full[s], rb * TMA_ROWS);
        __syncwarp();
        if (issuer) bulk_g2s(STAGE(s) + row * TPB, row_ptr(g, t0), rb, &full[s]);
    };
    // persistent 1-D grid: block b takes the (tile, group) items w = b, b + gridDim.x, ... with w = xt + xtiles * g
    size_t g = blockIdx.x / xtiles;
    int64_t xt = (int64_t)(blockIdx.x % xtiles);
    const size_t dg = gridDim.x / xtiles; const int64_t dx = (int64_t)(gridDim.x % xtiles);   // the stride, split once
    if (threadIdx.x < 32 && g < groups) issue_load(g, xt * TPB, 0);
    uint32_t phases = 0;   // bit s = parity the next wait on full[s] expects
    for (int s = 0; g < groups; s ^= 1) {
        size_t gn = g + dg; int64_t xn = xt + dx;
        if (xn >= (int64_t)xtiles) { xn -= xtiles; ++gn; }
        if (threadIdx.x < 32 && gn < groups) {
            // the other stage was stored by the previous trip: its shared-memory source must have been read before it is refilled
            if (issuer) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
            issue_load(gn, xn * TPB, s ^ 1);
        }
        mbar_wait(&full[s], (phases >> s) & 1u);
        phases ^= 1u << s;
        const int64_t t0 = xt * TPB;
        double2* col = STAGE(s) + threadIdx.x;
        if ((int64_t)threadIdx.x < npad - t0) {
            double2 x[4][4], y[4][4];
#pragma unroll
            for (int i = 0; i < 16; ++i) x[i & 3][i >> 2] = col[i * TPB];
#pragma unroll
            for (int c = 0; c < 4; ++c)
#pragma unroll
                for (int r = 0; r < 4; ++r) {
                    double2 acc = make_double2(0, 0);
#pragma unroll
                    for (int m = 0; m < 4; ++m) cfma(acc, a.u[r + m * 4], x[m][c]);
                    y[r][c] = acc;
                }
#pragma unroll
            for (int r = 0; r < 4; ++r)
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                    double2 acc = make_double2(0, 0);
#pragma unroll
                    for (int m = 0; m < 4; ++m) { double2 u = a.u[c + m * 4]; acc.x += y[r][m].x * u.x + y[r][m].y * u.y; acc.y += y[r][m].y * u.x - y[r][m].x * u.y; }
                    col[(r + 4 * c) * TPB] = acc;
                }
        }
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // generic-proxy writes -> visible to the bulk-copy engine
        __syncthreads();
        if (issuer) {
            bulk_s2g(row_ptr(g, t0), STAGE(s) + row * TPB, bytes_of(t0));
            asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        }
        g = gn; xt = xn;
    }
    if (issuer) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
#undef STAGE
}

static void sort_ints(int* v, int n) { for (int i = 1; i < n; ++i) for (int j = i; j > 0 && v[j] < v[j - 1]; --j) { int s = v[j]; v[j] = v[j - 1]; v[j - 1] = s; } }

int qsbd_apply(qsb_state* st, const int* idx, int k, const double* U, const qsb_mask* mask) {
    qsb_ctx* ctx = st->ctx;
    GateArgs a; memset(&a, 0, sizeof(a));
    int M = 1 << k;
    for (int i = 0; i < M * M; ++i) a.u[i] = make_double2(U[2 * i], U[2 * i + 1]);
    for (int i = 0; i < k; ++i) a.pos[i] = idx[i];
    const uint32_t* mw = mask ? mask->w : nullptr;
    int n = st->n;
    if (ctx->rec) {
        TabOp* o = qsb_rec_push(ctx, DOP_APPLY); o->n = n; o->k = k; o->q[0] = idx[0]; o->q[1] = k > 1 ? idx[1] : 0;
        o->s[0] = st->vbase; o->m[0] = mask ? mask->vplane : -1; o->obs = qsb_rec_daux(ctx, U, 2 * M * M);
        qsb_account(ctx, qsb_is_ketlike(st->kind) ? "k_apply_ket" : "k_apply_dm", (uint64_t)32 * ((uint64_t)1 << (qsb_is_ketlike(st->kind) ? n : 2 * n)) * ctx->ntraj);
        ctx->ops += ctx->ntraj;
        return QSB_OK;
    }
    if (qsb_is_ketlike(st->kind)) {
        for (int i = 0; i < k; ++i) a.sorted[i] = idx[i];
        sort_ints(a.sorted, k);
        size_t groups = (size_t)1 << (n - k);
        Grid2 g = plan_grid(ctx, groups);
        QSB_PRE(ctx, "k_apply_ket");
        if (k == 1) k_apply_ket<1><<<g.grid, TPB, 0, ctx->stream>>>((double2*)st->data, ctx->npad, groups, a, mw);
        else k_apply_ket<2><<<g.grid, TPB, 0, ctx->stream>>>((double2*)st->data, ctx->npad, groups, a, mw);
        qsb_account(ctx, "k_apply_ket", (uint64_t)32 * ((uint64_t)1 << n) * ctx->ntraj);
    } else {
        for (int i = 0; i < k; ++i) { a.sorted[i] = idx[i]; a.sorted[k + i] = idx[i] + n; }
        sort_ints(a.sorted, 2 * k);
        size_t groups = (size_t)1 << (2 * n - 2 * k);
        Grid2 g = plan_grid(ctx, groups);
        QSB_PRE(ctx, "k_apply_dm");
        const char* tma_env = getenv("QSB_APPLY_TMA");   // experiment switch, read per call so that a test can compare both paths
        const int use_tma = tma_env ? atoi(tma_env) : 0;
        if (k == 1) k_apply_dm<1><<<g.grid, TPB, 0, ctx->stream>>>((double2*)st->data, ctx->npad, groups, n, a, mw);
        else if (use_tma && !mw) {
            const int smem = 2 * TMA_ROWS * TPB * (int)sizeof(double2);
            static int configured[64];
            if (ctx->device < 64 && !configured[ctx->device]) { QSB_CUDA(cudaFuncSetAttribute(k_apply_dm2_tma, cudaFuncAttributeMaxDynamicSharedMemorySize, smem)); configured[ctx->device] = 1; }
            // persistent 1-D grid: `use_tma` blocks of 64 KB per SM (2 measured best), every block walks its share of the (tile, group) items
            size_t items = (size_t)((ctx->npad + TPB - 1) / TPB) * groups, nb = (size_t)ctx->sm_count * (use_tma > 3 ? 3 : use_tma);
            if (nb > items) nb = items;
            k_apply_dm2_tma<<<(unsigned)nb, TPB, smem, ctx->stream>>>((double2*)st->data, ctx->npad, groups, n, a);
        }
        else k_apply_dm<2><<<g.grid, TPB, 0, ctx->stream>>>((double2*)st->data, ctx->npad, groups, n, a, mw);
        qsb_account(ctx, "k_apply_dm", (uint64_t)32 * ((uint64_t)1 << (2 * n)) * ctx->ntraj);
    }
    QSB_LAUNCH_CHECK(ctx);
    ctx->ops += ctx->ntraj;
    return QSB_OK;
}

// =============================================================================================
// uptotime! on a density matrix: closed-form single-qubit channels (a5)
//   e00' = a e00 + b e11 ; e11' = c e11 + dd e00 ; e01' = f e01 ; e10' = f e10
// =============================================================================================
struct ChanArgs { double a, b, c, dd, f; int q, n; int offdiag_only; };

__global__ void k_channel_dm(double2* d, int64_t npad, size_t groups, ChanArgs p) {
    int64_t t = blockIdx.x * (int64_t)TPB + threadIdx.x;
    if (t >= npad) return;
    size_t rbit = (size_t)1 << p.q, cbit = (size_t)1 << (p.q + p.n);
    for (size_t g = blockIdx.y; g < groups; g += gridDim.y) {
        size_t base = qsb_insert0(qsb_insert0(g, p.q), p.q + p.n);
        double2* p01 = d + (base | cbit) * npad + t;  // row bit 0, col bit 1
        double2* p10 = d + (base | rbit) * npad + t;
        if (p.offdiag_only) {
            double2 e01 = *p01, e10 = *p10;
            *p01 = cscale(e01, p.f); *p10 = cscale(e10, p.f);
        } else {
            double2* p00 = d + base * npad + t;
            double2* p11 = d + (base | rbit | cbit) * npad + t;
            double2 e00 = *p00, e01 = *p01, e10 = *p10, e11 = *p11;
            *p00 = make_double2(p.a * e00.x + p.b * e11.x, p.a * e00.y + p.b * e11.y);
            *p11 = make_double2(p.c * e11.x + p.dd * e00.x, p.c * e11.y + p.dd * e00.y);
            *p01 = cscale(e01, p.f); *p10 = cscale(e10, p.f);
        }
    }
}

// ---- reduced single-qubit density matrix (shared by measurement and sampled Kraus) ----------------
// partial[(s*4+v)*npad + t], v: 0 rho00, 1 rho11, 2 Re rho01, 3 Im rho01
template <bool DM>
__global__ void k_rdm1(const double2* __restrict__ d, int64_t npad, int n, int q, size_t range, int nsplit, double* __restrict__ partial) {
    int64_t t = blockIdx.x * (int64_t)TPB + threadIdx.x;
    if (t >= npad) return;
    int s = blockIdx.y;
    size_t lo = range * s / nsplit, hi = range * (s + 1) / nsplit;
    size_t bit = (size_t)1 << q;
    double m00 = 0, m11 = 0, mre = 0, mim = 0;
#pragma unroll 4
    for (size_t aidx = lo; aidx < hi; ++aidx) {
        size_t b0 = qsb_insert0(aidx, q), b1 = b0 | bit;
        if (DM) {
            double2 e00 = d[(b0 + (b0 << n)) * npad + t];
            double2 e11 = d[(b1 + (b1 << n)) * npad + t];
            double2 e01 = d[(b0 + (b1 << n)) * npad + t];
            m00 += e00.x; m11 += e11.x; mre += e01.x; mim += e01.y;
        } else {
            double2 p0 = d[b0 * npad + t], p1 = d[b1 * npad + t];
            m00 += p0.x * p0.x + p0.y * p0.y; m11 += p1.x * p1.x + p1.y * p1.y;
            double2 c = cmulc(p0, p1);
            mre += c.x; mim += c.y;
        }
    }
    size_t o = (size_t)s * 4 * npad + t;
    partial[o] = m00; partial[o + npad] = m11; partial[o + 2 * npad] = mre; partial[o + 3 * npad] = mim;
}
static int launch_rdm1(qsb_state* st, int q, int* nsplit_out) {
    qsb_ctx* ctx = st->ctx;
    size_t range = (size_t)1 << (st->n - 1);
    int nsplit = plan_split(ctx, range);
    // layout: [partials nsplit*4*npad][scale npad][sel npad (as double-sized slots for alignment)]
    QSB_TRY(qsb_ensure_scratch(ctx, ((size_t)nsplit * 4 + 2) * ctx->npad * sizeof(double)));
    dim3 grid((unsigned)((ctx->npad + TPB - 1) / TPB), nsplit, 1);
    QSB_PRE(ctx, "k_rdm1");
    if (st->kind == QSB_DM) k_rdm1<true><<<grid, TPB, 0, ctx->stream>>>((const double2*)st->data, ctx->npad, st->n, q, range, nsplit, ctx->scratch);
    else k_rdm1<false><<<grid, TPB, 0, ctx->stream>>>((const double2*)st->data, ctx->npad, st->n, q, range, nsplit, ctx->scratch);
    QSB_LAUNCH_CHECK(ctx);
    // reads the ket once (dm: the three touched entries of each 2x2 block of qubit q on the diagonal), writes 4 partial sums per split
    qsb_account(ctx, "k_rdm1", ((uint64_t)16 * (st->kind == QSB_DM ? 3 * range : 2 * range) + (uint64_t)32 * nsplit) * ctx->ntraj);
    *nsplit_out = nsplit;
    return QSB_OK;
}

// ---- sampled Kraus branch on a trajectory ket (a6) ---------------------------------------------------
__global__ void k_mc_finalize(const double* __restrict__ partial, int nsplit, int64_t npad, KrausArgs ka,
                              uint64_t seed, uint64_t traj0, uint64_t draw_id, double* __restrict__ scale, int* __restrict__ sel) {
    int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (t >= npad) return;
    double m00 = 0, m11 = 0, mre = 0, mim = 0;
    for (int s = 0; s < nsplit; ++s) {
        size_t o = (size_t)s * 4 * npad + t;
        m00 += partial[o]; m11 += partial[o + npad]; mre += partial[o + 2 * npad]; mim += partial[o + 3 * npad];
    }
    // P_k = sum_r | K[r][0] psi0 + K[r][1] psi1 |^2 with M[s][s'] = sum psi_s conj(psi_s')
    double P[4], total = 0;
    for (int k = 0; k < ka.nk; ++k) {
        double acc = 0;
        for (int r = 0; r < 2; ++r) {
            double2 k0 = ka.k[k][r], k1 = ka.k[k][r + 2];
            double2 x = cmulc(k0, k1);  // K[r][0] conj(K[r][1])
            acc += (k0.x * k0.x + k0.y * k0.y) * m00 + (k1.x * k1.x + k1.y * k1.y) * m11 + 2.0 * (x.x * mre - x.y * mim);
        }
        P[k] = acc; total += acc;
    }
    double thr = qsb_draw(seed, traj0 + (uint64_t)t, draw_id) * total, cum = 0;
    int choice = ka.nk - 1;
    for (int k = 0; k < ka.nk; ++k) { cum += P[k]; if (cum > thr) { choice = k; break; } }
    sel[t] = choice;
    scale[t] = 1.0 / sqrt(P[choice]);
}
__global__ void k_apply_kraus_sel(double2* d, int64_t npad, size_t groups, int q, KrausArgs ka, const double* __restrict__ scale, const int* __restrict__ sel) {
    int64_t t = blockIdx.x * (int64_t)TPB + threadIdx.x;
    if (t >= npad) return;
    int k = sel[t];
    double sc = scale[t];
    double2 k00 = cscale(ka.k[k][0], sc), k10 = cscale(ka.k[k][1], sc), k01 = cscale(ka.k[k][2], sc), k11 = cscale(ka.k[k][3], sc);
    size_t bit = (size_t)1 << q;
    for (size_t g = blockIdx.y; g < groups; g += gridDim.y) {
        size_t b0 = qsb_insert0(g, q);
        double2 p0 = d[b0 * npad + t], p1 = d[(b0 | bit) * npad + t];
        d[b0 * npad + t] = cadd(cmul(k00, p0), cmul(k01, p1));
        d[(b0 | bit) * npad + t] = cadd(cmul(k10, p0), cmul(k11, p1));
    }
}

// krausops tables, uptotime.jl:106-173 (column-major 2x2: K[r + 2c])
static int host_kraus(int bg, const double* prm, double dt, KrausArgs* ka) {
    memset(ka, 0, sizeof(*ka));
    auto set = [&](int k, double r00, double r10, double r01, double r11) {
        ka->k[k][0] = make_double2(r00, 0); ka->k[k][1] = make_double2(r10, 0); ka->k[k][2] = make_double2(r01, 0); ka->k[k][3] = make_double2(r11, 0);
    };
    if (bg == QSB_BG_T1) {
        double p = exp(-dt / prm[0]);
        set(0, 0, 0, sqrt(1 - p), 0); set(1, 1, 0, 0, sqrt(p)); ka->nk = 2;
    } else if (bg == QSB_BG_T2) {
        double p = 1 - exp(-dt / prm[0]);
        set(0, sqrt(1 - p / 2), 0, 0, sqrt(1 - p / 2)); set(1, sqrt(p / 2), 0, 0, -sqrt(p / 2)); ka->nk = 2;
    } else if (bg == QSB_BG_DEPOL) {
        double p = 1 - exp(-dt / prm[0]);
        double a = sqrt(1 - 3 * p / 4), b = sqrt(p) / 2;
        set(0, a, 0, 0, a); set(1, 0, b, b, 0);
        ka->k[2][0] = make_double2(0, 0); ka->k[2][1] = make_double2(0, b); ka->k[2][2] = make_double2(0, -b); ka->k[2][3] = make_double2(0, 0);
        set(3, b, 0, 0, -b); ka->nk = 4;
    } else {
        double t1 = prm[0], t2 = prm[1];
        double p = exp(-dt / t1);
        double tphi_inv = 1 / t2 - 1 / (2 * t1);
        if (tphi_inv <= 0) { set(0, 0, 0, sqrt(1 - p), 0); set(1, 1, 0, 0, sqrt(p)); ka->nk = 2; }
        else {
            double pphi = 1 - exp(-dt * tphi_inv);
            double f0 = sqrt(1 - pphi / 2), f1 = sqrt(pphi / 2);
            // [F*E for F in dephase for E in T1]: F0*E0, F0*E1, F1*E0, F1*E1 ; F1 = f1*Z
            set(0, 0, 0, f0 * sqrt(1 - p), 0); set(1, f0, 0, 0, f0 * sqrt(p));
            set(2, 0, 0, f1 * sqrt(1 - p), 0); set(3, f1, 0, 0, -f1 * sqrt(p)); ka->nk = 4;
        }
    }
    return QSB_OK;
}

// One sampled branch of a single-qubit Kraus set on a trajectory ket (uptotime.jl:20-42): probabilities from the reduced
// 2x2 density matrix of the qubit, one draw, K_sel psi / |K_sel psi|.  Also the MCKet branch of qsb_evolve (qsb_evolve.cu).
int qsbd_kraus_sampled(qsb_state* st, int idx, const KrausArgs& ka, uint64_t draw_id) {
    qsb_ctx* ctx = st->ctx;
    const int n = st->n;
    if (ctx->rec) QSB_FAIL(QSB_ERR_UNSUPPORTED, "trajectory kets (QuantumMCRepr) cannot be recorded");
    int nsplit;
    QSB_TRY(launch_rdm1(st, idx, &nsplit));
    double* scale = ctx->scratch + (size_t)nsplit * 4 * ctx->npad;
    int* sel = (int*)(scale + ctx->npad);
    QSB_PRE(ctx, "k_mc_finalize");
    k_mc_finalize<<<(unsigned)((ctx->npad + 255) / 256), 256, 0, ctx->stream>>>(ctx->scratch, nsplit, ctx->npad, ka, ctx->seed, (uint64_t)ctx->traj_offset, draw_id, scale, sel);
    QSB_LAUNCH_CHECK(ctx);
    size_t groups = (size_t)1 << (n - 1);
    Grid2 g = plan_grid(ctx, groups);
    QSB_PRE(ctx, "k_apply_kraus_sel");
    k_apply_kraus_sel<<<g.grid, TPB, 0, ctx->stream>>>((double2*)st->data, ctx->npad, groups, idx, ka, scale, sel);
    QSB_LAUNCH_CHECK(ctx);
    qsb_account(ctx, "k_apply_kraus_sel", (uint64_t)32 * ((uint64_t)1 << n) * ctx->ntraj); ctx->ops += ctx->ntraj; ctx->draws++;
    return QSB_OK;
}

int qsbd_channel(qsb_state* st, int idx, int bg, const double* prm, double dt, uint64_t draw_id) {
    qsb_ctx* ctx = st->ctx;
    int n = st->n;
    if (st->kind == QSB_MCKET) {
        KrausArgs ka;
        host_kraus(bg, prm, dt, &ka);
        return qsbd_kraus_sampled(st, idx, ka, draw_id);
    }
    if (st->kind == QSB_KET) QSB_TRY(qsbd_promote_dm(st));  // uptotime.jl:1-4
    ChanArgs p; memset(&p, 0, sizeof(p));
    p.q = idx; p.n = n;
    if (bg == QSB_BG_T1) {
        double e = exp(-dt / prm[0]);
        p.a = 1; p.b = 1 - e; p.c = e; p.dd = 0; p.f = sqrt(e);
    } else if (bg == QSB_BG_T2) {
        double pp = 1 - exp(-dt / prm[0]);
        p.a = 1; p.b = 0; p.c = 1; p.dd = 0; p.f = 1 - pp; p.offdiag_only = 1;
    } else if (bg == QSB_BG_DEPOL) {
        double pp = 1 - exp(-dt / prm[0]);
        p.a = 1 - pp / 2; p.b = pp / 2; p.c = 1 - pp / 2; p.dd = pp / 2; p.f = 1 - pp;
    } else {
        double e = exp(-dt / prm[0]);
        double tphi_inv = 1 / prm[1] - 1 / (2 * prm[0]);
        p.a = 1; p.b = 1 - e; p.c = e; p.dd = 0; p.f = sqrt(e);
        if (tphi_inv > 0) { double pphi = 1 - exp(-dt * tphi_inv); p.f *= 1 - pphi; }
    }
    if (ctx->rec) {
        const double prm5[6] = {p.a, p.b, p.c, p.dd, p.f, (double)p.offdiag_only};
        TabOp* o = qsb_rec_push(ctx, DOP_CHANNEL); o->n = n; o->q[0] = idx; o->s[0] = st->vbase; o->iv[0] = p.offdiag_only; o->obs = qsb_rec_daux(ctx, prm5, 6);
        qsb_account(ctx, "k_channel_dm", (uint64_t)(p.offdiag_only ? 16 : 32) * ((uint64_t)1 << (2 * n)) * ctx->ntraj); ctx->ops += ctx->ntraj;
        return QSB_OK;
    }
    size_t groups = (size_t)1 << (2 * n - 2);
    Grid2 g = plan_grid(ctx, groups);
    QSB_PRE(ctx, "k_channel_dm");
    k_channel_dm<<<g.grid, TPB, 0, ctx->stream>>>((double2*)st->data, ctx->npad, groups, p);
    QSB_LAUNCH_CHECK(ctx);
    qsb_account(ctx, "k_channel_dm", (uint64_t)(p.offdiag_only ? 16 : 32) * ((uint64_t)1 << (2 * n)) * ctx->ntraj); ctx->ops += ctx->ntraj;
    return QSB_OK;
}

// =============================================================================================
// project_traceout! (a8): probabilities from the reduced 2x2 density matrix, one draw, project + shrink
// =============================================================================================
struct MeasArgs { double2 psi[2][2]; };
__global__ void k_measure_finalize(const double* __restrict__ partial, int nsplit, int64_t npad, MeasArgs ma, int is_ket, int check_norm,
                                   uint64_t seed, uint64_t traj0, uint64_t draw_id, const uint32_t* __restrict__ alive,
                                   double* __restrict__ scale, int* __restrict__ sel, uint32_t* __restrict__ outcome2,
                                   long long* __restrict__ counts, int* __restrict__ flags) {
    int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (t >= npad) return;
    double m00 = 0, m11 = 0, mre = 0, mim = 0;
    for (int s = 0; s < nsplit; ++s) {
        size_t o = (size_t)s * 4 * npad + t;
        m00 += partial[o]; m11 += partial[o + npad]; mre += partial[o + 2 * npad]; mim += partial[o + 3 * npad];
    }
    double P[2];
#pragma unroll
    for (int j = 0; j < 2; ++j) {
        double2 a0 = ma.psi[j][0], a1 = ma.psi[j][1];
        double2 x = cconjmul(a0, a1);  // conj(psi0) psi1
        // <psi|M|psi> = |a0|^2 M00 + |a1|^2 M11 + 2 Re( conj(a0) M01 a1 )
        P[j] = (a0.x * a0.x + a0.y * a0.y) * m00 + (a1.x * a1.x + a1.y * a1.y) * m11 + 2.0 * (x.x * mre - x.y * mim);
    }
    bool al = (alive[t >> 5] >> (t & 31)) & 1u;
    double u = qsb_draw(seed, traj0 + (uint64_t)t, draw_id);
    int j = (P[0] >= u) ? 0 : 1;  // findfirst(>=(rand()), cumsum(P)), quantumoptics.jl:54,61
    sel[t] = j;
    scale[t] = is_ket ? 1.0 / sqrt(P[j]) : 1.0 / P[j];
    if (check_norm && al && fabs(P[0] + P[1] - 1.0) > 1.5e-8) atomicOr(flags, 1);
    uint32_t w2 = __ballot_sync(0xffffffffu, al && j == 1);
    uint32_t wa = __ballot_sync(0xffffffffu, al);
    if ((threadIdx.x & 31) == 0) {
        outcome2[t >> 5] = w2;
        if (counts) {
            int na = __popc(wa), n2 = __popc(w2);
            if (na) atomicAdd((unsigned long long*)&counts[0], (unsigned long long)na);
            if (na - n2) atomicAdd((unsigned long long*)&counts[1], (unsigned long long)(na - n2));
            if (n2) atomicAdd((unsigned long long*)&counts[2], (unsigned long long)n2);
        }
    }
}
template <bool DM>
__global__ void k_project(double2* __restrict__ out, const double2* __restrict__ d, int64_t npad, int n, int q, size_t groups, MeasArgs ma,
                          const double* __restrict__ scale, const int* __restrict__ sel) {
    int64_t t = blockIdx.x * (int64_t)TPB + threadIdx.x;
    if (t >= npad) return;
    int j = sel[t];
    double sc = scale[t];
    double2 c0 = make_double2(ma.psi[j][0].x, -ma.psi[j][0].y), c1 = make_double2(ma.psi[j][1].x, -ma.psi[j][1].y);  // conj(psi)
    size_t bit = (size_t)1 << q;
    if (DM) {
        double2 p0 = ma.psi[j][0], p1 = ma.psi[j][1];
        size_t Dn = (size_t)1 << (n - 1);
        for (size_t g = blockIdx.y; g < groups; g += gridDim.y) {
            size_t a = g & (Dn - 1), b = g >> (n - 1);
            size_t ra = qsb_insert0(a, q), cb = qsb_insert0(b, q);
            double2 e00 = d[(ra + (cb << n)) * npad + t], e10 = d[((ra | bit) + (cb << n)) * npad + t];
            double2 e01 = d[(ra + ((cb | bit) << n)) * npad + t], e11 = d[((ra | bit) + ((cb | bit) << n)) * npad + t];
            double2 col0 = cadd(cmul(c0, e00), cmul(c1, e10));  // <psi| rho[:, (b,0)]
            double2 col1 = cadd(cmul(c0, e01), cmul(c1, e11));
            double2 v = cadd(cmul(col0, p0), cmul(col1, p1));
            out[g * npad + t] = cscale(v, sc);
        }
    } else {
        for (size_t g = blockIdx.y; g < groups; g += gridDim.y) {
            size_t b0 = qsb_insert0(g, q);
            double2 v = cadd(cmul(c0, d[b0 * npad + t]), cmul(c1, d[(b0 | bit) * npad + t]));
            out[g * npad + t] = cscale(v, sc);
        }
    }
}

static int fetch_counts(qsb_ctx* ctx, int64_t* counts_host) {
    long long h[3];
    QSB_CUDA(cudaMemcpyAsync(h, ctx->counts_tmp, sizeof(h), cudaMemcpyDeviceToHost, ctx->stream));
    QSB_TRY(qsb_ctx_sync(ctx));
    counts_host[0] = h[0]; counts_host[1] = h[1]; counts_host[2] = h[2];
    return QSB_OK;
}

int qsbd_measure(qsb_state* st, int idx, const double* kets, uint64_t draw_id, qsb_mask* outcome2, int64_t* counts_host) {
    qsb_ctx* ctx = st->ctx;
    int n = st->n;
    bool ket = qsb_is_ketlike(st->kind);
    MeasArgs ma;
    for (int j = 0; j < 2; ++j) for (int a = 0; a < 2; ++a) ma.psi[j][a] = make_double2(kets[4 * j + 2 * a], kets[4 * j + 2 * a + 1]);
    if (ctx->rec) {
        // the qubit is projected out IN PLACE: the state keeps its base, the tail of its block is recycled
        TabOp* o = qsb_rec_push(ctx, DOP_MEASURE); o->n = n; o->q[0] = idx; o->draw = draw_id; o->obs = qsb_rec_daux(ctx, kets, 8);
        o->s[0] = st->vbase; o->al[0] = st->valive; o->m[0] = outcome2->vplane;
        uint64_t e_in = ket ? ((uint64_t)1 << n) : dm_elems(n);
        qsb_account(ctx, "k_project", (uint64_t)16 * (e_in + e_in / (ket ? 2 : 4)) * ctx->ntraj); ctx->ops += ctx->ntraj;
        qsb_rec_shrink_to(ctx, 1, st->vbase, n > 1 ? 1 << (2 * n - 2) : 1);
        st->n = n - 1; st->bytes = n > 1 ? qsb_state_bytes(ctx, st->kind, n - 1) : 0;
        rec_relocate_small(ctx, st);
        return QSB_OK;
    }
    int nsplit;
    QSB_TRY(launch_rdm1(st, idx, &nsplit));
    double* scale = ctx->scratch + (size_t)nsplit * 4 * ctx->npad;
    int* sel = (int*)(scale + ctx->npad);
    if (counts_host) QSB_CUDA(cudaMemsetAsync(ctx->counts_tmp, 0, 3 * sizeof(long long), ctx->stream));
    QSB_PRE(ctx, "k_measure_finalize");
    k_measure_finalize<<<(unsigned)((ctx->npad + 255) / 256), 256, 0, ctx->stream>>>(ctx->scratch, nsplit, ctx->npad, ma, ket ? 1 : 0, 1,
        ctx->seed, (uint64_t)ctx->traj_offset, draw_id, st->alive, scale, sel, outcome2->w, counts_host ? ctx->counts_tmp : nullptr, ctx->dev_flags);
    QSB_LAUNCH_CHECK(ctx);
    uint64_t elems_in = ket ? ((uint64_t)1 << n) : ((uint64_t)1 << (2 * n));
    if (n == 1) {
        // quantumoptics.jl:48-55: the state is consumed ("j, nothing")
        QSB_TRY(qsb_dev_free(ctx, st->data));
        st->data = nullptr; st->bytes = 0; st->n = 0;
        QSB_TRY(qsb_dev_alloc(ctx, &st->data, 16));
    } else {
        size_t nbytes = qsb_state_bytes(ctx, st->kind, n - 1);
        void* nd;
        QSB_TRY(qsb_dev_alloc(ctx, &nd, nbytes));
        size_t groups = ket ? (size_t)1 << (n - 1) : (size_t)1 << (2 * n - 2);
        Grid2 g = plan_grid(ctx, groups);
        QSB_PRE(ctx, "k_project");
        if (ket) k_project<false><<<g.grid, TPB, 0, ctx->stream>>>((double2*)nd, (const double2*)st->data, ctx->npad, n, idx, groups, ma, scale, sel);
        else k_project<true><<<g.grid, TPB, 0, ctx->stream>>>((double2*)nd, (const double2*)st->data, ctx->npad, n, idx, groups, ma, scale, sel);
        QSB_LAUNCH_CHECK(ctx);
        QSB_TRY(qsb_dev_free(ctx, st->data));
        st->data = nd; st->bytes = nbytes; st->n = n - 1;
    }
    qsb_account(ctx, "k_project", (uint64_t)16 * (elems_in + elems_in / (ket ? 2 : 4)) * ctx->ntraj); ctx->ops += ctx->ntraj;
    if (counts_host) QSB_TRY(fetch_counts(ctx, counts_host));
    return QSB_OK;
}

// =============================================================================================
// traceout! (a10)
// =============================================================================================
__global__ void k_ptrace_dm(double2* __restrict__ out, const double2* __restrict__ d, int64_t npad, int n, int q, size_t groups) {
    int64_t t = blockIdx.x * (int64_t)TPB + threadIdx.x;
    if (t >= npad) return;
    size_t bit = (size_t)1 << q, Dn = (size_t)1 << (n - 1);
    for (size_t g = blockIdx.y; g < groups; g += gridDim.y) {
        size_t a = g & (Dn - 1), b = g >> (n - 1);
        size_t ra = qsb_insert0(a, q), cb = qsb_insert0(b, q);
        out[g * npad + t] = cadd(d[(ra + (cb << n)) * npad + t], d[((ra | bit) + ((cb | bit) << n)) * npad + t]);
    }
}
__global__ void k_ptrace_ket(double2* __restrict__ out, const double2* __restrict__ d, int64_t npad, int n, int q, size_t groups) {
    int64_t t = blockIdx.x * (int64_t)TPB + threadIdx.x;
    if (t >= npad) return;
    size_t bit = (size_t)1 << q, Dn = (size_t)1 << (n - 1);
    for (size_t g = blockIdx.y; g < groups; g += gridDim.y) {
        size_t a = g & (Dn - 1), b = g >> (n - 1);
        size_t ra = qsb_insert0(a, q), cb = qsb_insert0(b, q);
        out[g * npad + t] = cadd(cmulc(d[ra * npad + t], d[cb * npad + t]), cmulc(d[(ra | bit) * npad + t], d[(cb | bit) * npad + t]));
    }
}
int qsbd_ptrace(qsb_state* st, int idx, uint64_t draw_id) {
    qsb_ctx* ctx = st->ctx;
    int n = st->n;
    if (st->kind == QSB_MCKET) {
        // mcket.jl:46-54: sampled projection on the canonical basis, outcome discarded
        qsb_mask* tmp;
        QSB_TRY(qsb_mask_alloc(ctx, &tmp));
        const double zk[8] = {1, 0, 0, 0, 0, 0, 1, 0};
        int rc = qsbd_measure(st, idx, zk, draw_id, tmp, nullptr);
        qsb_mask_free(tmp);
        ctx->draws++;
        return rc;
    }
    size_t nbytes = qsb_state_bytes(ctx, QSB_DM, n - 1);
    if (ctx->rec) {
        TabOp* o = qsb_rec_push(ctx, DOP_PTRACE); o->n = n; o->q[0] = idx; o->s[0] = st->vbase;
        if (st->kind == QSB_DM) qsb_account(ctx, "k_ptrace_dm", (uint64_t)16 * (dm_elems(n) / 2 + dm_elems(n - 1)) * ctx->ntraj);
        else qsb_account(ctx, "k_ptrace_ket", (uint64_t)16 * (((uint64_t)1 << n) + dm_elems(n - 1)) * ctx->ntraj);
        qsb_rec_shrink_to(ctx, 1, st->vbase, 1 << (2 * n - 2));
        st->bytes = nbytes; st->n = n - 1; st->kind = QSB_DM;
        ctx->ops += ctx->ntraj;
        rec_relocate_small(ctx, st);
        return QSB_OK;
    }
    void* nd;
    QSB_TRY(qsb_dev_alloc(ctx, &nd, nbytes));
    size_t groups = (size_t)1 << (2 * n - 2);
    Grid2 g = plan_grid(ctx, groups);
    if (st->kind == QSB_DM) {
        QSB_PRE(ctx, "k_ptrace_dm");
        k_ptrace_dm<<<g.grid, TPB, 0, ctx->stream>>>((double2*)nd, (const double2*)st->data, ctx->npad, n, idx, groups);
        qsb_account(ctx, "k_ptrace_dm", (uint64_t)16 * (((uint64_t)1 << (2 * n)) / 2 + ((uint64_t)1 << (2 * n - 2))) * ctx->ntraj);
    } else {
        QSB_PRE(ctx, "k_ptrace_ket");
        k_ptrace_ket<<<g.grid, TPB, 0, ctx->stream>>>((double2*)nd, (const double2*)st->data, ctx->npad, n, idx, groups);
        qsb_account(ctx, "k_ptrace_ket", (uint64_t)16 * (((uint64_t)1 << n) + ((uint64_t)1 << (2 * n - 2))) * ctx->ntraj);
    }
    QSB_LAUNCH_CHECK(ctx);
    QSB_TRY(qsb_dev_free(ctx, st->data));
    st->data = nd; st->bytes = nbytes; st->n = n - 1; st->kind = QSB_DM;
    ctx->ops += ctx->ntraj;
    return QSB_OK;
}

// =============================================================================================
// observable (a11): per-trajectory expectation values + ensemble statistics
// =============================================================================================
struct ExpArgs {
    int mode;          // 0 Pauli string, 1 sparse projector, 2 k-qubit operator
    int n, k;
    int is_ket;
    size_t xmask, zmask;      // mode 0
    int nnz;                  // mode 1: entries (index, amplitude) in obs buffer
    int pos[8];               // mode 2: operator factor -> bit position
    int sorted[8];
    const void* buf;          // device: mode 1 -> {size_t idx; double2 amp}[nnz]; mode 2 -> double2 O[4^k] column-major
};
struct SparseAmp { unsigned long long idx; double2 amp; };

__global__ void k_expect(const double2* __restrict__ d, int64_t npad, ExpArgs a, size_t range, int nsplit, double* __restrict__ partial) {
    int64_t t = blockIdx.x * (int64_t)TPB + threadIdx.x;
    if (t >= npad) return;
    int s = blockIdx.y;
    size_t lo = range * s / nsplit, hi = range * (s + 1) / nsplit;
    double re = 0, im = 0;
    int n = a.n;
    if (a.mode == 0) {
#pragma unroll 4
        for (size_t c = lo; c < hi; ++c) {
            double sgn = (__popcll((unsigned long long)(c & a.zmask)) & 1) ? -1.0 : 1.0;
            if (a.is_ket) { double2 v = cconjmul(d[(c ^ a.xmask) * npad + t], d[c * npad + t]); re += sgn * v.x; im += sgn * v.y; }
            else { double2 v = d[(c + ((c ^ a.xmask) << n)) * npad + t]; re += sgn * v.x; im += sgn * v.y; }
        }
    } else if (a.mode == 1) {
        const SparseAmp* sp = (const SparseAmp*)a.buf;
        if (a.is_ket) {  // amp = sum conj(psi[r]) phi[r]  (a' * b), squared at finalize
            for (size_t i = lo; i < hi; ++i) { double2 v = cconjmul(d[sp[i].idx * npad + t], sp[i].amp); re += v.x; im += v.y; }
        } else {         // sum_{i,j} conj(phi_i) rho[r_i, r_j] phi_j ; item = i*nnz + j
            for (size_t it = lo; it < hi; ++it) {
                size_t i = it / a.nnz, j = it % a.nnz;
                double2 v = cmul(cconjmul(sp[i].amp, d[(sp[i].idx + (sp[j].idx << n)) * npad + t]), sp[j].amp);
                re += v.x; im += v.y;
            }
        }
    } else {
        const double2* O = (const double2*)a.buf;
        int K = 1 << a.k;
        for (size_t g = lo; g < hi; ++g) {
            size_t base = g;
            for (int i = 0; i < a.k; ++i) base = qsb_insert0(base, a.sorted[i]);
            for (int i = 0; i < K; ++i) {
                size_t ei = base;
                for (int b = 0; b < a.k; ++b) ei |= (size_t)((i >> b) & 1) << a.pos[b];
                for (int j = 0; j < K; ++j) {
                    double2 o = O[i + j * K];
                    if (o.x == 0.0 && o.y == 0.0) continue;
                    size_t ej = base;
                    for (int b = 0; b < a.k; ++b) ej |= (size_t)((j >> b) & 1) << a.pos[b];
                    double2 v;
                    if (a.is_ket) v = cmul(cconjmul(d[ei * npad + t], o), d[ej * npad + t]);  // conj(psi_i) O_ij psi_j
                    else v = cmul(o, d[(ej + (ei << n)) * npad + t]);                         // O_ij rho[j,i]
                    re += v.x; im += v.y;
                }
            }
        }
    }
    partial[(size_t)s * 2 * npad + t] = re;
    partial[(size_t)s * 2 * npad + npad + t] = im;
}

// sums the split partials, applies the global factor, writes per-trajectory values and per-block statistics
__global__ void k_expect_finalize(const double* __restrict__ partial, int nsplit, int64_t npad, double2 factor, int square_abs,
                                  const uint32_t* __restrict__ alive, double* __restrict__ per_traj, double* __restrict__ blockstats) {
    int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    double re = 0, im = 0; bool al = false;
    if (t < npad) {
        for (int s = 0; s < nsplit; ++s) { re += partial[(size_t)s * 2 * npad + t]; im += partial[(size_t)s * 2 * npad + npad + t]; }
        if (square_abs) { re = re * re + im * im; im = 0; }
        double2 v = cmul(make_double2(re, im), factor);
        re = v.x; im = v.y;
        al = (alive[t >> 5] >> (t & 31)) & 1u;
        if (!al) { re = 0; im = 0; }
        if (per_traj) { per_traj[2 * t] = re; per_traj[2 * t + 1] = im; }
    }
    double v[4] = {al ? 1.0 : 0.0, re, re * re, im};
    __shared__ double sm[4][8];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        double x = v[i];
        for (int o = 16; o; o >>= 1) x += __shfl_xor_sync(0xffffffffu, x, o);
        if ((threadIdx.x & 31) == 0) sm[i][threadIdx.x >> 5] = x;
    }
    __syncthreads();
    if (threadIdx.x < 4) {
        double x = 0;
        for (int w = 0; w < (int)(blockDim.x >> 5); ++w) x += sm[threadIdx.x][w];
        blockstats[(size_t)blockIdx.x * 4 + threadIdx.x] = x;
    }
}
__global__ void k_stats_final(const double* __restrict__ blockstats, int nblocks, double* __restrict__ out) {
    __shared__ double sm[4][256];
    double acc[4] = {0, 0, 0, 0};
    for (int b = threadIdx.x; b < nblocks; b += 256)
        for (int i = 0; i < 4; ++i) acc[i] += blockstats[(size_t)b * 4 + i];
    for (int i = 0; i < 4; ++i) sm[i][threadIdx.x] = acc[i];
    __syncthreads();
    for (int s = 128; s; s >>= 1) {
        if ((int)threadIdx.x < s) for (int i = 0; i < 4; ++i) sm[i][threadIdx.x] += sm[i][threadIdx.x + s];
        __syncthreads();
    }
    if (threadIdx.x < 4) out[threadIdx.x] = sm[threadIdx.x][0];
}

int qsbd_expect(qsb_state* st, const int* idx, int k, int obs_kind, const void* obs, int obs_sign,
                double* per_traj, double* stats_dev, double* stats_host) {
    qsb_ctx* ctx = st->ctx;
    int n = st->n;
    bool ket = qsb_is_ketlike(st->kind);
    if (ctx->rec) {
        // recorded form (the state is held as a dm whatever its logical kind): iv[0] = mode, f[0..1] = global factor,
        //   mode 0  Pauli string      u[0] = xmask, u[1] = zmask
        //   mode 1  full projector    aux = nnz x {index, re, im}, iv[1] = nnz        value = sum_ij conj(phi_i) rho[r_i, r_j] phi_j
        //   mode 2  k-qubit operator  aux = {pos[0..k) padded to even} ++ O (4^k complex, column-major)
        TabOp* o = qsb_rec_push(ctx, DOP_EXPECT);
        o->n = n; o->k = k; o->s[0] = st->vbase; o->al[0] = st->valive; o->stats = stats_dev; o->per_traj = per_traj;
        double fr = obs_sign ? -1.0 : 1.0, fi = 0.0;
        uint64_t algo;
        if (obs_kind == QSB_OBS_PAULI) {
            const uint8_t* codes = (const uint8_t*)obs;
            int ny = 0;
            for (int m = 0; m < k; ++m) {
                if (codes[m] > 3) QSB_FAIL(QSB_ERR_BAD_ARG, "bad Pauli code %d", codes[m]);
                if (codes[m] == 1 || codes[m] == 2) o->u[0] |= 1u << idx[m];
                if (codes[m] == 3 || codes[m] == 2) o->u[1] |= 1u << idx[m];
                if (codes[m] == 2) ++ny;
            }
            static const double ipw[4][2] = {{1, 0}, {0, 1}, {-1, 0}, {0, -1}};
            fi = ipw[ny & 3][1] * fr; fr = ipw[ny & 3][0] * fr;
            o->iv[0] = 0;
            algo = (uint64_t)16 * ((uint64_t)1 << n) * (ket ? 2 : 1);
        } else if (obs_kind == QSB_OBS_PROJ_KET && k == n) {
            const double* psi = (const double*)obs;
            std::vector<double> sp;
            for (size_t r = 0; r < ((size_t)1 << n); ++r) {
                size_t pp = 0;
                for (int m = 0; m < k; ++m) pp |= ((r >> idx[m]) & 1) << m;
                if (psi[2 * pp] != 0.0 || psi[2 * pp + 1] != 0.0) { sp.push_back((double)r); sp.push_back(psi[2 * pp]); sp.push_back(psi[2 * pp + 1]); }
            }
            if (sp.empty()) QSB_FAIL(QSB_ERR_BAD_ARG, "projector ket is zero");
            o->iv[0] = 1; o->iv[1] = (int)(sp.size() / 3); o->obs = qsb_rec_daux(ctx, sp.data(), (int)sp.size());
            algo = (uint64_t)16 * (ket ? sp.size() / 3 : (sp.size() / 3) * (sp.size() / 3));
        } else if (obs_kind == QSB_OBS_OP || obs_kind == QSB_OBS_PROJ_KET) {
            size_t K = (size_t)1 << k;
            std::vector<double> pay((k + 1) / 2 * 2, 0.0);
            for (int i = 0; i < k; ++i) pay[i] = (double)idx[i];
            if (obs_kind == QSB_OBS_OP) pay.insert(pay.end(), (const double*)obs, (const double*)obs + 2 * K * K);
            else { std::vector<double> O = outer_ri((const double*)obs, K); pay.insert(pay.end(), O.begin(), O.end()); }
            o->iv[0] = 2; o->obs = qsb_rec_daux(ctx, pay.data(), (int)pay.size());
            algo = (uint64_t)16 * (ket ? ((uint64_t)1 << n) : ((uint64_t)1 << (n + k)));
        } else {
            QSB_FAIL(QSB_ERR_UNSUPPORTED, "observable kind %d is not defined for dense states", obs_kind);
        }
        o->f[0] = fr; o->f[1] = fi;
        qsb_rec_note_stats(ctx, stats_dev);
        qsb_account(ctx, "k_expect", algo * ctx->ntraj); ctx->ops += ctx->ntraj;
        return QSB_OK;
    }
    ExpArgs a; memset(&a, 0, sizeof(a));
    a.n = n; a.k = k; a.is_ket = ket;
    size_t range;
    double2 factor = make_double2(obs_sign ? -1.0 : 1.0, 0.0);
    int square_abs = 0;
    uint64_t algo;
    if (obs_kind == QSB_OBS_PAULI) {
        const uint8_t* codes = (const uint8_t*)obs;
        int ny = 0;
        for (int m = 0; m < k; ++m) {
            if (codes[m] > 3) QSB_FAIL(QSB_ERR_BAD_ARG, "bad Pauli code %d", codes[m]);
            if (codes[m] == 1 || codes[m] == 2) a.xmask |= (size_t)1 << idx[m];
            if (codes[m] == 3 || codes[m] == 2) a.zmask |= (size_t)1 << idx[m];
            if (codes[m] == 2) ++ny;
        }
        // P|c> = i^ny (-1)^{popc(c & zmask)} |c ^ xmask>
        static const double2 ipow[4] = {{1, 0}, {0, 1}, {-1, 0}, {0, -1}};
        double2 f = ipow[ny & 3];
        // ket path accumulates conj(psi[c^x]) psi[c] * sgn(c) = <psi|P|psi> / i^ny ; dm path rho[c, c^x] sgn(c)
        factor = make_double2(f.x * factor.x, f.y * factor.x);
        a.mode = 0;
        range = (size_t)1 << n;
        algo = (uint64_t)16 * ((uint64_t)1 << n) * (ket ? 2 : 1);
    } else if (obs_kind == QSB_OBS_PROJ_KET && k == n) {
        // quantumoptics.jl:31-41: full-state projector; psi factor m sits on subsystem idx[m]
        size_t D = (size_t)1 << n;
        const double* psi = (const double*)obs;
        std::vector<SparseAmp> sp;
        for (size_t r = 0; r < D; ++r) {
            size_t p = 0;
            for (int m = 0; m < k; ++m) p |= ((r >> idx[m]) & 1) << m;
            if (psi[2 * p] != 0.0 || psi[2 * p + 1] != 0.0) { SparseAmp e; e.idx = r; e.amp = make_double2(psi[2 * p], psi[2 * p + 1]); sp.push_back(e); }
        }
        if (sp.empty()) QSB_FAIL(QSB_ERR_BAD_ARG, "projector ket is zero");
        void* dev;
        QSB_TRY(qsb_upload(ctx, sp.data(), sp.size() * sizeof(SparseAmp), &dev));
        a.mode = 1; a.nnz = (int)sp.size(); a.buf = dev;
        range = ket ? sp.size() : sp.size() * sp.size();
        square_abs = ket ? 1 : 0;
        algo = (uint64_t)16 * range;
    } else if (obs_kind == QSB_OBS_OP || obs_kind == QSB_OBS_PROJ_KET) {
        if (k > 6) QSB_FAIL(QSB_ERR_UNSUPPORTED, "dense observables on more than 6 subsystems are not implemented");
        size_t K = (size_t)1 << k;
        std::vector<double> O(2 * K * K);
        if (obs_kind == QSB_OBS_OP) memcpy(O.data(), obs, O.size() * sizeof(double));
        else {  // outer product |psi><psi| (the `else` branch at quantumoptics.jl:38-40)
            const double* psi = (const double*)obs;
            for (size_t c = 0; c < K; ++c) for (size_t r = 0; r < K; ++r) {
                O[2 * (r + c * K)] = psi[2 * r] * psi[2 * c] + psi[2 * r + 1] * psi[2 * c + 1];
                O[2 * (r + c * K) + 1] = psi[2 * r + 1] * psi[2 * c] - psi[2 * r] * psi[2 * c + 1];
            }
        }
        void* dev;
        QSB_TRY(qsb_upload(ctx, O.data(), O.size() * sizeof(double), &dev));
        a.mode = 2; a.buf = dev;
        for (int i = 0; i < k; ++i) { a.pos[i] = idx[i]; a.sorted[i] = idx[i]; }
        sort_ints(a.sorted, k);
        range = (size_t)1 << (n - k);
        algo = (uint64_t)16 * (ket ? ((uint64_t)1 << n) : ((uint64_t)1 << (n + k)));
    } else {
        QSB_FAIL(QSB_ERR_UNSUPPORTED, "observable kind %d is not defined for dense states", obs_kind);
    }
    int nsplit = plan_split(ctx, range);
    unsigned fblocks = (unsigned)((ctx->npad + 255) / 256);
    size_t need = ((size_t)nsplit * 2 * ctx->npad + (size_t)fblocks * 4) * sizeof(double);
    QSB_TRY(qsb_ensure_scratch(ctx, need));
    dim3 grid((unsigned)((ctx->npad + TPB - 1) / TPB), nsplit, 1);
    QSB_PRE(ctx, "k_expect");
    k_expect<<<grid, TPB, 0, ctx->stream>>>((const double2*)st->data, ctx->npad, a, range, nsplit, ctx->scratch);
    QSB_LAUNCH_CHECK(ctx);
    double* blockstats = ctx->scratch + (size_t)nsplit * 2 * ctx->npad;
    QSB_PRE(ctx, "k_expect_finalize");
    k_expect_finalize<<<fblocks, 256, 0, ctx->stream>>>(ctx->scratch, nsplit, ctx->npad, factor, square_abs, st->alive, per_traj, blockstats);
    QSB_LAUNCH_CHECK(ctx);
    double* sd = stats_dev ? stats_dev : ctx->stats_tmp;
    if (stats_dev || stats_host) {
        QSB_PRE(ctx, "k_stats_final");
        k_stats_final<<<1, 256, 0, ctx->stream>>>(blockstats, (int)fblocks, sd);
        QSB_LAUNCH_CHECK(ctx);
    }
    qsb_account(ctx, "k_expect", algo * ctx->ntraj); ctx->ops += ctx->ntraj;
    if (stats_host) {
        QSB_CUDA(cudaMemcpyAsync(stats_host, sd, 4 * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
        QSB_TRY(qsb_ctx_sync(ctx));
    }
    return QSB_OK;
}

// =============================================================================================
// qsb_expect_pauli_static: the Pauli-string observable for callers that must not allocate
// =============================================================================================
// qsb_expect keeps its reduction partials in ctx->scratch, which qsb_ensure_scratch frees and re-allocates from the
// stream's memory pool whenever a larger observable arrives.  Two kinds of caller cannot live with that:
//   - a host process whose GPU memory already belongs to somebody else's allocator (CUDA.jl's pool or torch's caching
//     allocator holding nearly all of the HBM for a big density matrix): an allocation made behind its back can fail, or
//     make the driver trim the other pool, which stalls every stream of the device;
//   - a control loop with a latency bound per call (an observable evaluated between two events of a discrete-event
//     run): growing the scratch buffer is a free + an allocation in the middle of the loop.
// This variant does the same two passes (k_expect writes one partial sum per (split, trajectory), k_expect_finalize
// folds them) with the partials in a file-scope device array of fixed size.  Cost: every GPU on which this library's
// code gets loaded holds QSB_STATIC_SCRATCH_BYTES (4 MiB) for the life of the process, used or not.
//
// Sharing rules.  There is ONE array per device, shared by every context and every call.  The call leaves its partials
// behind; nothing in it has to start at zero, because k_expect stores (never accumulates into) every slot that
// k_expect_finalize reads, and k_expect_finalize stores every block-statistics slot that k_stats_final reads -- so no
// clearing pass is needed and stale contents of an earlier, larger call are never read.  Two calls whose kernels can
// overlap on the GPU (two contexts on different streams, or one context whose stream was swapped between the calls)
// write the same slots and RACE: issue them on one stream or order them with an event.
#define QSB_STATIC_SCRATCH_DOUBLES (QSB_STATIC_SCRATCH_BYTES / sizeof(double))
static __device__ __align__(256) double g_static_scratch[QSB_STATIC_SCRATCH_DOUBLES];

// In host code `g_static_scratch` is only a handle for the runtime's symbol table, not an address a kernel can use: the
// device address differs per GPU and has to be asked for with that GPU current.  Looked up on the first call per device
// and kept; that first call may make the runtime load this module onto the device, so it should not be made while a
// stream capture is open.
#define QSB_STATIC_MAX_DEVICES 64
static double* g_static_scratch_dev[QSB_STATIC_MAX_DEVICES];
static int static_scratch_address(const qsb_ctx* ctx, double** out) {
    if (ctx->device < 0 || ctx->device >= QSB_STATIC_MAX_DEVICES)
        QSB_FAIL(QSB_ERR_UNSUPPORTED, "qsb_expect_pauli_static: device ordinal %d is beyond the %d devices the address table covers", ctx->device, QSB_STATIC_MAX_DEVICES);
    double* p = __atomic_load_n(&g_static_scratch_dev[ctx->device], __ATOMIC_ACQUIRE);
    if (!p) {
        qsb_use_device(ctx);
        void* sym = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&sym, g_static_scratch);
        if (e != cudaSuccess || !sym) {
            qsb_set_error("qsb_expect_pauli_static: cudaGetSymbolAddress of the built-in scratch array failed on device %d: %s", ctx->device, cudaGetErrorString(e));
            cudaGetLastError();
            return QSB_ERR_CUDA;
        }
        p = (double*)sym;   // every thread that gets here computes the same address, so a plain publish is enough
        __atomic_store_n(&g_static_scratch_dev[ctx->device], p, __ATOMIC_RELEASE);
    }
    *out = p;
    return QSB_OK;
}

extern "C" int qsb_expect_pauli_static(qsb_state* st, const int* idx, int k, const uint8_t* codes, int obs_sign,
                                       double* per_traj_dev, double* stats_dev) {
    if (!st || !idx || !codes) QSB_FAIL(QSB_ERR_BAD_ARG, "qsb_expect_pauli_static: null argument");
    qsb_ctx* ctx = st->ctx;
    if (ctx->rec) QSB_FAIL(QSB_ERR_UNSUPPORTED, "qsb_expect_pauli_static cannot be recorded into a program (recorded observables keep their partials on chip; use qsb_expect)");
    QSB_TRY(qsb_handle_ok(ctx, st->vbase >= 0, st->gen));
    if (!qsb_is_dense(st->kind)) QSB_FAIL(QSB_ERR_UNSUPPORTED, "qsb_expect_pauli_static is defined for ket / density-matrix states (a tableau expectation needs no scratch: use qsb_expect)");
    if (!per_traj_dev && !stats_dev) QSB_FAIL(QSB_ERR_BAD_ARG, "qsb_expect_pauli_static: pass per_traj_dev, stats_dev or both (there is no host-visible output)");
    const int n = st->n;
    if (k < 1 || k > n) QSB_FAIL(QSB_ERR_BAD_INDEX, "%d indices given for a state of %d subsystems", k, n);
    ExpArgs a; memset(&a, 0, sizeof(a));
    a.n = n; a.k = k; a.is_ket = qsb_is_ketlike(st->kind); a.mode = 0;
    int ny = 0;
    for (int m = 0; m < k; ++m) {
        if (idx[m] < 0 || idx[m] >= n) QSB_FAIL(QSB_ERR_BAD_INDEX, "subsystem index %d out of range for a state of %d subsystems", idx[m] + 1, n);
        for (int b = 0; b < m; ++b) if (idx[m] == idx[b]) QSB_FAIL(QSB_ERR_BAD_INDEX, "subsystem index %d repeated", idx[m] + 1);
        if (codes[m] > 3) QSB_FAIL(QSB_ERR_BAD_ARG, "bad Pauli code %d", codes[m]);
        if (codes[m] == 1 || codes[m] == 2) a.xmask |= (size_t)1 << idx[m];
        if (codes[m] == 3 || codes[m] == 2) a.zmask |= (size_t)1 << idx[m];
        if (codes[m] == 2) ++ny;
    }
    // same global factor as qsbd_expect: P|c> = i^ny (-1)^{popc(c & zmask)} |c ^ xmask>
    static const double2 ipow[4] = {{1, 0}, {0, 1}, {-1, 0}, {0, -1}};
    const double sgn = obs_sign ? -1.0 : 1.0;
    const double2 factor = make_double2(ipow[ny & 3].x * sgn, ipow[ny & 3].y * sgn);
    const size_t range = (size_t)1 << n;

    // layout inside the array: [partials 2 * nsplit * npad][block statistics 4 * fblocks].  The split count is the one
    // qsb_expect would pick, lowered until the partials fit; when even one split does not fit the ensemble is too large.
    const unsigned fblocks = (unsigned)((ctx->npad + 255) / 256);
    const size_t fixed = (size_t)fblocks * 4, per_split = (size_t)2 * ctx->npad;
    if (fixed + per_split > QSB_STATIC_SCRATCH_DOUBLES)
        QSB_FAIL(QSB_ERR_BAD_ARG, "qsb_expect_pauli_static: %lld trajectories need %zu bytes of reduction scratch but the built-in array holds %zu "
                 "(at most %zu trajectories per context); use qsb_expect, which allocates", (long long)ctx->ntraj,
                 (fixed + per_split) * sizeof(double), (size_t)QSB_STATIC_SCRATCH_BYTES, (size_t)QSB_STATIC_MAX_TRAJ);
    int nsplit = plan_split(ctx, range);
    const size_t fit = (QSB_STATIC_SCRATCH_DOUBLES - fixed) / per_split;
    if ((size_t)nsplit > fit) nsplit = (int)fit;

    double* scratch = nullptr;
    QSB_TRY(static_scratch_address(ctx, &scratch));
    double* blockstats = scratch + (size_t)nsplit * per_split;
    qsb_use_device(ctx);
    dim3 grid((unsigned)((ctx->npad + TPB - 1) / TPB), nsplit, 1);
    k_expect<<<grid, TPB, 0, ctx->stream>>>((const double2*)st->data, ctx->npad, a, range, nsplit, scratch);
    k_expect_finalize<<<fblocks, 256, 0, ctx->stream>>>(scratch, nsplit, ctx->npad, factor, 0, st->alive, per_traj_dev, blockstats);
    int launched = 2;
    if (stats_dev) { k_stats_final<<<1, 256, 0, ctx->stream>>>(blockstats, (int)fblocks, stats_dev); ++launched; }
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) QSB_FAIL(QSB_ERR_CUDA, "qsb_expect_pauli_static: kernel launch failed: %s", cudaGetErrorString(e));
    // plain counters only: the per-kernel profile rows of qsb_ctx_profile grow host containers and create events
    ctx->launches += launched;
    ctx->algo_bytes += (uint64_t)16 * range * (a.is_ket ? 2 : 1) * ctx->ntraj;
    ctx->ops += ctx->ntraj;
    return QSB_OK;
}
