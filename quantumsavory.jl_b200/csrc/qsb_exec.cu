// qsb_exec.cu -- recorded tableau programs: capture the op list of a lockstep schedule once, then execute
// the WHOLE list in one launch with every trajectory word's state resident in shared memory.
//
// Why: at the BASELINE configs the entire Clifford ensemble is a few MB and a per-op launch is ~31k
// threads of dependent word logic (3-38 us each, ~1700 launches per pass).  A thread only ever touches
// its own 32-trajectory word column of every plane, so consecutive ops need no grid synchronisation:
// one thread can run the complete op list on its column.  The arena (all live states and masks of the
// schedule, planes allocated/recycled at record time) is a [planes][blockDim] tile of shared memory.
#include <map>

#include "qsb_internal.cuh"
#include "qsb_tab_core.cuh"

struct qsb_recorder {
    std::vector<TabOp> ops;
    std::vector<TabObsArgs> obs;
    std::vector<double*> stats_ptrs;
    int next_plane = 0, high = 0;
    std::multimap<int, int> free_blocks;  // size -> base
};

struct qsb_program {
    qsb_ctx* ctx;
    TabOp* d_ops; int nops;
    TabObsArgs* d_obs; int nobs;
    double** d_stats; int nstats;
    int planes;
    uint32_t* arena;       // global-memory arena, only when the shared-memory tile does not fit
    int threads, smem_bytes;
};

int qsb_rec_alloc_planes(qsb_ctx* ctx, int nplanes) {
    qsb_recorder* R = ctx->rec;
    if (nplanes <= 0) nplanes = 1;
    auto it = R->free_blocks.lower_bound(nplanes);
    if (it != R->free_blocks.end()) {
        int size = it->first, base = it->second;
        R->free_blocks.erase(it);
        if (size > nplanes) R->free_blocks.insert({size - nplanes, base + nplanes});
        return base;
    }
    int base = R->next_plane;
    R->next_plane += nplanes;
    if (R->next_plane > R->high) R->high = R->next_plane;
    return base;
}
void qsb_rec_free_planes(qsb_ctx* ctx, int base, int nplanes) {
    if (!ctx->rec || base < 0) return;
    if (nplanes <= 0) nplanes = 1;
    qsb_recorder* R = ctx->rec;
    // coalesce with neighbours so that large composite states can be placed again
    for (bool merged = true; merged;) {
        merged = false;
        for (auto it = R->free_blocks.begin(); it != R->free_blocks.end(); ++it) {
            if (it->second + it->first == base) { base = it->second; nplanes += it->first; R->free_blocks.erase(it); merged = true; break; }
            if (base + nplanes == it->second) { nplanes += it->first; R->free_blocks.erase(it); merged = true; break; }
        }
    }
    if (base + nplanes == R->next_plane) R->next_plane = base;
    else R->free_blocks.insert({nplanes, base});
}
TabOp* qsb_rec_push(qsb_ctx* ctx, int type) {
    TabOp op;
    memset(&op, 0, sizeof(op));
    op.type = type;
    for (int i = 0; i < 6; ++i) { op.s[i] = -1; op.al[i] = -1; }
    for (int i = 0; i < 3; ++i) op.m[i] = -1;
    op.obs = -1;
    ctx->rec->ops.push_back(op);
    return &ctx->rec->ops.back();
}
int qsb_rec_push_obs(qsb_ctx* ctx, const void* a, size_t bytes) {
    TabObsArgs o;
    memcpy(&o, a, bytes < sizeof(o) ? bytes : sizeof(o));
    ctx->rec->obs.push_back(o);
    return (int)ctx->rec->obs.size() - 1;
}

// ---------------------------------------------------------------------------------------------
// the fused executor
// ---------------------------------------------------------------------------------------------
struct ExecArgs {
    const TabOp* ops; int nops;
    const TabObsArgs* obs;
    uint32_t* arena;          // nullptr -> shared-memory tile
    const uint32_t* valid;
    size_t W;                 // words in the ensemble
    uint64_t seed, traj0, draw_base;
};

__global__ void k_zero_stats(double* const* ptrs, int n) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) { double* p = ptrs[i]; p[0] = 0; p[1] = 0; p[2] = 0; p[3] = 0; }
}

__global__ void k_tab_exec(ExecArgs a) {
    extern __shared__ uint32_t tile[];
    const size_t wg = blockIdx.x * (size_t)blockDim.x + threadIdx.x;  // global word (trajectories 32*wg ..)
    const bool on = wg < a.W;
    uint32_t* base; size_t W, w;
    if (a.arena) { base = a.arena; W = a.W; w = on ? wg : 0; }
    else { base = tile; W = blockDim.x; w = threadIdx.x; }
    const uint32_t validw = on ? a.valid[wg] : 0u;
#define PL(idx) (base + (size_t)(idx) * W)
    for (int i = 0; i < a.nops; ++i) {
        const TabOp& op = a.ops[i];
        const uint64_t draw = op.draw + a.draw_base;
        switch (op.type) {
        case TOP_INIT:
            if (on) { tab_init_word(PL(op.s[0]), W, w, op.n, op.iv[0]); PL(op.al[0])[w] = validw; }
            break;
        case TOP_TENSOR:
            if (on) { tab_tensor_word(PL(op.s[0]), PL(op.s[1]), PL(op.s[2]), W, w, op.n, op.n2); PL(op.al[0])[w] = PL(op.al[1])[w] & PL(op.al[2])[w]; }
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
                uint32_t m[5] = {0, 0, 0, 0, 0};
                for (int l = 0; l < 32; ++l) {
                    double u = qsb_draw(a.seed, a.traj0 + (uint64_t)(wg * 32 + l), draw);
                    int sel = op.n2 - 1;
                    for (int j = 0; j < op.n2; ++j) if (op.f[j] > u) { sel = j; break; }
                    m[sel] |= 1u << l;
                }
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
        case TOP_GATE:
            if (on) {
                TabGateArgs g; g.n = op.n; g.k = op.k; g.q[0] = op.q[0]; g.q[1] = op.q[1]; g.lut = op.lut;
                for (int j = 0; j < 4; ++j) g.outmask[j] = op.outmask[j];
                tab_gate_word(PL(op.s[0]), W, w, g, op.m[0] >= 0 ? PL(op.m[0])[w] : 0xFFFFFFFFu);
            }
            break;
        case TOP_NOISE:
            if (on) {
                TabNoiseArgs g; g.n = op.n; g.q = op.q[0]; g.np = op.k;
                for (int j = 0; j < 3; ++j) { g.cum[j] = op.f[j]; g.op[j] = op.iv[j]; }
                tab_noise_word(PL(op.s[0]), W, w, wg, g, a.seed, a.traj0, draw);
            }
            break;
        case TOP_MEASURE:
            if (on) {
                uint32_t rnd;
                uint32_t outcome = tab_measure_word(PL(op.s[1]), PL(op.s[0]), W, w, wg, op.n, op.q[0], op.iv[0], op.iv[1], a.seed, a.traj0, draw, &rnd);
                uint32_t al = PL(op.al[1])[w];
                PL(op.m[0])[w] = outcome & al;
                if (op.m[1] >= 0) PL(op.m[1])[w] = rnd & al;
            }
            break;
        case TOP_PTRACE:
            if (on) tab_ptrace_word(PL(op.s[1]), PL(op.s[0]), W, w, op.n, op.q[0]);
            break;
        case TOP_FORCE:
            if (on) tab_force_word(PL(op.s[0]), W, w, op.n, op.u[0], op.u[1], op.iv[0]);
            break;
        case TOP_EXPECT: {
            double s_alive = 0, s_v = 0, s_v2 = 0;
            if (on) {
                const TabObsArgs& o = a.obs[op.obs];
                uint32_t cnt[8], cntm[8];
                tab_expect_word(PL(op.s[0]), W, w, o, cnt, cntm);
                uint32_t al = PL(op.al[0])[w];
                double inv = 1.0 / (double)o.nelem;
                for (int l = 0; l < 32; ++l) {
                    int plus = 0, minus = 0;
                    for (int b = 0; b < 8; ++b) { plus |= ((cnt[b] >> l) & 1) << b; minus |= ((cntm[b] >> l) & 1) << b; }
                    bool lane = (al >> l) & 1u;
                    double val = lane ? (double)(plus - minus) * inv : 0.0;
                    if (op.per_traj) { size_t t = wg * 32 + l; op.per_traj[2 * t] = val; op.per_traj[2 * t + 1] = 0.0; }
                    if (lane) { s_alive += 1.0; s_v += val; s_v2 += val * val; }
                }
            }
            if (op.stats) {  // whole warps reach this point together (the op list is uniform)
                for (int o = 16; o; o >>= 1) {
                    s_alive += __shfl_xor_sync(0xffffffffu, s_alive, o);
                    s_v += __shfl_xor_sync(0xffffffffu, s_v, o);
                    s_v2 += __shfl_xor_sync(0xffffffffu, s_v2, o);
                }
                if ((threadIdx.x & 31) == 0 && s_alive != 0.0) { atomicAdd(&op.stats[0], s_alive); atomicAdd(&op.stats[1], s_v); atomicAdd(&op.stats[2], s_v2); }
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
                PL(op.m[0])[w] = r & validw;
            }
            break;
        case TOP_MASK_RANDOM:
            if (on) {
                uint32_t r = 0;
                for (int l = 0; l < 32; ++l) if (qsb_draw(a.seed, a.traj0 + (uint64_t)(wg * 32 + l), draw) < op.f[0]) r |= 1u << l;
                PL(op.m[0])[w] = r & validw;
            }
            break;
        case TOP_KILL:
            if (on) PL(op.al[0])[w] &= ~PL(op.m[0])[w];
            break;
        case TOP_ALIVE_COPY:
            if (on) PL(op.m[0])[w] = PL(op.al[0])[w];
            break;
        }
    }
#undef PL
}

// ---------------------------------------------------------------------------------------------
// C-ABI: record / run
// ---------------------------------------------------------------------------------------------
extern "C" int qsb_record_begin(qsb_ctx* ctx) {
    if (!ctx) QSB_FAIL(QSB_ERR_BAD_ARG, "null context");
    if (ctx->rec) QSB_FAIL(QSB_ERR_BAD_ARG, "already recording");
    ctx->rec = new qsb_recorder();
    ctx->rec_gen++;
    return QSB_OK;
}

extern "C" int qsb_record_end(qsb_ctx* ctx, qsb_program** out) {
    if (!ctx || !out) QSB_FAIL(QSB_ERR_BAD_ARG, "null argument");
    if (!ctx->rec) QSB_FAIL(QSB_ERR_BAD_ARG, "not recording");
    qsb_recorder* R = ctx->rec;
    qsb_program* P = new qsb_program();
    memset(P, 0, sizeof(*P));
    P->ctx = ctx;
    P->nops = (int)R->ops.size(); P->nobs = (int)R->obs.size(); P->nstats = (int)R->stats_ptrs.size();
    P->planes = R->high > 0 ? R->high : 1;
    int rc = QSB_OK;
    auto up = [&](void** dst, const void* src, size_t bytes) {
        if (rc != QSB_OK) return;
        rc = qsb_dev_alloc(ctx, dst, bytes ? bytes : 16);
        if (rc == QSB_OK && bytes) {
            cudaError_t e = cudaMemcpyAsync(*dst, src, bytes, cudaMemcpyHostToDevice, ctx->stream);
            if (e != cudaSuccess) { qsb_set_error("program upload failed: %s", cudaGetErrorString(e)); rc = QSB_ERR_CUDA; }
        }
    };
    up((void**)&P->d_ops, R->ops.data(), R->ops.size() * sizeof(TabOp));
    up((void**)&P->d_obs, R->obs.data(), R->obs.size() * sizeof(TabObsArgs));
    up((void**)&P->d_stats, R->stats_ptrs.data(), R->stats_ptrs.size() * sizeof(double*));
    if (rc == QSB_OK) { cudaError_t e = cudaStreamSynchronize(ctx->stream); if (e != cudaSuccess) rc = QSB_ERR_CUDA; }  // host vectors die below
    // launch shape: the largest block whose [planes][threads] tile fits, preferring a single wave
    int dev_smem = 0;
    cudaDeviceGetAttribute(&dev_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, ctx->device);
    int best_t = 0; double best_cost = 1e300;
    for (int t = 32; t <= 256; t += 32) {
        size_t smem = (size_t)P->planes * t * sizeof(uint32_t);
        if (smem > (size_t)dev_smem) break;
        int per_sm = (int)((size_t)(228 * 1024) / (smem + 1024));
        if (per_sm < 1) per_sm = 1;
        if (per_sm * t > 2048) per_sm = 2048 / t;
        long blocks = (long)((ctx->nwords + t - 1) / t);
        long slots = (long)ctx->sm_count * per_sm;
        long waves = (blocks + slots - 1) / slots;
        // a wave costs about one thread's serial time; fewer resident warps per SM run faster per warp
        double cost = (double)waves * (1.0 + 0.02 * per_sm * t / 32.0);
        if (cost < best_cost) { best_cost = cost; best_t = t; }
    }
    if (best_t) { P->threads = best_t; P->smem_bytes = P->planes * best_t * (int)sizeof(uint32_t); P->arena = nullptr; }
    else {
        P->threads = 128; P->smem_bytes = 0;
        if (rc == QSB_OK) rc = qsb_dev_alloc(ctx, (void**)&P->arena, (size_t)P->planes * ctx->nwords * sizeof(uint32_t));
    }
    if (rc == QSB_OK && P->smem_bytes) {
        cudaError_t e = cudaFuncSetAttribute(k_tab_exec, cudaFuncAttributeMaxDynamicSharedMemorySize, P->smem_bytes);
        if (e != cudaSuccess) { qsb_set_error("cudaFuncSetAttribute(%d bytes) failed: %s", P->smem_bytes, cudaGetErrorString(e)); rc = QSB_ERR_CUDA; }
    }
    delete R;
    ctx->rec = nullptr;
    if (rc != QSB_OK) { delete P; return rc; }
    *out = P;
    return QSB_OK;
}

// traj_shift moves the whole ensemble to fresh global trajectory ids (a new batch), draw_base offsets every draw id.
extern "C" int qsb_program_run(qsb_program* P, uint64_t traj_shift, uint64_t draw_base) {
    if (!P) QSB_FAIL(QSB_ERR_BAD_ARG, "null program");
    qsb_ctx* ctx = P->ctx;
    if (ctx->rec) QSB_FAIL(QSB_ERR_BAD_ARG, "cannot run a program while recording");
    if (P->nstats) {
        QSB_PRE(ctx, "k_zero_stats");
        k_zero_stats<<<(P->nstats + 127) / 128, 128, 0, ctx->stream>>>(P->d_stats, P->nstats);
        QSB_LAUNCH_CHECK(ctx);
    }
    ExecArgs a;
    a.ops = P->d_ops; a.nops = P->nops; a.obs = P->d_obs; a.arena = P->arena; a.valid = ctx->valid; a.W = (size_t)ctx->nwords;
    a.seed = ctx->seed; a.traj0 = (uint64_t)ctx->traj_offset + traj_shift; a.draw_base = draw_base;
    unsigned blocks = (unsigned)((ctx->nwords + P->threads - 1) / P->threads);
    QSB_PRE(ctx, "k_tab_exec");
    k_tab_exec<<<blocks, P->threads, P->smem_bytes, ctx->stream>>>(a);
    QSB_LAUNCH_CHECK(ctx);
    return QSB_OK;
}

// info: {ops, arena planes, statistics rows, threads per block, shared bytes per block, 1 if the arena is in global memory}
extern "C" int qsb_program_info(const qsb_program* P, int64_t info[6]) {
    if (!P || !info) QSB_FAIL(QSB_ERR_BAD_ARG, "null argument");
    info[0] = P->nops; info[1] = P->planes; info[2] = P->nstats; info[3] = P->threads; info[4] = P->smem_bytes; info[5] = P->arena ? 1 : 0;
    return QSB_OK;
}

extern "C" int qsb_program_free(qsb_program* P) {
    if (!P) return QSB_OK;
    qsb_ctx* ctx = P->ctx;
    qsb_dev_free(ctx, P->d_ops); qsb_dev_free(ctx, P->d_obs); qsb_dev_free(ctx, P->d_stats);
    if (P->arena) qsb_dev_free(ctx, P->arena);
    delete P;
    return QSB_OK;
}

void qsb_rec_note_stats(qsb_ctx* ctx, double* p) { if (p) ctx->rec->stats_ptrs.push_back(p); }
