// qsb_core.cu -- contexts, handles, masks and the C-ABI dispatch of libqsb200.so
#include <stdarg.h>

#include "qsb_internal.cuh"

static thread_local char g_err[512] = "";

void qsb_set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

extern "C" const char* qsb_last_error(void) { return g_err; }
extern "C" int qsb_version(void) { return 100; }

// ---------------------------------------------------------------------------------------------
// per-kernel profile: CUDA events on the launching stream around every launch (bench.py's roofline)
// ---------------------------------------------------------------------------------------------
struct ProfRow { std::string name; uint64_t launches = 0, bytes = 0; double ms = 0; };
struct qsb_prof {
    std::vector<ProfRow> rows;
    std::vector<cudaEvent_t> pool;
    struct Pending { int row; cudaEvent_t a, b; };
    std::vector<Pending> pending;
    int cur_row = -1; cudaEvent_t cur_a = nullptr;
    int row_of(const char* name) {
        for (size_t i = 0; i < rows.size(); ++i) if (rows[i].name == name) return (int)i;
        ProfRow r; r.name = name; rows.push_back(r); return (int)rows.size() - 1;
    }
    cudaEvent_t get() {
        if (!pool.empty()) { cudaEvent_t e = pool.back(); pool.pop_back(); return e; }
        cudaEvent_t e; cudaEventCreate(&e); return e;
    }
    void drain() {
        for (auto& p : pending) {
            float ms = 0; cudaEventElapsedTime(&ms, p.a, p.b);
            rows[p.row].ms += ms; rows[p.row].launches++;
            pool.push_back(p.a); pool.push_back(p.b);
        }
        pending.clear();
    }
};
void qsb_prof_pre(qsb_ctx* ctx, const char* kernel) {
    qsb_prof* P = ctx->prof;
    P->cur_row = P->row_of(kernel);
    P->cur_a = P->get();
    cudaEventRecord(P->cur_a, ctx->stream);
}
void qsb_prof_post(qsb_ctx* ctx) {
    qsb_prof* P = ctx->prof;
    if (P->cur_row < 0) return;
    cudaEvent_t b = P->get();
    cudaEventRecord(b, ctx->stream);
    P->pending.push_back({P->cur_row, P->cur_a, b});
    P->cur_row = -1;
    if (P->pending.size() >= 4096) { cudaStreamSynchronize(ctx->stream); P->drain(); }
}
void qsb_account(qsb_ctx* ctx, const char* kernel, uint64_t bytes) {
    ctx->algo_bytes += bytes;
    if (ctx->prof) ctx->prof->rows[ctx->prof->row_of(kernel)].bytes += bytes;
}
extern "C" int qsb_ctx_profile(qsb_ctx* ctx, int enable) {
    if (!ctx) QSB_FAIL(QSB_ERR_BAD_ARG, "null context");
    if (enable && !ctx->prof) ctx->prof = new qsb_prof();
    if (!enable && ctx->prof) {
        cudaStreamSynchronize(ctx->stream);
        ctx->prof->drain();
        for (auto e : ctx->prof->pool) cudaEventDestroy(e);
        delete ctx->prof; ctx->prof = nullptr;
    }
    return QSB_OK;
}
extern "C" int qsb_ctx_profile_report(qsb_ctx* ctx, char* buf, size_t buflen) {
    if (!ctx || !buf || buflen < 4) QSB_FAIL(QSB_ERR_BAD_ARG, "bad argument");
    std::string s = "[";
    if (ctx->prof) {
        QSB_CUDA(cudaStreamSynchronize(ctx->stream));
        ctx->prof->drain();
        bool first = true;
        for (auto& r : ctx->prof->rows) {
            char line[256];
            snprintf(line, sizeof(line), "%s{\"kernel\":\"%s\",\"launches\":%llu,\"ms\":%.6f,\"algo_bytes\":%llu}", first ? "" : ",",
                     r.name.c_str(), (unsigned long long)r.launches, r.ms, (unsigned long long)r.bytes);
            s += line; first = false;
        }
    }
    s += "]";
    if (s.size() + 1 > buflen) QSB_FAIL(QSB_ERR_BAD_ARG, "report buffer too small (%zu needed)", s.size() + 1);
    memcpy(buf, s.c_str(), s.size() + 1);
    return QSB_OK;
}

// ---------------------------------------------------------------------------------------------
// allocation
// ---------------------------------------------------------------------------------------------
int qsb_dev_alloc(qsb_ctx* ctx, void** p, size_t bytes) {
    if (bytes == 0) bytes = 16;
    qsb_use_device(ctx);
    cudaError_t e = cudaMallocAsync(p, bytes, ctx->stream);
    if (e == cudaErrorMemoryAllocation) {
        // the pool keeps freed blocks (release threshold = max): wait for pending frees, give them back, retry once
        cudaGetLastError();
        cudaStreamSynchronize(ctx->stream);
        cudaMemPool_t pool;
        if (cudaDeviceGetDefaultMemPool(&pool, ctx->device) == cudaSuccess) cudaMemPoolTrimTo(pool, 0);
        e = cudaMallocAsync(p, bytes, ctx->stream);
    }
    if (e != cudaSuccess) {
        cudaGetLastError();
        size_t fr = 0, tot = 0;
        cudaMemGetInfo(&fr, &tot);
        qsb_set_error("cudaMallocAsync(%zu bytes) failed: %s (device has %zu of %zu bytes free)", bytes, cudaGetErrorString(e), fr, tot);
        return e == cudaErrorMemoryAllocation ? QSB_ERR_OOM : QSB_ERR_CUDA;
    }
    return QSB_OK;
}
int qsb_dev_free(qsb_ctx* ctx, void* p) {
    if (!p) return QSB_OK;
    qsb_use_device(ctx);
    QSB_CUDA(cudaFreeAsync(p, ctx->stream));
    return QSB_OK;
}
int qsb_ensure_scratch(qsb_ctx* ctx, size_t bytes) {
    if (ctx->scratch_bytes >= bytes) return QSB_OK;
    QSB_TRY(qsb_dev_free(ctx, ctx->scratch));
    ctx->scratch = nullptr; ctx->scratch_bytes = 0;
    size_t want = bytes < (1u << 20) ? (1u << 20) : bytes;
    QSB_TRY(qsb_dev_alloc(ctx, (void**)&ctx->scratch, want));
    ctx->scratch_bytes = want;
    return QSB_OK;
}
int qsb_upload(qsb_ctx* ctx, const void* host, size_t bytes, void** dev_out) {
    if (ctx->obs_bytes < bytes) {
        QSB_TRY(qsb_dev_free(ctx, ctx->obs_buf));
        ctx->obs_buf = nullptr; ctx->obs_bytes = 0;
        size_t want = bytes < 65536 ? 65536 : bytes;
        QSB_TRY(qsb_dev_alloc(ctx, &ctx->obs_buf, want));
        ctx->obs_bytes = want;
    }
    // pageable source: the runtime stages the bytes before returning, so `host` may be reused at once
    QSB_CUDA(cudaMemcpyAsync(ctx->obs_buf, host, bytes, cudaMemcpyHostToDevice, ctx->stream));
    *dev_out = ctx->obs_buf;
    return QSB_OK;
}
size_t qsb_state_bytes(const qsb_ctx* ctx, int kind, int n) {
    if (kind == QSB_TAB) return (size_t)qsb_tab_planes(n) * ctx->nwords * sizeof(uint32_t);
    size_t elems = qsb_is_ketlike(kind) ? ((size_t)1 << n) : ((size_t)1 << (2 * n));
    return elems * (size_t)ctx->npad * sizeof(double2);
}
int qsb_state_alloc(qsb_ctx* ctx, int kind, int n, qsb_state** out) {
    qsb_state* st = new qsb_state();
    st->ctx = ctx; st->kind = kind; st->n = n;
    st->bytes = qsb_state_bytes(ctx, kind, n);
    st->data = nullptr; st->alive = nullptr; st->vbase = st->valive = -1; st->gen = 0; st->vcap = 0;
    if (ctx->rec) {
        if (kind == QSB_MCKET) { delete st; QSB_FAIL(QSB_ERR_UNSUPPORTED, "QuantumMCRepr (MCKet) states cannot be recorded into a program"); }
        if (kind != QSB_TAB && n > QSB_DEXEC_MAXN) { delete st; QSB_FAIL(QSB_ERR_UNSUPPORTED, "dense states of more than %d subsystems cannot be recorded into a program (they do not fit on chip)", QSB_DEXEC_MAXN); }
        if (qsb_rec_set_family(ctx, kind != QSB_TAB) != QSB_OK) { delete st; return QSB_ERR_UNSUPPORTED; }
        // dense programs hold EVERY state as a density matrix (a ket is recorded as its projector): 4^n elements
        st->vbase = kind == QSB_TAB ? qsb_rec_alloc_planes(ctx, qsb_tab_planes(n)) : qsb_rec_alloc(ctx, 1, 1 << (2 * n));
        st->valive = qsb_rec_alloc_planes(ctx, 1);
        st->vcap = kind == QSB_TAB ? 0 : 1 << (2 * n);
        st->gen = ctx->rec_gen;
        ctx->live_states->insert(st);
        *out = st;
        return QSB_OK;
    }
    int rc = qsb_dev_alloc(ctx, &st->data, st->bytes);
    if (rc == QSB_OK) rc = qsb_dev_alloc(ctx, (void**)&st->alive, ctx->nwords * sizeof(uint32_t));
    if (rc != QSB_OK) { qsb_dev_free(ctx, st->data); delete st; return rc; }
    ctx->live_states->insert(st);
    *out = st;
    return QSB_OK;
}
int qsb_mask_alloc(qsb_ctx* ctx, qsb_mask** out) {
    qsb_mask* m = new qsb_mask();
    m->ctx = ctx; m->w = nullptr; m->vplane = -1; m->gen = 0;
    if (ctx->rec) {
        m->vplane = qsb_rec_alloc_planes(ctx, 1); m->gen = ctx->rec_gen;
        ctx->live_masks->insert(m);
        *out = m;
        return QSB_OK;
    }
    int rc = qsb_dev_alloc(ctx, (void**)&m->w, ctx->nwords * sizeof(uint32_t));
    if (rc != QSB_OK) { delete m; return rc; }
    ctx->live_masks->insert(m);
    *out = m;
    return QSB_OK;
}

// ---------------------------------------------------------------------------------------------
// small word kernels (masks)
// ---------------------------------------------------------------------------------------------
__global__ void k_valid_init(uint32_t* v, int64_t nwords, int64_t ntraj) {
    int64_t w = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (w >= nwords) return;
    int64_t lo = w * 32;
    uint32_t m = 0;
    if (lo + 32 <= ntraj) m = 0xFFFFFFFFu;
    else if (lo < ntraj) m = (1u << (ntraj - lo)) - 1u;
    v[w] = m;
}
__global__ void k_mask_fill(uint32_t* d, const uint32_t* valid, int64_t nwords, int fill) {
    int64_t w = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (w < nwords) d[w] = fill ? valid[w] : 0u;
}
__global__ void k_mask_logic(uint32_t* d, const uint32_t* a, const uint32_t* b, const uint32_t* valid, int64_t nwords, int op) {
    int64_t w = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (w >= nwords) return;
    uint32_t x = a[w], y = b ? b[w] : 0u, r;
    switch (op) {
        case QSB_MASK_AND: r = x & y; break;
        case QSB_MASK_OR: r = x | y; break;
        case QSB_MASK_XOR: r = x ^ y; break;
        case QSB_MASK_ANDNOT: r = x & ~y; break;
        case QSB_MASK_NOT: r = ~x; break;
        default: r = ~(x ^ y); break;
    }
    d[w] = r & valid[w];
}
__global__ void k_mask_count(const uint32_t* a, int64_t nwords, long long* out) {
    long long c = 0;
    for (int64_t w = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; w < nwords; w += (int64_t)gridDim.x * blockDim.x) c += __popc(a[w]);
    for (int o = 16; o; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
    if ((threadIdx.x & 31) == 0 && c) atomicAdd((unsigned long long*)out, (unsigned long long)c);
}
__global__ void k_mask_random(uint32_t* d, const uint32_t* valid, int64_t npad, uint64_t seed, uint64_t traj0, uint64_t draw_id, double p) {
    int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (t >= npad) return;
    bool b = qsb_draw(seed, traj0 + (uint64_t)t, draw_id) < p;
    uint32_t word = __ballot_sync(0xffffffffu, b);
    if ((threadIdx.x & 31) == 0) d[t >> 5] = word & valid[t >> 5];
}
__global__ void k_kill(uint32_t* alive, const uint32_t* lanes, int64_t nwords) {
    int64_t w = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (w < nwords) alive[w] &= ~lanes[w];
}

static inline unsigned wgrid(int64_t nwords) { return (unsigned)((nwords + 255) / 256); }

int qsb_handle_ok(const qsb_ctx* ctx, bool is_virtual, int gen) {
    if (ctx->rec) {
        if (!is_virtual) QSB_FAIL(QSB_ERR_BAD_ARG, "a handle created outside the recording cannot be used inside it");
        if (gen != ctx->rec_gen) QSB_FAIL(QSB_ERR_BAD_ARG, "handle belongs to an earlier recording");
    } else if (is_virtual) QSB_FAIL(QSB_ERR_BAD_ARG, "handle belongs to a recorded program and has no storage of its own");
    return QSB_OK;
}
#define ST_OK(st) QSB_TRY(qsb_handle_ok((st)->ctx, (st)->vbase >= 0, (st)->gen))
#define MK_OK(m) QSB_TRY(qsb_handle_ok((m)->ctx, (m)->vplane >= 0, (m)->gen))
#define NO_REC(ctx, what) do { if ((ctx)->rec) QSB_FAIL(QSB_ERR_UNSUPPORTED, what " is host-visible and cannot be recorded"); } while (0)

// ---------------------------------------------------------------------------------------------
// context
// ---------------------------------------------------------------------------------------------
extern "C" int qsb_ctx_create(int device, int64_t ntraj, int64_t traj_offset, uint64_t seed, void* cuda_stream, qsb_ctx** out) {
    if (!out || ntraj <= 0) QSB_FAIL(QSB_ERR_BAD_ARG, "qsb_ctx_create: ntraj must be positive");
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0) {
        cudaGetLastError();
        QSB_FAIL(QSB_ERR_CUDA, "qsb_ctx_create: no CUDA device available (%s); libqsb200 has no CPU fallback",
                 e == cudaSuccess ? "device count is 0" : cudaGetErrorString(e));
    }
    QSB_CUDA(cudaSetDevice(device));
    qsb_ctx* ctx = new qsb_ctx();
    memset(ctx, 0, sizeof(*ctx));
    ctx->live_states = new std::set<qsb_state*>();
    ctx->live_masks = new std::set<qsb_mask*>();
    ctx->device = device;
    ctx->ntraj = ntraj;
    ctx->npad = (ntraj + 31) / 32 * 32;
    ctx->nwords = ctx->npad / 32;
    ctx->traj_offset = traj_offset;
    ctx->seed = seed;
    if (cuda_stream) { ctx->stream = (cudaStream_t)cuda_stream; ctx->own_stream = false; }
    else { QSB_CUDA(cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking)); ctx->own_stream = true; }
    cudaDeviceProp prop;
    QSB_CUDA(cudaGetDeviceProperties(&prop, device));
    ctx->sm_count = prop.multiProcessorCount;
    cudaMemPool_t pool;
    QSB_CUDA(cudaDeviceGetDefaultMemPool(&pool, device));
    uint64_t thr = UINT64_MAX;
    QSB_CUDA(cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &thr));
    QSB_TRY(qsb_dev_alloc(ctx, (void**)&ctx->valid, ctx->nwords * sizeof(uint32_t)));
    QSB_TRY(qsb_dev_alloc(ctx, (void**)&ctx->dev_flags, 4 * sizeof(int)));
    QSB_TRY(qsb_dev_alloc(ctx, (void**)&ctx->stats_tmp, 8 * sizeof(double)));
    QSB_TRY(qsb_dev_alloc(ctx, (void**)&ctx->counts_tmp, 8 * sizeof(long long)));
    QSB_CUDA(cudaMemsetAsync(ctx->dev_flags, 0, 4 * sizeof(int), ctx->stream));
    QSB_PRE(ctx, "k_valid_init");
    k_valid_init<<<wgrid(ctx->nwords), 256, 0, ctx->stream>>>(ctx->valid, ctx->nwords, ctx->ntraj);
    QSB_LAUNCH_CHECK(ctx);
    *out = ctx;
    return QSB_OK;
}
extern "C" int qsb_ctx_destroy(qsb_ctx* ctx) {
    if (!ctx) return QSB_OK;
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    for (qsb_state* s : *ctx->live_states) { cudaFreeAsync(s->data, ctx->stream); cudaFreeAsync(s->alive, ctx->stream); delete s; }
    for (qsb_mask* m : *ctx->live_masks) { cudaFreeAsync(m->w, ctx->stream); delete m; }
    delete ctx->live_states; delete ctx->live_masks;
    cudaFreeAsync(ctx->valid, ctx->stream);
    cudaFreeAsync(ctx->dev_flags, ctx->stream);
    cudaFreeAsync(ctx->stats_tmp, ctx->stream);
    cudaFreeAsync(ctx->counts_tmp, ctx->stream);
    if (ctx->scratch) cudaFreeAsync(ctx->scratch, ctx->stream);
    if (ctx->obs_buf) cudaFreeAsync(ctx->obs_buf, ctx->stream);
    if (ctx->wave_bits) cudaFreeAsync(ctx->wave_bits, ctx->stream);
    if (ctx->wave_jobs) cudaFreeAsync(ctx->wave_jobs, ctx->stream);
    if (ctx->route_bins) cudaFreeAsync(ctx->route_bins, ctx->stream);
    cudaStreamSynchronize(ctx->stream);
    if (ctx->prof) qsb_ctx_profile(ctx, 0);
    cudaMemPool_t pool;
    if (cudaDeviceGetDefaultMemPool(&pool, ctx->device) == cudaSuccess) cudaMemPoolTrimTo(pool, 0);
    if (ctx->own_stream) cudaStreamDestroy(ctx->stream);
    delete ctx;
    return QSB_OK;
}
extern "C" int qsb_ctx_set_stream(qsb_ctx* ctx, void* s) {
    if (!ctx) QSB_FAIL(QSB_ERR_BAD_ARG, "null context");
    QSB_CUDA(cudaStreamSynchronize(ctx->stream));
    if (ctx->own_stream) { cudaStreamDestroy(ctx->stream); ctx->own_stream = false; }
    if (s) ctx->stream = (cudaStream_t)s;
    else { QSB_CUDA(cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking)); ctx->own_stream = true; }
    return QSB_OK;
}
extern "C" int qsb_ctx_sync(qsb_ctx* ctx) {
    if (!ctx) QSB_FAIL(QSB_ERR_BAD_ARG, "null context");
    qsb_use_device(ctx);
    int flags[4];
    QSB_CUDA(cudaMemcpyAsync(flags, ctx->dev_flags, sizeof(flags), cudaMemcpyDeviceToHost, ctx->stream));
    QSB_CUDA(cudaStreamSynchronize(ctx->stream));
    if (flags[0]) {
        cudaMemsetAsync(ctx->dev_flags, 0, 4 * sizeof(int), ctx->stream);
        QSB_FAIL(QSB_ERR_NOT_NORMALIZED, "State not normalized. Could be due to passing wrong state to `initialize!`");
    }
    return QSB_OK;
}
extern "C" int64_t qsb_ctx_ntraj(const qsb_ctx* ctx) { return ctx ? ctx->ntraj : 0; }
extern "C" int qsb_ctx_set_traj_offset(qsb_ctx* ctx, int64_t traj_offset) {
    if (!ctx || traj_offset < 0) QSB_FAIL(QSB_ERR_BAD_ARG, "qsb_ctx_set_traj_offset: bad argument");
    if (ctx->rec) QSB_FAIL(QSB_ERR_BAD_ARG, "the trajectory ids cannot change inside a recording (use traj_shift of qsb_program_run)");
    ctx->traj_offset = traj_offset;   // kernels read it at launch: stream order keeps earlier launches on the old ids
    return QSB_OK;
}
extern "C" int qsb_ctx_counters(const qsb_ctx* ctx, uint64_t c[4]) {
    if (!ctx) QSB_FAIL(QSB_ERR_BAD_ARG, "null context");
    c[0] = ctx->launches; c[1] = ctx->algo_bytes; c[2] = ctx->ops; c[3] = ctx->draws;
    return QSB_OK;
}
extern "C" int qsb_ctx_reset_counters(qsb_ctx* ctx) {
    if (!ctx) QSB_FAIL(QSB_ERR_BAD_ARG, "null context");
    ctx->launches = ctx->algo_bytes = ctx->ops = ctx->draws = 0;
    return QSB_OK;
}

// ---------------------------------------------------------------------------------------------
// masks
// ---------------------------------------------------------------------------------------------
extern "C" int qsb_mask_create(qsb_ctx* ctx, int fill, qsb_mask** out) {
    if (!ctx || !out) QSB_FAIL(QSB_ERR_BAD_ARG, "null argument");
    qsb_mask* m;
    QSB_TRY(qsb_mask_alloc(ctx, &m));
    if (ctx->rec) { TabOp* op = qsb_rec_push(ctx, TOP_MASK_FILL); op->m[0] = m->vplane; op->iv[0] = fill; *out = m; return QSB_OK; }
    QSB_PRE(ctx, "k_mask_fill");
    k_mask_fill<<<wgrid(ctx->nwords), 256, 0, ctx->stream>>>(m->w, ctx->valid, ctx->nwords, fill);
    QSB_LAUNCH_CHECK(ctx);
    *out = m;
    return QSB_OK;
}
extern "C" int qsb_mask_free(qsb_mask* m) {
    if (!m) return QSB_OK;
    m->ctx->live_masks->erase(m);
    if (m->vplane >= 0) { if (m->ctx->rec && m->gen == m->ctx->rec_gen) qsb_rec_free_planes(m->ctx, m->vplane, 1); delete m; return QSB_OK; }
    int rc = qsb_dev_free(m->ctx, m->w);
    delete m;
    return rc;
}
extern "C" int qsb_mask_logic(qsb_mask* dst, const qsb_mask* a, const qsb_mask* b, int op) {
    if (!dst || !a || (op != QSB_MASK_NOT && !b)) QSB_FAIL(QSB_ERR_BAD_ARG, "qsb_mask_logic: null operand");
    qsb_ctx* ctx = dst->ctx;
    MK_OK(dst); MK_OK(a); if (b) MK_OK(b);
    if (ctx->rec) { TabOp* o = qsb_rec_push(ctx, TOP_MASK_LOGIC); o->m[0] = dst->vplane; o->m[1] = a->vplane; o->m[2] = b ? b->vplane : -1; o->iv[0] = op; return QSB_OK; }
    QSB_PRE(ctx, "k_mask_logic");
    k_mask_logic<<<wgrid(ctx->nwords), 256, 0, ctx->stream>>>(dst->w, a->w, b ? b->w : nullptr, ctx->valid, ctx->nwords, op);
    QSB_LAUNCH_CHECK(ctx);
    return QSB_OK;
}
extern "C" int qsb_mask_count(const qsb_mask* m, int64_t* out) {
    if (!m || !out) QSB_FAIL(QSB_ERR_BAD_ARG, "null argument");
    qsb_ctx* ctx = m->ctx;
    NO_REC(ctx, "qsb_mask_count"); MK_OK(m);
    QSB_CUDA(cudaMemsetAsync(ctx->counts_tmp, 0, sizeof(long long), ctx->stream));
    unsigned g = wgrid(ctx->nwords); if (g > 1024) g = 1024;
    QSB_PRE(ctx, "k_mask_count");
    k_mask_count<<<g, 256, 0, ctx->stream>>>(m->w, ctx->nwords, ctx->counts_tmp);
    QSB_LAUNCH_CHECK(ctx);
    long long h = 0;
    QSB_CUDA(cudaMemcpyAsync(&h, ctx->counts_tmp, sizeof(h), cudaMemcpyDeviceToHost, ctx->stream));
    QSB_CUDA(cudaStreamSynchronize(ctx->stream));
    *out = h;
    return QSB_OK;
}
extern "C" int qsb_mask_to_host(const qsb_mask* m, uint32_t* words) {
    if (!m || !words) QSB_FAIL(QSB_ERR_BAD_ARG, "null argument");
    NO_REC(m->ctx, "qsb_mask_to_host"); MK_OK(m);
    QSB_CUDA(cudaMemcpyAsync(words, m->w, m->ctx->nwords * sizeof(uint32_t), cudaMemcpyDeviceToHost, m->ctx->stream));
    QSB_CUDA(cudaStreamSynchronize(m->ctx->stream));
    return QSB_OK;
}
extern "C" int qsb_mask_from_host(qsb_mask* m, const uint32_t* words) {
    if (!m || !words) QSB_FAIL(QSB_ERR_BAD_ARG, "null argument");
    qsb_ctx* ctx = m->ctx;
    NO_REC(ctx, "qsb_mask_from_host"); MK_OK(m);
    QSB_CUDA(cudaMemcpyAsync(m->w, words, ctx->nwords * sizeof(uint32_t), cudaMemcpyHostToDevice, ctx->stream));
    QSB_PRE(ctx, "k_mask_logic");
    k_mask_logic<<<wgrid(ctx->nwords), 256, 0, ctx->stream>>>(m->w, m->w, m->w, ctx->valid, ctx->nwords, QSB_MASK_AND);
    QSB_LAUNCH_CHECK(ctx);
    return QSB_OK;
}
extern "C" int qsb_mask_random(qsb_ctx* ctx, double p, uint64_t draw_id, qsb_mask** out) {
    if (!ctx || !out) QSB_FAIL(QSB_ERR_BAD_ARG, "null argument");
    qsb_mask* m;
    QSB_TRY(qsb_mask_alloc(ctx, &m));
    if (ctx->rec) { TabOp* o = qsb_rec_push(ctx, TOP_MASK_RANDOM); o->m[0] = m->vplane; o->f[0] = p; o->draw = draw_id; ctx->draws++; *out = m; return QSB_OK; }
    QSB_PRE(ctx, "k_mask_random");
    k_mask_random<<<(unsigned)((ctx->npad + 255) / 256), 256, 0, ctx->stream>>>(m->w, ctx->valid, ctx->npad, ctx->seed, (uint64_t)ctx->traj_offset, draw_id, p);
    QSB_LAUNCH_CHECK(ctx);
    ctx->draws++;
    *out = m;
    return QSB_OK;
}
extern "C" int qsb_state_alive(const qsb_state* st, qsb_mask** out) {
    if (!st || !out) QSB_FAIL(QSB_ERR_BAD_ARG, "null argument");
    ST_OK(st);
    qsb_mask* m;
    QSB_TRY(qsb_mask_alloc(st->ctx, &m));
    if (st->ctx->rec) { TabOp* o = qsb_rec_push(st->ctx, TOP_ALIVE_COPY); o->m[0] = m->vplane; o->al[0] = st->valive; *out = m; return QSB_OK; }
    QSB_CUDA(cudaMemcpyAsync(m->w, st->alive, st->ctx->nwords * sizeof(uint32_t), cudaMemcpyDeviceToDevice, st->ctx->stream));
    *out = m;
    return QSB_OK;
}
extern "C" int qsb_state_kill(qsb_state* st, const qsb_mask* lanes) {
    if (!st || !lanes) QSB_FAIL(QSB_ERR_BAD_ARG, "null argument");
    qsb_ctx* ctx = st->ctx;
    ST_OK(st); MK_OK(lanes);
    if (ctx->rec) { TabOp* o = qsb_rec_push(ctx, TOP_KILL); o->al[0] = st->valive; o->m[0] = lanes->vplane; return QSB_OK; }
    QSB_PRE(ctx, "k_kill");
    k_kill<<<wgrid(ctx->nwords), 256, 0, ctx->stream>>>(st->alive, lanes->w, ctx->nwords);
    QSB_LAUNCH_CHECK(ctx);
    return QSB_OK;
}

// ---------------------------------------------------------------------------------------------
// state lifetime + dispatch on kind
// ---------------------------------------------------------------------------------------------
static int check_idx(const qsb_state* st, const int* idx, int k) {
    if (!st) QSB_FAIL(QSB_ERR_BAD_ARG, "null state");
    ST_OK(st);
    if (k < 1 || k > st->n) QSB_FAIL(QSB_ERR_BAD_INDEX, "%d indices given for a state of %d subsystems", k, st->n);
    for (int a = 0; a < k; ++a) {
        if (idx[a] < 0 || idx[a] >= st->n) QSB_FAIL(QSB_ERR_BAD_INDEX, "subsystem index %d out of range for a state of %d subsystems", idx[a] + 1, st->n);
        for (int b = 0; b < a; ++b) if (idx[a] == idx[b]) QSB_FAIL(QSB_ERR_BAD_INDEX, "subsystem index %d repeated", idx[a] + 1);
    }
    return QSB_OK;
}

extern "C" int qsb_state_new(qsb_ctx* ctx, int kind, int nsub, qsb_state** out) {
    if (!ctx || !out || nsub < 1) QSB_FAIL(QSB_ERR_BAD_ARG, "qsb_state_new: bad argument");
    if (kind == QSB_TAB) { if (nsub > QSB_TAB_MAXN) QSB_FAIL(QSB_ERR_UNSUPPORTED, "tableaux are limited to %d qubits", QSB_TAB_MAXN); return qsbt_new(ctx, nsub, 1, out); }
    if (!qsb_is_dense(kind)) QSB_FAIL(QSB_ERR_BAD_ARG, "unknown state kind %d", kind);
    if (nsub > QSB_DENSE_MAXN) QSB_FAIL(QSB_ERR_UNSUPPORTED, "dense states are limited to %d qubits", QSB_DENSE_MAXN);
    return qsbd_new(ctx, kind, nsub, out);
}
extern "C" int qsb_state_from_host(qsb_ctx* ctx, int kind, int nsub, const double* data, qsb_state** out) {
    if (!ctx || !out || !data || nsub < 1) QSB_FAIL(QSB_ERR_BAD_ARG, "qsb_state_from_host: bad argument");
    if (!qsb_is_dense(kind)) QSB_FAIL(QSB_ERR_UNSUPPORTED, "qsb_state_from_host needs a dense kind; use qsb_tab_from_host");
    if (nsub > QSB_DENSE_MAXN) QSB_FAIL(QSB_ERR_UNSUPPORTED, "dense states are limited to %d qubits", QSB_DENSE_MAXN);
    return qsbd_from_host(ctx, kind, nsub, data, out);
}
extern "C" int qsb_tab_from_host(qsb_ctx* ctx, int nsub, int ngen, const uint8_t* xs, const uint8_t* zs, const uint8_t* signs, qsb_state** out) {
    if (!ctx || !out || nsub < 1 || ngen < 0 || ngen > nsub) QSB_FAIL(QSB_ERR_BAD_ARG, "qsb_tab_from_host: bad argument");
    if (nsub > QSB_TAB_MAXN) QSB_FAIL(QSB_ERR_UNSUPPORTED, "tableaux are limited to %d qubits", QSB_TAB_MAXN);
    return qsbt_from_host(ctx, nsub, ngen, xs, zs, signs, out);
}
extern "C" int qsb_state_select(qsb_ctx* ctx, const qsb_state* const* states, const double* weights, int m, uint64_t draw_id, qsb_state** out) {
    if (!ctx || !states || !weights || m < 1 || !out) QSB_FAIL(QSB_ERR_BAD_ARG, "qsb_state_select: bad argument");
    for (int i = 0; i < m; ++i) {
        if (!states[i] || states[i]->ctx != ctx) QSB_FAIL(QSB_ERR_BAD_ARG, "qsb_state_select: state %d is null or from another context", i);
        if (states[i]->kind != states[0]->kind || states[i]->n != states[0]->n) QSB_FAIL(QSB_ERR_UNSUPPORTED, "qsb_state_select: all states must share kind and size");
        ST_OK(states[i]);
    }
    ctx->draws++;
    return states[0]->kind == QSB_TAB ? qsbt_select(ctx, states, weights, m, draw_id, out) : qsbd_select(ctx, states, weights, m, draw_id, out);
}
extern "C" int qsb_state_copy(const qsb_state* st, qsb_state** out) {
    if (!st || !out) QSB_FAIL(QSB_ERR_BAD_ARG, "null argument");
    ST_OK(st);
    qsb_state* c;
    QSB_TRY(qsb_state_alloc(st->ctx, st->kind, st->n, &c));
    if (st->ctx->rec) {
        TabOp* o = qsb_rec_push(st->ctx, st->kind == QSB_TAB ? TOP_COPY : DOP_COPY); o->n = st->n; o->s[0] = c->vbase; o->s[1] = st->vbase; o->al[0] = c->valive; o->al[1] = st->valive;
        *out = c; return QSB_OK;
    }
    QSB_CUDA(cudaMemcpyAsync(c->data, st->data, st->bytes, cudaMemcpyDeviceToDevice, st->ctx->stream));
    QSB_CUDA(cudaMemcpyAsync(c->alive, st->alive, st->ctx->nwords * sizeof(uint32_t), cudaMemcpyDeviceToDevice, st->ctx->stream));
    *out = c;
    return QSB_OK;
}
extern "C" int qsb_state_free(qsb_state* st) {
    if (!st) return QSB_OK;
    st->ctx->live_states->erase(st);
    if (st->vbase >= 0) {
        if (st->ctx->rec && st->gen == st->ctx->rec_gen) { qsb_rec_free(st->ctx, st->kind == QSB_TAB ? 0 : 1, st->vbase); qsb_rec_free(st->ctx, 0, st->valive); }
        delete st; return QSB_OK;
    }
    int rc = qsb_dev_free(st->ctx, st->data);
    int rc2 = qsb_dev_free(st->ctx, st->alive);
    delete st;
    return rc ? rc : rc2;
}
extern "C" int qsb_state_kind(const qsb_state* st) { return st ? st->kind : -1; }
extern "C" int qsb_state_nsubsystems(const qsb_state* st) { return st ? st->n : -1; }
extern "C" int qsb_state_promote_dm(qsb_state* st) {
    if (!st) QSB_FAIL(QSB_ERR_BAD_ARG, "null state");
    if (st->kind == QSB_DM) return QSB_OK;
    if (!qsb_is_ketlike(st->kind)) QSB_FAIL(QSB_ERR_UNSUPPORTED, "dm() of a tableau is not defined");
    return qsbd_promote_dm(st);
}
extern "C" int qsb_state_to_host(const qsb_state* st, int64_t traj, double* out) {
    if (!st || !out) QSB_FAIL(QSB_ERR_BAD_ARG, "null argument");
    if (traj < 0 || traj >= st->ctx->ntraj) QSB_FAIL(QSB_ERR_BAD_INDEX, "trajectory %lld out of range", (long long)traj);
    NO_REC(st->ctx, "qsb_state_to_host"); ST_OK(st);
    if (!qsb_is_dense(st->kind)) QSB_FAIL(QSB_ERR_UNSUPPORTED, "use qsb_tab_to_host for tableaux");
    return qsbd_to_host(st, traj, out);
}
extern "C" int qsb_tab_to_host(const qsb_state* st, int64_t traj, uint8_t* xs, uint8_t* zs, uint8_t* signs, uint8_t* active) {
    if (!st) QSB_FAIL(QSB_ERR_BAD_ARG, "null argument");
    if (traj < 0 || traj >= st->ctx->ntraj) QSB_FAIL(QSB_ERR_BAD_INDEX, "trajectory %lld out of range", (long long)traj);
    NO_REC(st->ctx, "qsb_tab_to_host"); ST_OK(st);
    if (st->kind != QSB_TAB) QSB_FAIL(QSB_ERR_UNSUPPORTED, "not a tableau");
    return qsbt_to_host(st, traj, xs, zs, signs, active);
}

extern "C" int qsb_tensor(qsb_ctx* ctx, const qsb_state* const* states, int m, qsb_state** out) {
    if (!ctx || !states || m < 1 || !out) QSB_FAIL(QSB_ERR_BAD_ARG, "qsb_tensor: bad argument");
    for (int i = 0; i < m; ++i) { if (!states[i] || states[i]->ctx != ctx) QSB_FAIL(QSB_ERR_BAD_ARG, "qsb_tensor: state %d is null or from another context", i); ST_OK(states[i]); }
    bool any_tab = false, all_tab = true;
    for (int i = 0; i < m; ++i) { any_tab |= states[i]->kind == QSB_TAB; all_tab &= states[i]->kind == QSB_TAB; }
    if (any_tab && !all_tab) QSB_FAIL(QSB_ERR_UNSUPPORTED, "subsystemcompose has no method for a tableau and a dense state");
    if (m == 1) return qsb_state_copy(states[0], out);
    qsb_state* acc = nullptr;
    for (int i = 1; i < m; ++i) {
        const qsb_state* left = acc ? acc : states[0];
        qsb_state* next = nullptr;
        int rc = all_tab ? qsbt_tensor2(ctx, left, states[i], &next) : qsbd_tensor2(ctx, left, states[i], &next);
        if (acc) qsb_state_free(acc);
        if (rc != QSB_OK) return rc;
        acc = next;
    }
    *out = acc;
    return QSB_OK;
}
extern "C" int qsb_apply(qsb_state* st, const int* idx, int k, const double* U, const qsb_mask* mask) {
    QSB_TRY(check_idx(st, idx, k));
    if (!U) QSB_FAIL(QSB_ERR_BAD_ARG, "null operator");
    if (!qsb_is_dense(st->kind)) QSB_FAIL(QSB_ERR_UNSUPPORTED, "apply! of a dense operator on a tableau; lower the gate with qsb_apply_clifford");
    if (k > 2) QSB_FAIL(QSB_ERR_UNSUPPORTED, "dense operators on more than 2 subsystems are not implemented");
    return qsbd_apply(st, idx, k, U, mask);
}
extern "C" int qsb_apply_clifford(qsb_state* st, const int* idx, int k, const uint8_t* table, const qsb_mask* mask) {
    QSB_TRY(check_idx(st, idx, k));
    if (!table) QSB_FAIL(QSB_ERR_BAD_ARG, "null table");
    if (st->kind != QSB_TAB) QSB_FAIL(QSB_ERR_UNSUPPORTED, "qsb_apply_clifford needs a tableau state");
    if (k > 2) QSB_FAIL(QSB_ERR_UNSUPPORTED, "Clifford gates on more than 2 qubits are not implemented");
    return qsbt_apply(st, idx, k, table, mask);
}
extern "C" int qsb_channel(qsb_state* st, int idx, int bg, const double params[2], double dt, uint64_t draw_id) {
    QSB_TRY(check_idx(st, &idx, 1));
    if (!params || bg < 0 || bg > 3) QSB_FAIL(QSB_ERR_BAD_ARG, "qsb_channel: bad background");
    if (st->kind == QSB_TAB) return qsbt_channel(st, idx, bg, params, dt, draw_id);
    return qsbd_channel(st, idx, bg, params, dt, draw_id);
}
extern "C" int qsb_evolve(qsb_state* st, const int* idx, int k, const double* H_ri, int nl, const int* l_pos, const double* L_ri, double duration, uint64_t draw_id) {
    QSB_TRY(check_idx(st, idx, k));
    if (!qsb_is_dense(st->kind)) QSB_FAIL(QSB_ERR_UNSUPPORTED, "apply_noninstant! has no method for a tableau state");
    if (k > 2) QSB_FAIL(QSB_ERR_UNSUPPORTED, "Hamiltonians on more than 2 subsystems are not implemented");
    if (nl < 0 || nl > 8 || (nl > 0 && (!l_pos || !L_ri))) QSB_FAIL(QSB_ERR_BAD_ARG, "qsb_evolve: bad collapse operator list");
    for (int j = 0; j < nl; ++j) if (l_pos[j] < 0 || l_pos[j] >= k) QSB_FAIL(QSB_ERR_BAD_INDEX, "qsb_evolve: collapse operator %d acts on entry %d of %d indices", j, l_pos[j], k);
    if (!(duration >= 0)) QSB_FAIL(QSB_ERR_BAD_ARG, "qsb_evolve: negative duration");
    return qsbd_evolve(st, idx, k, H_ri, nl, l_pos, L_ri, duration, draw_id);
}
extern "C" int qsb_measure(qsb_state* st, int idx, int basis, const double* kets, uint64_t draw_id,
                           qsb_mask** outcome2, qsb_mask** was_random, int64_t* counts_host) {
    QSB_TRY(check_idx(st, &idx, 1));
    if (!outcome2) QSB_FAIL(QSB_ERR_BAD_ARG, "qsb_measure: outcome2 must not be null");
    qsb_ctx* ctx = st->ctx;
    if (ctx->rec && counts_host) QSB_FAIL(QSB_ERR_UNSUPPORTED, "qsb_measure with counts_host is host-visible and cannot be recorded");
    qsb_mask *o = nullptr, *r = nullptr;
    QSB_TRY(qsb_mask_alloc(ctx, &o));
    int rc;
    if (st->kind == QSB_TAB) {
        if (basis < QSB_BASIS_X || basis > QSB_BASIS_Z) { qsb_mask_free(o); QSB_FAIL(QSB_ERR_UNSUPPORTED, "Clifford measurements support the X, Y and Z bases only"); }
        if (was_random) { rc = qsb_mask_alloc(ctx, &r); if (rc) { qsb_mask_free(o); return rc; } }
        rc = qsbt_measure(st, idx, basis, draw_id, o, r, counts_host);
    } else {
        static const double s = 0.70710678118654752440;
        double std_kets[3][8] = {{s, 0, s, 0, s, 0, -s, 0}, {s, 0, 0, s, s, 0, 0, -s}, {1, 0, 0, 0, 0, 0, 1, 0}};
        const double* kk = kets;
        if (basis >= QSB_BASIS_X && basis <= QSB_BASIS_Z) kk = std_kets[basis - 1];
        else if (basis != QSB_BASIS_KETS || !kets) { qsb_mask_free(o); QSB_FAIL(QSB_ERR_BAD_ARG, "qsb_measure: bad basis"); }
        rc = qsbd_measure(st, idx, kk, draw_id, o, counts_host);
        if (was_random) *was_random = nullptr;
    }
    if (rc != QSB_OK) { qsb_mask_free(o); if (r) qsb_mask_free(r); return rc; }
    ctx->draws++;
    *outcome2 = o;
    if (was_random && st->kind == QSB_TAB) *was_random = r;
    return QSB_OK;
}
extern "C" int qsb_ptrace(qsb_state* st, int idx, uint64_t draw_id) {
    QSB_TRY(check_idx(st, &idx, 1));
    if (st->n == 1) QSB_FAIL(QSB_ERR_TRACE_ALL, "Partial trace can't be used to trace out all subsystems - use tr() instead.");
    if (st->kind == QSB_TAB) return qsbt_ptrace(st, idx);
    return qsbd_ptrace(st, idx, draw_id);
}
extern "C" int qsb_expect(qsb_state* st, const int* idx, int k, int obs_kind, const void* obs, int obs_sign,
                          double* per_traj, double* stats_dev, double* stats_host) {
    QSB_TRY(check_idx(st, idx, k));
    if (!obs) QSB_FAIL(QSB_ERR_BAD_ARG, "null observable");
    if (st->ctx->rec && stats_host) QSB_FAIL(QSB_ERR_UNSUPPORTED, "qsb_expect with stats_host is host-visible and cannot be recorded; pass stats_dev");
    if (st->kind == QSB_TAB) return qsbt_expect(st, idx, k, obs_kind, obs, obs_sign, per_traj, stats_dev, stats_host);
    return qsbd_expect(st, idx, k, obs_kind, obs, obs_sign, per_traj, stats_dev, stats_host);
}
