// qsb_exec.cu -- recorded tableau programs: capture the op list of a lockstep schedule once, then execute
// the WHOLE list in one launch with every trajectory word's state resident in shared memory.
//
// Why: at the BASELINE configs the entire Clifford ensemble is a few MB and a per-op launch is ~31k
// threads of dependent word logic (3-38 us each, ~1700 launches per pass).  A thread only ever touches
// its own 32-trajectory word column of every plane, so consecutive ops need no grid synchronisation:
// one thread can run the complete op list on its column.  The arena (all live states and masks of the
// schedule, planes allocated/recycled at record time) is a [planes][blockDim] tile of shared memory.
//
// Lanes per thread: this translation unit compiles the tableau core with 8-lane BYTE planes (QSB_PLANE_T), so an
// ensemble of N trajectories runs on N/8 threads instead of N/32 -- the bit logic is latency bound (dependent
// shared-memory round trips), and 10^6 trajectories are only 977 warps of 32-lane words on 148 SMs.
#define QSB_PLANE_T uint8_t
#define QSB_CORE_NS tabcore8
#include <map>

#include "qsb_internal.cuh"
#include "qsb_tab_core.cuh"

struct qsb_recorder {
    std::vector<TabOp> ops;
    std::vector<TabObsArgs> obs;
    std::vector<double*> stats_ptrs;
    int next_plane = 0, high = 0;
    std::multimap<int, int> free_blocks;  // size -> base
    uint64_t ops0 = 0, draws0 = 0;        // context counters at qsb_record_begin (recording itself executes nothing)
};

// one stochastic op of the program as the draw pass sees it
struct DrawSpec {
    uint64_t draw;
    int type, np, slot, pad;   // type: TOP_NOISE / TOP_SELECT / TOP_MEASURE / TOP_MASK_RANDOM
    double cum[5];
    int op[3];
};

struct qsb_program {
    qsb_ctx* ctx;
    DrawSpec* d_specs; int nspecs, nslots;
    uint8_t* d_aux; int naux, nfolded;   // constant tables of folded ops
    uint32_t* bits;        // [nslots][nwords] decision words written by k_draw_bits, read by k_tab_exec
    TabOp* d_ops; int nops;
    TabObsArgs* d_obs; int nobs;
    double** d_stats; int nstats;
    int planes;
    plane_t* arena;        // global-memory arena, only when the shared-memory tile does not fit
    int threads, smem_bytes;
    TabOp* h_ops;          // pinned host copy of the op list (qsb_program_upload re-sends it)
    uint64_t ops_per_run, draws_per_run, per_traj_rows;
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
    if (R->next_plane > R->high) {
        R->high = R->next_plane;
        if (getenv("QSB_REC_TRACE")) fprintf(stderr, "[rec] op %zu: high-water %d planes (block of %d)\n", R->ops.size(), R->high, nplanes);
    }
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
void qsb_rec_shrink(qsb_ctx* ctx, int base, int old_planes, int new_planes) {
    if (old_planes <= 0) old_planes = 1;
    if (new_planes <= 0) new_planes = 1;   // a 0-qubit state keeps one (unused) plane, like a fresh allocation would
    if (new_planes < old_planes) qsb_rec_free_planes(ctx, base + new_planes, old_planes - new_planes);
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
    const uint8_t* aux;
    plane_t* arena;           // nullptr -> shared-memory tile
    const plane_t* bits;      // decision bits of every stochastic op, [slot][W] (the draw pass's words seen lane-group wise)
    const plane_t* valid;
    size_t W;                 // lane groups (threads) in the ensemble = trajectories / QSB_LANES
    uint64_t seed, traj0, draw_base;
};

__global__ void k_zero_stats(double* const* ptrs, int n) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) { double* p = ptrs[i]; p[0] = 0; p[1] = 0; p[2] = 0; p[3] = 0; }
}

// Draw pass: ONE THREAD PER TRAJECTORY evaluates every stochastic op's Philox draw and its decision; a warp
// ballot packs the 32 decisions of a word, so the bit-sliced pass below never runs Philox (which a
// thread-per-word kernel would have to loop over its 32 lanes for, at 1/32 of the parallelism).
__global__ void __launch_bounds__(256) k_draw_bits(const DrawSpec* __restrict__ specs, int nspecs, uint32_t* __restrict__ bits,
                                                   size_t W, uint64_t seed, uint64_t traj0, uint64_t draw_base) {
    const size_t t = blockIdx.x * (size_t)blockDim.x + threadIdx.x;   // local trajectory; W*32 is a multiple of 32
    if (t >= W * 32) return;                                          // whole warps leave together
    const size_t wg = t >> 5;
    const unsigned lane = threadIdx.x & 31u;
    const uint64_t traj = traj0 + t;
    for (int i = 0; i < nspecs; ++i) {
        const DrawSpec& s = specs[i];
        const double u = qsb_draw(seed, traj, s.draw + draw_base);
        uint32_t d;
        if (s.type == TOP_NOISE) {
            d = 0;
            for (int j = 0; j < s.np; ++j) if (u > s.cum[j]) { int P = s.op[j]; d = ((P == 1 || P == 2) ? 1u : 0u) | ((P == 3 || P == 2) ? 2u : 0u); break; }
        } else if (s.type == TOP_SELECT) {
            d = (uint32_t)(s.np - 1);
            for (int j = 0; j < s.np; ++j) if (s.cum[j] > u) { d = (uint32_t)j; break; }
        } else if (s.type == TOP_MEASURE) {
            d = u >= 0.5 ? 1u : 0u;
        } else {
            d = u < s.cum[0] ? 1u : 0u;
        }
        const uint32_t b0 = __ballot_sync(0xffffffffu, d & 1u);
        uint32_t* dst = bits + (size_t)s.slot * W + wg;
        if (s.type == TOP_NOISE) {
            const uint32_t b1 = __ballot_sync(0xffffffffu, d & 2u);
            if (lane == 0) { dst[0] = b0; dst[W] = b1; }
        } else if (s.type == TOP_SELECT) {
            const uint32_t b1 = __ballot_sync(0xffffffffu, d & 2u), b2 = __ballot_sync(0xffffffffu, d & 4u);
            if (lane == 0) { dst[0] = b0; dst[W] = b1; dst[2 * W] = b2; }
        } else if (lane == 0) dst[0] = b0;
    }
}

// The tableau functions are force-inlined with a LITERAL qubit count so that their loops unroll and their
// per-row scratch arrays live in registers (a run-time n keeps them in local memory: 2.2 KB of stack per thread).
#define N_CASE(K, ...) case K: { constexpr int N = K; __VA_ARGS__; } break;
#define N_SWITCH(NV, ...) switch (NV) { N_CASE(1, __VA_ARGS__) N_CASE(2, __VA_ARGS__) N_CASE(3, __VA_ARGS__) N_CASE(4, __VA_ARGS__) \
                                        N_CASE(5, __VA_ARGS__) default: { const int N = (NV); __VA_ARGS__; } break; }

// T > 0: the arena is a [planes][T] tile of shared memory (block of T threads, compile-time stride);
// T == 0: the arena is in global memory with stride a.W (programs whose tile exceeds the shared memory of an SM).
template <int T>
__global__ void __launch_bounds__(T ? T : 128) k_tab_exec(ExecArgs a) {
    extern __shared__ plane_t tile[];
    const size_t wg = blockIdx.x * (size_t)blockDim.x + threadIdx.x;  // global lane group (trajectories QSB_LANES*wg ..)
    const bool on = wg < a.W;
    plane_t* const base = T ? tile : a.arena;
    const size_t W = T ? (size_t)T : a.W;
    const size_t w = T ? (size_t)threadIdx.x : (on ? wg : 0);
    const uint32_t validw = on ? a.valid[wg] : 0u;
#define PL(idx) (base + (size_t)(idx) * W)
    for (int i = 0; i < a.nops; ++i) {
        const TabOp& op = a.ops[i];
        switch (op.type) {
        case TOP_INIT:
            if (on) { N_SWITCH(op.n, tab_init_word(PL(op.s[0]), W, w, N, op.iv[0])) PL(op.al[0])[w] = validw; }
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
                const plane_t* d = a.bits + (size_t)op.dslot * a.W + wg;
                const uint32_t b0 = d[0], b1 = d[a.W], b2 = d[2 * a.W];   // bit planes of the selected index
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
                const plane_t* d = a.bits + (size_t)op.dslot * a.W + wg;
                const uint32_t b0 = d[0], b1 = d[a.W], b2 = d[2 * a.W];
                for (int j = 0; j < 5; ++j) m[j] = ((j & 1) ? b0 : ~b0) & ((j & 2) ? b1 : ~b1) & ((j & 4) ? b2 : ~b2);
                const uint8_t* tab = a.aux + op.obs;
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
                N_SWITCH(op.n, tab_gate_anf(PL(op.s[0]), W, w, N, op.k, op.q[0], op.q[1], op.outmask, op.u[0], gm))
            }
            break;
        case TOP_NOISE:
            if (on) {
                const plane_t* d = a.bits + (size_t)op.dslot * a.W + wg;
                const uint32_t fx = d[0], fz = d[a.W];
                N_SWITCH(op.n, tab_noise_apply(PL(op.s[0]), W, w, N, op.q[0], fx, fz))
            }
            break;
        case TOP_MEASURE:
            if (on) {
                uint32_t rnd;
                uint32_t outcome = 0; const uint32_t rb = a.bits[(size_t)op.dslot * a.W + wg];
                N_SWITCH(op.n, outcome = tab_measure_bits(PL(op.s[1]), PL(op.s[0]), W, w, N, op.q[0], op.iv[0], op.iv[1], rb, &rnd))
                uint32_t al = PL(op.al[1])[w];
                PL(op.m[0])[w] = outcome & al;
                if (op.m[1] >= 0) PL(op.m[1])[w] = rnd & al;
            }
            break;
        case TOP_PTRACE:
            if (on) { N_SWITCH(op.n, tab_ptrace_word(PL(op.s[1]), PL(op.s[0]), W, w, N, op.q[0])) }
            break;
        case TOP_FORCE:
            if (on) { N_SWITCH(op.n, tab_force_word(PL(op.s[0]), W, w, N, op.u[0], op.u[1], op.iv[0])) }
            break;
        case TOP_EXPECT: {
            // integer tallies (exact, order independent): alive lanes, sum of (plus - minus), sum of its square
            int n_alive = 0, num = 0, num2 = 0;
            int nelem = 1;
            if (on) {
                const TabObsArgs& o = a.obs[op.obs];
                nelem = o.nelem;
                uint32_t cnt[8], cntm[8];
                N_SWITCH(op.n, tab_expect_word_n(PL(op.s[0]), W, w, N, o, cnt, cntm))
                const uint32_t al = PL(op.al[0])[w];
                n_alive = __popc(al);
                if (nelem == 1 && !op.per_traj) {   // a single Pauli: each lane holds -1, 0 or +1
                    num = __popc(cnt[0] & al) - __popc(cntm[0] & al);
                    num2 = __popc((cnt[0] | cntm[0]) & al);
                } else {
                    const double inv = 1.0 / (double)nelem;
                    for (int l = 0; l < QSB_LANES; ++l) {
                        int plus = 0, minus = 0;
                        for (int b = 0; b < 8; ++b) { plus |= ((cnt[b] >> l) & 1) << b; minus |= ((cntm[b] >> l) & 1) << b; }
                        const bool lane = (al >> l) & 1u;
                        const int d = lane ? plus - minus : 0;
                        if (op.per_traj) { size_t t = wg * QSB_LANES + l; op.per_traj[2 * t] = (double)d * inv; op.per_traj[2 * t + 1] = 0.0; }
                        num += d; num2 += d * d;
                    }
                }
            }
            if (op.stats) {  // whole warps reach this point together (the op list is uniform)
                n_alive = __reduce_add_sync(0xffffffffu, n_alive);
                num = __reduce_add_sync(0xffffffffu, num);
                num2 = __reduce_add_sync(0xffffffffu, num2);
                nelem = __reduce_max_sync(0xffffffffu, nelem);
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
                PL(op.m[0])[w] = a.bits[(size_t)op.dslot * a.W + wg] & validw;
            }
            break;
        case TOP_KILL:
            if (on) PL(op.al[0])[w] = (plane_t)(PL(op.al[0])[w] & ~PL(op.m[0])[w]);
            break;
        case TOP_ALIVE_COPY:
            if (on) PL(op.m[0])[w] = PL(op.al[0])[w];
            break;
        }
    }
#undef PL
}

typedef void (*exec_fn)(ExecArgs);
static exec_fn exec_kernel(int threads, bool global_arena) {
    if (global_arena) return k_tab_exec<0>;
    switch (threads) {
        case 64: return k_tab_exec<64>;
        case 128: return k_tab_exec<128>;
        case 256: return k_tab_exec<256>;
        default: return nullptr;
    }
}

// ---------------------------------------------------------------------------------------------
// C-ABI: record / run
// ---------------------------------------------------------------------------------------------
extern "C" int qsb_record_begin(qsb_ctx* ctx) {
    if (!ctx) QSB_FAIL(QSB_ERR_BAD_ARG, "null context");
    if (ctx->rec) QSB_FAIL(QSB_ERR_BAD_ARG, "already recording");
    ctx->rec = new qsb_recorder();
    ctx->rec->ops0 = ctx->ops; ctx->rec->draws0 = ctx->draws;
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
    P->ops_per_run = ctx->ops - R->ops0; P->draws_per_run = ctx->draws - R->draws0;
    ctx->ops = R->ops0; ctx->draws = R->draws0;  // counted per qsb_program_run instead
    for (auto& o : R->ops) if (o.per_traj) P->per_traj_rows++;
    int rc = QSB_OK;
    // Constant folding: a mixture's components are built by INIT (+FORCE...) ops whose result is the same on every
    // lane.  Evaluate those on the host (the tableau core is host-callable), drop them, and let the SELECT pick each
    // plane from a per-plane table of "which components have this bit set" instead of reading component tableaux.
    std::vector<uint8_t> aux;
    {
        struct Pend { std::vector<int> ops; int alive, n; };
        std::map<int, Pend> pend;   // state base plane -> the ops that built it, untouched by anything else so far
        std::vector<plane_t> h((size_t)P->planes + 8, 0);
        for (int i = 0; i < (int)R->ops.size(); ++i) {
            TabOp& o = R->ops[i];
            if (o.type == TOP_INIT) { pend[o.s[0]] = Pend{{i}, o.al[0], o.n}; continue; }
            if (o.type == TOP_FORCE) { auto it = pend.find(o.s[0]); if (it != pend.end()) it->second.ops.push_back(i); continue; }
            bool fold = o.type == TOP_SELECT && qsb_tab_planes(o.n) <= 4096;
            if (fold) for (int j = 0; j < o.n2; ++j) { auto it = pend.find(o.s[1 + j]); if (it == pend.end() || it->second.alive != o.al[1 + j] || it->second.n != o.n) fold = false; }
            if (fold) {
                const int planes = qsb_tab_planes(o.n);
                std::vector<uint8_t> tab(planes, 0);
                for (int j = 0; j < o.n2 && fold; ++j) {
                    for (int k : pend[o.s[1 + j]].ops) {
                        const TabOp& c = R->ops[k];
                        if (c.type == TOP_INIT) tab_init_word(h.data() + c.s[0], 1, 0, c.n, c.iv[0]);
                        else tab_force_word(h.data() + c.s[0], 1, 0, c.n, c.u[0], c.u[1], c.iv[0]);
                    }
                    for (int p = 0; p < planes; ++p) {
                        plane_t v = h[(size_t)o.s[1 + j] + p];
                        if (v == (plane_t)QSB_FULL) tab[p] |= 1u << j; else if (v != 0) fold = false;   // not lane-uniform: keep the ops
                    }
                }
                if (fold) {
                    for (int j = 0; j < o.n2; ++j) { for (int k : pend[o.s[1 + j]].ops) R->ops[k].type = TOP_NOP; pend.erase(o.s[1 + j]); }
                    o.type = TOP_SELECT_CONST; o.obs = (int)aux.size();
                    aux.insert(aux.end(), tab.begin(), tab.end());
                    P->nfolded++;
                }
            }
            // any other use of a pending state (or of its alive plane) makes it a real state again
            for (auto it = pend.begin(); it != pend.end();) {
                bool used = false;
                for (int k = 0; k < 6; ++k) used |= (o.s[k] == it->first) || (o.al[k] == it->second.alive);
                it = used ? pend.erase(it) : std::next(it);
            }
        }
        // drop the NOPs
        std::vector<TabOp> kept;
        for (auto& o : R->ops) if (o.type != TOP_NOP) kept.push_back(o);
        R->ops.swap(kept);
        P->nops = (int)R->ops.size(); P->naux = (int)aux.size();
    }
    // stochastic ops -> draw specs + decision-word slots (noise: fx,fz; mixture: 3 index planes; others: 1)
    std::vector<DrawSpec> specs;
    for (auto& o : R->ops) {
        int need = o.type == TOP_NOISE ? 2 : (o.type == TOP_SELECT || o.type == TOP_SELECT_CONST) ? 3 : (o.type == TOP_MEASURE || o.type == TOP_MASK_RANDOM) ? 1 : 0;
        if (!need) continue;
        DrawSpec d; memset(&d, 0, sizeof(d));
        d.draw = o.draw; d.type = o.type == TOP_SELECT_CONST ? (int)TOP_SELECT : o.type; d.slot = P->nslots;
        d.np = o.type == TOP_NOISE ? o.k : d.type == TOP_SELECT ? o.n2 : 1;
        for (int j = 0; j < 5; ++j) d.cum[j] = o.f[j];
        for (int j = 0; j < 3; ++j) d.op[j] = o.iv[j];
        o.dslot = P->nslots; P->nslots += need;
        specs.push_back(d);
    }
    P->nspecs = (int)specs.size();
    if (!R->ops.empty()) {
        if (cudaMallocHost((void**)&P->h_ops, R->ops.size() * sizeof(TabOp)) != cudaSuccess) { cudaGetLastError(); qsb_set_error("pinned allocation of the op list failed"); rc = QSB_ERR_OOM; }
        else memcpy(P->h_ops, R->ops.data(), R->ops.size() * sizeof(TabOp));
    }
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
    up((void**)&P->d_specs, specs.data(), specs.size() * sizeof(DrawSpec));
    up((void**)&P->d_aux, aux.data(), aux.size());
    if (rc == QSB_OK) rc = qsb_dev_alloc(ctx, (void**)&P->bits, (size_t)(P->nslots ? P->nslots : 1) * ctx->nwords * sizeof(uint32_t));
    if (rc == QSB_OK) { cudaError_t e = cudaStreamSynchronize(ctx->stream); if (e != cudaSuccess) rc = QSB_ERR_CUDA; }  // host vectors die below
    // launch shape: the largest block whose [planes][threads] tile fits, preferring a single wave
    int dev_smem = 0;
    cudaDeviceGetAttribute(&dev_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, ctx->device);
    const long groups = (long)ctx->nwords * (32 / QSB_LANES);   // one thread per lane group
    int best_t = 0; double best_cost = 1e300;
    for (int t = 64; t <= 256; t *= 2) {
        size_t smem = (size_t)P->planes * t * sizeof(plane_t);
        if (smem > (size_t)dev_smem) break;
        int per_sm = (int)((size_t)(228 * 1024) / (smem + 1024));
        if (per_sm < 1) per_sm = 1;
        if (per_sm > 32) per_sm = 32;
        if (per_sm * t > 2048) per_sm = 2048 / t;
        long blocks = (groups + t - 1) / t;
        long slots = (long)ctx->sm_count * per_sm;
        long waves = (blocks + slots - 1) / slots;
        // a wave costs about one thread's serial time; more resident warps per SM stretch it only a little
        double cost = (double)waves * (1.0 + 0.01 * per_sm * t / 32.0);
        if (cost < best_cost) { best_cost = cost; best_t = t; }
    }
    if (best_t) { P->threads = best_t; P->smem_bytes = P->planes * best_t * (int)sizeof(plane_t); P->arena = nullptr; }
    else {
        P->threads = 128; P->smem_bytes = 0;
        if (rc == QSB_OK) rc = qsb_dev_alloc(ctx, (void**)&P->arena, (size_t)P->planes * groups * sizeof(plane_t));
    }
    if (rc == QSB_OK && P->smem_bytes) {
        cudaError_t e = cudaFuncSetAttribute(exec_kernel(P->threads, P->arena != nullptr), cudaFuncAttributeMaxDynamicSharedMemorySize, P->smem_bytes);
        if (e != cudaSuccess) { qsb_set_error("cudaFuncSetAttribute(%d bytes) failed: %s", P->smem_bytes, cudaGetErrorString(e)); rc = QSB_ERR_CUDA; }
    }
    delete R;
    ctx->rec = nullptr;
    if (rc != QSB_OK) { if (P->h_ops) cudaFreeHost(P->h_ops); delete P; return rc; }
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
    a.ops = P->d_ops; a.nops = P->nops; a.obs = P->d_obs; a.aux = P->d_aux; a.arena = P->arena; a.valid = (const plane_t*)ctx->valid; a.W = (size_t)ctx->nwords * (32 / QSB_LANES);
    a.seed = ctx->seed; a.traj0 = (uint64_t)ctx->traj_offset + traj_shift; a.draw_base = draw_base;
    a.bits = (const plane_t*)P->bits;
    if (P->nspecs) {
        QSB_PRE(ctx, "k_draw_bits");
        k_draw_bits<<<(unsigned)((ctx->nwords * 32 + 255) / 256), 256, 0, ctx->stream>>>(P->d_specs, P->nspecs, P->bits, (size_t)ctx->nwords, a.seed, a.traj0, draw_base);
        QSB_LAUNCH_CHECK(ctx);
        // decision words: written once by the draw pass, read once by the executor
        qsb_account(ctx, "k_draw_bits", (uint64_t)P->nslots * ctx->nwords * 4 + (uint64_t)P->nspecs * sizeof(DrawSpec));
    }
    unsigned blocks = (unsigned)((a.W + P->threads - 1) / P->threads);
    QSB_PRE(ctx, "k_tab_exec");
    exec_kernel(P->threads, P->arena != nullptr)<<<blocks, P->threads, P->smem_bytes, ctx->stream>>>(a);
    QSB_LAUNCH_CHECK(ctx);
    // compulsory HBM traffic of the fused launch: the op list and the valid words once, the statistics rows and
    // any per-trajectory observable rows written once -- the ensemble state itself never leaves the SM
    qsb_account(ctx, "k_tab_exec", (uint64_t)P->nops * sizeof(TabOp) + (uint64_t)ctx->nwords * 4 + (uint64_t)P->nstats * 32 + (uint64_t)P->nslots * ctx->nwords * 4 +
                                       P->per_traj_rows * (uint64_t)ctx->npad * 16);
    ctx->ops += P->ops_per_run; ctx->draws += P->draws_per_run;
    return QSB_OK;
}

// Re-sends the host op list to the device (asynchronous H2D on the context's stream): the op list is the
// program's only host-side input, so an end-to-end timing that starts from host buffers includes this copy.
extern "C" int qsb_program_upload(qsb_program* P, int64_t* bytes) {
    if (!P) QSB_FAIL(QSB_ERR_BAD_ARG, "null program");
    size_t b = (size_t)P->nops * sizeof(TabOp);
    if (b) QSB_CUDA(cudaMemcpyAsync(P->d_ops, P->h_ops, b, cudaMemcpyHostToDevice, P->ctx->stream));
    if (bytes) *bytes = (int64_t)b;
    return QSB_OK;
}

// info: {ops, arena planes, statistics rows, threads per block, shared bytes per block, 1 if the arena is in global memory,
//        register-state ops per run (all trajectories), draws per run}
extern "C" int qsb_program_info(const qsb_program* P, int64_t info[8]) {
    if (!P || !info) QSB_FAIL(QSB_ERR_BAD_ARG, "null argument");
    info[0] = P->nops; info[1] = P->planes; info[2] = P->nstats; info[3] = P->threads; info[4] = P->smem_bytes; info[5] = P->arena ? 1 : 0;
    info[6] = (int64_t)P->ops_per_run; info[7] = (int64_t)P->draws_per_run;
    return QSB_OK;
}

extern "C" int qsb_program_free(qsb_program* P) {
    if (!P) return QSB_OK;
    qsb_ctx* ctx = P->ctx;
    qsb_dev_free(ctx, P->d_ops); qsb_dev_free(ctx, P->d_obs); qsb_dev_free(ctx, P->d_stats);
    qsb_dev_free(ctx, P->d_specs); qsb_dev_free(ctx, P->bits); qsb_dev_free(ctx, P->d_aux);
    if (P->arena) qsb_dev_free(ctx, P->arena);
    if (P->h_ops) cudaFreeHost(P->h_ops);
    delete P;
    return QSB_OK;
}

void qsb_rec_note_stats(qsb_ctx* ctx, double* p) { if (p) ctx->rec->stats_ptrs.push_back(p); }
