// qsb_exec.cu -- recorded tableau programs: capture the op list of a lockstep schedule once, then execute
// the WHOLE list in one launch with every trajectory word's state resident in shared memory.
//
// Why: at the BASELINE configs the entire Clifford ensemble is a few MB and a per-op launch is ~31k
// threads of dependent word logic (3-38 us each, ~1700 launches per pass).  A thread only ever touches
// its own 32-trajectory word column of every plane, so consecutive ops need no grid synchronisation:
// one thread can run the complete op list on its column.  The arena (all live states and masks of the
// schedule, planes allocated/recycled at record time) is a [planes][blockDim] tile of shared memory.
//
// Lanes per thread (QSB_EXEC_PLANE_T): the executor is bound by the dependent shared-memory round trips of each op and by
// how many trajectories' tableaux fit in an SM's shared memory, so its run time barely depends on the plane width
// (8-lane bytes: 24 warps per SM at 0.54 issue utilisation, 3.52 ms; 32-lane words: 6 warps per SM and a quarter of the
// instructions, 3.23 ms per 10^6 firstgenrepeater trajectories).  Words are the default; -DQSB_EXEC_PLANE_T=uint8_t|uint16_t
// builds the narrower variants (QSB200_LIB selects such a build at run time).
#include <limits.h>
#include <math.h>
#include <algorithm>
#include <map>

#include "qsb_exec_kernel.cuh"
#include "qsb_dexec.cuh"

#define QSB_DRAW_CHUNK 64      // specs staged per shared-memory refill of the draw pass
// The draw pass's work list: one item per Philox block (= two consecutive draw ids), so a block is computed once and its two
// halves are consumed without any "is this block already in registers" test.  32 bytes = two 128-bit shared-memory loads.
enum { DK_NONE = 0, DK_NOISE = 1, DK_SELECT = 2, DK_MEASURE = 3, DK_MASK = 4 };
struct DrawItem {
    uint32_t blk;          // draw id >> 1 (before draw_base)
    uint32_t kinds;        // kind of half 0 | kind of half 1 << 8
    uint32_t spec[2];      // index into the DrawSpec table (exact thresholds: read only on the slow paths)
    uint32_t thr_hi[2];    // DK_NOISE: DrawSpec::thr_hi_any
    uint32_t slot[2];      // first decision slot
};
static_assert(sizeof(DrawItem) == 32, "DrawItem layout");
#define QSB_ITEM_CHUNK 64      // items staged per refill: at most 2 * 3 slots each

struct qsb_program {
    qsb_ctx* ctx;
    qsb_dprogram* dprog;   // non-null: a dense program (qsb_dexec.cu); the tableau fields below are unused except d_stats
    DrawSpec* d_specs; int nspecs, nslots, first_dslot;
    DrawItem* d_items; int nitems;
    uint8_t* d_aux; int naux, nfolded;   // constant tables of folded ops
    ExecOp* d_ops; int nops;
    TabObsArgs* d_obs; int nobs;
    double** d_stats; int nstats;
    int planes;
    plane_t* arena;        // global-memory arena, only when the shared-memory tile does not fit
    int threads, smem_bytes;
    ExecOp* h_ops;         // pinned host copy of the op list (qsb_program_upload re-sends it)
    uint64_t ops_per_run, draws_per_run, per_traj_rows;
    uint64_t table_bytes;  // SURVEY 8d per-op algorithmic bytes summed over the recorded ops (what the per-op path would account)
};

static int rec_alloc_raw(qsb_recorder* Rr, RecArena* R, int nplanes) {
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
        if (getenv("QSB_REC_TRACE")) fprintf(stderr, "[rec] op %zu: arena %d high-water %d (block of %d)\n", Rr->ops.size(), (int)(R - Rr->A), R->high, nplanes);
    }
    return base;
}
static void rec_free_raw(RecArena* R, int base, int nplanes) {
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
int qsb_rec_alloc(qsb_ctx* ctx, int arena, int nplanes) {
    qsb_recorder* Rr = ctx->rec;
    RecArena* R = &Rr->A[arena];
    if (nplanes <= 0) nplanes = 1;
    int base = rec_alloc_raw(Rr, R, nplanes);
    R->live[base] = (int)R->blks.size();
    R->blks.push_back({base, Rr->now(), INT_MAX, {{Rr->now(), nplanes}}, -1, false});
    return base;
}
void qsb_rec_free(qsb_ctx* ctx, int arena, int base) {
    if (!ctx->rec || base < 0) return;
    RecArena* R = &ctx->rec->A[arena];
    auto it = R->live.find(base);
    if (it == R->live.end()) return;
    RecArena::Blk& b = R->blks[it->second];
    b.t1 = ctx->rec->now();
    rec_free_raw(R, base, b.sizes.back().second);
    R->live.erase(it);
}
void qsb_rec_shrink_to(qsb_ctx* ctx, int arena, int base, int new_planes) {
    RecArena* R = &ctx->rec->A[arena];
    if (new_planes <= 0) new_planes = 1;   // a 0-qubit state keeps one (unused) unit, like a fresh allocation would
    auto it = R->live.find(base);
    if (it == R->live.end()) return;
    RecArena::Blk& b = R->blks[it->second];
    int cur = b.sizes.back().second;
    if (new_planes >= cur) return;
    b.sizes.push_back({ctx->rec->now(), new_planes});
    rec_free_raw(R, base + new_planes, cur - new_planes);
}
int qsb_rec_alloc_planes(qsb_ctx* ctx, int nplanes) { return qsb_rec_alloc(ctx, 0, nplanes); }
void qsb_rec_free_planes(qsb_ctx* ctx, int base, int nplanes) { (void)nplanes; qsb_rec_free(ctx, 0, base); }
void qsb_rec_shrink(qsb_ctx* ctx, int base, int old_planes, int new_planes) { (void)old_planes; qsb_rec_shrink_to(ctx, 0, base, new_planes); }
// a recording holds tableau states or dense states, never both (one executor per program)
int qsb_rec_set_family(qsb_ctx* ctx, bool dense) {
    qsb_recorder* R = ctx->rec;
    if ((dense && R->tab) || (!dense && R->dense)) { qsb_set_error("a recorded program cannot mix tableau and dense states"); return QSB_ERR_UNSUPPORTED; }
    if (dense) R->dense = true; else R->tab = true;
    return QSB_OK;
}
// constant payload of a dense op -> offset (in doubles, 2-aligned) into the program's aux table; identical payloads share storage
int qsb_rec_daux(qsb_ctx* ctx, const double* v, int n) {
    qsb_recorder* R = ctx->rec;
    std::vector<double> key(v, v + n);
    auto it = R->daux_index.find(key);
    if (it != R->daux_index.end()) return it->second;
    if (R->daux.size() & 1) R->daux.push_back(0.0);
    int off = (int)R->daux.size();
    R->daux.insert(R->daux.end(), key.begin(), key.end());
    R->daux_index[key] = off;
    return off;
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

__global__ void k_zero_stats(double* const* ptrs, int n) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) { double* p = ptrs[i]; p[0] = 0; p[1] = 0; p[2] = 0; p[3] = 0; }
}

// Draw pass: ONE THREAD PER TRAJECTORY evaluates every stochastic op's Philox draw and its decision; a warp
// ballot packs the 32 decisions of a word, so the bit-sliced pass below never runs Philox (which a
// thread-per-word kernel would have to loop over its 32 lanes for, at 1/32 of the parallelism).
// Like the executor it runs any number of (program, lane subset) jobs per launch: block b belongs to the job with the
// largest block0_draw <= b, thread t of a job draws for the trajectory whose lane index is lanes[t] (identity when null).
// Output layout bits[slot][wstride] (slot-major).  A block is 8 warps = 8 consecutive words = ONE 32-byte sector of every
// slot row: the decision words of a chunk of specs are collected in shared memory [slot][8] and written sector by sector.
// Noise specs take a fast path: "no error on any lane of this word" (the common case at small p) follows from the HIGH 32
// bits of the draw alone (hi < thr_hi_any on all 32 lanes); only otherwise are the exact 53-bit comparisons evaluated.
#define QSB_DRAW_MAXSLOTS (3 * QSB_DRAW_CHUNK)
__global__ void __launch_bounds__(256) k_draw_bits(WaveArgs a) {
    __shared__ __align__(16) DrawSpec sspec[QSB_DRAW_CHUNK];   // filled with 16-byte copies
    __shared__ __align__(16) uint32_t sbits[QSB_DRAW_MAXSLOTS][8];
    __shared__ ExecJob sjob;
    if (threadIdx.x == 0) {
        int lo = 0, hi = a.njobs - 1;
        while (lo < hi) { const int mid = (lo + hi + 1) >> 1; if (a.jobs[mid].block0_draw <= blockIdx.x) lo = mid; else hi = mid - 1; }
        sjob = a.jobs[lo];
    }
    __syncthreads();
    const DrawSpec* const specs = reinterpret_cast<const DrawSpec*>(sjob.specs);
    const int nspecs = sjob.nspecs;
    const size_t t = (blockIdx.x - sjob.block0_draw) * (size_t)blockDim.x + threadIdx.x;   // trajectory of the job
    const unsigned lane = threadIdx.x & 31u, wib = threadIdx.x >> 5;                      // warp in block = word in the sector
    const size_t word0 = (blockIdx.x - sjob.block0_draw) * (size_t)8;                      // first word of this block
    const uint64_t traj = a.traj0 + ((sjob.lanes && t < sjob.count) ? (uint64_t)__ldg(sjob.lanes + t) : (uint64_t)t);
    uint32_t blk[4];
    uint64_t have = ~0ull;   // Philox block currently in registers (two draws per block)
    for (int i0 = 0; i0 < nspecs; i0 += QSB_DRAW_CHUNK) {
        __syncthreads();     // the previous chunk's words have been written out
        const int nc = min(QSB_DRAW_CHUNK, nspecs - i0);
        for (int k = threadIdx.x; k < nc * (int)(sizeof(DrawSpec) / 16); k += blockDim.x)
            reinterpret_cast<uint4*>(sspec)[k] = __ldg(reinterpret_cast<const uint4*>(specs + i0) + k);
        __syncthreads();
        const int slot0 = sspec[0].slot;
        for (int i = 0; i < nc; ++i) {
            const DrawSpec& s = sspec[i];
            const uint64_t id = s.draw + a.draw_base;
            if ((id >> 1) != have) { have = id >> 1; qsb_philox_block(a.seed, traj, have, blk); }
            const uint32_t xlo = (id & 1) ? blk[2] : blk[0], xhi = (id & 1) ? blk[3] : blk[1];
            uint32_t* const out = &sbits[s.slot - slot0][wib];
            const int type = s.type;
            if (type == TOP_NOISE) {
                if (!__any_sync(0xffffffffu, xhi >= s.thr_hi_any)) { if (lane == 0) { out[0] = 0u; out[8] = 0u; } continue; }   // nobody errs
                const uint64_t k = (((uint64_t)xhi << 32) | xlo) >> 11;   // u = k / 2^53
                uint32_t d = 0;
                for (int j = s.np - 1; j >= 0; --j) if (k > s.thr[j]) d = s.code[j];   // first matching branch wins (thr descends)
                const uint32_t fx = __ballot_sync(0xffffffffu, d & 1u), fz = __ballot_sync(0xffffffffu, d & 2u);
                if (lane == 0) { out[0] = fx; out[8] = fz; }
            } else if (type == TOP_SELECT) {
                const uint64_t k = (((uint64_t)xhi << 32) | xlo) >> 11;
                uint32_t d = (uint32_t)(s.np - 1);
                for (int j = s.np - 2; j >= 0; --j) if (k < s.thr[j]) d = (uint32_t)j;           // first j with u < cum[j]
                const uint32_t b0 = __ballot_sync(0xffffffffu, d & 1u), b1 = __ballot_sync(0xffffffffu, d & 2u), b2 = __ballot_sync(0xffffffffu, d & 4u);
                if (lane == 0) { out[0] = b0; out[8] = b1; out[16] = b2; }
            } else if (type == TOP_MEASURE) {
                const uint32_t b = __ballot_sync(0xffffffffu, xhi >> 31);     // !(u < 0.5)  <=>  top bit of the draw
                if (lane == 0) out[0] = b;
            } else {
                const uint64_t k = (((uint64_t)xhi << 32) | xlo) >> 11;
                const uint32_t b = __ballot_sync(0xffffffffu, k < s.thr[0]);  // random mask: bit = u < p
                if (lane == 0) out[0] = b;
            }
        }
        __syncthreads();
        // write the chunk: slot row r, words word0 .. word0 + 7 = one aligned 32-byte sector
        const DrawSpec& last = sspec[nc - 1];
        const int nsl = last.slot - slot0 + (last.type == TOP_NOISE ? 2 : last.type == TOP_SELECT ? 3 : 1);
        for (int k = threadIdx.x; k < nsl * 8; k += blockDim.x) {
            const size_t wd = word0 + (k & 7);
            if (wd < sjob.wstride) sjob.bits[(size_t)(slot0 + (k >> 3)) * sjob.wstride + wd] = sbits[k >> 3][k & 7];
        }
    }
}

// The same pass over the item list (one Philox block per trip, both halves consumed in place): the default whenever
// draw_base is even, i.e. whenever the pairing of draw ids into blocks is the one the items were built for.
// Philox4x32-10 with the round keys taken from the kernel arguments (constant bank): 20 IMAD.WIDE + 20 LOP3 per block
__device__ __forceinline__ void philox_rk(const uint32_t (&rk)[20], uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t* out) {
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        const uint64_t p0 = (uint64_t)0xD2511F53u * c0, p1 = (uint64_t)0xCD9E8D57u * c2;
        const uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ rk[2 * r], n2 = (uint32_t)(p0 >> 32) ^ c3 ^ rk[2 * r + 1];
        c0 = n0; c1 = (uint32_t)p1; c2 = n2; c3 = (uint32_t)p0;
    }
    out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}
__global__ void __launch_bounds__(256) k_draw_items(const __grid_constant__ WaveArgs a) {
    __shared__ __align__(16) DrawItem sitem[QSB_ITEM_CHUNK];
    __shared__ __align__(16) uint32_t sbits[6 * QSB_ITEM_CHUNK][8];
    __shared__ ExecJob sjob;
    if (threadIdx.x == 0) {
        int lo = 0, hi = a.njobs - 1;
        while (lo < hi) { const int mid = (lo + hi + 1) >> 1; if (a.jobs[mid].block0_draw <= blockIdx.x) lo = mid; else hi = mid - 1; }
        sjob = a.jobs[lo];
    }
    __syncthreads();
    const DrawItem* const items = reinterpret_cast<const DrawItem*>(sjob.items);
    const DrawSpec* const specs = reinterpret_cast<const DrawSpec*>(sjob.specs);
    const int nitems = sjob.nitems;
    const size_t t = (blockIdx.x - sjob.block0_draw) * (size_t)blockDim.x + threadIdx.x;   // trajectory of the job
    const unsigned lane = threadIdx.x & 31u, wib = threadIdx.x >> 5;                      // warp in block = word in the sector
    const size_t word0 = (blockIdx.x - sjob.block0_draw) * (size_t)8;
    const uint64_t traj = a.traj0 + ((sjob.lanes && t < sjob.count) ? (uint64_t)__ldg(sjob.lanes + t) : (uint64_t)t);
    const uint64_t blk_base = a.draw_base >> 1;
    auto nslots_of = [](uint32_t kind) { return kind == DK_NOISE ? 2u : kind == DK_SELECT ? 3u : kind == DK_NONE ? 0u : 1u; };
    for (int i0 = 0; i0 < nitems; i0 += QSB_ITEM_CHUNK) {
        __syncthreads();     // the previous chunk's words have been written out
        const int nc = min(QSB_ITEM_CHUNK, nitems - i0);
        for (int k = threadIdx.x; k < nc * 2; k += blockDim.x) reinterpret_cast<uint4*>(sitem)[k] = __ldg(reinterpret_cast<const uint4*>(items + i0) + k);
        __syncthreads();
        const uint32_t slot0 = (sitem[0].kinds & 0xff) ? sitem[0].slot[0] : sitem[0].slot[1];
        uint32_t* const mine = &sbits[0][wib] - (size_t)slot0 * 8;   // word `wib` of slot s lives at mine[s * 8]
#pragma unroll 1
        for (int i = 0; i < nc; ++i) {
            const uint4 lo4 = reinterpret_cast<const uint4*>(sitem)[2 * i], hi4 = reinterpret_cast<const uint4*>(sitem)[2 * i + 1];
            uint32_t blk[4];
            const uint64_t bno = blk_base + lo4.x;
            philox_rk(a.rk, (uint32_t)traj, (uint32_t)(traj >> 32), (uint32_t)bno, (uint32_t)(bno >> 32), blk);
            if (lo4.y == (DK_NOISE | (DK_NOISE << 8))) {   // both halves are noise draws (two thirds of the items): no per-half dispatch
                const bool e0 = __any_sync(0xffffffffu, blk[1] >= hi4.x), e1 = __any_sync(0xffffffffu, blk[3] >= hi4.y);
                if (!(e0 | e1)) {
                    if (lane == 0) { uint32_t* o0 = mine + (size_t)hi4.z * 8; uint32_t* o1 = mine + (size_t)hi4.w * 8; o0[0] = 0u; o0[8] = 0u; o1[0] = 0u; o1[8] = 0u; }
                    continue;
                }
            }
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                const uint32_t kind = (lo4.y >> (8 * h)) & 0xffu;
                if (kind == DK_NONE) continue;
                const uint32_t xlo = h ? blk[2] : blk[0], xhi = h ? blk[3] : blk[1];
                const uint32_t thr_hi = h ? hi4.y : hi4.x, slot = h ? hi4.w : hi4.z, sidx = h ? lo4.w : lo4.z;
                uint32_t* const out = mine + (size_t)slot * 8;
                if (kind == DK_NOISE) {
                    if (!__any_sync(0xffffffffu, xhi >= thr_hi)) { if (lane == 0) { out[0] = 0u; out[8] = 0u; } continue; }   // nobody errs
                    const DrawSpec* sp = specs + sidx;
                    const uint64_t k = (((uint64_t)xhi << 32) | xlo) >> 11;   // u = k / 2^53
                    const int np = __ldg(&sp->np);
                    uint32_t d = 0;
                    for (int j = np - 1; j >= 0; --j) if (k > __ldg(&sp->thr[j])) d = __ldg(&sp->code[j]);   // first matching branch wins (thr descends)
                    const uint32_t fx = __ballot_sync(0xffffffffu, d & 1u), fz = __ballot_sync(0xffffffffu, d & 2u);
                    if (lane == 0) { out[0] = fx; out[8] = fz; }
                } else if (kind == DK_MEASURE) {
                    const uint32_t b = __ballot_sync(0xffffffffu, xhi >> 31);     // !(u < 0.5)  <=>  top bit of the draw
                    if (lane == 0) out[0] = b;
                } else if (kind == DK_SELECT) {
                    const DrawSpec* sp = specs + sidx;
                    const uint64_t k = (((uint64_t)xhi << 32) | xlo) >> 11;
                    const int np = __ldg(&sp->np);
                    uint32_t d = (uint32_t)(np - 1);
                    for (int j = np - 2; j >= 0; --j) if (k < __ldg(&sp->thr[j])) d = (uint32_t)j;           // first j with u < cum[j]
                    const uint32_t b0 = __ballot_sync(0xffffffffu, d & 1u), b1 = __ballot_sync(0xffffffffu, d & 2u), b2 = __ballot_sync(0xffffffffu, d & 4u);
                    if (lane == 0) { out[0] = b0; out[8] = b1; out[16] = b2; }
                } else {
                    const uint64_t k = (((uint64_t)xhi << 32) | xlo) >> 11;
                    const uint32_t b = __ballot_sync(0xffffffffu, k < __ldg(&(specs + sidx)->thr[0]));  // random mask: bit = u < p
                    if (lane == 0) out[0] = b;
                }
            }
        }
        __syncthreads();
        const DrawItem& last = sitem[nc - 1];
        const uint32_t kb = (last.kinds >> 8) & 0xff, ka = last.kinds & 0xff;
        const uint32_t end = kb ? last.slot[1] + nslots_of(kb) : last.slot[0] + nslots_of(ka);
        const int nsl = (int)(end - slot0);
        for (int k = threadIdx.x; k < nsl * 8; k += blockDim.x) {
            const size_t wd = word0 + (k & 7);
            if (wd < sjob.wstride) sjob.bits[(size_t)(slot0 + (k >> 3)) * sjob.wstride + wd] = sbits[k >> 3][k & 7];
        }
    }
}

// the two instantiations live in qsb_exec_inst.cu (one object each, csrc/Makefile)
cudaError_t qsb_exec_launch_0(const WaveArgs&, unsigned, int, int, cudaStream_t);
cudaError_t qsb_exec_launch_1(const WaveArgs&, unsigned, int, int, cudaStream_t);
qsb_exec_launch_fn qsb_exec_launcher(int variant) { return variant ? qsb_exec_launch_1 : qsb_exec_launch_0; }

// Launch shape of the shared-memory executor for a set of jobs (words[j] = 32-trajectory words of job j, `planes` = the
// largest arena among them).  A block runs for (ops x latency of an op body) whatever its width, and the tile capacity of
// an SM fixes how many words are resident, so a small tail wave costs as much as a full one (10^6 firstgenrepeater
// trajectories on 96-thread blocks: 1.1 waves, 3.22 ms; 909 312 trajectories, exactly one wave: 1.67 ms).  Choose the
// fewest waves, then the fewest blocks per SM, then the smallest block (32..256 threads, partial warps allowed) that
// covers every job.  Returns 0 when not even a 32-thread tile fits (the arena then lives in global memory).
static int qsb_exec_shape(const qsb_ctx* ctx, const uint32_t* words, int njobs, int planes, long* blocks_out) {
    int dev_smem = 0, sm_smem = 0, sms = 0;
    cudaDeviceGetAttribute(&dev_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, ctx->device);
    cudaDeviceGetAttribute(&sm_smem, cudaDevAttrMaxSharedMemoryPerMultiprocessor, ctx->device);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, ctx->device);
    const size_t per_thread = (size_t)qsb_exec_stride(planes) * sizeof(plane_t), fixed = 1024 + 2 * QSB_EXEC_CHUNK * sizeof(ExecOp) + sizeof(ExecJob);
    if (per_thread * 32 > (size_t)dev_smem) return 0;
    auto blocks_of = [&](int t) { long b = 0; for (int j = 0; j < njobs; ++j) b += ((long)words[j] + t - 1) / t; return b; };
    const char* force_t = getenv("QSB_EXEC_THREADS");   // tuning aid: pin the block size
    if (force_t && atoi(force_t) >= 32 && atoi(force_t) <= 256 && per_thread * atoi(force_t) <= (size_t)dev_smem) {
        if (blocks_out) *blocks_out = blocks_of(atoi(force_t));
        return atoi(force_t);
    }
    const int force_b = getenv("QSB_EXEC_BLOCKS_PER_SM") ? atoi(getenv("QSB_EXEC_BLOCKS_PER_SM")) : 0;   // tuning aid
    long total = 0; for (int j = 0; j < njobs; ++j) total += words[j];
    for (long w = 1;; ++w) {
        for (int b = 1; b <= 8; ++b) {
            if (force_b && b != force_b) continue;
            long tmax = ((long)sm_smem / b - (long)fixed) / (long)per_thread;
            if (tmax > 256) tmax = 256;
            if ((size_t)tmax * per_thread > (size_t)dev_smem) tmax = (long)(dev_smem / per_thread);
            if (tmax < 32) continue;
            const long cap = (long)sms * b * w;     // blocks the device holds in w waves at b blocks per SM
            if (blocks_of((int)tmax) > cap) continue;
            int lo = 32, hi = (int)tmax;             // blocks_of is non-increasing in t: smallest block that still fits
            while (lo < hi) { const int mid = (lo + hi) / 2; if (blocks_of(mid) <= cap) hi = mid; else lo = mid + 1; }
            if (blocks_out) *blocks_out = blocks_of(lo);
            return lo;
        }
        if (w > total + 1) break;   // cannot happen: one word per block always fits eventually
    }
    return 32;
}

// Offline re-packing of one arena.  The online allocator had to number units while the schedule was still
// unfolding; now every block's lifetime is known, blocks only the folded ops touched can be dropped, and
// the rest is placed big-blocks-first at the lowest offset free during its whole lifetime.  A smaller arena
// means more resident threads per SM (the tile is [units][threads] of shared memory).
// use_s / use_alm: which op references (state bases / alive+mask planes) live in this arena.
static void repack_arena(qsb_recorder* R, int ai, bool use_s, bool use_alm, int& planes) {
    typedef RecArena::Blk Blk;
    std::vector<Blk>& B = R->A[ai].blks;
    const int nops = (int)R->ops.size();
    auto size_at = [](const Blk& b, int t) { int sz = b.sizes[0].second; for (auto& e : b.sizes) if (e.first <= t) sz = e.second; return sz; };
    auto find_blk = [&](int plane, int t) -> int {
        for (int k = 0; k < (int)B.size(); ++k) if (B[k].t0 <= t && t < B[k].t1 && plane >= B[k].base && plane < B[k].base + size_at(B[k], t)) return k;
        return -1;
    };
    for (int i = 0; i < nops; ++i) {
        const TabOp& o = R->ops[i];
        if (o.type == TOP_NOP) continue;
        for (int k = 0; k < 6; ++k) { if (use_s && o.s[k] >= 0) { int b = find_blk(o.s[k], i); if (b >= 0) B[b].used = true; } if (use_alm && o.al[k] >= 0) { int b = find_blk(o.al[k], i); if (b >= 0) B[b].used = true; } }
        for (int k = 0; k < 3; ++k) if (use_alm && o.m[k] >= 0) { int b = find_blk(o.m[k], i); if (b >= 0) B[b].used = true; }
    }
    auto conflict = [&](const Blk& x, int xo, const Blk& y, int yo) {
        int lo = std::max(x.t0, y.t0), hi = std::min(x.t1, y.t1);
        if (lo >= hi) return -1;
        // sizes only shrink: the pair overlaps at some common time iff it does at the first common time
        // or at any later step of either staircase; check every breakpoint inside [lo, hi)
        std::vector<int> ts{lo};
        for (auto& e : x.sizes) if (e.first > lo && e.first < hi) ts.push_back(e.first);
        for (auto& e : y.sizes) if (e.first > lo && e.first < hi) ts.push_back(e.first);
        int need = -1;   // lowest offset for x that clears y at every conflicting time, -1: no conflict
        for (int t : ts) { int xs = size_at(x, t), ys = size_at(y, t); if (xo < yo + ys && yo < xo + xs) need = std::max(need, yo + ys); }
        return need;
    };
    std::vector<int> order;
    for (int k = 0; k < (int)B.size(); ++k) if (B[k].used) order.push_back(k);
    // several placement orders, keep the tightest: long-lived blocks first (temporaries stack on top),
    // big blocks first, and allocation order
    auto life = [&](int k) { return (long)std::min(B[k].t1, nops) - B[k].t0; };
    auto place = [&](const std::vector<int>& ord, std::vector<int>& nb) {
        std::vector<int> placed;
        int pk = 0;
        for (int k : ord) {
            int off = 0;
            for (bool moved = true; moved;) {
                moved = false;
                for (int j : placed) { int need = conflict(B[k], off, B[j], nb[j]); if (need >= 0) { off = need; moved = true; } }
            }
            nb[k] = off; placed.push_back(k);
            pk = std::max(pk, off + B[k].sizes[0].second);
        }
        return pk;
    };
    int peak = INT_MAX;
    for (int variant = 0; variant < 4; ++variant) {
        std::vector<int> ord = order, nb(B.size(), -1);
        if (variant == 0) std::stable_sort(ord.begin(), ord.end(), [&](int a, int b) { return life(a) > life(b); });
        if (variant == 1) std::stable_sort(ord.begin(), ord.end(), [&](int a, int b) { return B[a].sizes[0].second > B[b].sizes[0].second; });
        if (variant == 2) std::stable_sort(ord.begin(), ord.end(), [&](int a, int b) { return life(a) * B[a].sizes[0].second > life(b) * B[b].sizes[0].second; });
        int pk = place(ord, nb);
        if (getenv("QSB_REC_TRACE")) fprintf(stderr, "[rec] packing order %d: %d planes\n", variant, pk);
        if (pk < peak) { peak = pk; for (int k : order) B[k].nbase = nb[k]; }
    }
    if (order.empty()) peak = 0;
    if (peak > 0 && peak <= planes && !getenv("QSB_NO_REPACK")) {
        auto remap = [&](int& plane, int t) { if (plane < 0) return; int b = find_blk(plane, t); if (b >= 0 && B[b].used) plane = plane - B[b].base + B[b].nbase; };
        for (int i = 0; i < nops; ++i) {
            TabOp& o = R->ops[i];
            if (o.type == TOP_NOP) continue;
            for (int k = 0; k < 6; ++k) { if (use_s) remap(o.s[k], i); if (use_alm) remap(o.al[k], i); }
            for (int k = 0; k < 3; ++k) if (use_alm) remap(o.m[k], i);
        }
        if (getenv("QSB_REC_TRACE")) fprintf(stderr, "[rec] arena re-packed: %d -> %d planes (%zu blocks)\n", planes, peak, order.size());
        planes = peak;
    } else if (getenv("QSB_REC_TRACE")) fprintf(stderr, "[rec] arena NOT re-packed: online %d planes, offline %d\n", planes, peak);
    if (getenv("QSB_REC_TRACE") && nops > 0) {
        int best = 0, bt = 0;
        for (int t = 0; t < nops; ++t) { int sum = 0; for (int k : order) if (B[k].t0 <= t && t < B[k].t1) sum += size_at(B[k], t); if (sum > best) { best = sum; bt = t; } }
        fprintf(stderr, "[rec] live-plane lower bound %d at op %d (type %d):", best, bt, R->ops[bt].type);
        for (int k : order) if (B[k].t0 <= bt && bt < B[k].t1) fprintf(stderr, " %d[%d,%d)", size_at(B[k], bt), B[k].t0, B[k].t1 == INT_MAX ? -1 : B[k].t1);
        fprintf(stderr, "\n");
    }

}

// ---------------------------------------------------------------------------------------------
// C-ABI: record / run
// ---------------------------------------------------------------------------------------------
extern "C" int qsb_record_begin(qsb_ctx* ctx) {
    if (!ctx) QSB_FAIL(QSB_ERR_BAD_ARG, "null context");
    if (ctx->rec) QSB_FAIL(QSB_ERR_BAD_ARG, "already recording");
    ctx->rec = new qsb_recorder();
    ctx->rec->ops0 = ctx->ops; ctx->rec->draws0 = ctx->draws; ctx->rec->bytes0 = ctx->algo_bytes;
    ctx->rec_gen++;
    return QSB_OK;
}

extern "C" int qsb_record_end(qsb_ctx* ctx, qsb_program** out) {
    if (!ctx || !out) QSB_FAIL(QSB_ERR_BAD_ARG, "null argument");
    if (!ctx->rec) QSB_FAIL(QSB_ERR_BAD_ARG, "not recording");
    qsb_use_device(ctx);
    qsb_recorder* R = ctx->rec;
    qsb_program* P = new qsb_program();
    memset(P, 0, sizeof(*P));
    P->ctx = ctx;
    P->nops = (int)R->ops.size(); P->nobs = (int)R->obs.size(); P->nstats = (int)R->stats_ptrs.size();
    P->planes = R->A[0].high > 0 ? R->A[0].high : 1;
    P->ops_per_run = ctx->ops - R->ops0; P->draws_per_run = ctx->draws - R->draws0;
    P->table_bytes = ctx->algo_bytes - R->bytes0;
    ctx->ops = R->ops0; ctx->draws = R->draws0; ctx->algo_bytes = R->bytes0;  // counted per qsb_program_run instead
    for (auto& o : R->ops) if (o.per_traj) P->per_traj_rows++;
    int rc = QSB_OK;
    if (R->dense) {
        int nmasks = R->A[0].high > 0 ? R->A[0].high : 1, elems = R->A[1].high > 0 ? R->A[1].high : 1;
        repack_arena(R, 0, false, true, nmasks);
        repack_arena(R, 1, true, false, elems);
        rc = qsb_dexec_build(ctx, R, elems, nmasks, &P->dprog);
        if (rc == QSB_OK) {
            rc = qsb_dev_alloc(ctx, (void**)&P->d_stats, R->stats_ptrs.empty() ? 16 : R->stats_ptrs.size() * sizeof(double*));
            if (rc == QSB_OK && !R->stats_ptrs.empty()) {
                if (cudaMemcpyAsync(P->d_stats, R->stats_ptrs.data(), R->stats_ptrs.size() * sizeof(double*), cudaMemcpyHostToDevice, ctx->stream) != cudaSuccess ||
                    cudaStreamSynchronize(ctx->stream) != cudaSuccess) { qsb_set_error("program upload failed"); rc = QSB_ERR_CUDA; }
            }
        }
        P->planes = elems; P->threads = QSB_DEXEC_THREADS;
        delete R;
        ctx->rec = nullptr;
        if (rc != QSB_OK) { if (P->dprog) qsb_dexec_free(P->dprog); if (P->d_stats) qsb_dev_free(ctx, P->d_stats); delete P; return rc; }
        *out = P;
        return QSB_OK;
    }
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
        repack_arena(R, 0, true, true, P->planes);
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
        for (int j = 0; j < 5; ++j) {
            const double c = o.f[j], scaled = ldexp(c < 0 ? 0.0 : (c > 1 ? 1.0 : c), 53);
            d.thr[j] = o.type == TOP_NOISE ? (uint64_t)floor(scaled) : (uint64_t)ceil(scaled);
        }
        if (o.type == TOP_MEASURE) d.thr[0] = 1ull << 52;   // u >= 0.5  <=>  !(k < 2^52)
        if (o.type == TOP_NOISE) {
            d.thr_any = ~0ull;
            for (int j = 0; j < d.np; ++j) {
                if (d.thr[j] < d.thr_any) d.thr_any = d.thr[j];
                d.code[j] = (uint8_t)(((o.iv[j] == 1 || o.iv[j] == 2) ? 1 : 0) | ((o.iv[j] == 3 || o.iv[j] == 2) ? 2 : 0));
            }
            // k > thr_any  <=>  x >= (thr_any + 1) << 11 for the 64 draw bits x (k = x >> 11); thr_any + 1 = 2^53 can never be reached
            const uint64_t t1 = d.thr_any + 1;
            d.thr_hi_any = (d.np == 0 || t1 >= (1ull << 53)) ? 0xffffffffu : (uint32_t)((t1 << 11) >> 32);
            if (d.np == 0) d.thr_any = ~0ull;
        }
        o.dslot = P->nslots; P->nslots += need;
        specs.push_back(d);
    }
    P->nspecs = (int)specs.size();
    std::vector<DrawItem> items;
    for (size_t i = 0; i < specs.size(); ++i) {
        const DrawSpec& d = specs[i];
        const uint32_t blk = (uint32_t)(d.draw >> 1), half = (uint32_t)(d.draw & 1);
        if (items.empty() || items.back().blk != blk || ((items.back().kinds >> (8 * half)) & 0xff)) {
            DrawItem it; memset(&it, 0, sizeof(it)); it.blk = blk; items.push_back(it);
        }
        DrawItem& it = items.back();
        const uint32_t kind = d.type == TOP_NOISE ? DK_NOISE : d.type == TOP_SELECT ? DK_SELECT : d.type == TOP_MEASURE ? DK_MEASURE : DK_MASK;
        it.kinds |= kind << (8 * half); it.spec[half] = (uint32_t)i; it.thr_hi[half] = d.thr_hi_any; it.slot[half] = (uint32_t)d.slot;
    }
    P->nitems = (int)items.size();
    P->first_dslot = -1;
    {   // every stochastic op names the slots of the next one: the executor fetches decision words one stochastic op ahead
        int next = -1;
        for (size_t i = R->ops.size(); i-- > 0;) {
            TabOp& o = R->ops[i];
            const bool stoch = o.type == TOP_NOISE || o.type == TOP_SELECT || o.type == TOP_SELECT_CONST || o.type == TOP_MEASURE || o.type == TOP_MASK_RANDOM;
            o.lut = (uint32_t)next;          // host-only field, free at this point: carries next_dslot into the ExecOp
            if (stoch) next = o.dslot;
        }
        P->first_dslot = next;
    }
    // the kernel's compact op records
    std::vector<ExecOp> xops(R->ops.size());
    for (size_t i = 0; i < R->ops.size() && rc == QSB_OK; ++i) {
        const TabOp& o = R->ops[i];
        ExecOp& x = xops[i];
        memset(&x, 0, sizeof(x));
        if (o.n > 255 || o.n2 > 255 || P->planes > 32767) { qsb_set_error("program too large for the executor's op encoding"); rc = QSB_ERR_UNSUPPORTED; break; }
        x.type = (uint8_t)o.type; x.n = (uint8_t)o.n; x.n2 = (uint8_t)o.n2; x.k = (uint8_t)o.k;
        x.q[0] = (uint8_t)o.q[0]; x.q[1] = (uint8_t)o.q[1]; x.iv[0] = (uint8_t)o.iv[0]; x.iv[1] = (uint8_t)o.iv[1];
        for (int j = 0; j < 4; ++j) x.outmask[j] = o.outmask[j];
        x.u[0] = o.u[0]; x.u[1] = o.u[1];
        for (int j = 0; j < 6; ++j) { x.s[j] = (int16_t)o.s[j]; x.al[j] = (int16_t)o.al[j]; }
        for (int j = 0; j < 3; ++j) x.m[j] = (int16_t)o.m[j];
        x.obs = o.obs; x.dslot = o.dslot; x.next_dslot = (int32_t)o.lut; x.stats = o.stats; x.per_traj = o.per_traj;
        if (o.type == TOP_EXPECT) {   // single-Pauli observables carry their arguments in the record
            const TabObsArgs& oa = R->obs[o.obs];
            x.k = oa.nelem == 1 ? 1 : 0;
            if (x.k) { x.u[0] = oa.px[0]; x.u[1] = oa.pz[0]; x.iv[0] = oa.sign[0]; }
        }
    }
    if (!xops.empty() && rc == QSB_OK) {
        if (cudaMallocHost((void**)&P->h_ops, xops.size() * sizeof(ExecOp)) != cudaSuccess) { cudaGetLastError(); qsb_set_error("pinned allocation of the op list failed"); rc = QSB_ERR_OOM; }
        else memcpy(P->h_ops, xops.data(), xops.size() * sizeof(ExecOp));
    }
    auto up = [&](void** dst, const void* src, size_t bytes) {
        if (rc != QSB_OK) return;
        rc = qsb_dev_alloc(ctx, dst, bytes ? bytes : 16);
        if (rc == QSB_OK && bytes) {
            cudaError_t e = cudaMemcpyAsync(*dst, src, bytes, cudaMemcpyHostToDevice, ctx->stream);
            if (e != cudaSuccess) { qsb_set_error("program upload failed: %s", cudaGetErrorString(e)); rc = QSB_ERR_CUDA; }
        }
    };
    up((void**)&P->d_ops, xops.data(), xops.size() * sizeof(ExecOp));
    up((void**)&P->d_obs, R->obs.data(), R->obs.size() * sizeof(TabObsArgs));
    up((void**)&P->d_stats, R->stats_ptrs.data(), R->stats_ptrs.size() * sizeof(double*));
    up((void**)&P->d_specs, specs.data(), specs.size() * sizeof(DrawSpec));
    up((void**)&P->d_items, items.data(), items.size() * sizeof(DrawItem));
    up((void**)&P->d_aux, aux.data(), aux.size());
    if (rc == QSB_OK) { cudaError_t e = cudaStreamSynchronize(ctx->stream); if (e != cudaSuccess) rc = QSB_ERR_CUDA; }  // host vectors die below
    // launch shape of the full-ensemble replay (reported by qsb_program_info; subsets and waves are shaped at run time)
    const uint32_t full_words = (uint32_t)ctx->nwords;
    long shape_blocks = 0;
    int t = qsb_exec_shape(ctx, &full_words, 1, P->planes, &shape_blocks);
    // Shared memory or global memory for the arena?  The shared-memory executor is latency bound: a block runs for (ops x latency of
    // an op body) whatever its width, so its run time is (waves x that), and the waves follow from how many word columns fit in an
    // SM's shared memory.  When the ensemble needs many waves (repeatergrid: 1341 planes -> 42 columns per SM -> 50 waves for 10^7
    // lanes) the same kernel over a plane-major arena in global memory wins: every word is a thread, 48 warps per SM hide the L2
    // latency of the plane accesses, and the traffic is what the per-op path would move anyway (measured: 270 -> 93 ms).
    if (t) {
        int sm_smem = 0, sms = 1;
        cudaDeviceGetAttribute(&sm_smem, cudaDevAttrMaxSharedMemoryPerMultiprocessor, ctx->device);
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, ctx->device);
        long resident = sm_smem / ((long)qsb_exec_stride(P->planes) * t * (long)sizeof(plane_t) + 4096);
        if (resident < 1) resident = 1;
        const double waves = (double)shape_blocks / ((double)sms * (double)resident);
        const char* g = getenv("QSB_EXEC_GLOBAL");          // tuning aid: 1 forces the global-memory arena, 0 the shared-memory tile
        if (g ? atoi(g) != 0 : waves > 4.0) t = 0;
    }
    if (t) { P->threads = t; P->smem_bytes = qsb_exec_stride(P->planes) * t * (int)sizeof(plane_t); P->arena = nullptr; }
    else {
        P->threads = 128; P->smem_bytes = 0;
        if (rc == QSB_OK) rc = qsb_dev_alloc(ctx, (void**)&P->arena, (size_t)P->planes * ctx->nwords * sizeof(plane_t));
    }
    delete R;
    ctx->rec = nullptr;
    if (rc != QSB_OK) {   // release whatever was uploaded before the failure (stream ordered; null pointers are skipped)
        qsb_dev_free(ctx, P->d_ops); qsb_dev_free(ctx, P->d_obs); qsb_dev_free(ctx, P->d_stats);
        qsb_dev_free(ctx, P->d_specs); qsb_dev_free(ctx, P->d_items); qsb_dev_free(ctx, P->d_aux);
        if (P->arena) qsb_dev_free(ctx, P->arena);
        if (P->h_ops) cudaFreeHost(P->h_ops);
        delete P;
        return rc;
    }
    *out = P;
    return QSB_OK;
}

// ---------------------------------------------------------------------------------------------
// run: one launch pair (draw pass + executor) over any number of (program, lane subset) jobs
// ---------------------------------------------------------------------------------------------
struct WaveJobIn { qsb_program* P; const uint32_t* lanes; int64_t count; };

static int wave_buffer(qsb_ctx* ctx, void** buf, size_t* have, size_t need) {
    if (*have >= need) return QSB_OK;
    QSB_TRY(qsb_dev_free(ctx, *buf));   // stream ordered: launches already queued keep the old block until they finish
    *buf = nullptr; *have = 0;
    size_t want = need + need / 4;
    QSB_TRY(qsb_dev_alloc(ctx, buf, want));
    *have = want;
    return QSB_OK;
}

static int run_wave(qsb_ctx* ctx, const WaveJobIn* in, int n, uint64_t traj_shift, uint64_t draw_base, int flags) {
    if (ctx->rec) QSB_FAIL(QSB_ERR_BAD_ARG, "cannot run a program while recording");
    qsb_use_device(ctx);
    for (int j = 0; j < n; ++j) {
        if (!in[j].P || in[j].P->ctx != ctx) QSB_FAIL(QSB_ERR_BAD_ARG, "job %d: null program or program of another context", j);
        if (in[j].count < 0 || in[j].count > ctx->ntraj) QSB_FAIL(QSB_ERR_BAD_ARG, "job %d: %lld lanes requested from an ensemble of %lld", j, (long long)in[j].count, (long long)ctx->ntraj);
        if ((in[j].P->dprog != nullptr) != (in[0].P->dprog != nullptr)) QSB_FAIL(QSB_ERR_UNSUPPORTED, "a launch runs tableau programs or dense programs, not both");
    }
    if (!(flags & QSB_RUN_KEEP_STATS)) {
        for (int j = 0; j < n; ++j) {
            qsb_program* P = in[j].P;
            if (!P->nstats) continue;
            QSB_PRE(ctx, "k_zero_stats");
            k_zero_stats<<<(P->nstats + 127) / 128, 128, 0, ctx->stream>>>(P->d_stats, P->nstats);
            QSB_LAUNCH_CHECK(ctx);
        }
    }
    if (in[0].P->dprog) {   // dense programs: one launch per (program, lane subset) job -- a tile of trajectories is a block's whole working set
        // The cohorts of a tree level are independent and mostly small (a few tiles each, far fewer blocks than SMs), so their
        // launches are dealt round-robin to a pool of side streams that fork from and join back into the context's stream.
        const int NS = 8;
        static thread_local cudaStream_t pool[64][NS];
        static thread_local bool pool_ok[64];
        const bool fan = n > 1 && !ctx->prof && ctx->device >= 0 && ctx->device < 64;
        cudaEvent_t ev = nullptr;
        if (fan) {
            if (!pool_ok[ctx->device]) { for (int k = 0; k < NS; ++k) QSB_CUDA(cudaStreamCreateWithFlags(&pool[ctx->device][k], cudaStreamNonBlocking)); pool_ok[ctx->device] = true; }
            QSB_CUDA(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
            QSB_CUDA(cudaEventRecord(ev, ctx->stream));
            for (int k = 0; k < NS && k < n; ++k) QSB_CUDA(cudaStreamWaitEvent(pool[ctx->device][k], ev, 0));
        }
        for (int j = 0; j < n; ++j) {
            qsb_program* P = in[j].P;
            QSB_TRY(qsb_dexec_run(P->dprog, in[j].lanes, in[j].count, traj_shift, draw_base, fan ? pool[ctx->device][j % NS] : nullptr));
            ctx->ops += P->ops_per_run / (uint64_t)ctx->ntraj * (uint64_t)in[j].count; ctx->draws += P->draws_per_run;
        }
        if (fan) {
            for (int k = 0; k < NS && k < n; ++k) { QSB_CUDA(cudaEventRecord(ev, pool[ctx->device][k])); QSB_CUDA(cudaStreamWaitEvent(ctx->stream, ev, 0)); }
            QSB_CUDA(cudaEventDestroy(ev));
        }
        return QSB_OK;
    }
    // A program whose arena lives in global memory (too many planes for a shared-memory tile, or too many waves) cannot share a
    // launch: the global-arena instantiation is shaped for one job.  A level of a cohort tree may mix both kinds, so such jobs go
    // out one launch pair each, on the same stream, and the shared-memory jobs keep their common launch.  Statistics rows were
    // zeroed above where asked, hence KEEP_STATS below.
    if (n > 1) {
        int with_arena = 0;
        for (int j = 0; j < n; ++j) with_arena += in[j].P->arena != nullptr;
        if (with_arena) {
            std::vector<WaveJobIn> tiled;
            for (int j = 0; j < n; ++j) {
                if (in[j].P->arena) QSB_TRY(run_wave(ctx, &in[j], 1, traj_shift, draw_base, flags | QSB_RUN_KEEP_STATS));
                else tiled.push_back(in[j]);
            }
            if (!tiled.empty()) QSB_TRY(run_wave(ctx, tiled.data(), (int)tiled.size(), traj_shift, draw_base, flags | QSB_RUN_KEEP_STATS));
            return QSB_OK;
        }
    }
    std::vector<ExecJob> jobs(n);
    std::vector<uint32_t> words(n);
    size_t bits_words = 0;
    int planes = 1, nspecs_any = 0;
    for (int j = 0; j < n; ++j) {
        qsb_program* P = in[j].P;
        words[j] = (uint32_t)((in[j].count + 31) / 32);
        bits_words += (size_t)(P->nslots + 4) * ((words[j] + 7u) & ~7u);   // +4 rows: the executor's read-ahead may run past the last slot
        planes = std::max(planes, P->planes);
        nspecs_any |= P->nspecs;
    }
    QSB_TRY(wave_buffer(ctx, (void**)&ctx->wave_bits, &ctx->wave_bits_bytes, (bits_words + 32) * sizeof(uint32_t)));
    QSB_TRY(wave_buffer(ctx, &ctx->wave_jobs, &ctx->wave_jobs_bytes, (size_t)n * sizeof(ExecJob)));
    const bool global_arena = n == 1 && in[0].P->arena != nullptr;   // several jobs per launch always take the shared-memory tile
    long blocks = 0;
    int threads;
    if (global_arena) { threads = 128; blocks = (words[0] + 127) / 128; }
    else if (n == 1 && !in[0].lanes && in[0].count == ctx->ntraj && !in[0].P->arena) { threads = in[0].P->threads; blocks = ((long)words[0] + threads - 1) / threads; }
    else threads = qsb_exec_shape(ctx, words.data(), n, planes, &blocks);
    if (!threads) QSB_FAIL(QSB_ERR_UNSUPPORTED, "a program of this launch does not fit in shared memory");
    size_t boff = 0;
    uint32_t be = 0, bd = 0;
    for (int j = 0; j < n; ++j) {
        qsb_program* P = in[j].P;
        ExecJob& J = jobs[j];
        J.ops = P->d_ops; J.obs = P->d_obs; J.aux = P->d_aux; J.specs = P->d_specs; J.lanes = in[j].lanes;
        J.items = P->d_items; J.nitems = P->nitems; J.pad_items = 0;
        J.bits = ctx->wave_bits + boff; J.arena = (uint32_t*)P->arena;
        J.nops = P->nops; J.nspecs = P->nspecs; J.nslots = P->nslots; J.planes = P->planes;
        J.wstride = (words[j] + 7u) & ~7u; J.first_dslot = P->first_dslot;
        J.count = (uint32_t)in[j].count; J.words = words[j];
        J.block0_exec = be; J.block0_draw = bd;
        boff += (size_t)(P->nslots + 4) * J.wstride;
        be += (words[j] + threads - 1) / threads;
        bd += P->nspecs ? (uint32_t)(((size_t)words[j] * 32 + 255) / 256) : 0u;
    }
    // pageable source: the runtime stages the bytes before returning, so the host vector may die at once
    QSB_CUDA(cudaMemcpyAsync(ctx->wave_jobs, jobs.data(), (size_t)n * sizeof(ExecJob), cudaMemcpyHostToDevice, ctx->stream));
    WaveArgs a;
    a.jobs = (const ExecJob*)ctx->wave_jobs; a.njobs = n;
    a.seed = ctx->seed; a.traj0 = (uint64_t)ctx->traj_offset + traj_shift; a.draw_base = draw_base;
    a.tile_stride = qsb_exec_stride(planes);
    for (int r = 0; r < 10; ++r) { a.rk[2 * r] = (uint32_t)ctx->seed + (uint32_t)r * 0x9E3779B9u; a.rk[2 * r + 1] = (uint32_t)(ctx->seed >> 32) + (uint32_t)r * 0xBB67AE85u; }
    if (bd) {
        QSB_PRE(ctx, "k_draw_bits");
        if (draw_base & 1) k_draw_bits<<<bd, 256, 0, ctx->stream>>>(a);   // odd offset: the items' pairing of draws into blocks does not apply
        else k_draw_items<<<bd, 256, 0, ctx->stream>>>(a);
        QSB_LAUNCH_CHECK(ctx);
    }
    if (be) {
        QSB_PRE(ctx, "k_tab_exec");
        cudaError_t e = qsb_exec_launcher(global_arena ? 0 : 1)(a, be, threads, global_arena ? 0 : qsb_exec_stride(planes) * threads * (int)sizeof(plane_t), ctx->stream);
        if (e != cudaSuccess) QSB_FAIL(QSB_ERR_CUDA, "k_tab_exec launch failed: %s", cudaGetErrorString(e));
        QSB_LAUNCH_CHECK(ctx);
    }
    for (int j = 0; j < n; ++j) {
        qsb_program* P = in[j].P;
        const uint64_t nw = words[j], share_num = (uint64_t)in[j].count, share_den = (uint64_t)ctx->ntraj;
        // decision words: written once by the draw pass, read once by the executor
        if (P->nspecs) qsb_account(ctx, "k_draw_bits", (uint64_t)P->nslots * nw * 4 + (uint64_t)P->nspecs * sizeof(DrawSpec));
        // compulsory HBM traffic of the fused launch: the op list once, the statistics rows and any per-trajectory observable
        // rows written once, the decision words read once -- the ensemble state itself never leaves the SM
        qsb_account(ctx, "k_tab_exec", (uint64_t)P->nops * sizeof(ExecOp) + (uint64_t)P->nstats * 32 + (uint64_t)P->nslots * nw * 4 +
                                           P->per_traj_rows * (uint64_t)nw * 32 * 16);
        ctx->ops += P->ops_per_run / share_den * share_num; ctx->draws += P->draws_per_run;
    }
    ctx->wave_threads = threads; ctx->wave_blocks = (int)be;
    return QSB_OK;
}

// traj_shift moves the whole ensemble to fresh global trajectory ids (a new batch), draw_base offsets every draw id.
extern "C" int qsb_program_run(qsb_program* P, uint64_t traj_shift, uint64_t draw_base) {
    if (!P) QSB_FAIL(QSB_ERR_BAD_ARG, "null program");
    WaveJobIn in{P, nullptr, P->ctx->ntraj};
    return run_wave(P->ctx, &in, 1, traj_shift, draw_base, 0);
}

extern "C" int qsb_program_run_lanes(qsb_program* P, const uint32_t* lanes_dev, int64_t count, uint64_t traj_shift, uint64_t draw_base, int flags) {
    if (!P) QSB_FAIL(QSB_ERR_BAD_ARG, "null program");
    WaveJobIn in{P, lanes_dev, lanes_dev ? count : P->ctx->ntraj};
    return run_wave(P->ctx, &in, 1, traj_shift, draw_base, flags);
}

extern "C" int qsb_wave_run(qsb_ctx* ctx, qsb_program* const* progs, const uint32_t* const* lanes_dev, const int64_t* counts, int njobs,
                            uint64_t traj_shift, uint64_t draw_base, int flags) {
    if (!ctx || !progs || !counts || njobs < 1) QSB_FAIL(QSB_ERR_BAD_ARG, "qsb_wave_run: bad argument");
    std::vector<WaveJobIn> in(njobs);
    for (int j = 0; j < njobs; ++j) in[j] = WaveJobIn{progs[j], lanes_dev ? lanes_dev[j] : nullptr, (lanes_dev && lanes_dev[j]) ? counts[j] : ctx->ntraj};
    return run_wave(ctx, in.data(), njobs, traj_shift, draw_base, flags);
}

// Cohort forking.  The lanes of `lanes` leave the lockstep ensemble at this point of the schedule: their
// classical history differs from the one the host is following.  They stop counting in every later statistic of this run
// and codes_dev[lane index] (nullable; int32 per trajectory of the context) receives `code`, so that the host can re-run
// each retired lane under the program of its own history (qsb_wave_run with lane subsets).  Recorded programs only.
extern "C" int qsb_ctx_retire(qsb_ctx* ctx, const qsb_mask* lanes, int32_t code, int32_t* codes_dev) {
    if (!ctx || !lanes) QSB_FAIL(QSB_ERR_BAD_ARG, "qsb_ctx_retire: null argument");
    if (!ctx->rec) QSB_FAIL(QSB_ERR_UNSUPPORTED, "qsb_ctx_retire needs a recording: cohorts are re-run as recorded programs over lane subsets");
    QSB_TRY(qsb_handle_ok(ctx, lanes->vplane >= 0, lanes->gen));
    TabOp* o = qsb_rec_push(ctx, TOP_RETIRE);
    o->m[0] = lanes->vplane; o->u[0] = (uint32_t)code; o->per_traj = reinterpret_cast<double*>(codes_dev);
    return QSB_OK;
}

// Re-sends the host op list to the device (asynchronous H2D on the context's stream): the op list is the
// program's only host-side input, so an end-to-end timing that starts from host buffers includes this copy.
extern "C" int qsb_program_upload(qsb_program* P, int64_t* bytes) {
    if (!P) QSB_FAIL(QSB_ERR_BAD_ARG, "null program");
    if (P->dprog) return qsb_dexec_upload(P->dprog, bytes);
    qsb_use_device(P->ctx);
    size_t b = (size_t)P->nops * sizeof(ExecOp);
    if (b) QSB_CUDA(cudaMemcpyAsync(P->d_ops, P->h_ops, b, cudaMemcpyHostToDevice, P->ctx->stream));
    if (bytes) *bytes = (int64_t)b;
    return QSB_OK;
}

// info: {ops, arena planes, statistics rows, threads per block, shared bytes per block, 1 if the arena is in global memory,
//        register-state ops per run (all trajectories), draws per run}
extern "C" int qsb_program_info(const qsb_program* P, int64_t info[8]) {
    if (!P || !info) QSB_FAIL(QSB_ERR_BAD_ARG, "null argument");
    info[0] = P->nops; info[1] = P->planes; info[2] = P->nstats; info[3] = P->threads; info[4] = P->smem_bytes; info[5] = P->arena ? 1 : 0;
    if (P->dprog) qsb_dexec_info(P->dprog, info);
    info[2] = P->nstats; info[6] = (int64_t)P->ops_per_run; info[7] = (int64_t)P->draws_per_run;
    return QSB_OK;
}

// Raw records of a tableau program as the executor reads them (layouts: csrc/qsb_exec_kernel.cuh -- a verification aid, not a stable
// ABI): what = 0 ExecOp[], 1 DrawSpec[], 2 TabObsArgs[], 3 aux bytes, 4 statistics row pointers (device addresses, row order),
// 5 int32 {planes, nslots, nops, nspecs, nobs, naux, nstats, first_dslot}.  *needed receives the size; data is copied when buf holds it.
// Host-visible.  Used by the bit-sliced CPU replay of bench.py / tests (oracle/qsb_bitsliced.cpp), never by the product.
extern "C" int qsb_program_export(const qsb_program* P, int what, void* buf, size_t buflen, size_t* needed) {
    if (!P || !needed) QSB_FAIL(QSB_ERR_BAD_ARG, "null argument");
    if (P->dprog) QSB_FAIL(QSB_ERR_UNSUPPORTED, "qsb_program_export covers tableau programs");
    qsb_use_device(P->ctx);
    const int32_t meta[8] = {P->planes, P->nslots, P->nops, P->nspecs, P->nobs, P->naux, P->nstats, P->first_dslot};
    const void* src = nullptr; size_t n = 0; bool dev = true;
    switch (what) {
        case 0: src = P->d_ops; n = (size_t)P->nops * sizeof(ExecOp); break;
        case 1: src = P->d_specs; n = (size_t)P->nspecs * sizeof(DrawSpec); break;
        case 2: src = P->d_obs; n = (size_t)P->nobs * sizeof(TabObsArgs); break;
        case 3: src = P->d_aux; n = (size_t)P->naux; break;
        case 4: src = P->d_stats; n = (size_t)P->nstats * sizeof(double*); break;
        case 5: src = meta; n = sizeof(meta); dev = false; break;
        default: QSB_FAIL(QSB_ERR_BAD_ARG, "qsb_program_export: unknown section %d", what);
    }
    *needed = n;
    if (!buf || buflen < n || n == 0) return QSB_OK;
    if (!dev) { memcpy(buf, src, n); return QSB_OK; }
    QSB_CUDA(cudaMemcpyAsync(buf, src, n, cudaMemcpyDeviceToHost, P->ctx->stream));
    QSB_CUDA(cudaStreamSynchronize(P->ctx->stream));
    return QSB_OK;
}

extern "C" int qsb_program_table_bytes(const qsb_program* P, uint64_t* bytes) {
    if (!P || !bytes) QSB_FAIL(QSB_ERR_BAD_ARG, "null argument");
    *bytes = P->table_bytes;
    return QSB_OK;
}

extern "C" int qsb_program_free(qsb_program* P) {
    if (!P) return QSB_OK;
    qsb_ctx* ctx = P->ctx;
    if (P->dprog) { qsb_dexec_free(P->dprog); qsb_dev_free(ctx, P->d_stats); delete P; return QSB_OK; }
    qsb_dev_free(ctx, P->d_ops); qsb_dev_free(ctx, P->d_obs); qsb_dev_free(ctx, P->d_stats);
    qsb_dev_free(ctx, P->d_specs); qsb_dev_free(ctx, P->d_items); qsb_dev_free(ctx, P->d_aux);
    if (P->arena) qsb_dev_free(ctx, P->arena);
    if (P->h_ops) cudaFreeHost(P->h_ops);
    delete P;
    return QSB_OK;
}

void qsb_rec_note_stats(qsb_ctx* ctx, double* p) { if (p) ctx->rec->stats_ptrs.push_back(p); }

// ---------------------------------------------------------------------------------------------
// cohort routing: group the lanes a level retired by (program, decision) without leaving the device
// ---------------------------------------------------------------------------------------------
// key of a retired lane = node[lane] * maxdec + code[lane]; bins[key] counts them, bins[nkeys] counts the censored lanes
__global__ void k_cohort_count(const int32_t* __restrict__ codes, const int32_t* __restrict__ node, int64_t ntraj, int maxdec, int32_t censor_base,
                               int nkeys, int32_t* __restrict__ bins) {
    for (int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; t < ntraj; t += (int64_t)gridDim.x * blockDim.x) {
        const int32_t c = codes[t];
        if (c < 0) continue;
        if (c >= censor_base) { atomicAdd(&bins[nkeys], 1); continue; }
        const int key = node[t] * maxdec + c;
        if (key < nkeys) atomicAdd(&bins[key], 1);
    }
}
// lanes_out[offset[key] + i] = the i-th lane of the key (order inside a cohort is arbitrary: lanes are independent and every
// row a program writes is indexed by lane); node[lane] <- the program the lane continues in; codes reset for the next level
__global__ void k_cohort_scatter(int32_t* __restrict__ codes, int32_t* __restrict__ node, int64_t ntraj, int maxdec, int32_t censor_base, int nkeys,
                                 const int32_t* __restrict__ offset, const int32_t* __restrict__ child, int32_t* __restrict__ cursor, uint32_t* __restrict__ lanes_out) {
    for (int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; t < ntraj; t += (int64_t)gridDim.x * blockDim.x) {
        const int32_t c = codes[t];
        if (c < 0) continue;
        codes[t] = -1;
        if (c >= censor_base) continue;
        const int key = node[t] * maxdec + c;
        if (key >= nkeys || child[key] < 0) continue;          // no program for this history (budget spent): the lane is dropped
        lanes_out[offset[key] + atomicAdd(&cursor[key], 1)] = (uint32_t)t;
        node[t] = child[key];
    }
}

// counts_host[key] for key < nnodes * maxdec, counts_host[nnodes * maxdec] = censored lanes.  Host-visible (one D2H + sync per level).
extern "C" int qsb_cohort_count(qsb_ctx* ctx, const int32_t* codes_dev, const int32_t* node_dev, int nnodes, int maxdec, int32_t censor_base, int32_t* counts_host) {
    if (!ctx || !codes_dev || !node_dev || !counts_host || nnodes < 1 || maxdec < 1) QSB_FAIL(QSB_ERR_BAD_ARG, "qsb_cohort_count: bad argument");
    const int nkeys = nnodes * maxdec;
    QSB_TRY(wave_buffer(ctx, (void**)&ctx->route_bins, &ctx->route_bins_bytes, (size_t)(3 * nkeys + 4) * sizeof(int32_t)));
    QSB_CUDA(cudaMemsetAsync(ctx->route_bins, 0, (size_t)(nkeys + 1) * sizeof(int32_t), ctx->stream));
    QSB_PRE(ctx, "k_cohort_count");
    k_cohort_count<<<ctx->sm_count * 8, 256, 0, ctx->stream>>>(codes_dev, node_dev, ctx->ntraj, maxdec, censor_base, nkeys, ctx->route_bins);
    QSB_LAUNCH_CHECK(ctx);
    QSB_CUDA(cudaMemcpyAsync(counts_host, ctx->route_bins, (size_t)(nkeys + 1) * sizeof(int32_t), cudaMemcpyDeviceToHost, ctx->stream));
    QSB_CUDA(cudaStreamSynchronize(ctx->stream));
    return QSB_OK;
}
// offsets_host / child_host: per key, where its lanes start in lanes_out_dev and which program they continue in (-1: dropped).
// Must follow qsb_cohort_count with the same nnodes / maxdec (it reuses the device scratch).
extern "C" int qsb_cohort_scatter(qsb_ctx* ctx, int32_t* codes_dev, int32_t* node_dev, int nnodes, int maxdec, int32_t censor_base,
                                  const int32_t* offsets_host, const int32_t* child_host, uint32_t* lanes_out_dev) {
    if (!ctx || !codes_dev || !node_dev || !offsets_host || !child_host || !lanes_out_dev) QSB_FAIL(QSB_ERR_BAD_ARG, "qsb_cohort_scatter: bad argument");
    const int nkeys = nnodes * maxdec;
    if (ctx->route_bins_bytes < (size_t)(3 * nkeys + 4) * sizeof(int32_t)) QSB_FAIL(QSB_ERR_BAD_ARG, "qsb_cohort_scatter without a matching qsb_cohort_count");
    int32_t* off = ctx->route_bins + nkeys + 1; int32_t* ch = off + nkeys;
    QSB_CUDA(cudaMemcpyAsync(off, offsets_host, (size_t)nkeys * sizeof(int32_t), cudaMemcpyHostToDevice, ctx->stream));   // pageable: staged before returning
    QSB_CUDA(cudaMemcpyAsync(ch, child_host, (size_t)nkeys * sizeof(int32_t), cudaMemcpyHostToDevice, ctx->stream));
    QSB_CUDA(cudaMemsetAsync(ctx->route_bins, 0, (size_t)nkeys * sizeof(int32_t), ctx->stream));    // the bins become the per-key cursors
    QSB_PRE(ctx, "k_cohort_scatter");
    k_cohort_scatter<<<ctx->sm_count * 8, 256, 0, ctx->stream>>>(codes_dev, node_dev, ctx->ntraj, maxdec, censor_base, nkeys, off, ch, ctx->route_bins, lanes_out_dev);
    QSB_LAUNCH_CHECK(ctx);
    return QSB_OK;
}
