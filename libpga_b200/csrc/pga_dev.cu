// pga_dev.cu -- device layer of the B200-native libpga engine (sm_100a only).
//
// Implements the internal C-ABI of pga_dev.h: device buffers, streams, events,
// kernel dispatch, ranking, migration payloads and the dlopen'ed NCCL transport.
// The host logic above it is plain C (pga_host.c).
//
// Data layout in HBM, per population (reference: src/pga.cu:37-46, :107-111):
//   genes[2]  [n][stride] 32-bit words, ping-pong; stride = row words rounded up
//             to 4 so every row is 16-byte aligned and contiguous (hooks take a
//             contiguous float[len]; TMA bulk copies and 128-bit loads need 16 B)
//   score[2]  [n] float, ping-pong (the fused kernel reads the parents' scores
//             and writes the children's)
//   best[4]   ring of 64-bit (order key, ~index) maxima, one slot per scoring launch
// The reference's third full-size buffer, rand[n][len], does not exist here.
#include <cuda_runtime.h>
#include <dlfcn.h>
#include <nccl.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "../../include/pga_b200.h"
#include "pga_dev.h"
#include "pga_kernels.cuh"

using namespace pgak;

// libpga's error convention (src/pga.cu:25-33)
#define CCE(ans) gpu_assert((ans), __FILE__, __LINE__)
static inline void gpu_assert(cudaError_t code, const char* file, int line)
{
    if (code != cudaSuccess) {
        fprintf(stderr, "GPUassert: %s %s %d\n", cudaGetErrorString(code), file, line);
        exit((int)code);
    }
}

struct pgadev_ctx {
    int device;
    int sm_count;
    cudaStream_t own_stream, stream, side;
    int route_side;  // ranking and packing go to the side stream (emigrant selection overlaps the next generation)
    cudaEvent_t ev_main, ev_side, ev_packed, t0, t1;
    unsigned long long launches;
    unsigned long long d2h_bytes;  // bytes copied device -> host by this solver
    ncclComm_t comm;
    int hook_threads;  // $PGA_HOOK_THREADS override, 0 = automatic
    // tuning knobs, settable from the environment for on-device experiments
    uint32_t fast_blocks_per_sm;  // $PGA_FAST_BLOCKS_PER_SM: grid cap of the fast kernels
    int fast_blocks_per_sm_set;
    int force_generic;            // $PGA_FORCE_GENERIC=1: skip the compile-time tiled kernels
    int no_hook_split;            // $PGA_NO_HOOK_SPLIT=1: generations with hooks are always one launch
    int no_coop;                  // $PGA_NO_COOP=1: rank with the 5-launch select instead of the cooperative kernel
    uint32_t reserve_now;         // SMs' worth of blocks the current launch leaves free (pgadev_step.reserve_sms)
    // early-stop check (pgax_set_target): best score of the local populations
    float* check_dev;
    float* check_host;  // pinned
    cudaEvent_t ev_check;
};

struct pgadev_pop {
    uint64_t n;
    uint32_t len;
    int enc;
    uint32_t row_words;  // host-visible words per row
    uint32_t stride;     // device words per row (multiple of 4)
    uint32_t group;      // lanes per child on the fast path
    uint32_t* genes[2];
    float* score[2];
    int gidx, sidx;
    unsigned long long* best;
    uint32_t best_seq;
    int32_t* sel;
    int capture;
    // ranking scratch
    TopkState2* topk[2];  // one set per index list: emigrants may be ranked on the side stream
    TopkOne* coop[2];
    unsigned long long coop_arrivals[2];  // barrier arrivals consumed so far by the cooperative select
    uint32_t coop_calls[2];               // its histograms alternate between two copies
    uint32_t *blk_gt[2], *blk_eq[2];
    uint32_t nb;
    uint32_t* idx[2];
    float* idx_score[2];  // scores of the listed individuals (pgadev_download_top_scores)
    uint64_t idx_cap[2];
    // hook path: fused launch or crossover+mutate then evaluate (see launch_hooks_fused)
    int hook_mode, hook_trials;
    float hook_ms[2];
    // migration boxes
    uint32_t *outbox, *inbox;
    uint64_t box_cap;
    // host <-> device staging for padded rows
    uint32_t* stage;
    size_t stage_bytes;
};

#define LAUNCH_CHECK(ctx)              \
    do {                               \
        CCE(cudaPeekAtLastError());    \
        (ctx)->launches += 1;          \
    } while (0)

// every call works on the solver's own device, whatever the caller made current since pga_init
static inline void use_device(pgadev_ctx* c)
{
    int d = -1;
    if (cudaGetDevice(&d) == cudaSuccess && d != c->device) CCE(cudaSetDevice(c->device));
}

static uint32_t grid_for(uint64_t work_items, uint32_t per_block, uint32_t cap)
{
    uint64_t g = (work_items + per_block - 1) / per_block;
    if (g < 1) g = 1;
    if (g > cap) g = cap;
    return (uint32_t)g;
}

// ---- context ------------------------------------------------------------------

extern "C" pgadev_ctx* pgadev_ctx_create(void)
{
    pgadev_ctx* c = (pgadev_ctx*)calloc(1, sizeof(*c));
    if (!c) return NULL;
    CCE(cudaGetDevice(&c->device));
    cudaDeviceProp prop;
    CCE(cudaGetDeviceProperties(&prop, c->device));
    c->sm_count = prop.multiProcessorCount;
    // the side stream (emigrant ranking, packing, NCCL) gets the highest priority so its
    // small kernels are placed ahead of the next generation's blocks
    int prio_lo = 0, prio_hi = 0;
    CCE(cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi));
    CCE(cudaStreamCreateWithPriority(&c->own_stream, cudaStreamNonBlocking, prio_lo));
    CCE(cudaStreamCreateWithPriority(&c->side, cudaStreamNonBlocking, prio_hi));
    c->stream = c->own_stream;
    CCE(cudaEventCreateWithFlags(&c->ev_main, cudaEventDisableTiming));
    CCE(cudaEventCreateWithFlags(&c->ev_side, cudaEventDisableTiming));
    CCE(cudaEventCreateWithFlags(&c->ev_packed, cudaEventDisableTiming));
    CCE(cudaEventCreate(&c->t0));
    CCE(cudaEventCreate(&c->t1));
    CCE(cudaEventCreateWithFlags(&c->ev_check, cudaEventDisableTiming));
    const char* ht = getenv("PGA_HOOK_THREADS");
    c->hook_threads = ht ? atoi(ht) : 0;
    const char* e;
    c->fast_blocks_per_sm = (e = getenv("PGA_FAST_BLOCKS_PER_SM")) && atoi(e) > 0 ? (uint32_t)atoi(e) : 16u;
    c->fast_blocks_per_sm_set = e && atoi(e) > 0;
    c->force_generic = (e = getenv("PGA_FORCE_GENERIC")) ? atoi(e) : 0;
    c->no_hook_split = (e = getenv("PGA_NO_HOOK_SPLIT")) ? atoi(e) : 0;
    c->no_coop = (e = getenv("PGA_NO_COOP")) ? atoi(e) : 0;
    const char* stack = getenv("PGA_STACK_BYTES");  // device stack for deep user hooks
    if (stack && atoi(stack) > 0) CCE(cudaDeviceSetLimit(cudaLimitStackSize, (size_t)atoi(stack)));
    return c;
}

static void nccl_destroy(pgadev_ctx* c);

extern "C" void pgadev_ctx_destroy(pgadev_ctx* c)
{
    if (!c) return;
    nccl_destroy(c);
    CCE(cudaEventDestroy(c->ev_main));
    CCE(cudaEventDestroy(c->ev_side));
    CCE(cudaEventDestroy(c->ev_packed));
    CCE(cudaEventDestroy(c->t0));
    CCE(cudaEventDestroy(c->t1));
    CCE(cudaEventDestroy(c->ev_check));
    if (c->check_dev) CCE(cudaFree(c->check_dev));
    if (c->check_host) CCE(cudaFreeHost(c->check_host));
    CCE(cudaStreamDestroy(c->own_stream));
    CCE(cudaStreamDestroy(c->side));
    free(c);
}

extern "C" void pgadev_ctx_set_stream(pgadev_ctx* c, void* s) { c->stream = s ? (cudaStream_t)s : c->own_stream; }

extern "C" void pgadev_ctx_sync(pgadev_ctx* c)
{
    use_device(c);
    CCE(cudaStreamSynchronize(c->stream));
    CCE(cudaStreamSynchronize(c->side));
}

extern "C" unsigned long long pgadev_ctx_launches(pgadev_ctx* c) { return c->launches; }
extern "C" unsigned long long pgadev_ctx_d2h_bytes(pgadev_ctx* c) { return c->d2h_bytes; }
extern "C" void pgadev_timer_start(pgadev_ctx* c) { use_device(c); CCE(cudaEventRecord(c->t0, c->stream)); }
extern "C" void pgadev_timer_stop(pgadev_ctx* c) { use_device(c); CCE(cudaEventRecord(c->t1, c->stream)); }
extern "C" float pgadev_timer_ms(pgadev_ctx* c)
{
    float ms = 0.f;
    CCE(cudaEventSynchronize(c->t1));
    CCE(cudaEventElapsedTime(&ms, c->t0, c->t1));
    return ms;
}

// ---- $PGA_TRACE=1: where the time of a run goes -------------------------------------------------------
// Events on the main stream at named points of the generation / migration schedule; the gaps
// between consecutive marks are summed per (from, to) pair and printed to stderr when the run
// ends.  Costs nothing when the variable is not set.
#define TRACE_MAX 8192
static struct {
    int enabled, init;
    cudaEvent_t ev[TRACE_MAX];
    const char* label[TRACE_MAX];
    int count;
} g_trace;

extern "C" void pgadev_trace_mark(pgadev_ctx* c, const char* label)
{
    if (!g_trace.init) {
        const char* e = getenv("PGA_TRACE");
        g_trace.enabled = e && atoi(e) > 0;
        g_trace.init = 1;
    }
    if (!g_trace.enabled || g_trace.count >= TRACE_MAX) return;
    if (!g_trace.ev[g_trace.count]) CCE(cudaEventCreate(&g_trace.ev[g_trace.count]));
    CCE(cudaEventRecord(g_trace.ev[g_trace.count], c->stream));
    g_trace.label[g_trace.count++] = label;
}

extern "C" void pgadev_trace_report(pgadev_ctx* c)
{
    if (!g_trace.enabled || g_trace.count < 2) { g_trace.count = 0; return; }
    CCE(cudaStreamSynchronize(c->stream));
    struct { const char *a, *b; double ms; int n; } agg[64];
    int na = 0;
    for (int i = 1; i < g_trace.count; ++i) {
        float ms = 0.f;
        CCE(cudaEventElapsedTime(&ms, g_trace.ev[i - 1], g_trace.ev[i]));
        int j = 0;
        while (j < na && !(agg[j].a == g_trace.label[i - 1] && agg[j].b == g_trace.label[i])) ++j;
        if (j == na) { if (na == 64) continue; agg[na].a = g_trace.label[i - 1]; agg[na].b = g_trace.label[i]; agg[na].ms = 0; agg[na].n = 0; ++na; }
        agg[j].ms += ms; agg[j].n += 1;
    }
    for (int j = 0; j < na; ++j)
        fprintf(stderr, "pga-trace: %-22s -> %-22s %6d x %9.2f us\n", agg[j].a, agg[j].b, agg[j].n, 1e3 * agg[j].ms / agg[j].n);
    g_trace.count = 0;
}

// ---- populations ------------------------------------------------------------------

extern "C" pgadev_pop* pgadev_pop_create(pgadev_ctx* c, uint64_t size, uint32_t genome_len, int encoding)
{
    use_device(c);
    pgadev_pop* p = (pgadev_pop*)calloc(1, sizeof(*p));
    if (!p) return NULL;
    p->n = size;
    p->len = genome_len;
    p->enc = encoding;
    p->row_words = encoding == PGAX_GENES_BITS ? (genome_len + 31u) / 32u : genome_len;
    p->stride = (p->row_words + 3u) & ~3u;
    const uint32_t nblk = p->stride >> 2;
    p->group = 1;
    while (p->group < nblk && p->group < 32u) p->group <<= 1;
    const size_t gene_bytes = (size_t)(size ? size : 1) * p->stride * 4u;
    const size_t score_bytes = (size_t)(size ? size : 1) * sizeof(float);
    // reference: four cudaMallocs, three of them population-sized (src/pga.cu:108-111)
    for (int i = 0; i < 2; ++i) {
        if (cudaMalloc((void**)&p->genes[i], gene_bytes) != cudaSuccess ||
            cudaMalloc((void**)&p->score[i], score_bytes) != cudaSuccess) {
            cudaGetLastError();
            pgadev_pop_destroy(c, p);
            return NULL;
        }
        CCE(cudaMemsetAsync(p->score[i], 0, score_bytes, c->stream));
        // next_gen starts zeroed: a crossover hook that does not write every gene sees this
        // row (the reference leaves it uninitialised, src/pga.cu:109, :302)
        CCE(cudaMemsetAsync(p->genes[i], 0, gene_bytes, c->stream));
    }
    CCE(cudaMalloc((void**)&p->best, 4 * sizeof(unsigned long long)));
    CCE(cudaMemsetAsync(p->best, 0, 4 * sizeof(unsigned long long), c->stream));
    return p;
}

extern "C" void pgadev_pop_destroy(pgadev_ctx* c, pgadev_pop* p)
{
    use_device(c);
    if (!p) return;
    for (int i = 0; i < 2; ++i) {
        if (p->genes[i]) CCE(cudaFree(p->genes[i]));
        if (p->score[i]) CCE(cudaFree(p->score[i]));
        if (p->idx[i]) CCE(cudaFree(p->idx[i]));
        if (p->idx_score[i]) CCE(cudaFree(p->idx_score[i]));
    }
    if (p->best) CCE(cudaFree(p->best));
    if (p->sel) CCE(cudaFree(p->sel));
    for (int i = 0; i < 2; ++i) {
        if (p->topk[i]) CCE(cudaFree(p->topk[i]));
        if (p->coop[i]) CCE(cudaFree(p->coop[i]));
        if (p->blk_gt[i]) CCE(cudaFree(p->blk_gt[i]));
        if (p->blk_eq[i]) CCE(cudaFree(p->blk_eq[i]));
    }
    if (p->outbox) CCE(cudaFree(p->outbox));
    if (p->inbox) CCE(cudaFree(p->inbox));
    if (p->stage) CCE(cudaFree(p->stage));
    free(p);
}

extern "C" uint32_t pgadev_pop_row_words(pgadev_pop* p) { return p->row_words; }

static DrawCtx make_draw(uint64_t seed, uint32_t island, uint32_t epoch)
{
    DrawCtx d;
    d.k0 = (uint32_t)seed;
    d.k1 = (uint32_t)(seed >> 32);
    d.epoch = epoch;
    d.island = island;
    return d;
}

extern "C" void pgadev_pop_init_random(pgadev_ctx* c, pgadev_pop* p, uint64_t seed, uint32_t island)
{
    use_device(c);
    if (p->n == 0) return;
    const uint64_t blocks16 = p->n * (p->stride >> 2);
    const uint32_t grid = grid_for(blocks16, 256, (uint32_t)c->sm_count * 16u);
    const DrawCtx d = make_draw(seed, island, 0);
    if (p->enc == PGAX_GENES_BITS)
        init_population<ENC_BITS><<<grid, 256, 0, c->stream>>>(p->genes[p->gidx], (uint32_t)p->n, p->len, p->stride, d);
    else
        init_population<ENC_F32><<<grid, 256, 0, c->stream>>>(p->genes[p->gidx], (uint32_t)p->n, p->len, p->stride, d);
    LAUNCH_CHECK(c);
}

extern "C" void pgadev_pop_capture_selection(pgadev_ctx* c, pgadev_pop* p, int enable)
{
    use_device(c);
    if (enable && !p->sel) CCE(cudaMalloc((void**)&p->sel, (size_t)(p->n ? p->n : 1) * 2 * sizeof(int32_t)));
    p->capture = enable;
}

extern "C" void pgadev_swap_genes(pgadev_pop* p) { p->gidx ^= 1; }
extern "C" void pgadev_swap_scores(pgadev_pop* p) { p->sidx ^= 1; }

// ---- kernel dispatch ----------------------------------------------------------------

#define SMEM_LIMIT (227u * 1024u)
#define MAX_DEVICES 64

template <uint32_t PH>
static int launch_hooks(pgadev_ctx* c, pgadev_pop* p, const GenArgs& a)
{
    const bool want_rand = ((PH & PH_CROSSOVER) && a.xov) || ((PH & PH_MUTATE) && a.mut);
    const bool stage_parents = (PH & PH_CROSSOVER) && a.xov;  // only a crossover hook needs parents in shared memory
    const uint32_t rows = 1u + (stage_parents ? 2u : 0u) + (want_rand ? 1u : 0u);
    const size_t per_thread = (size_t)rows * p->stride * 4u;
    const size_t limit = SMEM_LIMIT - 64;  // static shared memory of the kernel
    uint32_t T = 128;
    while (T > 32 && 128 + T * per_thread > limit) T >>= 1;
    if (128 + T * per_thread > limit) return PGAX_ERANGE;
    if (c->hook_threads == 32 || c->hook_threads == 64 || c->hook_threads == 128) {
        if ((uint32_t)c->hook_threads < T) T = (uint32_t)c->hook_threads;
    } else {
        // enough blocks to give every SM four, when the population allows it
        // (C1, 40000 x 100: 16.4 us per generation with 64 threads against 18.5 with 128)
        while (T > 32 && (p->n + T - 1) / T < 4ull * (uint64_t)c->sm_count) T >>= 1;
    }
    const size_t smem_final = 128 + T * per_thread;
    CCE(cudaFuncSetAttribute(generation_hooks<PH>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_final));
    const uint32_t grid = (uint32_t)((p->n + T - 1) / T);
    generation_hooks<PH><<<grid, T, smem_final, c->stream>>>(a);
    LAUNCH_CHECK(c);
    return 0;
}

// A generation with a crossover hook stages four rows per thread in shared memory (child,
// both parents, the rand row), which leaves 4 warps per SM.  When the objective hook is the
// expensive part (test3: 10^4 compare iterations per evaluation) it pays to score the
// children in a second launch that holds one row per thread and so runs 4x the threads:
// at 2^20 x 100 the fused launch takes 20.2 ms, crossover+mutate then evaluate 16.5 ms.
// Which is faster depends on the user's hooks, so the first two generations are timed
// (one each way) and the population keeps the winner.  Results are identical either way.
static int launch_hooks_split(pgadev_ctx* c, pgadev_pop* p, const GenArgs& a)
{
    int rc = launch_hooks<PH_CROSSOVER | PH_MUTATE>(c, p, a);
    if (rc != 0) return rc;
    GenArgs e = a;
    e.cur = a.nxt;  // the evaluate phase scores "current_gen": point it at the children
    return launch_hooks<PH_EVALUATE>(c, p, e);
}

static int launch_hooks_fused(pgadev_ctx* c, pgadev_pop* p, const GenArgs& a)
{
    const bool candidate = a.xov != nullptr && a.objective == OBJ_HOOK && !c->no_hook_split && p->n >= 4096;
    if (!candidate) return launch_hooks<PH_FUSED>(c, p, a);
    if (p->hook_mode == 1) return launch_hooks<PH_FUSED>(c, p, a);
    if (p->hook_mode == 2) return launch_hooks_split(c, p, a);
    // a stream that is being captured cannot be timed or waited on: keep the single launch
    cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
    if (cudaStreamIsCapturing(c->stream, &cap) != cudaSuccess || cap != cudaStreamCaptureStatusNone) {
        cudaGetLastError();
        return launch_hooks<PH_FUSED>(c, p, a);
    }
    // undecided: the first four generations go fused, split, fused, split.  The first pair
    // is not timed (it pays the module load and the first-launch set-up of each form), the
    // second pair is, and the faster form is kept.  Every generation is launched exactly
    // once, so a hook that reads its child row sees what it would see without the trial.
    const int trial = p->hook_trials;
    const bool split = (trial & 1) != 0, timed = trial >= 2;
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    if (timed) {
        CCE(cudaEventCreate(&e0));
        CCE(cudaEventCreate(&e1));
        CCE(cudaEventRecord(e0, c->stream));
    }
    const int rc = split ? launch_hooks_split(c, p, a) : launch_hooks<PH_FUSED>(c, p, a);
    if (timed) {
        CCE(cudaEventRecord(e1, c->stream));
        CCE(cudaEventSynchronize(e1));
        CCE(cudaEventElapsedTime(&p->hook_ms[trial & 1], e0, e1));
        CCE(cudaEventDestroy(e0));
        CCE(cudaEventDestroy(e1));
    }
    p->hook_trials += 1;
    if (p->hook_trials == 4) p->hook_mode = p->hook_ms[1] < p->hook_ms[0] ? 2 : 1;
    return rc;
}

template <int ENC, uint32_t PH, int OBJ>
static int launch_fast_inst(pgadev_ctx* c, pgadev_pop* p, const GenArgs& a)
{
    const uint32_t grid = grid_for(p->n, 256, (uint32_t)c->sm_count * 16u);
    generation_fast<ENC, PH, OBJ><<<grid, 256, 0, c->stream>>>(a);
    LAUNCH_CHECK(c);
    return 0;
}

// Tile batch and occupancy target per shape, from the B200 sweeps in
// profiles/r1_sweeps.md: packed bits run best with 16 parent loads in flight per
// lane at 2 blocks/SM (the group width hardly matters: 8 lanes x 8 tiles 71.9 us,
// 4 lanes x 2 blocks x 4 tiles 73.1 us, 8 lanes x 4 tiles 82 us per OneMax generation),
// f32 rows (more ALU work per byte) with 8 at 3 blocks/SM.
template <int ENC, int LOGG, int NB>
struct TiledShape {
    static constexpr int G = 1 << LOGG;
    static constexpr int TB = ENC == ENC_BITS ? (NB == 1 ? (G < 8 ? G : 8) : (NB == 2 ? (G < 4 ? G : 4) : 2)) : (NB >= 2 ? 2 : (G < 4 ? G : 4));
    static constexpr int MINB = ENC == ENC_BITS ? 2 : 3;
};

template <int ENC, int OBJ, int LOGG, int NB, bool EXACT>
static int launch_tiled_one(pgadev_ctx* c, pgadev_pop* p, const GenArgs& a)
{
    constexpr int TB = TiledShape<ENC, LOGG, NB>::TB, MINB = TiledShape<ENC, LOGG, NB>::MINB;
    // exactly one resident wave of 8 * MINB warps per SM (a partial second wave costs a full
    // kernel tail).  While a migration is in flight the launch holds `reserve_now` SMs back
    // and takes the one-block-per-SM form, which leaves those SMs empty for the side stream.
    const uint32_t per_sm = c->fast_blocks_per_sm_set ? c->fast_blocks_per_sm : (uint32_t)MINB;
    if (c->reserve_now == 0) {
        const uint32_t grid = grid_for(p->n, 256, (uint32_t)c->sm_count * per_sm);
        generation_fused_tiled<ENC, OBJ, LOGG, NB, TB, EXACT, MINB, false><<<grid, 256, 0, c->stream>>>(a);
    } else {
        const uint32_t sms = (uint32_t)c->sm_count - (c->reserve_now < (uint32_t)c->sm_count / 2u ? c->reserve_now : (uint32_t)c->sm_count / 2u);
        const uint32_t grid = grid_for(p->n, 256u * MINB, sms);
        generation_fused_tiled<ENC, OBJ, LOGG, NB, TB, EXACT, MINB, true><<<grid, 256 * MINB, 0, c->stream>>>(a);
    }
    LAUNCH_CHECK(c);
    return 0;
}

template <int ENC, int OBJ, int LOGG, int NB>
static int launch_tiled_inst(pgadev_ctx* c, pgadev_pop* p, const GenArgs& a)
{
    // EXACT: every lane of a group owns exactly NB blocks and (f32) no gene is padding
    const bool exact = (p->stride >> 2) == (uint32_t)NB << LOGG && (ENC == ENC_BITS || p->len == p->stride);
    return exact ? launch_tiled_one<ENC, OBJ, LOGG, NB, true>(c, p, a) : launch_tiled_one<ENC, OBJ, LOGG, NB, false>(c, p, a);
}

// Shapes with a compile-time tiling: G = min(32, pow2ceil(blocks)) lanes per child (the
// oracle's canonical group for f32 scores), NB = blocks per lane up to 4.  Anything
// larger takes the generic kernel.
template <int ENC, int OBJ>
static int launch_tiled(pgadev_ctx* c, pgadev_pop* p, const GenArgs& a)
{
    const uint32_t nblk = p->stride >> 2;
    if (c->force_generic) return 1;
    uint32_t G = p->group;
    const uint32_t NBr = (nblk + G - 1) / G;
#define TL(L, N) return launch_tiled_inst<ENC, OBJ, L, N>(c, p, a)
    if (NBr == 1) {
        switch (G) {
            case 1: TL(0, 1);
            case 2: TL(1, 1);
            case 4: TL(2, 1);
            case 8: TL(3, 1);
            case 16: TL(4, 1);
            default: TL(5, 1);
        }
    }
    if (NBr == 2) {
        switch (G) {
            case 1: TL(0, 2);
            case 2: TL(1, 2);
            case 4: TL(2, 2);
            case 8: TL(3, 2);
            case 16: TL(4, 2);
            default: TL(5, 2);
        }
    }
    if (G == 32 && NBr == 3) TL(5, 3);
    if (G == 32 && NBr == 4) TL(5, 4);
#undef TL
    return 1;
}

template <int ENC, int OBJ>
static int launch_fast_ph(pgadev_ctx* c, pgadev_pop* p, const GenArgs& a, uint32_t ph)
{
    if (ph == PH_FUSED && a.tk == 2u) {  // the compile-time tilings are built for the reference's size-2 tournament
        int rc = launch_tiled<ENC, OBJ>(c, p, a);
        if (rc <= 0) return rc;
    }
    switch (ph) {
        case PH_FUSED: return launch_fast_inst<ENC, PH_FUSED, OBJ>(c, p, a);
        case PH_EVALUATE: return launch_fast_inst<ENC, PH_EVALUATE, OBJ>(c, p, a);
        case PH_CROSSOVER: return launch_fast_inst<ENC, PH_CROSSOVER, OBJ_SUM>(c, p, a);
        case PH_MUTATE: return launch_fast_inst<ENC, PH_MUTATE, OBJ_SUM>(c, p, a);
        default: return PGAX_EINVAL;
    }
}

extern "C" int pgadev_launch(pgadev_ctx* c, pgadev_pop* p, const pgadev_step* st)
{
    use_device(c);
    if (p->n == 0) return 0;
    const uint32_t ph = st->phases;
    if (ph != PH_FUSED && ph != PH_EVALUATE && ph != PH_CROSSOVER && ph != PH_MUTATE) return PGAX_EINVAL;

    c->reserve_now = st->reserve_sms;
    GenArgs a;
    memset(&a, 0, sizeof(a));
    a.cur = p->genes[p->gidx];
    a.nxt = p->genes[p->gidx ^ 1];
    a.score_in = p->score[p->sidx];
    a.score_out = ph == PH_FUSED ? p->score[p->sidx ^ 1] : p->score[p->sidx];
    a.best = p->best;
    a.best_slot = p->best_seq & 3u;
    a.sel = p->capture ? p->sel : nullptr;
    a.n = (uint32_t)p->n;
    a.nf = (float)p->n;
    a.len = p->len;
    a.lenf = (float)p->len;
    a.stride = p->stride;
    a.group = p->group;
    a.tk = st->tournament ? st->tournament : 2u;
    a.draw = make_draw(st->seed, st->island, st->epoch);
    a.objective = st->objective;
    a.obj = (obj_fn)st->obj_hook;
    a.mut = (mutate_fn)st->mutate_hook;
    a.xov = (crossover_fn)st->crossover_hook;

    const bool uses_xov = (ph & PH_CROSSOVER) && a.xov, uses_mut = (ph & PH_MUTATE) && a.mut;
    const bool scores = (ph & PH_EVALUATE) != 0;
    int rc;
    if (p->enc == PGAX_GENES_BITS) {
        // packed bits: default operators and the bit objective only
        if (uses_xov || uses_mut) return PGAX_ENOSYS;
        if (scores && st->objective != PGAX_OBJ_ONEMAX) return PGAX_ENOSYS;
        rc = launch_fast_ph<ENC_BITS, OBJ_ONEMAX>(c, p, a, ph);
    } else {
        if (scores && st->objective == PGAX_OBJ_ONEMAX) return PGAX_ENOSYS;
        const bool hook_obj = scores && st->objective == PGAX_OBJ_HOOK;
        if (uses_xov || uses_mut || hook_obj) {
            if (hook_obj && !a.obj) return PGAX_EINVAL;
            switch (ph) {
                case PH_FUSED: rc = launch_hooks_fused(c, p, a); break;
                case PH_EVALUATE: rc = launch_hooks<PH_EVALUATE>(c, p, a); break;
                case PH_CROSSOVER: rc = launch_hooks<PH_CROSSOVER>(c, p, a); break;
                default: rc = launch_hooks<PH_MUTATE>(c, p, a); break;
            }
        } else {
            switch (scores ? st->objective : PGAX_OBJ_SUM) {
                case PGAX_OBJ_SUM: rc = launch_fast_ph<ENC_F32, OBJ_SUM>(c, p, a, ph); break;
                case PGAX_OBJ_RASTRIGIN: rc = launch_fast_ph<ENC_F32, OBJ_RASTRIGIN>(c, p, a, ph); break;
                case PGAX_OBJ_SPHERE: rc = launch_fast_ph<ENC_F32, OBJ_SPHERE>(c, p, a, ph); break;
                default: return PGAX_EINVAL;
            }
        }
    }
    if (rc == 0 && scores) p->best_seq += 1;
    return rc;
}

// ---- best individual ---------------------------------------------------------------------

extern "C" void pgadev_best(pgadev_ctx* c, pgadev_pop* p, int reduced_in_kernel, uint64_t* index, float* score)
{
    use_device(c);
    if (!reduced_in_kernel) {
        const uint32_t grid = grid_for(p->n, 256 * 8, (uint32_t)c->sm_count * 8u);
        reduce_best<<<grid, 256, 0, c->stream>>>(p->score[p->sidx], (uint32_t)p->n, p->best, p->best_seq & 3u);
        LAUNCH_CHECK(c);
        p->best_seq += 1;
    }
    unsigned long long key = 0;
    float first = 0.f;
    CCE(cudaMemcpyAsync(&key, p->best + ((p->best_seq - 1u) & 3u), sizeof(key), cudaMemcpyDeviceToHost, c->stream));
    c->d2h_bytes += sizeof(key) + 2 * sizeof(float);
    CCE(cudaMemcpyAsync(&first, p->score[p->sidx], sizeof(float), cudaMemcpyDeviceToHost, c->stream));
    CCE(cudaStreamSynchronize(c->stream));
    // src/pga.cu:221-229 never displaces a NaN sitting at index 0
    uint64_t idx = first != first ? 0u : (uint64_t)(0xFFFFFFFFu - (uint32_t)(key & 0xFFFFFFFFull));
    float s = first;
    if (idx != 0) {
        CCE(cudaMemcpyAsync(&s, p->score[p->sidx] + idx, sizeof(float), cudaMemcpyDeviceToHost, c->stream));
        CCE(cudaStreamSynchronize(c->stream));
    }
    *index = idx;
    *score = s;
}

// ---- host <-> device copies -----------------------------------------------------------------

static void copy_rows(pgadev_ctx* c, pgadev_pop* p, uint32_t* dev_rows, void* host, bool upload)
{
    if (p->n == 0) return;
    const size_t host_bytes = (size_t)p->n * p->row_words * 4u;
    if (p->row_words == p->stride) {
        if (upload) CCE(cudaMemcpyAsync(dev_rows, host, host_bytes, cudaMemcpyHostToDevice, c->stream));
        else { CCE(cudaMemcpyAsync(host, dev_rows, host_bytes, cudaMemcpyDeviceToHost, c->stream)); c->d2h_bytes += host_bytes; }
        CCE(cudaStreamSynchronize(c->stream));
        return;
    }
    // unpadded staging copy of the rows, kept for the next transfer (freed with the population)
    if (p->stage_bytes < host_bytes) {
        if (p->stage) CCE(cudaFree(p->stage));
        CCE(cudaMalloc((void**)&p->stage, host_bytes));
        p->stage_bytes = host_bytes;
    }
    uint32_t* tmp = p->stage;
    const uint32_t grid = grid_for(p->n * p->stride, 256, (uint32_t)c->sm_count * 16u);
    if (upload) {
        CCE(cudaMemcpyAsync(tmp, host, host_bytes, cudaMemcpyHostToDevice, c->stream));
        repack_rows<<<grid, 256, 0, c->stream>>>(tmp, dev_rows, p->n, p->row_words, p->stride);
        LAUNCH_CHECK(c);
    } else {
        repack_rows<<<grid, 256, 0, c->stream>>>(dev_rows, tmp, p->n, p->stride, p->row_words);
        LAUNCH_CHECK(c);
        CCE(cudaMemcpyAsync(host, tmp, host_bytes, cudaMemcpyDeviceToHost, c->stream));
        c->d2h_bytes += host_bytes;
    }
    CCE(cudaStreamSynchronize(c->stream));
}

extern "C" void pgadev_upload_rows(pgadev_ctx* c, pgadev_pop* p, const void* host, int next_gen)
{
    use_device(c);
    copy_rows(c, p, p->genes[p->gidx ^ (next_gen ? 1 : 0)], (void*)host, true);
}

extern "C" void pgadev_download_rows(pgadev_ctx* c, pgadev_pop* p, void* host, int next_gen)
{
    use_device(c);
    copy_rows(c, p, p->genes[p->gidx ^ (next_gen ? 1 : 0)], host, false);
}

extern "C" void pgadev_download_row(pgadev_ctx* c, pgadev_pop* p, uint64_t index, void* host_words)
{
    use_device(c);
    CCE(cudaMemcpyAsync(host_words, p->genes[p->gidx] + index * p->stride, (size_t)p->row_words * 4u, cudaMemcpyDeviceToHost,
                        c->stream));
    c->d2h_bytes += (size_t)p->row_words * 4u;
    CCE(cudaStreamSynchronize(c->stream));
}

extern "C" void pgadev_upload_scores(pgadev_ctx* c, pgadev_pop* p, const float* host)
{
    use_device(c);
    CCE(cudaMemcpyAsync(p->score[p->sidx], host, p->n * sizeof(float), cudaMemcpyHostToDevice, c->stream));
    CCE(cudaStreamSynchronize(c->stream));
}

extern "C" void pgadev_download_scores(pgadev_ctx* c, pgadev_pop* p, float* host)
{
    use_device(c);
    CCE(cudaMemcpyAsync(host, p->score[p->sidx], p->n * sizeof(float), cudaMemcpyDeviceToHost, c->stream));
    c->d2h_bytes += p->n * sizeof(float);
    CCE(cudaStreamSynchronize(c->stream));
}

extern "C" int pgadev_download_selection(pgadev_ctx* c, pgadev_pop* p, int32_t* host_pairs)
{
    use_device(c);
    if (!p->sel) return PGAX_EINVAL;
    CCE(cudaMemcpyAsync(host_pairs, p->sel, p->n * 2 * sizeof(int32_t), cudaMemcpyDeviceToHost, c->stream));
    c->d2h_bytes += p->n * 2 * sizeof(int32_t);
    CCE(cudaStreamSynchronize(c->stream));
    return 0;
}

// ---- ranking ------------------------------------------------------------------------------

static void ensure_index_list(pgadev_ctx* c, pgadev_pop* p, int slot, uint64_t k)
{
    if (p->idx_cap[slot] >= k) return;
    pgadev_ctx_sync(c);
    if (p->idx[slot]) CCE(cudaFree(p->idx[slot]));
    if (p->idx_score[slot]) CCE(cudaFree(p->idx_score[slot]));
    CCE(cudaMalloc((void**)&p->idx[slot], k * sizeof(uint32_t)));
    CCE(cudaMalloc((void**)&p->idx_score[slot], k * sizeof(float)));
    p->idx_cap[slot] = k;
}

static cudaStream_t work_stream(pgadev_ctx* c) { return c->route_side ? c->side : c->stream; }

// The whole selection as one cooperative launch of one 1024-thread block per SM, when E scores
// per thread (E = 8, 16 or 32) cover the population.
template <int E>
static bool select_top_coop_e(pgadev_ctx* c, pgadev_pop* p, uint32_t k, int worst, int slot, cudaStream_t s)
{
    static int occupancy[MAX_DEVICES];  // per device ordinal; -1 = the kernel does not fit
    int blocks_per_sm = occupancy[c->device % MAX_DEVICES];
    if (blocks_per_sm == 0) {
        CCE(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&blocks_per_sm, topk_select_coop<E>, (int)TOPK1_THREADS, 0));
        if (blocks_per_sm < 1) blocks_per_sm = -1;
        occupancy[c->device % MAX_DEVICES] = blocks_per_sm;
    }
    if (blocks_per_sm < 0) return false;
    const uint64_t per_block = (uint64_t)TOPK1_THREADS * E;
    const uint64_t grid = (p->n + per_block - 1) / per_block;
    if (grid > (uint64_t)c->sm_count || grid > 160u) return false;
    if (!p->coop[slot]) {
        CCE(cudaMalloc((void**)&p->coop[slot], sizeof(TopkOne)));
        CCE(cudaMemsetAsync(p->coop[slot], 0, sizeof(TopkOne), s));
        p->coop_arrivals[slot] = 0;
        p->coop_calls[slot] = 0;
    }
    const float* score = p->score[p->sidx];
    uint32_t n = (uint32_t)p->n;
    TopkOne* st = p->coop[slot];
    uint32_t* out = p->idx[slot];
    unsigned long long base = p->coop_arrivals[slot];
    uint32_t parity = p->coop_calls[slot] & 1u;
    void* args[] = { (void*)&score, (void*)&n, (void*)&k, (void*)&worst, (void*)&st, (void*)&out, (void*)&base, (void*)&parity };
    CCE(cudaLaunchCooperativeKernel((const void*)topk_select_coop<E>, dim3((unsigned)grid), dim3(TOPK1_THREADS), args, 0, s));
    c->launches += 1;
    p->coop_arrivals[slot] += 3ull * grid;  // one grid barrier per radix pass
    p->coop_calls[slot] += 1;
    return true;
}

static bool select_top_coop(pgadev_ctx* c, pgadev_pop* p, uint32_t k, int worst, int slot, cudaStream_t s)
{
    if (c->no_coop) return false;
    const uint64_t threads = (uint64_t)c->sm_count * TOPK1_THREADS;
    if (p->n <= threads * 8u) return select_top_coop_e<8>(c, p, k, worst, slot, s);
    if (p->n <= threads * 16u) return select_top_coop_e<16>(c, p, k, worst, slot, s);
    return select_top_coop_e<32>(c, p, k, worst, slot, s);
}

extern "C" void pgadev_route_to_side(pgadev_ctx* c, int on) { c->route_side = on; }

// five launches: three radix passes (the last block of each picks the digit), the per-chunk
// counts (the last block scans them), the stable compaction.  grid_cap = 0: one block per chunk.
static void launch_select_5(pgadev_ctx* c, const float* score, uint32_t n, uint32_t k, int worst, TopkState2* st, uint32_t* bg,
                            uint32_t* be, uint32_t* out, uint32_t nb, uint32_t grid_cap, cudaStream_t s)
{
    const uint32_t grid = grid_cap && grid_cap < nb ? grid_cap : nb;
    topk2_histogram<21, 11><<<grid, 256, 0, s>>>(score, n, worst, k, st, nb);
    topk2_histogram<10, 11><<<grid, 256, 0, s>>>(score, n, worst, k, st, nb);
    topk2_histogram<0, 10><<<grid, 256, 0, s>>>(score, n, worst, k, st, nb);
    topk2_count_scan<<<grid, TOPK_BLOCK, 0, s>>>(score, n, worst, st, bg, be, nb);
    topk2_emit<<<grid, TOPK_BLOCK, 0, s>>>(score, n, worst, st, bg, be, out, nb);
    if (c) {
        CCE(cudaPeekAtLastError());
        c->launches += 5;
    }
}

extern "C" void pgadev_select_top(pgadev_ctx* c, pgadev_pop* p, uint64_t k, int worst, int slot)
{
    use_device(c);
    if (k == 0 || p->n == 0) return;
    if (k > p->n) k = p->n;
    if (!p->topk[slot]) {
        p->nb = (uint32_t)((p->n + TOPK_CHUNK - 1) / TOPK_CHUNK);
        CCE(cudaMalloc((void**)&p->topk[slot], sizeof(TopkState2)));
        CCE(cudaMemset(p->topk[slot], 0, sizeof(TopkState2)));
        CCE(cudaMalloc((void**)&p->blk_gt[slot], (size_t)p->nb * sizeof(uint32_t)));
        CCE(cudaMalloc((void**)&p->blk_eq[slot], (size_t)p->nb * sizeof(uint32_t)));
    }
    ensure_index_list(c, p, slot, k);
    cudaStream_t s = work_stream(c);
    TopkState2* st = p->topk[slot];
    uint32_t *bg = p->blk_gt[slot], *be = p->blk_eq[slot];
    const float* score = p->score[p->sidx];
    const uint32_t n = (uint32_t)p->n;
    if (select_top_coop(c, p, (uint32_t)k, worst, slot, s)) return;
    launch_select_5(c, score, n, (uint32_t)k, worst, st, bg, be, p->idx[slot], p->nb, 0u, s);
}

extern "C" void pgadev_download_top(pgadev_ctx* c, pgadev_pop* p, int slot, uint64_t k, uint32_t* host_idx)
{
    use_device(c);
    CCE(cudaMemcpyAsync(host_idx, p->idx[slot], k * sizeof(uint32_t), cudaMemcpyDeviceToHost, c->stream));
    c->d2h_bytes += k * sizeof(uint32_t);
    CCE(cudaStreamSynchronize(c->stream));
}

// scores of the k listed individuals only: k floats cross PCIe, not the whole score array
extern "C" void pgadev_download_top_scores(pgadev_ctx* c, pgadev_pop* p, int slot, uint64_t k, float* host_scores)
{
    use_device(c);
    if (k == 0) return;
    gather_scores<<<grid_for(k, 256, (uint32_t)c->sm_count * 8u), 256, 0, c->stream>>>(p->score[p->sidx], p->idx[slot], (uint32_t)k,
                                                                                     p->idx_score[slot]);
    LAUNCH_CHECK(c);
    CCE(cudaMemcpyAsync(host_scores, p->idx_score[slot], k * sizeof(float), cudaMemcpyDeviceToHost, c->stream));
    c->d2h_bytes += k * sizeof(float);
    CCE(cudaStreamSynchronize(c->stream));
}

// ---- migration payloads -----------------------------------------------------------------------

static void ensure_boxes(pgadev_ctx* c, pgadev_pop* p, uint64_t k)
{
    if (p->box_cap >= k) return;
    pgadev_ctx_sync(c);
    if (p->outbox) CCE(cudaFree(p->outbox));
    if (p->inbox) CCE(cudaFree(p->inbox));
    const size_t bytes = (size_t)k * (p->stride + 1u) * 4u;
    CCE(cudaMalloc((void**)&p->outbox, bytes));
    CCE(cudaMalloc((void**)&p->inbox, bytes));
    p->box_cap = k;
}

static size_t box_words(const pgadev_pop* p, uint64_t k) { return (size_t)k * (p->stride + 1u); }

extern "C" void pgadev_reserve_boxes(pgadev_ctx* c, pgadev_pop* p, uint64_t k) { use_device(c); ensure_boxes(c, p, k); }

extern "C" void pgadev_pack_emigrants(pgadev_ctx* c, pgadev_pop* p, uint64_t k)
{
    use_device(c);
    ensure_boxes(c, p, k);
    const uint32_t grid = grid_for(k * (p->stride >> 2), 256, (uint32_t)c->sm_count * 8u);
    pack_rows<<<grid, 256, 0, work_stream(c)>>>(p->genes[p->gidx], p->score[p->sidx], p->idx[0], (uint32_t)k, p->stride, p->outbox, (uint32_t)p->n);
    LAUNCH_CHECK(c);
}

extern "C" void pgadev_transfer_local(pgadev_ctx* c, pgadev_pop* src, pgadev_pop* dst, uint64_t k)
{
    use_device(c);
    CCE(cudaMemcpyAsync(dst->inbox, src->outbox, box_words(src, k) * 4u, cudaMemcpyDeviceToDevice, c->side));
}

extern "C" void pgadev_unpack_immigrants(pgadev_ctx* c, pgadev_pop* p, uint64_t k)
{
    use_device(c);
    const uint32_t grid = grid_for(k * (p->stride >> 2), 256, (uint32_t)c->sm_count * 8u);
    unpack_rows<<<grid, 256, 0, c->stream>>>(p->genes[p->gidx], p->score[p->sidx], p->idx[1], (uint32_t)k, p->stride, p->inbox, (uint32_t)p->n);
    LAUNCH_CHECK(c);

}

extern "C" void pgadev_side_after_main(pgadev_ctx* c)
{
    use_device(c);
    CCE(cudaEventRecord(c->ev_main, c->stream));
    CCE(cudaStreamWaitEvent(c->side, c->ev_main, 0));
}

// the emigrants of this exchange have left the population buffers (side stream)
extern "C" void pgadev_mark_packed(pgadev_ctx* c) { CCE(cudaEventRecord(c->ev_packed, c->side)); }
// the main stream may overwrite those buffers only after that
extern "C" void pgadev_main_wait_packed(pgadev_ctx* c) { CCE(cudaStreamWaitEvent(c->stream, c->ev_packed, 0)); }

extern "C" void pgadev_main_after_side(pgadev_ctx* c)
{
    use_device(c);
    CCE(cudaEventRecord(c->ev_side, c->side));
    CCE(cudaStreamWaitEvent(c->stream, c->ev_side, 0));
}

// ---- NCCL, loaded on demand ----------------------------------------------------------------------
// The reference's link line is "-lpga -lcurand" (test/Makefile:3): nothing may add
// -lnccl to it, so the transport is resolved with dlopen the first time a
// multi-process island layout asks for it.

struct nccl_api {
    void* handle;
    ncclResult_t (*GetUniqueId)(ncclUniqueId*);
    ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int);
    ncclResult_t (*CommDestroy)(ncclComm_t);
    ncclResult_t (*Send)(const void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t);
    ncclResult_t (*Recv)(void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t);
    ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t);
    ncclResult_t (*GroupStart)(void);
    ncclResult_t (*GroupEnd)(void);
    const char* (*GetErrorString)(ncclResult_t);
};
static nccl_api g_nccl;

static int nccl_load(void)
{
    if (g_nccl.handle) return 0;
    const char* names[] = { getenv("PGA_NCCL_LIB"), "libnccl.so.2", "libnccl.so" };
    void* h = NULL;
    for (int i = 0; i < 3 && !h; ++i)
        if (names[i] && *names[i]) h = dlopen(names[i], RTLD_NOW | RTLD_GLOBAL);
    if (!h) return PGAX_ENCCL;
#define SYM(field, name)                                  \
    *(void**)(&g_nccl.field) = dlsym(h, name);            \
    if (!g_nccl.field) { dlclose(h); return PGAX_ENCCL; }
    SYM(GetUniqueId, "ncclGetUniqueId")
    SYM(CommInitRank, "ncclCommInitRank")
    SYM(CommDestroy, "ncclCommDestroy")
    SYM(Send, "ncclSend")
    SYM(Recv, "ncclRecv")
    SYM(AllReduce, "ncclAllReduce")
    SYM(GroupStart, "ncclGroupStart")
    SYM(GroupEnd, "ncclGroupEnd")
    SYM(GetErrorString, "ncclGetErrorString")
#undef SYM
    g_nccl.handle = h;
    return 0;
}

#define NCE(ans) nccl_assert((ans), __FILE__, __LINE__)
static inline void nccl_assert(ncclResult_t r, const char* file, int line)
{
    if (r != ncclSuccess) {
        fprintf(stderr, "GPUassert: NCCL %s %s %d\n", g_nccl.GetErrorString ? g_nccl.GetErrorString(r) : "error", file, line);
        exit((int)r);
    }
}

static void nccl_destroy(pgadev_ctx* c)
{
    if (c->comm && g_nccl.handle) g_nccl.CommDestroy(c->comm);
    c->comm = NULL;
}

extern "C" int pgadev_nccl_unique_id(void* out128)
{
    if (!out128) return PGAX_EINVAL;
    if (nccl_load() != 0) return PGAX_ENCCL;
    ncclUniqueId id;
    static_assert(sizeof(ncclUniqueId) == PGAX_UNIQUE_ID_BYTES, "ncclUniqueId size");
    NCE(g_nccl.GetUniqueId(&id));
    memcpy(out128, &id, sizeof(id));
    return 0;
}

extern "C" int pgadev_nccl_init(pgadev_ctx* c, int rank, int world, const void* id128)
{
    if (!id128) return PGAX_EINVAL;
    if (nccl_load() != 0) return PGAX_ENCCL;
    ncclUniqueId id;
    memcpy(&id, id128, sizeof(id));
    nccl_destroy(c);
    CCE(cudaSetDevice(c->device));
    NCE(g_nccl.CommInitRank(&c->comm, world, id, rank));
    return 0;
}

extern "C" void pgadev_nccl_group_begin(pgadev_ctx* c) { (void)c; NCE(g_nccl.GroupStart()); }
extern "C" void pgadev_nccl_group_end(pgadev_ctx* c) { (void)c; NCE(g_nccl.GroupEnd()); }

extern "C" void pgadev_nccl_send_outbox(pgadev_ctx* c, pgadev_pop* src, uint64_t k, int peer)
{
    NCE(g_nccl.Send(src->outbox, box_words(src, k), ncclUint32, peer, c->comm, c->side));
}

extern "C" void pgadev_nccl_recv_inbox(pgadev_ctx* c, pgadev_pop* dst, uint64_t k, int peer)
{
    NCE(g_nccl.Recv(dst->inbox, box_words(dst, k), ncclUint32, peer, c->comm, c->side));
}

// ---- early stop: best score of all local populations, read back without stalling ------------------
// The generation kernel already reduced each population's best into its ring slot; one
// small launch turns the slots into a float, an optional all-reduce takes the maximum over
// the ranks, and 4 bytes travel to pinned host memory behind an event.  The host looks at
// them one generation later (pgadev_check_wait), so the copy overlaps the next generation.
extern "C" void pgadev_check_begin(pgadev_ctx* c, pgadev_pop** pops, const int* reduced_in_kernel, unsigned count, int all_ranks)
{
    use_device(c);
    if (!c->check_dev) {
        CCE(cudaMalloc((void**)&c->check_dev, sizeof(float)));
        CCE(cudaMallocHost((void**)&c->check_host, sizeof(float)));
    }
    BestSlots slots;
    memset(&slots, 0, sizeof(slots));
    for (unsigned i = 0; i < count && i < 10u; ++i) {
        pgadev_pop* p = pops[i];
        if (p->n == 0) continue;
        if (!reduced_in_kernel[i]) {
            const uint32_t grid = grid_for(p->n, 256 * 8, (uint32_t)c->sm_count * 8u);
            reduce_best<<<grid, 256, 0, c->stream>>>(p->score[p->sidx], (uint32_t)p->n, p->best, p->best_seq & 3u);
            LAUNCH_CHECK(c);
            p->best_seq += 1;
        }
        slots.slot[slots.count++] = p->best + ((p->best_seq - 1u) & 3u);
    }
    best_of_slots<<<1, 32, 0, c->stream>>>(slots, c->check_dev);
    LAUNCH_CHECK(c);
    if (all_ranks && c->comm) NCE(g_nccl.AllReduce(c->check_dev, c->check_dev, 1, ncclFloat, ncclMax, c->comm, c->stream));
    CCE(cudaMemcpyAsync(c->check_host, c->check_dev, sizeof(float), cudaMemcpyDeviceToHost, c->stream));
    c->d2h_bytes += sizeof(float);
    CCE(cudaEventRecord(c->ev_check, c->stream));
}

extern "C" float pgadev_check_wait(pgadev_ctx* c)
{
    CCE(cudaEventSynchronize(c->ev_check));
    return *c->check_host;
}

// ---- ranking without any allocation ----------------------------------------------------------------
//
// pgax_select_top_noalloc: the k best / k worst of a score array, exactly as
// pgax_select_top computes them, for callers that must not (or cannot) reach the
// CUDA allocator while they work:
//   * programs that handed the whole of HBM to their own pool or arena at start-up,
//     where a cudaMalloc from inside a library either fails or eats into memory the
//     owner believes it has;
//   * steady-state loops (one selection per migration, thousands of migrations) that
//     cannot afford cudaMalloc/cudaFree, because both synchronise the whole device
//     and stall every other stream for the duration.
// pgax_select_top sizes its histogram, counters and index list with cudaMalloc on
// first use.  Here the caller brings the input and the output, and the only other
// memory the selection needs -- the radix histogram, the running prefix and the
// per-block counts -- lives in the fixed array below, which is part of the library
// image: every GPU that loads the library carries these 72 KB until the process ends.
//
// The array bounds the problem: at most PGAX_NOALLOC_MAX_INDIVIDUALS scores
// (2^24, i.e. 8192 blocks of 2048).  Larger inputs are refused, not truncated.
//
// `noalloc_scratch` below is a DEVICE object.  In host code its name is only a
// handle for the runtime, not an address a kernel could dereference, so the device
// address is looked up once per device with cudaGetSymbolAddress (with that device
// current) and that pointer is what the kernels receive.
#define NOALLOC_STATE_BYTES 8448u  /* TopkState2 (2048-bin histogram + running state), padded to 256 */
#define NOALLOC_MAX_BLOCKS ((PGAX_NOALLOC_MAX_INDIVIDUALS + TOPK_CHUNK - 1u) / TOPK_CHUNK)
#define NOALLOC_SCRATCH_BYTES (NOALLOC_STATE_BYTES + 2u * NOALLOC_MAX_BLOCKS * 4u)
static_assert(sizeof(TopkState2) <= NOALLOC_STATE_BYTES, "scratch layout");

static __device__ __align__(256) unsigned char noalloc_scratch[NOALLOC_SCRATCH_BYTES];

#define NOALLOC_MAX_DEVICES 64
static unsigned char* g_noalloc_ptr[NOALLOC_MAX_DEVICES];  // device address of the array, per device
static int g_noalloc_sms[NOALLOC_MAX_DEVICES];

extern "C" int pgax_select_top_noalloc(const float* dev_scores, unsigned long n, unsigned long k, int worst,
                                       uint32_t* dev_indices, void* cuda_stream)
{
    if (dev_scores == NULL || dev_indices == NULL) return PGAX_EINVAL;
    if (n > PGAX_NOALLOC_MAX_INDIVIDUALS) {
        fprintf(stderr, "libpga: pgax_select_top_noalloc: %lu scores exceed the fixed scratch (limit %u); use pgax_select_top\n", n,
                (unsigned)PGAX_NOALLOC_MAX_INDIVIDUALS);
        return PGAX_ERANGE;
    }
    if (k > n) k = n;
    if (k == 0) return PGAX_OK;

    int dev = -1;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= NOALLOC_MAX_DEVICES) return PGAX_ECUDA;
    if (g_noalloc_ptr[dev] == NULL) {
        // first call on this device: resolve the array's device address (this is the
        // set-up step; make this first call before any stream capture begins)
        void* ptr = NULL;
        const cudaError_t e = cudaGetSymbolAddress(&ptr, noalloc_scratch);
        if (e != cudaSuccess || ptr == NULL) {
            fprintf(stderr, "libpga: pgax_select_top_noalloc: cudaGetSymbolAddress failed: %s\n", cudaGetErrorString(e));
            return PGAX_ECUDA;
        }
        int sms = 0;
        if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || sms < 1) return PGAX_ECUDA;
        g_noalloc_sms[dev] = sms;
        g_noalloc_ptr[dev] = (unsigned char*)ptr;
    }
    unsigned char* base = g_noalloc_ptr[dev];
    TopkState2* st = reinterpret_cast<TopkState2*>(base);
    uint32_t* blk_gt = reinterpret_cast<uint32_t*>(base + NOALLOC_STATE_BYTES);
    uint32_t* blk_eq = blk_gt + NOALLOC_MAX_BLOCKS;
    cudaStream_t s = (cudaStream_t)cuda_stream;
    const uint32_t n32 = (uint32_t)n, nb = (uint32_t)((n + TOPK_CHUNK - 1) / TOPK_CHUNK);

    // The array still holds whatever the previous call left (histogram, prefix, ticket, the
    // per-chunk counts).  The selection kernels leave the histogram and the ticket clean when
    // they run to completion, but a call that was cut short would not have: the state is
    // cleared explicitly, and the count pass overwrites all nb chunk counts before they are read.
    topk2_reset<<<1, 256, 0, s>>>(st);
    launch_select_5(NULL, dev_scores, n32, (uint32_t)k, worst ? 1 : 0, st, blk_gt, blk_eq, dev_indices, nb,
                    (uint32_t)g_noalloc_sms[dev] * 8u, s);
    return cudaPeekAtLastError() == cudaSuccess ? PGAX_OK : PGAX_ECUDA;
}
