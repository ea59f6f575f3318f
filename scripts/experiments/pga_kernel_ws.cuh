// pga_kernel_ws.cuh -- the fused generation as a warp-specialised pipeline (sm_100a).
//
// What the register-tiled kernel (generation_fused_tiled) cannot do is keep parent rows
// in flight while it computes: a warp alternates between "16 loads outstanding, nothing
// to do" and "800 instructions of Philox, crossover and popcount, nothing outstanding",
// and with 119 registers only 16 warps fit an SM.  ncu on that kernel: no unit above 55 %,
// 4.5 of every 10.6 stall cycles on the long scoreboard, and the same 66-70 ns per 1 000
// children whether the population is L2-resident (2^18 rows) or not (2^22 rows), while the
// bare access pattern needs 39-52 ns (profiles/r2_sweeps.md).
//
// Here the two jobs belong to different warps of one persistent block per SM:
//
//   producer warps   lane j draws Philox block 0 of child j, runs both tournaments (four
//                    score gathers in flight, issued one work item ahead), and fetches the
//                    64 parent rows of the 32 children into a shared-memory stage with
//                    TMA: cp.async.bulk.tensor.2d ... tile::gather4, four independent row
//                    indices per instruction (SASS UTMALDG.2D.GATHER4), completion counted
//                    in bytes on the stage's "full" mbarrier.  It also leaves the
//                    children's mutation draws in the stage.
//   consumer warps   wait on "full", read both parents from shared memory, draw the
//                    crossover mask, mutate, score (popcount / canonical f32 reduction),
//                    store the child rows and scores, reduce the best key, and hand the
//                    stage back through its "empty" mbarrier.
//
// Each consumer warp owns a two-stage ring fed by one producer warp; a producer serves
// CPP consumers round-robin.  Up to 2 * consumers stages (192 KB for 128-byte rows) are in
// flight per SM and no register holds row data while it travels.
//
// Results are identical to every other generation kernel: the same draws, the same
// tournament rule (src/pga.cu:280-292), the same crossover / mutation (:127-143), the same
// canonical score reduction.  Replaces the reference's __g_crossover / __g_mutate /
// __g_evaluate passes (src/pga.cu:250-262, :294-317, :333-347).
#pragma once

#include <cuda.h>

#include "pga_kernels.cuh"

namespace pgak {

__device__ __forceinline__ void mbar_arrive(uint64_t* bar)
{
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_addr(bar)) : "memory");
}

// four rows of a [n][stride] tensor, chosen by index, into 4 consecutive row slots of shared memory
__device__ __forceinline__ void tma_gather4(void* smem_dst, const CUtensorMap* map, int32_t r0, int32_t r1, int32_t r2, int32_t r3, uint64_t* bar)
{
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.tile::gather4.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4, %5, %6}], [%7];" ::"r"(
            smem_addr(smem_dst)),
        "l"(map), "r"(0), "r"(r0), "r"(r1), "r"(r2), "r"(r3), "r"(smem_addr(bar))
        : "memory");
}

// one work item of a producer: the selection of 32 children, gathers still in flight
struct WsSelection {
    uint32_t c0, c1, c2, c3;  // tournament candidates
    float s0, s1, s2, s3;     // their scores
    uint32_t mut;             // bit 31: mutate, bit 30: value (bits), bits 0..29: gene index
    uint32_t val;             // f32 genes: the mutation value (draw 2), as bits
};

template <int ENC>
__device__ __forceinline__ WsSelection ws_select_begin(const GenArgs& a, uint32_t child)
{
    WsSelection d;
    const bool have = child < a.n;
    const uint4 w = draw_block(a.draw, STREAM_ROW, child, 0);
    const float u0 = to_uniform(w.x), u1 = to_uniform(w.y), u2 = to_uniform(w.z), u3 = to_uniform(w.w);
    d.c0 = have ? draw_index(u0, a.nf, a.n) : 0u;
    d.c1 = have ? draw_index(u1, a.nf, a.n) : 0u;
    d.c2 = have ? draw_index(u2, a.nf, a.n) : 0u;
    d.c3 = have ? draw_index(u3, a.nf, a.n) : 0u;
    d.s0 = __ldg(a.score_in + d.c0);
    d.s1 = __ldg(a.score_in + d.c1);
    d.s2 = __ldg(a.score_in + d.c2);
    d.s3 = __ldg(a.score_in + d.c3);
    const bool m_on = have && u1 <= PGA_MUTATION_CHANCE;  // src/pga.cu:127-133
    const uint32_t m_idx = draw_index(u0, a.lenf, a.len);
    const uint32_t m_val = ENC == ENC_BITS ? (u2 > 0.5f ? 1u : 0u) : 0u;
    d.mut = (m_on ? 0x80000000u : 0u) | (m_val << 30) | m_idx;
    d.val = ENC == ENC_F32 ? __float_as_uint(u2) : 0u;
    return d;
}

// Shared-memory layout (dynamic): per consumer warp, 2 stages of STAGE bytes, then per
// consumer 2 x 32 mutation words (+ 2 x 32 mutation values for f32 genes), then the barriers.
template <int ENC, int OBJ, int LOGG, int NPROD, int CPP>
__global__ void __launch_bounds__((NPROD + NPROD * CPP) * 32, 1)
    generation_fused_ws(const __grid_constant__ GenArgs a, const __grid_constant__ CUtensorMap cur_map)
{
    constexpr uint32_t G = 1u << LOGG, CPT = 32u / G, ROWB = G * 16u, STAGE = 64u * ROWB;
    constexpr uint32_t NCONS = NPROD * CPP, NWARPS = NPROD + NCONS;
    static_assert(LOGG >= 1 && LOGG <= 3, "rows of 32 .. 128 bytes");
    extern __shared__ __align__(128) unsigned char smem_raw[];
    __shared__ unsigned long long warp_best[NWARPS];
    __shared__ __align__(8) uint64_t full_bar[NCONS][2];
    __shared__ __align__(8) uint64_t empty_bar[NCONS][2];
    __shared__ uint32_t meta[NCONS][2][32];
    __shared__ uint32_t meta_val[ENC == ENC_F32 ? NCONS : 1][2][32];

    const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    const uint32_t total = (a.n + 31u) >> 5;        // groups of 32 children
    const uint32_t stride = gridDim.x * NCONS;      // consumers in the grid
    // TMA destinations need 128-byte alignment; dynamic shared memory only promises 16
    unsigned char* const stages = smem_raw + ((128u - (smem_addr(smem_raw) & 127u)) & 127u);

    if (threadIdx.x == 0) {
        for (uint32_t c = 0; c < NCONS; ++c)
            for (uint32_t s = 0; s < 2; ++s) {
                mbar_init(&full_bar[c][s], 1);
                mbar_init(&empty_bar[c][s], 1);
            }
        if (blockIdx.x == 0) a.best[(a.best_slot + 1u) & 3u] = 0ull;
    }
    __syncthreads();

    unsigned long long key = 0ull;

    if (warp < NPROD) {
        // ---------------- producer ----------------
        // work items in round-robin order over this producer's consumers: item i belongs to
        // consumer (i % CPP), its k-th group, k = i / CPP
        const uint32_t cons0 = warp * CPP;
        auto item_group = [&](uint32_t i, uint32_t& cons, uint32_t& k) -> uint32_t {
            cons = cons0 + i % CPP;
            k = i / CPP;
            return blockIdx.x * NCONS + cons + k * stride;  // group index, may be >= total
        };
        // number of items: until every consumer has run out of groups
        uint32_t first = blockIdx.x * NCONS + cons0;
        const uint32_t kmax = first < total ? (total - first + stride - 1u) / stride : 0u;  // groups of the first consumer (the most)
        const uint32_t items = kmax * CPP;

        uint32_t cons = 0, k = 0;
        uint32_t g = items ? item_group(0, cons, k) : 0u;
        WsSelection nx;
        if (items && g < total) nx = ws_select_begin<ENC>(a, g * 32u + lane);
        for (uint32_t i = 0; i < items; ++i) {
            const uint32_t g_now = g, cons_now = cons, k_now = k;
            const WsSelection cur = nx;
            if (i + 1u < items) {
                g = item_group(i + 1u, cons, k);
                if (g < total) nx = ws_select_begin<ENC>(a, g * 32u + lane);  // its gathers fly while this item is finished
            }
            if (g_now >= total) continue;  // this consumer has no k-th group
            const uint32_t child = g_now * 32u + lane;
            const uint32_t p1 = cur.s0 < cur.s1 ? cur.c1 : cur.c0;  // src/pga.cu:286-290
            const uint32_t p2 = cur.s2 < cur.s3 ? cur.c3 : cur.c2;
            if (a.sel && child < a.n) { a.sel[2 * (size_t)child] = (int32_t)p1; a.sel[2 * (size_t)child + 1] = (int32_t)p2; }
            const uint32_t s = k_now & 1u;
            mbar_wait(&empty_bar[cons_now][s], ((k_now >> 1) & 1u) ^ 1u);
            meta[cons_now][s][lane] = cur.mut;
            if (ENC == ENC_F32) meta_val[ENC == ENC_F32 ? cons_now : 0][s][lane] = cur.val;
            __syncwarp();
            if (lane == 0) mbar_arrive_expect_tx(&full_bar[cons_now][s], STAGE);
            __syncwarp();
            // lane l < 16: both parents of children 2l and 2l+1 = row slots 4l .. 4l+3 of the stage
            const uint32_t q0 = __shfl_sync(0xFFFFFFFFu, p1, (2u * lane) & 31u), q1 = __shfl_sync(0xFFFFFFFFu, p2, (2u * lane) & 31u);
            const uint32_t q2 = __shfl_sync(0xFFFFFFFFu, p1, (2u * lane + 1u) & 31u), q3 = __shfl_sync(0xFFFFFFFFu, p2, (2u * lane + 1u) & 31u);
            if (lane < 16u) {
                unsigned char* st = stages + ((size_t)cons_now * 2u + s) * STAGE;
                tma_gather4(st + lane * 4u * ROWB, &cur_map, (int32_t)q0, (int32_t)q1, (int32_t)q2, (int32_t)q3, &full_bar[cons_now][s]);
            }
        }
    } else {
        // ---------------- consumer ----------------
        const uint32_t cons = warp - NPROD;
        const uint32_t gl = lane & (G - 1u), grp = lane >> LOGG;
        uint4* __restrict__ nxt = reinterpret_cast<uint4*>(a.nxt);
        const uint32_t first = blockIdx.x * NCONS + cons;
        for (uint32_t k = 0, g = first; g < total; ++k, g += stride) {
            const uint32_t s = k & 1u, cbase = g * 32u;
            const unsigned char* st = stages + ((size_t)cons * 2u + s) * STAGE;
            mbar_wait(&full_bar[cons][s], (k >> 1) & 1u);
            const uint32_t my_mut = meta[cons][s][lane];
            const uint32_t my_val = ENC == ENC_F32 ? meta_val[ENC == ENC_F32 ? cons : 0][s][lane] : 0u;
            const uint32_t any_mut = __ballot_sync(0xFFFFFFFFu, (my_mut >> 31) != 0u);
            float my_score = 0.f;
#pragma unroll
            for (uint32_t tile = 0; tile < G; ++tile) {
                const uint32_t c = tile * CPT + grp;  // child slot 0..31 of the group
                const uint32_t child = cbase + c;
                const uint4 xx = *reinterpret_cast<const uint4*>(st + (2u * c + 0u) * ROWB + gl * 16u);
                const uint4 yy = *reinterpret_cast<const uint4*>(st + (2u * c + 1u) * ROWB + gl * 16u);
                uint4 o;
                if (ENC == ENC_BITS) {
                    const uint4 m = draw_block(a.draw, STREAM_MASK, child, gl);
                    o.x = (xx.x & m.x) | (yy.x & ~m.x);
                    o.y = (xx.y & m.y) | (yy.y & ~m.y);
                    o.z = (xx.z & m.z) | (yy.z & ~m.z);
                    o.w = (xx.w & m.w) | (yy.w & ~m.w);
                } else {
                    const uint4 r = draw_block(a.draw, STREAM_ROW, child, gl);
                    o.x = draw_above_half(r.x) ? xx.x : yy.x;  // src/pga.cu:137-141
                    o.y = draw_above_half(r.y) ? xx.y : yy.y;
                    o.z = draw_above_half(r.z) ? xx.z : yy.z;
                    o.w = draw_above_half(r.w) ? xx.w : yy.w;
                }
                if (any_mut) {  // warp-uniform; 72 % of groups have no mutation at all
                    const uint32_t cm = __shfl_sync(0xFFFFFFFFu, my_mut, c);
                    const uint32_t cv = ENC == ENC_F32 ? __shfl_sync(0xFFFFFFFFu, my_val, c) : ((cm >> 30) & 1u);
                    const uint32_t c_idx = cm & 0x3FFFFFFFu;
                    const uint32_t c_blk = ENC == ENC_BITS ? (c_idx >> 7) : (c_idx >> 2);
                    if ((cm >> 31) && c_blk == gl) {  // src/pga.cu:127-133
                        if (ENC == ENC_BITS) {
                            const uint32_t word = (c_idx >> 5) & 3u, bit = 1u << (c_idx & 31u);
                            uint32_t* cw = word == 0 ? &o.x : word == 1 ? &o.y : word == 2 ? &o.z : &o.w;
                            *cw = cv ? (*cw | bit) : (*cw & ~bit);
                        } else {
                            const uint32_t e = c_idx & 3u;
                            if (e == 0) o.x = cv; else if (e == 1) o.y = cv; else if (e == 2) o.z = cv; else o.w = cv;
                        }
                    }
                }
                float sc;
                if (ENC == ENC_BITS) {
                    uint32_t iacc = __popc(o.x) + __popc(o.y) + __popc(o.z) + __popc(o.w);
#pragma unroll
                    for (uint32_t off = G >> 1; off >= 1u; off >>= 1) iacc += __shfl_xor_sync(0xFFFFFFFFu, iacc, off);
                    sc = __uint2float_rn(iacc);
                } else {
                    float facc = 0.f;
                    facc = obj_accumulate<OBJ>(facc, __uint_as_float(o.x));
                    facc = obj_accumulate<OBJ>(facc, __uint_as_float(o.y));
                    facc = obj_accumulate<OBJ>(facc, __uint_as_float(o.z));
                    facc = obj_accumulate<OBJ>(facc, __uint_as_float(o.w));
#pragma unroll
                    for (uint32_t off = G >> 1; off >= 1u; off >>= 1) facc = __fadd_rn(facc, __shfl_xor_sync(0xFFFFFFFFu, facc, off));
                    sc = obj_finalize<OBJ>(facc, a.len);
                }
                if (child < a.n) store_row_block(nxt + (size_t)child * G + gl, o);
                const float v = __shfl_sync(0xFFFFFFFFu, sc, (lane & (CPT - 1u)) << LOGG);
                if ((lane >> (5 - LOGG)) == tile) my_score = v;
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(&empty_bar[cons][s]);  // the stage may be refilled
            const uint32_t mine = cbase + lane;
            if (mine < a.n) {
                a.score_out[mine] = my_score;
                const unsigned long long kk = best_key(my_score, mine);
                key = kk > key ? kk : key;
            }
        }
    }
    block_best_commit(key, a.best + a.best_slot, warp_best);
}

// ---- the same pipeline inside every warp ------------------------------------------------------
//
// Measured on B200 (profiles/r2_sweeps.md): the producer/consumer split above is correct
// but slower than the register-tiled kernel (77.9 against 72.0 us per OneMax generation):
// a stage that is being consumed, or that waits for its producer to come round, holds
// shared memory without holding a request in flight, and a producer serialises the
// selections of three consumers.  generation_fused_pipe gives every warp its own two
// stages and lets it run three iterations at once:
//     iteration i+2   Philox block 0 drawn, the four score gathers issued
//     iteration i+1   tournaments resolved, both parents of its 32 children requested with
//                     sixteen tile::gather4 TMA instructions into the warp's other stage
//     iteration i     the eight crossover masks drawn FIRST (they depend on nothing that
//                     is in flight), then the warp waits for its stage, combines, mutates,
//                     scores and stores
// so the rows of iteration i+1 travel while the 550 Philox instructions of iteration i
// issue, and nothing in flight lives in a register.
template <int ENC, int OBJ, int LOGG, int WARPS>
__global__ void __launch_bounds__(WARPS * 32, 1)
    generation_fused_pipe(const __grid_constant__ GenArgs a, const __grid_constant__ CUtensorMap cur_map)
{
    constexpr uint32_t G = 1u << LOGG, CPT = 32u / G, ROWB = G * 16u, STAGE = 64u * ROWB;
    static_assert(LOGG >= 1 && LOGG <= 3, "rows of 32 .. 128 bytes");
    // The kernel has no static shared memory: dynamic shared memory starts right behind the
    // static arrays (16-byte aligned whatever the declaration asks for), and the stages are TMA
    // destinations.  With nothing in front of it the buffer starts at offset 0 of the block's
    // window; the barriers and the best-key scratch sit behind the stages.
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    uint64_t (*full_bar)[2] = reinterpret_cast<uint64_t (*)[2]>(smem_raw + (size_t)WARPS * 2u * STAGE);
    unsigned long long* warp_best = reinterpret_cast<unsigned long long*>(smem_raw + (size_t)WARPS * 2u * STAGE + (size_t)WARPS * 16u);

    const uint32_t lane = threadIdx.x & 31u, w = threadIdx.x >> 5;
    const uint32_t gl = lane & (G - 1u), grp = lane >> LOGG;
    unsigned char* const my = smem_raw + (size_t)w * 2u * STAGE;
    uint4* __restrict__ nxt = reinterpret_cast<uint4*>(a.nxt);

    if (lane == 0) {
        mbar_init(&full_bar[w][0], 1);
        mbar_init(&full_bar[w][1], 1);
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) a.best[(a.best_slot + 1u) & 3u] = 0ull;
    __syncwarp();

    const uint32_t total = (a.n + 31u) >> 5;  // groups of 32 children
    const uint32_t gwarp = blockIdx.x * WARPS + w, nwarps = gridDim.x * WARPS;
    const uint32_t iters = gwarp < total ? (total - gwarp + nwarps - 1u) / nwarps : 0u;
    unsigned long long key = 0ull;

    // request both parents of the 32 children of iteration `it` into stage it & 1
    auto issue_rows = [&](const WsSelection& d, uint32_t it) {
        const uint32_t child = (gwarp + it * nwarps) * 32u + lane;
        const uint32_t p1 = d.s0 < d.s1 ? d.c1 : d.c0;  // src/pga.cu:286-290
        const uint32_t p2 = d.s2 < d.s3 ? d.c3 : d.c2;
        if (a.sel && child < a.n) { a.sel[2 * (size_t)child] = (int32_t)p1; a.sel[2 * (size_t)child + 1] = (int32_t)p2; }
        const uint32_t s = it & 1u;
        // the stage was last read by this warp's generic-proxy loads; the copies below write it through the async proxy
        fence_async_smem();
        __syncwarp();
        if (lane == 0) mbar_arrive_expect_tx(&full_bar[w][s], STAGE);
        __syncwarp();
        const uint32_t q0 = __shfl_sync(0xFFFFFFFFu, p1, (2u * lane) & 31u), q1 = __shfl_sync(0xFFFFFFFFu, p2, (2u * lane) & 31u);
        const uint32_t q2 = __shfl_sync(0xFFFFFFFFu, p1, (2u * lane + 1u) & 31u), q3 = __shfl_sync(0xFFFFFFFFu, p2, (2u * lane + 1u) & 31u);
        // lane l < 16: both parents of children 2l and 2l+1 = row slots 4l .. 4l+3 of the stage
        if (lane < 16u) tma_gather4(my + s * STAGE + lane * 4u * ROWB, &cur_map, (int32_t)q0, (int32_t)q1, (int32_t)q2, (int32_t)q3, &full_bar[w][s]);
    };

    WsSelection nx;          // selection whose gathers are in flight
    uint32_t mut0 = 0, val0 = 0, mut1 = 0, val1 = 0;  // mutation draws of iteration it / it+1
    if (iters > 0) {
        nx = ws_select_begin<ENC>(a, gwarp * 32u + lane);
        mut0 = nx.mut; val0 = nx.val;
        issue_rows(nx, 0);
        if (iters > 1) nx = ws_select_begin<ENC>(a, (gwarp + nwarps) * 32u + lane);
    }

    for (uint32_t it = 0; it < iters; ++it) {
        const uint32_t cbase = (gwarp + it * nwarps) * 32u;
        if (it + 1u < iters) {
            mut1 = nx.mut; val1 = nx.val;
            issue_rows(nx, it + 1u);
            if (it + 2u < iters) nx = ws_select_begin<ENC>(a, (gwarp + (it + 2u) * nwarps) * 32u + lane);
        }
        // the masks of this iteration, while its rows (and the next iteration's) travel
        uint4 m[G];
#pragma unroll
        for (uint32_t tile = 0; tile < G; ++tile)
            m[tile] = draw_block(a.draw, ENC == ENC_BITS ? STREAM_MASK : STREAM_ROW, cbase + tile * CPT + grp, gl);
        // pin the masks in front of the wait (the optimiser would otherwise sink them to their uses behind it)
#pragma unroll
        for (uint32_t tile = 0; tile < G; ++tile) asm volatile("" ::"r"(m[tile].x), "r"(m[tile].y), "r"(m[tile].z), "r"(m[tile].w));
        const uint32_t any_mut = __ballot_sync(0xFFFFFFFFu, (mut0 >> 31) != 0u);

        const uint32_t s = it & 1u;
        const unsigned char* st = my + s * STAGE;
        mbar_wait(&full_bar[w][s], (it >> 1) & 1u);
        float my_score = 0.f;
#pragma unroll
        for (uint32_t tile = 0; tile < G; ++tile) {
            const uint32_t c = tile * CPT + grp;  // child slot 0..31 of the group
            const uint32_t child = cbase + c;
            const uint4 xx = *reinterpret_cast<const uint4*>(st + (2u * c + 0u) * ROWB + gl * 16u);
            const uint4 yy = *reinterpret_cast<const uint4*>(st + (2u * c + 1u) * ROWB + gl * 16u);
            uint4 o;
            if (ENC == ENC_BITS) {
                o.x = (xx.x & m[tile].x) | (yy.x & ~m[tile].x);
                o.y = (xx.y & m[tile].y) | (yy.y & ~m[tile].y);
                o.z = (xx.z & m[tile].z) | (yy.z & ~m[tile].z);
                o.w = (xx.w & m[tile].w) | (yy.w & ~m[tile].w);
            } else {
                o.x = draw_above_half(m[tile].x) ? xx.x : yy.x;  // src/pga.cu:137-141
                o.y = draw_above_half(m[tile].y) ? xx.y : yy.y;
                o.z = draw_above_half(m[tile].z) ? xx.z : yy.z;
                o.w = draw_above_half(m[tile].w) ? xx.w : yy.w;
            }
            if (any_mut) {  // warp-uniform; 72 % of groups have no mutation at all
                const uint32_t cm = __shfl_sync(0xFFFFFFFFu, mut0, c);
                const uint32_t cv = ENC == ENC_F32 ? __shfl_sync(0xFFFFFFFFu, val0, c) : ((cm >> 30) & 1u);
                const uint32_t c_idx = cm & 0x3FFFFFFFu;
                const uint32_t c_blk = ENC == ENC_BITS ? (c_idx >> 7) : (c_idx >> 2);
                if ((cm >> 31) && c_blk == gl) {  // src/pga.cu:127-133
                    if (ENC == ENC_BITS) {
                        const uint32_t word = (c_idx >> 5) & 3u, bit = 1u << (c_idx & 31u);
                        o.x = word == 0 ? (cv ? (o.x | bit) : (o.x & ~bit)) : o.x;
                        o.y = word == 1 ? (cv ? (o.y | bit) : (o.y & ~bit)) : o.y;
                        o.z = word == 2 ? (cv ? (o.z | bit) : (o.z & ~bit)) : o.z;
                        o.w = word == 3 ? (cv ? (o.w | bit) : (o.w & ~bit)) : o.w;
                    } else {
                        const uint32_t e = c_idx & 3u;
                        if (e == 0) o.x = cv; else if (e == 1) o.y = cv; else if (e == 2) o.z = cv; else o.w = cv;
                    }
                }
            }
            float sc;
            if (ENC == ENC_BITS) {
                uint32_t iacc = __popc(o.x) + __popc(o.y) + __popc(o.z) + __popc(o.w);
#pragma unroll
                for (uint32_t off = G >> 1; off >= 1u; off >>= 1) iacc += __shfl_xor_sync(0xFFFFFFFFu, iacc, off);
                sc = __uint2float_rn(iacc);
            } else {
                float facc = 0.f;
                facc = obj_accumulate<OBJ>(facc, __uint_as_float(o.x));
                facc = obj_accumulate<OBJ>(facc, __uint_as_float(o.y));
                facc = obj_accumulate<OBJ>(facc, __uint_as_float(o.z));
                facc = obj_accumulate<OBJ>(facc, __uint_as_float(o.w));
#pragma unroll
                for (uint32_t off = G >> 1; off >= 1u; off >>= 1) facc = __fadd_rn(facc, __shfl_xor_sync(0xFFFFFFFFu, facc, off));
                sc = obj_finalize<OBJ>(facc, a.len);
            }
            if (child < a.n) store_row_block(nxt + (size_t)child * G + gl, o);
            const float v = __shfl_sync(0xFFFFFFFFu, sc, (lane & (CPT - 1u)) << LOGG);
            if ((lane >> (5 - LOGG)) == tile) my_score = v;
        }
        const uint32_t mine = cbase + lane;
        if (mine < a.n) {
            a.score_out[mine] = my_score;
            const unsigned long long kk = best_key(my_score, mine);
            key = kk > key ? kk : key;
        }
        mut0 = mut1; val0 = val1;
    }
    block_best_commit(key, a.best + a.best_slot, warp_best);
}

}  // namespace pgak
