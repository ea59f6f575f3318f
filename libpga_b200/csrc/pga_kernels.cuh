// pga_kernels.cuh -- the generation kernels of the B200-native libpga engine.
//
// One launch = one pass over a population.  With all three phases enabled it is
// the fused generation: tournament selection, crossover, mutation, scoring of the
// child and the best-of-population reduction, reading each parent row once and
// writing each child row once.  It replaces the reference's four kernels and its
// cuRAND pass (src/pga.cu:81-86, :99-101, :250-262, :294-317, :333-347) and the
// host argmax of pga_get_best (src/pga.cu:218-236).
//
// Two families:
//   generation_hooks  - user hooks are device function pointers with a serial,
//                       one-thread-per-individual ABI (include/pga.h:46-48); rows are
//                       staged in shared memory with TMA bulk copies and handed to
//                       the hooks there.
//   generation_fast   - default operators + built-in objective, no indirect calls:
//                       a group of lanes per child, 128-bit coalesced row accesses,
//                       everything in registers.
#pragma once

#include "pga_device_utils.cuh"

namespace pgak {

typedef float (*obj_fn)(float*, unsigned);
typedef void (*mutate_fn)(float*, float*, unsigned);
typedef void (*crossover_fn)(float*, float*, float*, float*, unsigned);

enum : uint32_t { PH_CROSSOVER = 1u, PH_MUTATE = 2u, PH_EVALUATE = 4u, PH_FUSED = 7u };
enum : int { ENC_F32 = 0, ENC_BITS = 1 };

#define PGA_MUTATION_CHANCE 0.01f /* src/pga.cu:128 */

struct GenArgs {
    const uint32_t* cur;    // [n][stride] current generation
    uint32_t* nxt;          // [n][stride] next generation
    const float* score_in;  // scores of cur (selection)
    float* score_out;       // scores written by the EVALUATE phase
    unsigned long long* best;  // ring of 4 best-keys
    uint32_t best_slot;        // slot this launch reduces into
    int32_t* sel;              // optional [n][2] capture of the tournament winners
    uint32_t n;
    float nf;        // (float)n, the multiplier of the index draw
    uint32_t len;    // genes per genome (bits for ENC_BITS)
    float lenf;      // (float)len
    uint32_t stride; // 32-bit words per stored row, multiple of 4
    uint32_t group;  // fast path: lanes per child (power of two <= 32)
    uint32_t tk;     // tournament size (candidates per parent); 2 = the reference's TOURNAMENT
    DrawCtx draw;
    int objective;
    obj_fn obj;
    mutate_fn mut;
    crossover_fn xov;
};

// ---- shared pieces ----------------------------------------------------------

// size-2 tournament (src/pga.cu:280-292): the second candidate wins only if
// strictly better, so ties and NaNs keep the first
__device__ __forceinline__ uint32_t tournament(const float* __restrict__ score, float ua, float ub, float nf, uint32_t n)
{
    const uint32_t a = draw_index(ua, nf, n), b = draw_index(ub, nf, n);
    PGA_ASSERT(a < n && b < n);
    const float sa = __ldg(score + a), sb = __ldg(score + b);
    return sa < sb ? b : a;
}

// Tournament of K candidates on draws first .. first+K-1 of the child's rand row:
// src/pga.cu:280-292 with TOURNAMENT_POPULATION = K (a later candidate wins only if
// strictly better than the best so far).  The reference hard-wires K = 2; other sizes are
// the selection schemes of include/pga_b200.h (PGAX_TOURNAMENT_OF).
__device__ __forceinline__ uint32_t tournament_of(const DrawCtx& d, const float* __restrict__ score, uint32_t child, uint32_t first,
                                                  uint32_t K, float nf, uint32_t n)
{
    uint32_t best = 0, have = 0xFFFFFFFFu;
    float sbest = 0.f;
    uint4 w = make_uint4(0, 0, 0, 0);
    for (uint32_t i = 0; i < K; ++i) {
        const uint32_t j = first + i;
        if ((j >> 2) != have) { have = j >> 2; w = draw_block(d, STREAM_ROW, child, have); }
        const uint32_t x = (j & 3u) == 0 ? w.x : (j & 3u) == 1 ? w.y : (j & 3u) == 2 ? w.z : w.w;
        const uint32_t c = draw_index(to_uniform(x), nf, n);
        PGA_ASSERT(c < n);
        const float sc = __ldg(score + c);
        if (i == 0 || sbest < sc) { best = c; sbest = sc; }
    }
    return best;
}

// per-block best: warp shuffles, then one atomic per block
__device__ __forceinline__ void block_best_commit(unsigned long long key, unsigned long long* slot,
                                                  unsigned long long* warp_scratch)
{
    key = warp_max_u64(key);
    const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5, nwarps = (blockDim.x + 31u) >> 5;
    if (lane == 0) warp_scratch[warp] = key;
    __syncthreads();
    if (warp == 0) {
        unsigned long long k = lane < nwarps ? warp_scratch[lane] : 0ull;
        k = warp_max_u64(k);
        if (lane == 0 && k != 0ull) atomicMax(slot, k);
    }
}

// ---- hook path ----------------------------------------------------------------
//
// Block of T threads <-> T consecutive children.  Shared memory holds, per thread,
// the child row, plus the rand row when a hook wants one, plus -- when the user
// supplied a crossover hook, which must be handed two contiguous parent rows -- both
// parents.  Those parent rows arrive by one TMA bulk copy each, issued by the thread
// that ran the tournament.  With the default crossover the parents are never staged:
// each warp reads them straight from global memory with coalesced 128-bit loads while
// it builds its 32 children, which leaves four times as many threads resident.  The
// block's child rows leave by ONE bulk store, because consecutive children are
// contiguous in next_gen.
template <uint32_t PH>
__global__ void __launch_bounds__(128) generation_hooks(const __grid_constant__ GenArgs a)
{
    extern __shared__ __align__(128) unsigned char smem_raw[];
    __shared__ unsigned long long warp_best[4];
    uint64_t* bar = reinterpret_cast<uint64_t*>(smem_raw);
    float* rows = reinterpret_cast<float*>(smem_raw + 128);

    const uint32_t T = blockDim.x, t = threadIdx.x, lane = t & 31u, warp = t >> 5;
    const uint32_t S = a.stride;
    const bool want_rand = ((PH & PH_CROSSOVER) && a.xov != nullptr) || ((PH & PH_MUTATE) && a.mut != nullptr);

    const bool stage_parents = (PH & PH_CROSSOVER) && a.xov != nullptr;
    float* s_child = rows;
    float* s_p1 = s_child + (size_t)T * S;
    float* s_p2 = stage_parents ? s_p1 + (size_t)T * S : s_p1;
    float* s_rnd = stage_parents ? s_p2 + (size_t)T * S : s_p1;

    const uint32_t base = blockIdx.x * T;
    const uint32_t nact = min(T, a.n - base);
    const bool active = t < nact;
    const uint32_t i = base + t;
    const uint32_t row_bytes = S * 4u;

    if (t == 0) {
        mbar_init(bar, 1);
        if ((PH & PH_EVALUATE) && blockIdx.x == 0) a.best[(a.best_slot + 1u) & 3u] = 0ull;
    }
    __syncthreads();

    // draws 0..3 of the child's rand row: selection and the default mutation
    float u0 = 0.f, u1 = 1.f, u2 = 0.f, u3 = 0.f;
    if (active && (PH & (PH_CROSSOVER | PH_MUTATE))) {
        const uint4 w = draw_block(a.draw, STREAM_ROW, i, 0);
        u0 = to_uniform(w.x); u1 = to_uniform(w.y); u2 = to_uniform(w.z); u3 = to_uniform(w.w);
    }

    uint32_t p1 = 0, p2 = 0;
    if (PH & PH_CROSSOVER) {
        if (stage_parents && t == 0) {
            // a crossover hook gets what the reference hands it (src/pga.cu:302): the child
            // row starts as the row next_gen holds at that index (the generation before
            // last), so a hook that leaves genes unwritten yields the same child here
            mbar_arrive_expect_tx(bar, nact * 3u * row_bytes);
            bulk_load(s_child, a.nxt + (size_t)base * S, nact * row_bytes, bar);
        }
        if (active) {
            if (a.tk == 2u) {
                p1 = tournament(a.score_in, u0, u1, a.nf, a.n);  // src/pga.cu:306
                p2 = tournament(a.score_in, u2, u3, a.nf, a.n);  // src/pga.cu:307
            } else {
                p1 = tournament_of(a.draw, a.score_in, i, 0u, a.tk, a.nf, a.n);
                p2 = tournament_of(a.draw, a.score_in, i, a.tk, a.tk, a.nf, a.n);
            }
            if (a.sel) { a.sel[2 * (size_t)i] = (int32_t)p1; a.sel[2 * (size_t)i + 1] = (int32_t)p2; }
            if (stage_parents) {
                bulk_load(s_p1 + (size_t)t * S, a.cur + (size_t)p1 * S, row_bytes, bar);
                bulk_load(s_p2 + (size_t)t * S, a.cur + (size_t)p2 * S, row_bytes, bar);
            }
        }
    } else {
        // the block's own rows are contiguous: one copy
        const uint32_t* src = (PH & PH_MUTATE) ? a.nxt : a.cur;
        if (t == 0) {
            mbar_arrive_expect_tx(bar, nact * row_bytes);
            bulk_load(s_child, src + (size_t)base * S, nact * row_bytes, bar);
        }
    }

    // rand rows for the hooks, generated while the copies are in flight; each warp
    // serves its own 32 children, lanes across Philox blocks
    if (want_rand) {
        for (uint32_t c = 0; c < 32u; ++c) {
            const uint32_t tt = warp * 32u + c;
            if (tt >= nact) break;
            for (uint32_t b = lane; b < (S >> 2); b += 32u) {
                const uint4 w = draw_block(a.draw, STREAM_ROW, base + tt, b);
                *reinterpret_cast<float4*>(s_rnd + (size_t)tt * S + 4u * b) =
                    make_float4(to_uniform(w.x), to_uniform(w.y), to_uniform(w.z), to_uniform(w.w));
            }
        }
    }

    if (!(PH & PH_CROSSOVER) || stage_parents) mbar_wait(bar, 0);
    __syncwarp();

    if (PH & PH_CROSSOVER) {
        if (a.xov != nullptr) {
            if (active) {
                float* c = s_child + (size_t)t * S;
                a.xov(s_p1 + (size_t)t * S, s_p2 + (size_t)t * S, c, s_rnd + (size_t)t * S, a.len);  // src/pga.cu:305
                for (uint32_t j = a.len; j < S; ++j) c[j] = 0.f;
            }
        } else {
            // default uniform crossover (src/pga.cu:135-143), cooperative per warp: four
            // children at a time, lanes across 16-byte blocks, parents read from global
            const uint4* __restrict__ cur4 = reinterpret_cast<const uint4*>(a.cur);
            const uint32_t nblk = S >> 2;
            for (uint32_t c0 = 0; c0 < 32u; c0 += 4u) {
                if (warp * 32u + c0 >= nact) break;  // warp-uniform
                uint32_t q1[4], q2[4];
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    q1[u] = __shfl_sync(0xFFFFFFFFu, p1, c0 + u);
                    q2[u] = __shfl_sync(0xFFFFFFFFu, p2, c0 + u);
                }
                for (uint32_t b = lane; b < nblk; b += 32u) {
                    uint4 x[4], y[4];
#pragma unroll
                    for (int u = 0; u < 4; ++u) {
                        x[u] = load_row_block<false>(cur4 + (size_t)q1[u] * nblk + b);
                        y[u] = load_row_block<false>(cur4 + (size_t)q2[u] * nblk + b);
                    }
#pragma unroll
                    for (int u = 0; u < 4; ++u) {
                        const uint32_t tt = warp * 32u + c0 + u;
                        if (tt < nact) {
                            uint4 w;
                            if (want_rand) {
                                const float4 r = *reinterpret_cast<const float4*>(s_rnd + (size_t)tt * S + 4u * b);
                                // the rand row holds the uniforms themselves: compare as floats
                                w.x = r.x > 0.5f ? 0xFFFFFFFFu : 0u; w.y = r.y > 0.5f ? 0xFFFFFFFFu : 0u;
                                w.z = r.z > 0.5f ? 0xFFFFFFFFu : 0u; w.w = r.w > 0.5f ? 0xFFFFFFFFu : 0u;
                            } else {
                                w = draw_block(a.draw, STREAM_ROW, base + tt, b);
                            }
                            uint4 o;
                            o.x = draw_above_half(w.x) ? x[u].x : y[u].x;
                            o.y = draw_above_half(w.y) ? x[u].y : y[u].y;
                            o.z = draw_above_half(w.z) ? x[u].z : y[u].z;
                            o.w = draw_above_half(w.w) ? x[u].w : y[u].w;
                            *reinterpret_cast<uint4*>(s_child + (size_t)tt * S + 4u * b) = o;
                        }
                    }
                }
            }
        }
        __syncwarp();
    }

    if (PH & PH_MUTATE) {
        if (active) {
            float* c = s_child + (size_t)t * S;
            if (a.mut != nullptr) {
                a.mut(c, s_rnd + (size_t)t * S, a.len);  // src/pga.cu:341
            } else if (u1 <= PGA_MUTATION_CHANCE) {      // src/pga.cu:127-133
                PGA_ASSERT(draw_index(u0, a.lenf, a.len) < a.len);
                c[draw_index(u0, a.lenf, a.len)] = u2;
            }
        }
        __syncwarp();
    }

    if (PH & PH_EVALUATE) {
        unsigned long long key = 0ull;
        if (active) {
            float* c = s_child + (size_t)t * S;
            const float s = a.objective == OBJ_HOOK ? a.obj(c, a.len) : objective_serial(a.objective, c, a.len);  // src/pga.cu:258
            a.score_out[i] = s;
            key = best_key(s, i);
        }
        block_best_commit(key, a.best + a.best_slot, warp_best);
    }

    if (PH & (PH_CROSSOVER | PH_MUTATE)) {
        fence_async_smem();
        __syncthreads();
        if (t == 0) {
            bulk_store(a.nxt + (size_t)base * S, s_child, nact * row_bytes);
            bulk_store_commit_and_wait();
        }
    }
}

// ---- fast path ------------------------------------------------------------------
//
// A warp takes 32 consecutive children per iteration.  Lane j runs the
// tournaments of child base+j (four independent score gathers in flight per
// lane).  The rows are then processed in G tiles of 32/G children: a group of G
// lanes owns one child and walks its 16-byte blocks gl, gl+G, ... with 128-bit
// loads from both parents and a 128-bit store of the child; consecutive groups
// write consecutive rows, so a tile's stores are one contiguous span.
//
// f32 scores follow the oracle's canonical reduction (lane gl accumulates its
// blocks in order, xor butterfly over the group) and are therefore bit-exact.
template <int ENC, uint32_t PH, int OBJ>
__global__ void __launch_bounds__(256) generation_fast(const __grid_constant__ GenArgs a)
{
    __shared__ unsigned long long warp_best[8];
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t gwarp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const uint32_t nwarps = (gridDim.x * blockDim.x) >> 5;
    const uint32_t S = a.stride, nblk = S >> 2;
    const uint32_t G = a.group, cpt = 32u / G;
    const uint32_t gl = lane & (G - 1u), grp = lane / G;
    const uint4* __restrict__ cur = reinterpret_cast<const uint4*>(a.cur);
    uint4* __restrict__ nxt = reinterpret_cast<uint4*>(a.nxt);

    if ((PH & PH_EVALUATE) && blockIdx.x == 0 && threadIdx.x == 0) a.best[(a.best_slot + 1u) & 3u] = 0ull;

    unsigned long long key = 0ull;

    for (uint64_t cbase64 = (uint64_t)gwarp * 32u; cbase64 < a.n; cbase64 += (uint64_t)nwarps * 32u) {
        const uint32_t cbase = (uint32_t)cbase64;
        const uint32_t mine = cbase + lane;
        const bool have = mine < a.n;

        uint32_t p1 = 0, p2 = 0, m_idx = 0, m_val = 0;
        bool m_on = false;
        if ((PH & (PH_CROSSOVER | PH_MUTATE)) && have) {
            const uint4 w = draw_block(a.draw, STREAM_ROW, mine, 0);
            const float u0 = to_uniform(w.x), u1 = to_uniform(w.y), u2 = to_uniform(w.z), u3 = to_uniform(w.w);
            if (PH & PH_CROSSOVER) {
                if (a.tk == 2u) {
                    p1 = tournament(a.score_in, u0, u1, a.nf, a.n);
                    p2 = tournament(a.score_in, u2, u3, a.nf, a.n);
                } else {
                    p1 = tournament_of(a.draw, a.score_in, mine, 0u, a.tk, a.nf, a.n);
                    p2 = tournament_of(a.draw, a.score_in, mine, a.tk, a.tk, a.nf, a.n);
                }
                if (a.sel) { a.sel[2 * (size_t)mine] = (int32_t)p1; a.sel[2 * (size_t)mine + 1] = (int32_t)p2; }
            }
            if (PH & PH_MUTATE) {
                m_on = u1 <= PGA_MUTATION_CHANCE;
                m_idx = draw_index(u0, a.lenf, a.len);
                m_val = ENC == ENC_BITS ? (u2 > 0.5f ? 1u : 0u) : __float_as_uint(u2);
            }
        }

        float my_score = 0.f;
        for (uint32_t tile = 0; tile < G; ++tile) {
            const uint32_t src_lane = tile * cpt + grp;
            const uint32_t child = cbase + src_lane;
            const bool valid = child < a.n;
            const uint32_t q1 = __shfl_sync(0xFFFFFFFFu, p1, src_lane);
            const uint32_t q2 = __shfl_sync(0xFFFFFFFFu, p2, src_lane);
            const bool c_on = __shfl_sync(0xFFFFFFFFu, (int)m_on, src_lane) != 0;
            const uint32_t c_idx = __shfl_sync(0xFFFFFFFFu, m_idx, src_lane);
            const uint32_t c_val = __shfl_sync(0xFFFFFFFFu, m_val, src_lane);
            // the 16-byte block that holds the mutated gene
            const uint32_t c_blk = ENC == ENC_BITS ? (c_idx >> 7) : (c_idx >> 2);

            float facc = 0.f;
            uint32_t iacc = 0;
            if (valid) {
                for (uint32_t b = gl; b < nblk; b += G) {
                    uint4 c;
                    if (PH & PH_CROSSOVER) {
                        PGA_ASSERT(q1 < a.n && q2 < a.n && b < nblk);
                        const uint4 x = __ldg(cur + (size_t)q1 * nblk + b);
                        const uint4 y = __ldg(cur + (size_t)q2 * nblk + b);
                        if (ENC == ENC_BITS) {
                            const uint4 m = draw_block(a.draw, STREAM_MASK, child, b);
                            c.x = (x.x & m.x) | (y.x & ~m.x);
                            c.y = (x.y & m.y) | (y.y & ~m.y);
                            c.z = (x.z & m.z) | (y.z & ~m.z);
                            c.w = (x.w & m.w) | (y.w & ~m.w);
                        } else {
                            const uint4 w = draw_block(a.draw, STREAM_ROW, child, b);
                            c.x = to_uniform(w.x) > 0.5f ? x.x : y.x;  // src/pga.cu:137-141
                            c.y = to_uniform(w.y) > 0.5f ? x.y : y.y;
                            c.z = to_uniform(w.z) > 0.5f ? x.z : y.z;
                            c.w = to_uniform(w.w) > 0.5f ? x.w : y.w;
                        }
                    } else {
                        // a pass without crossover works on the row itself: the mutate
                        // step edits next_gen in place, the evaluate step reads current_gen
                        c = (PH & PH_MUTATE) ? nxt[(size_t)child * nblk + b] : __ldg(cur + (size_t)child * nblk + b);
                    }
                    bool touched = false;
                    if ((PH & PH_MUTATE) && c_on && c_blk == b) {  // src/pga.cu:127-133
                        touched = true;
                        if (ENC == ENC_BITS) {
                            const uint32_t word = (c_idx >> 5) & 3u, bit = 1u << (c_idx & 31u);
                            uint32_t* cw = word == 0 ? &c.x : word == 1 ? &c.y : word == 2 ? &c.z : &c.w;
                            *cw = c_val ? (*cw | bit) : (*cw & ~bit);
                        } else {
                            const uint32_t e = c_idx & 3u;
                            if (e == 0) c.x = c_val; else if (e == 1) c.y = c_val; else if (e == 2) c.z = c_val; else c.w = c_val;
                        }
                    }
                    if (PH & PH_EVALUATE) {
                        if (ENC == ENC_BITS) {
                            iacc += __popc(c.x) + __popc(c.y) + __popc(c.z) + __popc(c.w);
                        } else {
                            const uint32_t j = 4u * b;
                            if (j + 0 < a.len) facc = obj_accumulate<OBJ>(facc, __uint_as_float(c.x));
                            if (j + 1 < a.len) facc = obj_accumulate<OBJ>(facc, __uint_as_float(c.y));
                            if (j + 2 < a.len) facc = obj_accumulate<OBJ>(facc, __uint_as_float(c.z));
                            if (j + 3 < a.len) facc = obj_accumulate<OBJ>(facc, __uint_as_float(c.w));
                        }
                    }
                    if (PH & PH_CROSSOVER)
                        nxt[(size_t)child * nblk + b] = c;
                    else if ((PH & PH_MUTATE) && touched)
                        nxt[(size_t)child * nblk + b] = c;
                }
            }
            if (PH & PH_EVALUATE) {
                float s;
                if (ENC == ENC_BITS) {
                    for (uint32_t off = G >> 1; off >= 1u; off >>= 1) iacc += __shfl_xor_sync(0xFFFFFFFFu, iacc, off);
                    s = __uint2float_rn(iacc);
                } else {
                    for (uint32_t off = G >> 1; off >= 1u; off >>= 1)
                        facc = __fadd_rn(facc, __shfl_xor_sync(0xFFFFFFFFu, facc, off));
                    s = obj_finalize<OBJ>(facc, a.len);
                }
                // hand the score to the lane that ran this child's tournaments
                const float v = __shfl_sync(0xFFFFFFFFu, s, (lane % cpt) * G);
                if (lane / cpt == tile) my_score = v;
            }
        }
        if ((PH & PH_EVALUATE) && have) {
            a.score_out[mine] = my_score;
            const unsigned long long k = best_key(my_score, mine);
            key = k > key ? k : key;
        }
    }
    if (PH & PH_EVALUATE) block_best_commit(key, a.best + a.best_slot, warp_best);
}

// ---- fused generation, tiled -----------------------------------------------------
//
// The same algorithm as generation_fast<.., PH_FUSED, ..> with the shape known at
// compile time: G = 2^LOGG lanes per child, NB 16-byte blocks per lane, TB tiles
// batched so that 2*TB*NB independent 128-bit parent loads are in flight per lane
// before any of them is consumed.  EXACT says every lane owns exactly NB blocks
// (row blocks == NB*G), which removes the per-block bounds test.  Children past
// the end of the population are computed from row 0 and only their stores are
// suppressed, so the loop body has no divergent region around the shuffles.
// This is the kernel the roofline numbers are quoted on.
//
// ONEBLOCK: the same 8 * MINB warps per SM as ONE block instead of MINB blocks of 8 warps, so
// that a grid of (SMs - r) blocks leaves r SMs completely empty.  Used for the generation
// that runs beside a migration: the packing and the NCCL send/recv kernels on the side
// stream need whole SMs (two half-filled SMs do not hold an NCCL block; one empty SM does).
// It costs ~1 % (72.6 against 72.0 us per OneMax generation, 552 against 527 for
// Rastrigin), so every other generation keeps the MINB-blocks form.
template <int ENC, int OBJ, int LOGG, int NB, int TB, bool EXACT, int MINB, bool ONEBLOCK>
__global__ void __launch_bounds__(ONEBLOCK ? 256 * MINB : 256, ONEBLOCK ? 1 : MINB) generation_fused_tiled(const __grid_constant__ GenArgs a)
{
    constexpr uint32_t G = 1u << LOGG, CPT = 32u / G;
    constexpr bool KEEP_L2 = (uint32_t)NB * G * 16u < 1024u;  // rows under 1 KiB: see load_row_block
    __shared__ unsigned long long warp_best[8 * MINB];
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t gwarp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const uint32_t nwarps = (gridDim.x * blockDim.x) >> 5;
    const uint32_t nblk = EXACT ? (uint32_t)NB * G : a.stride >> 2;
    const uint32_t gl = lane & (G - 1u), grp = lane >> LOGG;
    const uint4* __restrict__ cur = reinterpret_cast<const uint4*>(a.cur);
    uint4* __restrict__ nxt = reinterpret_cast<uint4*>(a.nxt);

    if (blockIdx.x == 0 && threadIdx.x == 0) a.best[(a.best_slot + 1u) & 3u] = 0ull;

    unsigned long long key = 0ull;

    for (uint64_t cbase64 = (uint64_t)gwarp * 32u; cbase64 < a.n; cbase64 += (uint64_t)nwarps * 32u) {
        const uint32_t cbase = (uint32_t)cbase64;
        const uint32_t mine = cbase + lane;
        const bool have = mine < a.n;

        // lane j: tournaments and mutation draw of child cbase + j
        uint32_t p1 = 0, p2 = 0, m_idx = 0, m_val = 0;
        bool m_on = false;
        if (have) {
            const uint4 w = draw_block(a.draw, STREAM_ROW, mine, 0);
            const float u0 = to_uniform(w.x), u1 = to_uniform(w.y), u2 = to_uniform(w.z), u3 = to_uniform(w.w);
            p1 = tournament(a.score_in, u0, u1, a.nf, a.n);
            p2 = tournament(a.score_in, u2, u3, a.nf, a.n);
            if (a.sel) { a.sel[2 * (size_t)mine] = (int32_t)p1; a.sel[2 * (size_t)mine + 1] = (int32_t)p2; }
            m_on = u1 <= PGA_MUTATION_CHANCE;
            m_idx = draw_index(u0, a.lenf, a.len);
            m_val = ENC == ENC_BITS ? (u2 > 0.5f ? 1u : 0u) : __float_as_uint(u2);
        }
        const uint32_t any_mut = __ballot_sync(0xFFFFFFFFu, m_on);  // 72 % of warps: none

        float my_score = 0.f;
#pragma unroll 1
        for (uint32_t tile0 = 0; tile0 < G; tile0 += TB) {
            uint4 x[TB][NB], y[TB][NB];
            // issue every parent load of the batch
#pragma unroll
            for (int tb = 0; tb < TB; ++tb) {
                const uint32_t src_lane = (tile0 + tb) * CPT + grp;
                const uint32_t q1 = __shfl_sync(0xFFFFFFFFu, p1, src_lane);
                const uint32_t q2 = __shfl_sync(0xFFFFFFFFu, p2, src_lane);
                PGA_ASSERT(q1 < a.n && q2 < a.n);
                const uint4* r1 = cur + (size_t)q1 * nblk + gl;
                const uint4* r2 = cur + (size_t)q2 * nblk + gl;
#pragma unroll
                for (int nb = 0; nb < NB; ++nb) {
                    if (EXACT || gl + (uint32_t)nb * G < nblk) {
                        x[tb][nb] = load_row_block<KEEP_L2>(r1 + nb * G);
                        y[tb][nb] = load_row_block<KEEP_L2>(r2 + nb * G);
                    }
                }
            }
            // consume
#pragma unroll
            for (int tb = 0; tb < TB; ++tb) {
                const uint32_t src_lane = (tile0 + tb) * CPT + grp;
                const uint32_t child = cbase + src_lane;
                const bool valid = child < a.n;
                uint32_t c_idx = 0, c_val = 0;
                if (any_mut) {  // warp-uniform
                    c_idx = __shfl_sync(0xFFFFFFFFu, m_idx, src_lane);
                    c_val = __shfl_sync(0xFFFFFFFFu, m_val, src_lane);
                }
                const bool c_on = (any_mut >> src_lane) & 1u;
                const uint32_t c_blk = ENC == ENC_BITS ? (c_idx >> 7) : (c_idx >> 2);
                uint4* out = nxt + (size_t)child * nblk + gl;
                float facc = 0.f;
                uint32_t iacc = 0;
#pragma unroll
                for (int nb = 0; nb < NB; ++nb) {
                    const uint32_t b = gl + (uint32_t)nb * G;
                    if (EXACT || b < nblk) {
                        const uint4 xx = x[tb][nb], yy = y[tb][nb];
                        uint4 c;
                        if (ENC == ENC_BITS) {
                            const uint4 m = draw_block(a.draw, STREAM_MASK, child, b);
                            c.x = (xx.x & m.x) | (yy.x & ~m.x);
                            c.y = (xx.y & m.y) | (yy.y & ~m.y);
                            c.z = (xx.z & m.z) | (yy.z & ~m.z);
                            c.w = (xx.w & m.w) | (yy.w & ~m.w);
                        } else {
                            const uint4 w = draw_block(a.draw, STREAM_ROW, child, b);
                            c.x = draw_above_half(w.x) ? xx.x : yy.x;  // src/pga.cu:137-141
                            c.y = draw_above_half(w.y) ? xx.y : yy.y;
                            c.z = draw_above_half(w.z) ? xx.z : yy.z;
                            c.w = draw_above_half(w.w) ? xx.w : yy.w;
                        }
                        if (c_on && c_blk == b) {  // src/pga.cu:127-133; rare
                            if (ENC == ENC_BITS) {
                                const uint32_t word = (c_idx >> 5) & 3u, bit = 1u << (c_idx & 31u);
                                uint32_t* cw = word == 0 ? &c.x : word == 1 ? &c.y : word == 2 ? &c.z : &c.w;
                                *cw = c_val ? (*cw | bit) : (*cw & ~bit);
                            } else {
                                const uint32_t e = c_idx & 3u;
                                if (e == 0) c.x = c_val; else if (e == 1) c.y = c_val; else if (e == 2) c.z = c_val; else c.w = c_val;
                            }
                        }
                        if (ENC == ENC_BITS) {
                            iacc += __popc(c.x) + __popc(c.y) + __popc(c.z) + __popc(c.w);
                        } else {
                            const uint32_t j = 4u * b;
                            if (EXACT || j + 0 < a.len) facc = obj_accumulate<OBJ>(facc, __uint_as_float(c.x));
                            if (EXACT || j + 1 < a.len) facc = obj_accumulate<OBJ>(facc, __uint_as_float(c.y));
                            if (EXACT || j + 2 < a.len) facc = obj_accumulate<OBJ>(facc, __uint_as_float(c.z));
                            if (EXACT || j + 3 < a.len) facc = obj_accumulate<OBJ>(facc, __uint_as_float(c.w));
                        }
                        PGA_ASSERT(!valid || (child < a.n && b < (a.stride >> 2)));
                        if (valid) store_row_block(out + nb * G, c);
                    }
                }
                float s;
                if (ENC == ENC_BITS) {
#pragma unroll
                    for (uint32_t off = G >> 1; off >= 1u; off >>= 1) iacc += __shfl_xor_sync(0xFFFFFFFFu, iacc, off);
                    s = __uint2float_rn(iacc);
                } else {
#pragma unroll
                    for (uint32_t off = G >> 1; off >= 1u; off >>= 1)
                        facc = __fadd_rn(facc, __shfl_xor_sync(0xFFFFFFFFu, facc, off));
                    s = obj_finalize<OBJ>(facc, a.len);
                }
                const float v = __shfl_sync(0xFFFFFFFFu, s, (lane & (CPT - 1u)) << LOGG);
                if ((lane >> (5 - LOGG)) == tile0 + (uint32_t)tb) my_score = v;
            }
        }
        if (have) {
            a.score_out[mine] = my_score;
            const unsigned long long k = best_key(my_score, mine);
            key = k > key ? k : key;
        }
    }
    block_best_commit(key, a.best + a.best_slot, warp_best);
}

// ---- population set-up (src/pga.cu:81-86, :116-117): genes := rand row of epoch 0 ----
template <int ENC>
__global__ void init_population(uint32_t* rows, uint32_t n, uint32_t len, uint32_t stride, DrawCtx d)
{
    const uint32_t nblk = stride >> 2;
    const uint64_t total = (uint64_t)n * nblk;
    for (uint64_t k = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; k < total; k += (uint64_t)gridDim.x * blockDim.x) {
        const uint32_t i = (uint32_t)(k / nblk), b = (uint32_t)(k % nblk);
        uint4 w = draw_block(d, ENC == ENC_BITS ? STREAM_INITBITS : STREAM_ROW, i, b);
        if (ENC == ENC_BITS) {
            uint32_t out[4];
            for (int q = 0; q < 4; ++q) {
                const uint32_t lo = (4u * b + q) * 32u;  // first bit of this word
                uint32_t v = q == 0 ? w.x : q == 1 ? w.y : q == 2 ? w.z : w.w;
                if (lo >= len) v = 0u;
                else if (len - lo < 32u) v &= (1u << (len - lo)) - 1u;
                out[q] = v;
            }
            w = make_uint4(out[0], out[1], out[2], out[3]);
        } else {
            uint32_t out[4] = { w.x, w.y, w.z, w.w };
            for (int q = 0; q < 4; ++q)
                out[q] = (4u * b + q < len) ? __float_as_uint(to_uniform(out[q])) : 0u;
            w = make_uint4(out[0], out[1], out[2], out[3]);
        }
        reinterpret_cast<uint4*>(rows)[k] = w;
    }
}

// ---- stand-alone best reduction (after uploads and migrations) -----------------
__global__ void __launch_bounds__(256) reduce_best(const float* __restrict__ score, uint32_t n, unsigned long long* best,
                                                   uint32_t slot)
{
    __shared__ unsigned long long warp_best[8];
    if (blockIdx.x == 0 && threadIdx.x == 0) best[(slot + 1u) & 3u] = 0ull;
    unsigned long long key = 0ull;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) {
        const unsigned long long k = best_key(__ldg(score + i), (uint32_t)i);
        key = k > key ? k : key;
    }
    block_best_commit(key, best + slot, warp_best);
}

// ---- best score over several populations' slots (early-stop check) ------------------
struct BestSlots {
    const unsigned long long* slot[10];  // MAX_POPULATIONS
    uint32_t count;
};

// inverse of order_key: the score a best-key was built from (an empty slot or NaN gives -inf)
__device__ __forceinline__ float score_of_key(unsigned long long key)
{
    const uint32_t k = (uint32_t)(key >> 32);
    if (k == 0u) return -INFINITY;
    return __uint_as_float((k & 0x80000000u) ? (k ^ 0x80000000u) : ~k);
}

__global__ void best_of_slots(const BestSlots s, float* out)
{
    if (threadIdx.x == 0) {
        float best = -INFINITY;
        for (uint32_t i = 0; i < s.count; ++i) best = fmaxf(best, score_of_key(*s.slot[i]));
        *out = best;
    }
}

// ---- top-k radix select ------------------------------------------------------------
//
// Picks the k best (or k worst) individuals under the engine's total order and
// lists them in ascending index order.  Radix passes over the 32-bit order key find
// the k-th key T; a stable compaction then takes every key beyond T and as many keys
// equal to T as are still needed (lowest indices first for "best", highest first for
// "worst").  Histogram updates are aggregated per warp with __match_any_sync before
// they touch shared memory.
#define TOPK_BLOCK 256u
#define TOPK_ROUNDS 8u
#define TOPK_CHUNK (TOPK_BLOCK * TOPK_ROUNDS)

__device__ __forceinline__ uint32_t select_key(float s, int worst)
{
    const uint32_t k = order_key(s);
    return worst ? ~k : k;
}

// ---- the selection in five launches ----------------------------------------------------
//
// Three radix passes of 11 + 11 + 10 bits; the last block to finish a pass (ticket counter after
// a __threadfence) picks the digit, so no separate pick launch; the last block of the
// count pass scans the per-block counts.  State is left clean for the next call.
#define TOPK2_BINS 2048u
struct TopkState2 {
    uint32_t hist[TOPK2_BINS];
    uint32_t prefix, mask, remaining;
    uint32_t threshold, need_eq, total_eq, total_gt;
    uint32_t ticket;
};

// block-wide: find the highest bin d with (count of bins above d) + h[d] >= rem.
// h[] lives in shared memory (nbins <= 2048, 256 threads).  Returns d and the count above it.
__device__ __forceinline__ void pick_digit(const uint32_t* h, uint32_t nbins, uint32_t rem, uint32_t* out_digit, uint32_t* out_above,
                                           uint32_t* scratch /* >= 264 words */)
{
    const uint32_t t = threadIdx.x, per = nbins / 256u;  // bins per thread, descending order: thread 0 owns the top bins
    uint32_t mine = 0;
    for (uint32_t j = 0; j < per; ++j) mine += h[nbins - 1u - (t * per + j)];
    // inclusive scan of `mine` over threads (warp shuffles + 8 warp totals)
    uint32_t inc = mine;
    const uint32_t lane = t & 31u, warp = t >> 5;
    for (int off = 1; off < 32; off <<= 1) {
        const uint32_t o = __shfl_up_sync(0xFFFFFFFFu, inc, off);
        if ((int)lane >= off) inc += o;
    }
    if (lane == 31u) scratch[warp] = inc;
    __syncthreads();
    uint32_t before_warp = 0;
    for (uint32_t w2 = 0; w2 < warp; ++w2) before_warp += scratch[w2];
    inc += before_warp;
    const uint32_t excl = inc - mine;  // keys in bins owned by lower-numbered threads (= higher bins)
    if (excl < rem && inc >= rem) {    // the crossing happens inside this thread's bins; exactly one thread
        uint32_t above = excl;
        for (uint32_t j = 0; j < per; ++j) {
            const uint32_t d = nbins - 1u - (t * per + j);
            if (above + h[d] >= rem) { *out_digit = d; *out_above = above; break; }
            above += h[d];
        }
    }
    __syncthreads();
}

template <uint32_t SHIFT, uint32_t BITS>
__global__ void __launch_bounds__(256) topk2_histogram(const float* __restrict__ score, uint32_t n, int worst, uint32_t k, TopkState2* st, uint32_t nb)
{
    constexpr uint32_t NBINS = 1u << BITS;
    __shared__ uint32_t h[TOPK2_BINS];
    __shared__ uint32_t scratch[16];
    __shared__ uint32_t is_last, s_digit, s_above;
    for (uint32_t i = threadIdx.x; i < NBINS; i += 256u) h[i] = 0u;
    __syncthreads();
    const bool first = SHIFT + BITS == 32u;  // the first pass starts from an empty prefix
    const uint32_t prefix = first ? 0u : st->prefix, mask = first ? 0u : st->mask;
    const uint32_t lane = threadIdx.x & 31u;
    // 2048-score chunks, grid-stride: the grid may be far smaller than the number of chunks
    // (a ranking that runs beside a generation kernel only gets the SMs that kernel left free).
    // All eight loads of a thread are issued before the first is used (the ballots below
    // would otherwise serialise eight round trips).
    for (uint32_t chunk = blockIdx.x; chunk < nb; chunk += gridDim.x) {
        const uint64_t base = (uint64_t)chunk * TOPK_CHUNK;
        float v[TOPK_ROUNDS];
#pragma unroll
        for (uint32_t r = 0; r < TOPK_ROUNDS; ++r) {
            const uint64_t i = base + r * TOPK_BLOCK + threadIdx.x;
            v[r] = i < n ? __ldg(score + i) : 0.f;
        }
#pragma unroll
        for (uint32_t r = 0; r < TOPK_ROUNDS; ++r) {
            const uint64_t i = base + r * TOPK_BLOCK + threadIdx.x;
            const uint32_t key = select_key(v[r], worst);
            const bool hit = i < n && (key & mask) == prefix;
            const uint32_t digit = (key >> SHIFT) & (NBINS - 1u);
            const uint32_t voters = __ballot_sync(0xFFFFFFFFu, hit);
            if (hit) {  // one shared-memory atomic per distinct digit in the warp
                const uint32_t peers = __match_any_sync(voters, digit);
                if ((peers & ((1u << lane) - 1u)) == 0u) atomicAdd(&h[digit], (uint32_t)__popc(peers));
            }
        }
    }
    __syncthreads();
    for (uint32_t i = threadIdx.x; i < NBINS; i += 256u)
        if (h[i]) atomicAdd(&st->hist[i], h[i]);
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) is_last = atomicAdd(&st->ticket, 1u) == gridDim.x - 1u;
    __syncthreads();
    if (!is_last) return;
    // last block: all histograms have landed; choose the digit of this pass
    __threadfence();
    for (uint32_t i = threadIdx.x; i < NBINS; i += 256u) {
        h[i] = __ldcg(&st->hist[i]);
        st->hist[i] = 0u;  // clean for the next pass / next call
    }
    __syncthreads();
    const uint32_t rem = first ? k : st->remaining;
    pick_digit(h, NBINS, rem, &s_digit, &s_above, scratch);
    if (threadIdx.x == 0) {
        const uint32_t d = s_digit, above = s_above;
        st->remaining = rem - above;
        st->prefix = prefix | (d << SHIFT);
        st->mask = mask | ((NBINS - 1u) << SHIFT);
        st->total_gt = (first ? 0u : st->total_gt) + above;
        if (SHIFT == 0u) {
            st->threshold = prefix | d;
            st->need_eq = rem - above;
            st->total_eq = h[d];
        }
        st->ticket = 0u;
    }
}

// per-block counts of keys beyond / equal to the threshold; the last block scans them
__global__ void __launch_bounds__(256) topk2_count_scan(const float* __restrict__ score, uint32_t n, int worst, TopkState2* st,
                                                        uint32_t* block_gt, uint32_t* block_eq, uint32_t nb)
{
    __shared__ uint32_t s_gt, s_eq, is_last;
    __shared__ uint32_t wsum[2][8];
    __shared__ uint32_t carry[2];
    const uint32_t T = st->threshold;
    for (uint32_t chunk = blockIdx.x; chunk < nb; chunk += gridDim.x) {  // counts are per chunk, whatever the grid
        if (threadIdx.x == 0) { s_gt = 0u; s_eq = 0u; }
        __syncthreads();
        const uint64_t base = (uint64_t)chunk * TOPK_CHUNK;
        float v[TOPK_ROUNDS];
#pragma unroll
        for (uint32_t r = 0; r < TOPK_ROUNDS; ++r) {
            const uint64_t i = base + r * TOPK_BLOCK + threadIdx.x;
            v[r] = i < n ? __ldg(score + i) : 0.f;
        }
        uint32_t gt = 0u, eq = 0u;
#pragma unroll
        for (uint32_t r = 0; r < TOPK_ROUNDS; ++r) {
            const uint64_t i = base + r * TOPK_BLOCK + threadIdx.x;
            const uint32_t key = select_key(v[r], worst);
            gt += i < n && key > T;
            eq += i < n && key == T;
        }
        for (int off = 16; off >= 1; off >>= 1) {
            gt += __shfl_xor_sync(0xFFFFFFFFu, gt, off);
            eq += __shfl_xor_sync(0xFFFFFFFFu, eq, off);
        }
        if ((threadIdx.x & 31u) == 0u) { atomicAdd(&s_gt, gt); atomicAdd(&s_eq, eq); }
        __syncthreads();
        if (threadIdx.x == 0) {
            block_gt[chunk] = s_gt;
            block_eq[chunk] = s_eq;
        }
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        __threadfence();
        is_last = atomicAdd(&st->ticket, 1u) == gridDim.x - 1u;
    }
    __syncthreads();
    if (!is_last) return;
    __threadfence();
    // exclusive scan of both arrays in place, 256 entries at a time
    const uint32_t t = threadIdx.x, lane = t & 31u, warp = t >> 5;
    if (t == 0) { carry[0] = 0u; carry[1] = 0u; st->ticket = 0u; }
    __syncthreads();
    for (uint32_t b0 = 0; b0 < nb; b0 += 256u) {
        const uint32_t i = b0 + t;
        uint32_t v[2] = { i < nb ? __ldcg(block_gt + i) : 0u, i < nb ? __ldcg(block_eq + i) : 0u };
        uint32_t inc[2] = { v[0], v[1] };
        for (int q = 0; q < 2; ++q) {
            for (int off = 1; off < 32; off <<= 1) {
                const uint32_t o = __shfl_up_sync(0xFFFFFFFFu, inc[q], off);
                if ((int)lane >= off) inc[q] += o;
            }
            if (lane == 31u) wsum[q][warp] = inc[q];
        }
        __syncthreads();
        uint32_t tot[2];
        for (int q = 0; q < 2; ++q) {
            uint32_t before = carry[q];
            for (uint32_t w2 = 0; w2 < warp; ++w2) before += wsum[q][w2];
            const uint32_t excl = before + inc[q] - v[q];
            if (i < nb) (q == 0 ? block_gt : block_eq)[i] = excl;
            tot[q] = excl + v[q];
        }
        __syncthreads();
        if (t == 255u) { carry[0] = tot[0]; carry[1] = tot[1]; }
        __syncthreads();
    }
}

__global__ void __launch_bounds__(256) topk2_emit(const float* __restrict__ score, uint32_t n, int worst, const TopkState2* st,
                                                  const uint32_t* block_gt, const uint32_t* block_eq, uint32_t* out_idx, uint32_t nb)
{
    // Each warp owns 256 CONSECUTIVE scores of a chunk (8 sub-rounds of 32), so index
    // order is warp-major and a chunk needs one barrier: warp totals -> prefix.
    __shared__ uint32_t w_gt[8], w_eq[8];
    const uint32_t T = st->threshold, need = st->need_eq, total_eq = st->total_eq;
    // "best": the first `need` equal keys by index; "worst": the last `need`
    const uint32_t eq_lo = worst ? total_eq - need : 0u, eq_hi = worst ? total_eq : need;
    const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    for (uint32_t chunk = blockIdx.x; chunk < nb; chunk += gridDim.x) {
        const uint64_t base = (uint64_t)chunk * TOPK_CHUNK + warp * 256u;
        float v[TOPK_ROUNDS];
#pragma unroll
        for (uint32_t r = 0; r < TOPK_ROUNDS; ++r) {
            const uint64_t i = base + r * 32u + lane;
            v[r] = i < n ? __ldg(score + i) : 0.f;
        }
        uint32_t bg[TOPK_ROUNDS], be[TOPK_ROUNDS], tot_g = 0u, tot_e = 0u;
#pragma unroll
        for (uint32_t r = 0; r < TOPK_ROUNDS; ++r) {
            const uint64_t i = base + r * 32u + lane;
            const uint32_t key = select_key(v[r], worst);
            bg[r] = __ballot_sync(0xFFFFFFFFu, i < n && key > T);
            be[r] = __ballot_sync(0xFFFFFFFFu, i < n && key == T);
            tot_g += __popc(bg[r]);
            tot_e += __popc(be[r]);
        }
        if (lane == 0) { w_gt[warp] = tot_g; w_eq[warp] = tot_e; }
        __syncthreads();
        uint32_t run_g = block_gt[chunk], run_e = block_eq[chunk];
        for (uint32_t w2 = 0; w2 < warp; ++w2) { run_g += w_gt[w2]; run_e += w_eq[w2]; }
        const uint32_t lt_mask = (1u << lane) - 1u;
#pragma unroll
        for (uint32_t r = 0; r < TOPK_ROUNDS; ++r) {
            const uint64_t i = base + r * 32u + lane;
            const bool gt = (bg[r] >> lane) & 1u, eq = (be[r] >> lane) & 1u;
            const uint32_t gt_before = run_g + __popc(bg[r] & lt_mask), eq_before = run_e + __popc(be[r] & lt_mask);
            // position among the selected = selected elements with a lower index
            const uint32_t eq_sel_before = eq_before <= eq_lo ? 0u : (eq_before >= eq_hi ? eq_hi - eq_lo : eq_before - eq_lo);
            if (gt) out_idx[gt_before + eq_sel_before] = (uint32_t)i;
            else if (eq && eq_before >= eq_lo && eq_before < eq_hi) out_idx[gt_before + eq_sel_before] = (uint32_t)i;
            run_g += __popc(bg[r]);
            run_e += __popc(be[r]);
        }
        __syncthreads();  // the warp totals are rewritten for the next chunk
    }
}

// state of a ranking that starts from memory of unknown content (pgax_select_top_noalloc)
__global__ void topk2_reset(TopkState2* st)
{
    for (uint32_t i = threadIdx.x; i < TOPK2_BINS; i += blockDim.x) st->hist[i] = 0u;
    if (threadIdx.x == 0) {
        st->prefix = 0u; st->mask = 0u; st->remaining = 0u;
        st->threshold = 0u; st->need_eq = 0u; st->total_eq = 0u; st->total_gt = 0u; st->ticket = 0u;
    }
}

// ---- top-k select in ONE cooperative launch ------------------------------------------
//
// A migration is latency-bound: the five-launch select above spends ~7 us per launch on
// launch gaps, a fresh round trip to the scores and the last-block hand-over.  Here the
// whole selection is one kernel with one block of 1024 threads per SM, all resident
// (cooperative launch).  Every thread loads its E scores ONCE and keeps the keys in
// registers through the three radix passes and the compaction.  Every block derives the
// chosen digit itself from the global histogram, so no block has to publish a decision.
//
// Three grid barriers, one per radix pass.  There is no separate count pass: in the last pass every block
// also publishes, for each of its 1024 digits d, how many of its candidate keys have a
// digit >= d, and how many of its keys lie beyond the 22-bit prefix; once the digit is
// known each block reads two numbers per lower-numbered block and has its output offset.
// The state alternates between two copies of the histograms (parity of the call number):
// a call clears the copy the NEXT call will use, so no barrier is spent on cleaning up.
// ncu on the 512-block, four-barrier form this replaces: 34 us at 2^20 scores, 13 of 31
// stall cycles per instruction at the barriers.
#define TOPK1_THREADS 1024u
struct TopkOne {
    uint32_t hist[2][3][TOPK2_BINS];       // [call parity][pass][bin]
    unsigned long long arrive;             // monotonic barrier counter (never reset)
    uint32_t beyond[160];                  // per block: keys beyond the 22-bit prefix
    uint32_t suffix[160][1025];            // per block: candidate keys with last digit >= d
};

__device__ __forceinline__ void grid_barrier(unsigned long long* ctr, unsigned long long target)
{
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        atomicAdd(ctr, 1ull);
        while (*reinterpret_cast<volatile unsigned long long*>(ctr) < target) { }
        __threadfence();
    }
    __syncthreads();
}

// block-wide (1024 threads): highest bin d of h[0..nbins) with (count above d) + h[d] >= rem
__device__ __forceinline__ void pick_digit_1024(const uint32_t* h, uint32_t nbins, uint32_t rem, uint32_t* out_digit, uint32_t* out_above,
                                                uint32_t* scratch /* >= 32 words */)
{
    const uint32_t t = threadIdx.x, per = nbins / TOPK1_THREADS;  // descending: thread 0 owns the top bins
    uint32_t mine = 0;
    for (uint32_t j = 0; j < per; ++j) mine += h[nbins - 1u - (t * per + j)];
    uint32_t inc = mine;
    const uint32_t lane = t & 31u, warp = t >> 5;
    for (int off = 1; off < 32; off <<= 1) {
        const uint32_t o = __shfl_up_sync(0xFFFFFFFFu, inc, off);
        if ((int)lane >= off) inc += o;
    }
    if (lane == 31u) scratch[warp] = inc;
    __syncthreads();
    if (warp == 0) {  // exclusive scan of the 32 warp totals
        const uint32_t w = scratch[lane];
        uint32_t wi = w;
        for (int off = 1; off < 32; off <<= 1) {
            const uint32_t o = __shfl_up_sync(0xFFFFFFFFu, wi, off);
            if ((int)lane >= off) wi += o;
        }
        scratch[lane] = wi - w;
    }
    __syncthreads();
    inc += scratch[warp];
    const uint32_t excl = inc - mine;  // keys in bins owned by lower-numbered threads (= higher bins)
    if (excl < rem && inc >= rem) {    // the crossing happens inside this thread's bins; exactly one thread
        uint32_t above = excl;
        for (uint32_t j = 0; j < per; ++j) {
            const uint32_t d = nbins - 1u - (t * per + j);
            if (above + h[d] >= rem) { *out_digit = d; *out_above = above; break; }
            above += h[d];
        }
    }
    __syncthreads();
}

template <int E>
__global__ void __launch_bounds__(TOPK1_THREADS, 1) topk_select_coop(const float* __restrict__ score, uint32_t n, uint32_t k, int worst, TopkOne* st,
                                                                     uint32_t* out_idx, unsigned long long barrier_base, uint32_t parity)
{
    __shared__ uint32_t h[TOPK2_BINS];
    __shared__ uint32_t scratch[32];
    __shared__ uint32_t s_digit, s_above;
    __shared__ uint32_t w_gt[32], w_eq[32];
    __shared__ uint32_t s_before[2], s_beyond;
    const uint32_t t = threadIdx.x, lane = t & 31u, warp = t >> 5;
    // warp-major layout: each warp owns 32*E consecutive scores, so index order is
    // (block, warp, round, lane) and the compaction needs no block-wide scan per round
    const uint64_t base = ((uint64_t)blockIdx.x * 32u + warp) * (32u * E);
    uint32_t key[E];
#pragma unroll
    for (int r = 0; r < E; ++r) {
        const uint64_t i = base + (uint32_t)r * 32u + lane;
        key[r] = i < n ? select_key(__ldg(score + i), worst) : 0u;
    }
    // the copy of the histograms the next call will use is cleared now (it was last read two calls ago)
    if (blockIdx.x == 0)
        for (uint32_t i = t; i < 3u * TOPK2_BINS; i += TOPK1_THREADS) (&st->hist[parity ^ 1u][0][0])[i] = 0u;
    unsigned long long bar = barrier_base;

    uint32_t prefix = 0u, mask = 0u, rem = k, total_eq = 0u;
#pragma unroll
    for (int pass = 0; pass < 3; ++pass) {
        const uint32_t shift = pass == 0 ? 21u : (pass == 1 ? 10u : 0u);
        const uint32_t nbins = pass == 2 ? 1024u : 2048u;
        {
            for (uint32_t i = t; i < nbins; i += TOPK1_THREADS) h[i] = 0u;
            if (t == 0) s_beyond = 0u;
            __syncthreads();
            uint32_t beyond = 0u;
#pragma unroll
            for (int r = 0; r < E; ++r) {
                const uint64_t i = base + (uint32_t)r * 32u + lane;
                const bool in = i < n;
                const bool hit = in && (key[r] & mask) == prefix;
                if (pass == 2) beyond += in && (key[r] & mask) > prefix;
                const uint32_t digit = (key[r] >> shift) & (nbins - 1u);
                const uint32_t voters = __ballot_sync(0xFFFFFFFFu, hit);
                if (hit) {  // one shared-memory atomic per distinct digit in the warp
                    const uint32_t peers = __match_any_sync(voters, digit);
                    if ((peers & ((1u << lane) - 1u)) == 0u) atomicAdd(&h[digit], (uint32_t)__popc(peers));
                }
            }
            if (pass == 2) {
                for (int off = 16; off >= 1; off >>= 1) beyond += __shfl_xor_sync(0xFFFFFFFFu, beyond, off);
                if (lane == 0 && beyond) atomicAdd(&s_beyond, beyond);
            }
            __syncthreads();
            for (uint32_t i = t; i < nbins; i += TOPK1_THREADS)
                if (h[i]) atomicAdd(&st->hist[parity][pass][i], h[i]);
            if (pass == 2) {
                // suffix counts of this block's candidates: thread t owns digit t (1024 digits, 1024 threads)
                const uint32_t mine = h[t];
                uint32_t inc = mine;  // inclusive scan from the top digit down = reverse-ordered scan
                // scan in reversed thread order: use lane' = 31 - lane, warp' = 31 - warp
                for (int off = 1; off < 32; off <<= 1) {
                    const uint32_t o = __shfl_down_sync(0xFFFFFFFFu, inc, off);
                    if ((int)lane + off < 32) inc += o;
                }
                if (lane == 0) scratch[warp] = inc;  // total of the warp's 32 digits
                __syncthreads();
                uint32_t above_warps = 0u;
                for (uint32_t w2 = warp + 1u; w2 < 32u; ++w2) above_warps += scratch[w2];
                st->suffix[blockIdx.x][t] = inc + above_warps;  // candidates with digit >= t
                if (t == 0) {
                    st->suffix[blockIdx.x][1024] = 0u;
                    st->beyond[blockIdx.x] = s_beyond;
                }
            }
            bar += gridDim.x;
            grid_barrier(&st->arrive, bar);
            for (uint32_t i = t; i < nbins; i += TOPK1_THREADS) h[i] = __ldcg(&st->hist[parity][pass][i]);
            __syncthreads();
        }
        pick_digit_1024(h, nbins, rem, &s_digit, &s_above, scratch);
        const uint32_t d = s_digit, above = s_above;
        if (pass == 2) total_eq = h[d];
        prefix |= d << shift;
        mask |= (nbins - 1u) << shift;
        rem -= above;
        __syncthreads();
    }
    const uint32_t T = prefix, need = rem, dT = T & 1023u;

    // selected / equal elements in the blocks before this one, from what they published in the last pass
    uint32_t bg = 0u, be = 0u;
    if (t < blockIdx.x) {
        const uint32_t ge = __ldcg(&st->suffix[t][dT]), gt = __ldcg(&st->suffix[t][dT + 1u]);
        bg = __ldcg(&st->beyond[t]) + gt;
        be = ge - gt;
    }
    for (int off = 16; off >= 1; off >>= 1) {
        bg += __shfl_xor_sync(0xFFFFFFFFu, bg, off);
        be += __shfl_xor_sync(0xFFFFFFFFu, be, off);
    }
    if (t == 0) { s_before[0] = 0u; s_before[1] = 0u; }
    __syncthreads();
    if (lane == 0 && (bg | be)) { atomicAdd(&s_before[0], bg); atomicAdd(&s_before[1], be); }

    // counts of keys beyond / equal to the threshold, per warp
    uint32_t tot_g = 0u, tot_e = 0u;
#pragma unroll
    for (int r = 0; r < E; ++r) {
        const uint64_t i = base + (uint32_t)r * 32u + lane;
        tot_g += __popc(__ballot_sync(0xFFFFFFFFu, i < n && key[r] > T));
        tot_e += __popc(__ballot_sync(0xFFFFFFFFu, i < n && key[r] == T));
    }
    if (lane == 0) { w_gt[warp] = tot_g; w_eq[warp] = tot_e; }
    __syncthreads();
    uint32_t run_g = s_before[0], run_e = s_before[1];
    for (uint32_t w2 = 0; w2 < warp; ++w2) { run_g += w_gt[w2]; run_e += w_eq[w2]; }

    // "best": the first `need` equal keys by index; "worst": the last `need`
    const uint32_t eq_lo = worst ? total_eq - need : 0u, eq_hi = worst ? total_eq : need;
    const uint32_t lt_mask = (1u << lane) - 1u;
#pragma unroll
    for (int r = 0; r < E; ++r) {
        const uint64_t i = base + (uint32_t)r * 32u + lane;
        const bool gt = i < n && key[r] > T, eq = i < n && key[r] == T;
        const uint32_t mg = __ballot_sync(0xFFFFFFFFu, gt), me = __ballot_sync(0xFFFFFFFFu, eq);
        const uint32_t gt_before = run_g + __popc(mg & lt_mask), eq_before = run_e + __popc(me & lt_mask);
        const uint32_t eq_sel_before = eq_before <= eq_lo ? 0u : (eq_before >= eq_hi ? eq_hi - eq_lo : eq_before - eq_lo);
        PGA_ASSERT(!(gt || (eq && eq_before >= eq_lo && eq_before < eq_hi)) || gt_before + eq_sel_before < k);
        if (gt) out_idx[gt_before + eq_sel_before] = (uint32_t)i;
        else if (eq && eq_before >= eq_lo && eq_before < eq_hi) out_idx[gt_before + eq_sel_before] = (uint32_t)i;
        run_g += __popc(mg);
        run_e += __popc(me);
    }
}

// ---- migration payloads ---------------------------------------------------------
// box layout: k rows of `stride` words, then k scores
__global__ void pack_rows(const uint32_t* __restrict__ rows, const float* __restrict__ score, const uint32_t* __restrict__ idx,
                          uint32_t k, uint32_t stride, uint32_t* box, uint32_t n)
{
    const uint32_t nblk = stride >> 2;
    const uint64_t total = (uint64_t)k * nblk;
    float* box_score = reinterpret_cast<float*>(box + (size_t)k * stride);
    for (uint64_t e = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (uint64_t)gridDim.x * blockDim.x) {
        const uint32_t j = (uint32_t)(e / nblk), b = (uint32_t)(e % nblk);
        const uint32_t src = idx[j];
        PGA_ASSERT(src < n);
        reinterpret_cast<uint4*>(box)[e] = __ldg(reinterpret_cast<const uint4*>(rows) + (size_t)src * nblk + b);
        if (b == 0) box_score[j] = score[src];
    }
}

__global__ void unpack_rows(uint32_t* rows, float* score, const uint32_t* __restrict__ idx, uint32_t k, uint32_t stride,
                            const uint32_t* __restrict__ box, uint32_t n)
{
    const uint32_t nblk = stride >> 2;
    const uint64_t total = (uint64_t)k * nblk;
    const float* box_score = reinterpret_cast<const float*>(box + (size_t)k * stride);
    for (uint64_t e = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (uint64_t)gridDim.x * blockDim.x) {
        const uint32_t j = (uint32_t)(e / nblk), b = (uint32_t)(e % nblk);
        const uint32_t dst = idx[j];
        PGA_ASSERT(dst < n);
        reinterpret_cast<uint4*>(rows)[(size_t)dst * nblk + b] = __ldg(reinterpret_cast<const uint4*>(box) + e);
        if (b == 0) score[dst] = box_score[j];
    }
}

__global__ void gather_scores(const float* __restrict__ score, const uint32_t* __restrict__ idx, uint32_t k, float* out)
{
    for (uint32_t j = blockIdx.x * blockDim.x + threadIdx.x; j < k; j += gridDim.x * blockDim.x) out[j] = score[idx[j]];  // idx from the select: < n (asserted there)
}

// host rows are unpadded; device rows are padded to a multiple of 4 words
__global__ void repack_rows(const uint32_t* __restrict__ src, uint32_t* dst, uint64_t n, uint32_t src_stride, uint32_t dst_stride)
{
    const uint64_t total = n * dst_stride;
    for (uint64_t e = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (uint64_t)gridDim.x * blockDim.x) {
        const uint64_t i = e / dst_stride;
        const uint32_t j = (uint32_t)(e % dst_stride);
        dst[e] = j < src_stride ? src[i * src_stride + j] : 0u;
    }
}

}  // namespace pgak
