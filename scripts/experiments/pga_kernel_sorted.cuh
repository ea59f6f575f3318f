// pga_kernel_sorted.cuh -- the fused generation for long rows, children visited in parent order.
//
// With rows of 512 bytes and more the generation kernel is bound by DRAM, and a third of
// what it reads is redundant: under size-2 tournaments only 56.8 % of the individuals are
// ever chosen as FIRST parent, but a child-ordered kernel fetches a first-parent row per
// child (ncu on Rastrigin 2^20 x 256 f32: 3.19 GB moved per generation for 2.15 GB
// algorithmic, L2 hit rate 3.5 %).  The reference has the same shape and says so:
// src/pga.cu:304 "TODO: figure a way to select genomes from shared memory".
//
// Here the generation is split:
//   sorted_select    every child's two tournaments (src/pga.cu:280-292, :306-307) from the
//                    score array alone (4 MiB, L2-resident); writes the pair and counts the
//                    children per first parent
//   sorted_scan_*    exclusive prefix sum of those counts (a counting sort by first parent)
//   sorted_scatter   each child's record (child, p1, p2) goes to its slot: children that
//                    share a first parent are now adjacent
//   generation_sorted  a warp walks 32 consecutive records; the first parent's row is loaded
//                    only when it differs from the previous record's (it is still in
//                    registers otherwise), the second parent's row always; crossover,
//                    mutation, scoring and the best-key reduction as in every other kernel.
// Child i still lands in row i of next_gen with the draws of individual i, so results are
// bit-identical to the child-ordered kernels and to the oracle; only the order of the work
// changes.  The order of equal keys out of the scatter depends on atomics; nothing observable does.
#pragma once

#include "pga_kernels.cuh"

namespace pgak {

#define SORTED_SCAN_BLOCK 1024u
#define SORTED_SCAN_ITEMS 4u
#define SORTED_SCAN_CHUNK (SORTED_SCAN_BLOCK * SORTED_SCAN_ITEMS)  // counts per scan block; also the shift of the block-offset table

// pair[i] = the two tournament winners of child i; count[p1] += 1
__global__ void __launch_bounds__(256) sorted_select(const GenArgs a, uint2* __restrict__ pair, uint32_t* __restrict__ count)
{
    for (uint64_t i64 = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i64 < a.n; i64 += (uint64_t)gridDim.x * blockDim.x) {
        const uint32_t i = (uint32_t)i64;
        const uint4 w = draw_block(a.draw, STREAM_ROW, i, 0);
        const uint32_t p1 = tournament(a.score_in, to_uniform(w.x), to_uniform(w.y), a.nf, a.n);
        const uint32_t p2 = tournament(a.score_in, to_uniform(w.z), to_uniform(w.w), a.nf, a.n);
        if (a.sel) { a.sel[2 * (size_t)i] = (int32_t)p1; a.sel[2 * (size_t)i + 1] = (int32_t)p2; }
        pair[i] = make_uint2(p1, p2);
        atomicAdd(count + p1, 1u);
    }
}

// in place: count[] -> exclusive prefix within each chunk of SORTED_SCAN_CHUNK entries; block_total[b] = the chunk's sum
__global__ void __launch_bounds__(SORTED_SCAN_BLOCK) sorted_scan_local(uint32_t* __restrict__ count, uint32_t n, uint32_t* __restrict__ block_total)
{
    __shared__ uint32_t wsum[32];
    const uint32_t t = threadIdx.x, lane = t & 31u, warp = t >> 5;
    const uint64_t base = (uint64_t)blockIdx.x * SORTED_SCAN_CHUNK + (uint64_t)t * SORTED_SCAN_ITEMS;
    uint32_t v[SORTED_SCAN_ITEMS], sum = 0u;
#pragma unroll
    for (uint32_t k = 0; k < SORTED_SCAN_ITEMS; ++k) {
        v[k] = base + k < n ? count[base + k] : 0u;
        sum += v[k];
    }
    uint32_t inc = sum;
    for (int off = 1; off < 32; off <<= 1) {
        const uint32_t o = __shfl_up_sync(0xFFFFFFFFu, inc, off);
        if ((int)lane >= off) inc += o;
    }
    if (lane == 31u) wsum[warp] = inc;
    __syncthreads();
    if (warp == 0) {
        uint32_t w = wsum[lane], wi = w;
        for (int off = 1; off < 32; off <<= 1) {
            const uint32_t o = __shfl_up_sync(0xFFFFFFFFu, wi, off);
            if ((int)lane >= off) wi += o;
        }
        wsum[lane] = wi - w;  // exclusive over warps
        if (lane == 31u) block_total[blockIdx.x] = wi;
    }
    __syncthreads();
    uint32_t run = wsum[warp] + inc - sum;
#pragma unroll
    for (uint32_t k = 0; k < SORTED_SCAN_ITEMS; ++k) {
        if (base + k < n) count[base + k] = run;
        run += v[k];
    }
}

// one block: block_total[] -> exclusive prefix, in place
__global__ void __launch_bounds__(1024) sorted_scan_totals(uint32_t* __restrict__ block_total, uint32_t nb)
{
    __shared__ uint32_t wsum[32];
    __shared__ uint32_t carry;
    const uint32_t t = threadIdx.x, lane = t & 31u, warp = t >> 5;
    if (t == 0) carry = 0u;
    __syncthreads();
    for (uint32_t b0 = 0; b0 < nb; b0 += 1024u) {
        const uint32_t i = b0 + t;
        const uint32_t v = i < nb ? block_total[i] : 0u;
        uint32_t inc = v;
        for (int off = 1; off < 32; off <<= 1) {
            const uint32_t o = __shfl_up_sync(0xFFFFFFFFu, inc, off);
            if ((int)lane >= off) inc += o;
        }
        if (lane == 31u) wsum[warp] = inc;
        __syncthreads();
        uint32_t before = carry;
        for (uint32_t w2 = 0; w2 < warp; ++w2) before += wsum[w2];
        if (i < nb) block_total[i] = before + inc - v;
        __syncthreads();
        if (t == 1023u) carry = before + inc;
        __syncthreads();
    }
}

// rec[slot] = (child, p1, p2, 0); slot = offset of p1's run + the child's arrival number in it
__global__ void __launch_bounds__(256) sorted_scatter(const uint2* __restrict__ pair, uint32_t n, uint32_t* __restrict__ offset,
                                                      const uint32_t* __restrict__ block_offset, uint4* __restrict__ rec)
{
    for (uint64_t i64 = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i64 < n; i64 += (uint64_t)gridDim.x * blockDim.x) {
        const uint32_t i = (uint32_t)i64;
        const uint2 p = pair[i];
        const uint32_t slot = block_offset[p.x / SORTED_SCAN_CHUNK] + atomicAdd(offset + p.x, 1u);
        rec[slot] = make_uint4(i, p.x, p.y, 0u);
    }
}

// one child of the walk: crossover of the two rows held in registers, mutation, scoring, store
template <int ENC, int OBJ, int NB, bool EXACT>
__device__ __forceinline__ float sorted_child(const GenArgs& a, uint32_t child, bool valid, const uint4 (&x)[NB], const uint4 (&y)[NB],
                                              uint32_t nblk, uint32_t lane, bool c_on, uint32_t c_idx, uint32_t c_val, uint4* __restrict__ nxt)
{
    const uint32_t c_blk = ENC == ENC_BITS ? (c_idx >> 7) : (c_idx >> 2);
    float facc = 0.f;
    uint32_t iacc = 0;
#pragma unroll
    for (int nb = 0; nb < NB; ++nb) {
        const uint32_t b = lane + (uint32_t)nb * 32u;
        if (EXACT || b < nblk) {
            uint4 c;
            if (ENC == ENC_BITS) {
                const uint4 m = draw_block(a.draw, STREAM_MASK, child, b);
                c.x = (x[nb].x & m.x) | (y[nb].x & ~m.x);
                c.y = (x[nb].y & m.y) | (y[nb].y & ~m.y);
                c.z = (x[nb].z & m.z) | (y[nb].z & ~m.z);
                c.w = (x[nb].w & m.w) | (y[nb].w & ~m.w);
            } else {
                const uint4 w = draw_block(a.draw, STREAM_ROW, child, b);
                c.x = draw_above_half(w.x) ? x[nb].x : y[nb].x;  // src/pga.cu:137-141
                c.y = draw_above_half(w.y) ? x[nb].y : y[nb].y;
                c.z = draw_above_half(w.z) ? x[nb].z : y[nb].z;
                c.w = draw_above_half(w.w) ? x[nb].w : y[nb].w;
            }
            if (c_on && c_blk == b) {  // src/pga.cu:127-133; rare
                if (ENC == ENC_BITS) {
                    const uint32_t word = (c_idx >> 5) & 3u, bit = 1u << (c_idx & 31u);
                    c.x = word == 0 ? (c_val ? (c.x | bit) : (c.x & ~bit)) : c.x;
                    c.y = word == 1 ? (c_val ? (c.y | bit) : (c.y & ~bit)) : c.y;
                    c.z = word == 2 ? (c_val ? (c.z | bit) : (c.z & ~bit)) : c.z;
                    c.w = word == 3 ? (c_val ? (c.w | bit) : (c.w & ~bit)) : c.w;
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
            if (valid) store_row_block(nxt + (size_t)child * nblk + b, c);
        }
    }
    if (ENC == ENC_BITS) {
#pragma unroll
        for (uint32_t off = 16u; off >= 1u; off >>= 1) iacc += __shfl_xor_sync(0xFFFFFFFFu, iacc, off);
        return __uint2float_rn(iacc);
    }
#pragma unroll
    for (uint32_t off = 16u; off >= 1u; off >>= 1) facc = __fadd_rn(facc, __shfl_xor_sync(0xFFFFFFFFu, facc, off));
    return obj_finalize<OBJ>(facc, a.len);
}

// 32 lanes per child, NB 16-byte blocks per lane (rows of 16*32*(NB-1)+16 .. 512*NB bytes).
// One block per SM of 8*MINB warps, grid-stride over groups of 32 records.
template <int ENC, int OBJ, int NB, bool EXACT, int MINB>
__global__ void __launch_bounds__(256 * MINB, 1) generation_sorted(const __grid_constant__ GenArgs a, const uint4* __restrict__ rec)
{
    __shared__ unsigned long long warp_best[8 * MINB];
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t gwarp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const uint32_t nwarps = (gridDim.x * blockDim.x) >> 5;
    const uint32_t nblk = EXACT ? (uint32_t)NB * 32u : a.stride >> 2;
    const uint4* __restrict__ cur = reinterpret_cast<const uint4*>(a.cur);
    uint4* __restrict__ nxt = reinterpret_cast<uint4*>(a.nxt);

    if (blockIdx.x == 0 && threadIdx.x == 0) a.best[(a.best_slot + 1u) & 3u] = 0ull;

    unsigned long long key = 0ull;
    const uint32_t groups = (a.n + 31u) >> 5;
    for (uint32_t g = gwarp; g < groups; g += nwarps) {
        const uint32_t e = g * 32u + lane;
        const bool have = e < a.n;
        const uint4 r = have ? __ldg(rec + e) : make_uint4(0u, 0u, 0u, 0u);  // (child, first parent, second parent, -)
        // lane j: the mutation draws of record j's child (Philox block 0 of its rand row)
        uint32_t m_idx = 0, m_val = 0;
        bool m_on = false;
        if (have) {
            const uint4 w = draw_block(a.draw, STREAM_ROW, r.x, 0);
            const float u0 = to_uniform(w.x), u1 = to_uniform(w.y), u2 = to_uniform(w.z);
            m_on = u1 <= PGA_MUTATION_CHANCE;
            m_idx = draw_index(u0, a.lenf, a.len);
            m_val = ENC == ENC_BITS ? (u2 > 0.5f ? 1u : 0u) : __float_as_uint(u2);
        }
        const uint32_t any_mut = __ballot_sync(0xFFFFFFFFu, m_on);

        uint32_t prev = 0xFFFFFFFFu;  // first parent whose row xp holds
        uint4 xp[NB];
        float my_score = 0.f;
#pragma unroll 1
        for (uint32_t t = 0; t < 32u; t += 2u) {
            const uint32_t ca = __shfl_sync(0xFFFFFFFFu, r.x, t), cb = __shfl_sync(0xFFFFFFFFu, r.x, t + 1u);
            const uint32_t q1a = __shfl_sync(0xFFFFFFFFu, r.y, t), q1b = __shfl_sync(0xFFFFFFFFu, r.y, t + 1u);
            const uint32_t q2a = __shfl_sync(0xFFFFFFFFu, r.z, t), q2b = __shfl_sync(0xFFFFFFFFu, r.z, t + 1u);
            const bool va = g * 32u + t < a.n, vb = g * 32u + t + 1u < a.n;
            const bool la = q1a != prev, lb = q1b != q1a;  // warp-uniform: does the first parent change?
            uint4 xa[NB], ya[NB], xb[NB], yb[NB];
#pragma unroll
            for (int nb = 0; nb < NB; ++nb) {
                const uint32_t b = lane + (uint32_t)nb * 32u;
                if (EXACT || b < nblk) {
                    if (la) xa[nb] = load_row_block(cur + (size_t)q1a * nblk + b);
                    ya[nb] = load_row_block(cur + (size_t)q2a * nblk + b);
                    if (lb) xb[nb] = load_row_block(cur + (size_t)q1b * nblk + b);
                    yb[nb] = load_row_block(cur + (size_t)q2b * nblk + b);
                }
            }
            uint32_t ia = 0, ib = 0, wa = 0, wb = 0;
            if (any_mut) {  // warp-uniform
                ia = __shfl_sync(0xFFFFFFFFu, m_idx, t); ib = __shfl_sync(0xFFFFFFFFu, m_idx, t + 1u);
                wa = __shfl_sync(0xFFFFFFFFu, m_val, t); wb = __shfl_sync(0xFFFFFFFFu, m_val, t + 1u);
            }
            if (!la) {
#pragma unroll
                for (int nb = 0; nb < NB; ++nb) xa[nb] = xp[nb];
            }
            const float sa = sorted_child<ENC, OBJ, NB, EXACT>(a, ca, va, xa, ya, nblk, lane, (any_mut >> t) & 1u, ia, wa, nxt);
            if (!lb) {
#pragma unroll
                for (int nb = 0; nb < NB; ++nb) xb[nb] = xa[nb];
            }
            const float sb = sorted_child<ENC, OBJ, NB, EXACT>(a, cb, vb, xb, yb, nblk, lane, (any_mut >> (t + 1u)) & 1u, ib, wb, nxt);
#pragma unroll
            for (int nb = 0; nb < NB; ++nb) xp[nb] = xb[nb];
            prev = q1b;
            if (lane == t) my_score = sa;
            if (lane == t + 1u) my_score = sb;
        }
        if (have) {
            a.score_out[r.x] = my_score;
            const unsigned long long k = best_key(my_score, r.x);
            key = k > key ? k : key;
        }
    }
    block_best_commit(key, a.best + a.best_slot, warp_best);
}

}  // namespace pgak
