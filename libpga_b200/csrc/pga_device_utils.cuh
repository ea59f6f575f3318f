// pga_device_utils.cuh -- device-side building blocks of the generation kernels:
// the counter-based generator, the draw layout, sm_100a bulk-copy (TMA) wrappers,
// the ranking key and the built-in objectives.
//
// The draw layout replaces the reference's per-generation cuRAND buffer
// (src/pga.cu:99-101): nothing is stored, every draw is recomputed from
// (seed, island, epoch, individual, position) wherever it is needed.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

// Bounds checks for the debug build (make debug: -DPGA_DEBUG, libpga_b200_dbg.so).  The GPU
// pool has no compute-sanitizer, so every index that addresses a population row, a score or an
// output slot is asserted there; the release build compiles them away.
#ifdef PGA_DEBUG
#include <assert.h>
#define PGA_ASSERT(x) assert(x)
#else
#define PGA_ASSERT(x) ((void)0)
#endif

namespace pgak {

// ---- Philox4x32-10 --------------------------------------------------------
__device__ __forceinline__ uint4 philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3,
                                               uint32_t k0, uint32_t k1)
{
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        // one IMAD.WIDE.U32 per product: high and low halves together
        const unsigned long long m0 = (unsigned long long)0xD2511F53u * c0;
        const unsigned long long m1 = (unsigned long long)0xCD9E8D57u * c2;
        const uint32_t n0 = (uint32_t)(m1 >> 32) ^ c1 ^ k0;
        const uint32_t n2 = (uint32_t)(m0 >> 32) ^ c3 ^ k1;
        c0 = n0; c1 = (uint32_t)m1; c2 = n2; c3 = (uint32_t)m0;
        k0 += 0x9E3779B9u;
        k1 += 0xBB67AE85u;
    }
    return make_uint4(c0, c1, c2, c3);
}

enum : uint32_t { STREAM_ROW = 0u, STREAM_MASK = 1u, STREAM_INITBITS = 2u, STREAM_MIGRATE = 3u };

// the identity of a launch's draws: key = seed, counter words 2 and 3
struct DrawCtx {
    uint32_t k0, k1;  // seed
    uint32_t epoch;   // counter word 2
    uint32_t island;  // low 24 bits of counter word 3
};

// Philox block b (words 4b..4b+3) of `stream` for one individual
__device__ __forceinline__ uint4 draw_block(const DrawCtx& d, uint32_t stream, uint32_t individual, uint32_t b)
{
    return philox4x32_10(b, individual, d.epoch, (stream << 24) | (d.island & 0xFFFFFFu), d.k0, d.k1);
}

// uniform in [0,1) from the top 24 bits (exact)
__device__ __forceinline__ float to_uniform(uint32_t x)
{
    return __uint2float_rn(x >> 8) * 0x1.0p-24f;
}

// (int)(u * n), the index draw of src/pga.cu:284 and :129, clamped into range
__device__ __forceinline__ uint32_t draw_index(float u, float nf, uint32_t n)
{
    const uint32_t v = __float2uint_rz(__fmul_rn(u, nf));
    return v < n ? v : n - 1u;
}

// rand > 0.5 (src/pga.cu:137) on the raw word: (x >> 8) * 2^-24 > 0.5  <=>  x >= 0x80000100
__device__ __forceinline__ bool draw_above_half(uint32_t x)
{
    return x > 0x800000FFu;
}

// ---- population row accesses ---------------------------------------------------
// Parent rows: read-only path, no L1 allocation (a row is used once per child).
// L2 eviction-priority hints (createpolicy evict_last on the gathers, evict_first
// on the child stores, and the other combinations) were measured on B200 and made
// no difference to time or DRAM traffic (profiles/r1_sweeps.md), so none are used.
// Parent rows bypass L1 (each is read once per SM).  In L2 a bare L1::no_allocate load is
// treated as evict-first too -- it times exactly like an explicit L2::evict_first policy --
// which throws away what L2 can reuse: a winner is the parent of several children, and the
// generation being written turns into the one being read.  With KEEP the load names the
// normal L2 policy instead.  Measured per generation (profiles/r2_l2_hint_forms.log,
// r2_ab_row_loads.log): 128-byte rows 72.0 -> 66.0 us at 2^20, 144.8 -> 133.4 at 2^21,
// 677.8 -> 662.6 at 2^23; 512-byte rows 266.1 -> 251.4; 768-byte rows 538.4 -> 533.1;
// rows of 1 KiB go the other way (525.3 -> 538.1) and keep the bare form.
template <bool KEEP>
__device__ __forceinline__ uint4 load_row_block(const uint4* p)
{
    uint4 v;
    if (KEEP) {
        uint64_t pol;
        asm("createpolicy.fractional.L2::evict_normal.b64 %0, 1.0;" : "=l"(pol));
        asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.v4.u32 {%0,%1,%2,%3}, [%4], %5;"
                     : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
                     : "l"(p), "l"(pol));
    } else {
        asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
    }
    return v;
}

__device__ __forceinline__ void store_row_block(uint4* p, const uint4& v)
{
    asm volatile("st.global.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

// ---- ranking key: higher score first, then lower index; NaN below everything ----
__device__ __forceinline__ uint32_t order_key(float s)
{
    if (s != s) return 0u;
    s = s + 0.0f;  // -0 -> +0
    const uint32_t b = __float_as_uint(s);
    return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}

__device__ __forceinline__ unsigned long long best_key(float s, uint32_t index)
{
    return ((unsigned long long)order_key(s) << 32) | (unsigned long long)(0xFFFFFFFFu - index);
}

__device__ __forceinline__ unsigned long long warp_max_u64(unsigned long long v)
{
#pragma unroll
    for (int off = 16; off >= 1; off >>= 1) {
        const unsigned long long o = __shfl_xor_sync(0xFFFFFFFFu, v, off);
        v = o > v ? o : v;
    }
    return v;
}

// ---- sm_100a async bulk copies (TMA, SASS: UBLKCP) and mbarrier ------------
__device__ __forceinline__ uint32_t smem_addr(const void* p)
{
    return (uint32_t)__cvta_generic_to_shared(p);
}

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_addr(bar)), "r"(count) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}

__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_addr(bar)), "r"(bytes) : "memory");
}

__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity)
{
    uint32_t done;
    do {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(done)
            : "r"(smem_addr(bar)), "r"(parity)
            : "memory");
    } while (!done);
}

// global -> shared, completion counted in bytes on the mbarrier
__device__ __forceinline__ void bulk_load(void* smem_dst, const void* gmem_src, uint32_t bytes, uint64_t* bar)
{
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_addr(smem_dst)),
        "l"(gmem_src), "r"(bytes), "r"(smem_addr(bar))
        : "memory");
}

// shared -> global, tracked by the bulk async-group
__device__ __forceinline__ void bulk_store(void* gmem_dst, const void* smem_src, uint32_t bytes)
{
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gmem_dst), "r"(smem_addr(smem_src)),
                 "r"(bytes)
                 : "memory");
}

__device__ __forceinline__ void bulk_store_commit_and_wait()
{
    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
    asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
}

// make generic-proxy writes to shared memory visible to the async proxy
__device__ __forceinline__ void fence_async_smem()
{
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}

// ---- built-in objectives --------------------------------------------------
// Operation order is part of the specification ("canonical reduction
// shape"): explicit _rn intrinsics keep nvcc from contracting, so float scores are
// reproducible bit for bit on any conforming implementation.
enum : int { OBJ_HOOK = 0, OBJ_SUM = 1, OBJ_RASTRIGIN = 2, OBJ_SPHERE = 3, OBJ_ONEMAX = 4 };

__device__ __forceinline__ float cos2pi(float x)
{
    const float k = rintf(x);
    const float t = __fsub_rn(x, k);  // exact, |t| <= 0.5
    const float a = fabsf(t);
    const bool flip = a > 0.25f;
    const float r = flip ? __fsub_rn(0.5f, a) : a;
    const float s = __fmul_rn(r, r);
    float p = 0x1.f9d38ap+2f;
    p = __fmaf_rn(p, s, -0x1.a6d1f2p+4f);
    p = __fmaf_rn(p, s, 0x1.e1f506p+5f);
    p = __fmaf_rn(p, s, -0x1.55d3c8p+6f);
    p = __fmaf_rn(p, s, 0x1.03c1f0p+6f);
    p = __fmaf_rn(p, s, -0x1.3bd3ccp+4f);
    p = __fmaf_rn(p, s, 1.0f);
    return flip ? -p : p;
}

template <int OBJ>
__device__ __forceinline__ float obj_accumulate(float acc, float g)
{
    if (OBJ == OBJ_SUM) return __fadd_rn(acc, g);
    const float x = __fmul_rn(__fsub_rn(g, 0.5f), 10.24f);
    if (OBJ == OBJ_SPHERE) return __fmaf_rn(x, x, acc);
    const float c = cos2pi(x);
    const float term = __fmaf_rn(x, x, -__fmul_rn(10.0f, c));
    return __fadd_rn(acc, term);
}

template <int OBJ>
__device__ __forceinline__ float obj_finalize(float total, uint32_t len)
{
    if (OBJ == OBJ_SUM) return total;
    if (OBJ == OBJ_SPHERE) return -total;
    return -__fadd_rn(__uint2float_rn(10u * len), total);
}

__device__ __forceinline__ float obj_accumulate_rt(int obj, float acc, float g)
{
    switch (obj) {
        case OBJ_SUM: return obj_accumulate<OBJ_SUM>(acc, g);
        case OBJ_SPHERE: return obj_accumulate<OBJ_SPHERE>(acc, g);
        default: return obj_accumulate<OBJ_RASTRIGIN>(acc, g);
    }
}

// the canonical reduction evaluated by ONE thread (hook-path kernels that score
// with a built-in objective): G virtual lanes, blocks v, v+G, ..., xor butterfly
__device__ inline float objective_serial(int obj, const float* row, uint32_t len)
{
    const uint32_t nblk = (len + 3u) >> 2;
    uint32_t G = 1;
    while (G < nblk && G < 32u) G <<= 1;
    float part[32];
    for (uint32_t v = 0; v < G; ++v) {
        float acc = 0.0f;
        for (uint32_t b = v; b < nblk; b += G)
            for (uint32_t e = 0; e < 4u; ++e)
                if (4u * b + e < len) acc = obj_accumulate_rt(obj, acc, row[4u * b + e]);
        part[v] = acc;
    }
    for (uint32_t off = G >> 1; off >= 1u; off >>= 1) {
        float tmp[32];
        for (uint32_t v = 0; v < G; ++v) tmp[v] = __fadd_rn(part[v], part[v ^ off]);
        for (uint32_t v = 0; v < G; ++v) part[v] = tmp[v];
    }
    switch (obj) {
        case OBJ_SUM: return obj_finalize<OBJ_SUM>(part[0], len);
        case OBJ_SPHERE: return obj_finalize<OBJ_SPHERE>(part[0], len);
        default: return obj_finalize<OBJ_RASTRIGIN>(part[0], len);
    }
}

}  // namespace pgak
