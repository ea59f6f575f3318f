// Microbenchmark 2: per-warp double-buffered staging of parent rows in shared memory,
// (a) with one TMA bulk copy (cp.async.bulk) per row, (b) with 16-byte cp.async (LDGSTS).
// Same access pattern as microbench_gather.cu: child = rowA[ia] ^ rowB[ib], 128 B rows.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s %d\n", cudaGetErrorString(e), __LINE__); return 1; } } while (0)
__device__ __forceinline__ uint32_t mix(uint32_t x) { x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16; return x; }
__device__ __forceinline__ uint32_t sa(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// ROWB bytes per row; a warp handles 32 children per iteration; stage = 32*2*ROWB bytes
template <int ROWB, int WARPS, int MODE>
__global__ void __launch_bounds__(WARPS * 32) staged(const uint4* __restrict__ cur, uint4* __restrict__ nxt, uint32_t n, uint32_t salt)
{
    constexpr int NBLK = ROWB / 16;           // 16-byte blocks per row
    constexpr int G = NBLK < 32 ? NBLK : 32;  // lanes per child when consuming
    constexpr int CPT = 32 / G;
    constexpr int STAGE = 32 * 2 * ROWB;
    extern __shared__ __align__(128) unsigned char smem[];
    __shared__ __align__(8) uint64_t bars[WARPS][2];
    const uint32_t lane = threadIdx.x & 31u, w = threadIdx.x >> 5;
    unsigned char* my = smem + (size_t)w * 2 * STAGE;
    const uint32_t gwarp = blockIdx.x * WARPS + w, nwarps = gridDim.x * WARPS;
    if (MODE == 0 && lane == 0) {
        for (int s = 0; s < 2; ++s) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(sa(&bars[w][s])));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncwarp();
    const uint32_t iters = (n / 32 + nwarps - 1 - gwarp) / nwarps;  // n multiple of 32
    auto issue = [&](uint32_t it) {
        const uint32_t base = (gwarp + it * nwarps) * 32u;
        unsigned char* st = my + (it & 1) * STAGE;
        if (MODE == 0) {
            const uint32_t child = base + lane;
            const uint32_t a = mix(child * 2u + salt) % n, b = mix(child * 2u + 1u + salt) % n;
            if (lane == 0) asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(sa(&bars[w][it & 1])), "r"(STAGE) : "memory");
            __syncwarp();
            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(sa(st + (lane * 2 + 0) * ROWB)),
                         "l"(cur + (size_t)a * NBLK), "r"(ROWB), "r"(sa(&bars[w][it & 1])) : "memory");
            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(sa(st + (lane * 2 + 1) * ROWB)),
                         "l"(cur + (size_t)b * NBLK), "r"(ROWB), "r"(sa(&bars[w][it & 1])) : "memory");
        } else {
            // every lane copies 16 B; 32 lanes cover 512 B = (512/ROWB) rows per instruction
            constexpr int ROWS_PER_OP = 512 / ROWB > 0 ? 512 / ROWB : 1;
            for (int r0 = 0; r0 < 64; r0 += ROWS_PER_OP) {
                const uint32_t r = r0 + lane / NBLK;   // row slot 0..63 : child = r/2, parent = r&1
                const uint32_t child = base + (r >> 1);
                const uint32_t src = mix(child * 2u + (r & 1u) + salt) % n;
                asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(sa(st + r * ROWB + (lane % NBLK) * 16)),
                             "l"(cur + (size_t)src * NBLK + (lane % NBLK)) : "memory");
            }
            asm volatile("cp.async.commit_group;" ::: "memory");
        }
    };
    if (iters) issue(0);
    for (uint32_t it = 0; it < iters; ++it) {
        if (it + 1 < iters) {
            if (MODE == 0) asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            issue(it + 1);
        }
        if (MODE == 0) {
            uint32_t done;
            do {
                asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                             : "=r"(done) : "r"(sa(&bars[w][it & 1])), "r"((it >> 1) & 1u) : "memory");
            } while (!done);
        } else {
            if (it + 1 < iters) asm volatile("cp.async.wait_group 1;" ::: "memory");
            else asm volatile("cp.async.wait_group 0;" ::: "memory");
            __syncwarp();
        }
        const uint32_t base = (gwarp + it * nwarps) * 32u;
        const unsigned char* st = my + (it & 1) * STAGE;
        const uint32_t gl = lane % G, grp = lane / G;
        for (int tile = 0; tile < G; ++tile) {
            const uint32_t c = tile * CPT + grp;
            for (int b = gl; b < NBLK; b += G) {
                const uint4 x = *reinterpret_cast<const uint4*>(st + (c * 2 + 0) * ROWB + b * 16);
                const uint4 y = *reinterpret_cast<const uint4*>(st + (c * 2 + 1) * ROWB + b * 16);
                nxt[(size_t)(base + c) * NBLK + b] = make_uint4(x.x ^ y.x, x.y ^ y.y, x.z ^ y.z, x.w ^ y.w);
            }
        }
        __syncwarp();
    }
}

template <int ROWB, int WARPS, int MODE>
int run(uint32_t n, int blocks_per_sm)
{
    uint4 *a, *b;
    CK(cudaMalloc(&a, (size_t)n * ROWB)); CK(cudaMalloc(&b, (size_t)n * ROWB));
    CK(cudaMemset(a, 1, (size_t)n * ROWB)); CK(cudaMemset(b, 2, (size_t)n * ROWB));
    const size_t smem = (size_t)WARPS * 2 * 32 * 2 * ROWB;
    CK(cudaFuncSetAttribute(staged<ROWB, WARPS, MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int sms; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    const int grid = sms * blocks_per_sm;
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    for (int i = 0; i < 4; ++i) { staged<ROWB, WARPS, MODE><<<grid, WARPS * 32, smem>>>(a, b, n, i); uint4* t = a; a = b; b = t; }
    CK(cudaDeviceSynchronize());
    const int reps = 20;
    cudaEventRecord(e0);
    for (int i = 0; i < reps; ++i) { staged<ROWB, WARPS, MODE><<<grid, WARPS * 32, smem>>>(a, b, n, 100 + i); uint4* t = a; a = b; b = t; }
    cudaEventRecord(e1);
    CK(cudaDeviceSynchronize());
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    const double us = 1e3 * ms / reps;
    printf("%s rowB=%d warps/block=%d blocks/SM=%d smem/block=%zu : %8.1f us/pass  algorithmic(2 passes) %7.1f GB/s\n", MODE == 0 ? "TMA-bulk" : "cp.async",
           ROWB, WARPS, blocks_per_sm, smem, us, 2.0 * n * ROWB / us / 1e3);
    cudaFree(a); cudaFree(b);
    return 0;
}

int main()
{
    const uint32_t n = 1u << 20;
    run<128, 4, 0>(n, 1); run<128, 4, 0>(n, 2); run<128, 4, 0>(n, 3);
    run<128, 8, 0>(n, 1); run<128, 12, 0>(n, 1);
    run<128, 4, 1>(n, 1); run<128, 4, 1>(n, 2); run<128, 4, 1>(n, 3);
    run<128, 8, 1>(n, 1); run<128, 12, 1>(n, 1);
    run<1024, 1, 0>(n, 1); run<1024, 1, 0>(n, 2); run<1024, 1, 0>(n, 3);
    run<1024, 1, 1>(n, 1); run<1024, 1, 1>(n, 3);
    run<512, 2, 0>(n, 1); run<512, 2, 0>(n, 3);
    return 0;
}
