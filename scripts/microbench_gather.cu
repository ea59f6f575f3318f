// Microbenchmark: what does the parent-gather access pattern cost on this GPU, with
// no RNG or fitness work in the way?  child[i] = rowA[ia(i)] ^ rowB[ib(i)], indices
// pseudo-random over the whole population, rows of ROWB bytes, UNROLL children per
// lane-group in flight.  Prints requested GB/s (2 reads + 1 write per child).
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/mb scripts/microbench_gather.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s %d\n", cudaGetErrorString(e), __LINE__); return 1; } } while (0)

__device__ __forceinline__ uint32_t mix(uint32_t x) { x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16; return x; }

template <int LOGG, int NB, int UNROLL>
__global__ void __launch_bounds__(256) gather(const uint4* __restrict__ cur, uint4* __restrict__ nxt, uint32_t n, uint32_t salt)
{
    constexpr uint32_t G = 1u << LOGG, CPT = 32u / G, NBLK = G * NB;
    const uint32_t lane = threadIdx.x & 31u, gl = lane & (G - 1u), grp = lane >> LOGG;
    const uint32_t gwarp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, nwarps = (gridDim.x * blockDim.x) >> 5;
    for (uint64_t base = (uint64_t)gwarp * CPT * UNROLL; base < n; base += (uint64_t)nwarps * CPT * UNROLL) {
        uint4 x[UNROLL][NB], y[UNROLL][NB];
#pragma unroll
        for (int u = 0; u < UNROLL; ++u) {
            const uint32_t child = (uint32_t)base + u * CPT + grp;
            const uint32_t a = mix(child * 2u + salt) % n, b = mix(child * 2u + 1u + salt) % n;
#pragma unroll
            for (int nb = 0; nb < NB; ++nb) {
                x[u][nb] = __ldg(cur + (size_t)a * NBLK + gl + nb * G);
                y[u][nb] = __ldg(cur + (size_t)b * NBLK + gl + nb * G);
            }
        }
#pragma unroll
        for (int u = 0; u < UNROLL; ++u) {
            const uint32_t child = (uint32_t)base + u * CPT + grp;
            if (child < n)
#pragma unroll
                for (int nb = 0; nb < NB; ++nb) {
                    uint4 c = make_uint4(x[u][nb].x ^ y[u][nb].x, x[u][nb].y ^ y[u][nb].y, x[u][nb].z ^ y[u][nb].z, x[u][nb].w ^ y[u][nb].w);
                    nxt[(size_t)child * NBLK + gl + nb * G] = c;
                }
        }
    }
}

template <int LOGG, int NB, int UNROLL>
int run(uint32_t n, const char* name, int blocks_per_sm)
{
    constexpr size_t rowb = (size_t)(1u << LOGG) * NB * 16;
    uint4 *a, *b;
    CK(cudaMalloc(&a, n * rowb)); CK(cudaMalloc(&b, n * rowb));
    CK(cudaMemset(a, 1, n * rowb)); CK(cudaMemset(b, 2, n * rowb));
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    int sms; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    const int grid = sms * blocks_per_sm;
    for (int i = 0; i < 4; ++i) { gather<LOGG, NB, UNROLL><<<grid, 256>>>(a, b, n, i); uint4* t = a; a = b; b = t; }
    CK(cudaDeviceSynchronize());
    const int reps = 20;
    cudaEventRecord(e0);
    for (int i = 0; i < reps; ++i) { gather<LOGG, NB, UNROLL><<<grid, 256>>>(a, b, n, 100 + i); uint4* t = a; a = b; b = t; }
    cudaEventRecord(e1);
    CK(cudaDeviceSynchronize());
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    const double us = 1e3 * ms / reps;
    printf("%-28s rows=%u rowB=%zu unroll=%d bps=%d : %8.1f us/pass  requested(3 passes) %7.1f GB/s  algorithmic(2 passes) %7.1f GB/s\n", name, n,
           rowb, UNROLL, blocks_per_sm, us, 3.0 * n * rowb / us / 1e3, 2.0 * n * rowb / us / 1e3);
    cudaFree(a); cudaFree(b);
    return 0;
}

int main()
{
    const uint32_t n = 1u << 20;
    for (int bps : {2, 4, 8}) {
        run<3, 1, 1>(n, "128B rows", bps);
        run<3, 1, 2>(n, "128B rows", bps);
        run<3, 1, 4>(n, "128B rows", bps);
        run<3, 1, 8>(n, "128B rows", bps);
    }
    for (int bps : {2, 4, 8}) {
        run<5, 2, 1>(n, "1KiB rows", bps);
        run<5, 2, 2>(n, "1KiB rows", bps);
        run<5, 2, 4>(n, "1KiB rows", bps);
    }
    run<5, 1, 4>(1u << 22, "512B rows 2^22", 4);
    run<3, 1, 4>(1u << 22, "128B rows 2^22", 4);
    return 0;
}
