// Round-2 microbenchmarks: which unit bounds the OneMax generation kernel?
//
//   A  the parent-gather pattern (child = rowA[ia] ^ rowB[ib], 128 B rows) at population
//      sizes from L2-resident (2^17, 2^18) to DRAM-resident (2^20 .. 2^22): if ns/child does
//      not fall when everything fits in L2, the bound is the SM<->L2 path, not DRAM
//   B  read-only random 128 B row gathers from an L2-resident buffer: the practical L2->SM ceiling
//   C  random 4-byte gathers (one 32 B sector each) from a 4 MiB array: the score-gather rate
//   D  streaming 128 B row writes
//   E  the staged pattern of microbench_stage.cu with cp.async.bulk.tensor.2d ... tile::gather4
//      (four independent row indices per TMA instruction, sm_100 only), checked against
//      the plain-load result
//
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o scripts/microbench_r2.bin scripts/microbench_r2.cu
#include <cstdio>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <cuda.h>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at line %d\n", cudaGetErrorString(e), __LINE__); exit(1); } } while (0)

__device__ __forceinline__ uint32_t mix(uint32_t x) { x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16; return x; }
__device__ __forceinline__ uint32_t sa(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

static int g_sms = 148;

struct Timer {
    cudaEvent_t e0, e1;
    Timer() { cudaEventCreate(&e0); cudaEventCreate(&e1); }
    void start() { cudaEventRecord(e0); }
    double stop_us(int reps) { cudaEventRecord(e1); CK(cudaDeviceSynchronize()); float ms; cudaEventElapsedTime(&ms, e0, e1); return 1e3 * ms / reps; }
};

// ---- A: gather pattern ---------------------------------------------------------------------
template <int UNROLL>
__global__ void __launch_bounds__(256) gather(const uint4* __restrict__ cur, uint4* __restrict__ nxt, uint32_t n, uint32_t salt)
{
    constexpr uint32_t G = 8, CPT = 4, NBLK = 8;
    const uint32_t lane = threadIdx.x & 31u, gl = lane & (G - 1u), grp = lane >> 3;
    const uint32_t gwarp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, nwarps = (gridDim.x * blockDim.x) >> 5;
    for (uint64_t base = (uint64_t)gwarp * CPT * UNROLL; base < n; base += (uint64_t)nwarps * CPT * UNROLL) {
        uint4 x[UNROLL], y[UNROLL];
#pragma unroll
        for (int u = 0; u < UNROLL; ++u) {
            const uint32_t child = (uint32_t)base + u * CPT + grp;
            const uint32_t a = mix(child * 2u + salt) % n, b = mix(child * 2u + 1u + salt) % n;
            x[u] = __ldg(cur + (size_t)a * NBLK + gl);
            y[u] = __ldg(cur + (size_t)b * NBLK + gl);
        }
#pragma unroll
        for (int u = 0; u < UNROLL; ++u) {
            const uint32_t child = (uint32_t)base + u * CPT + grp;
            if (child < n) nxt[(size_t)child * NBLK + gl] = make_uint4(x[u].x ^ y[u].x, x[u].y ^ y[u].y, x[u].z ^ y[u].z, x[u].w ^ y[u].w);
        }
    }
}

static void run_gather(uint32_t n)
{
    uint4 *a, *b;
    CK(cudaMalloc(&a, (size_t)n * 128)); CK(cudaMalloc(&b, (size_t)n * 128));
    CK(cudaMemset(a, 1, (size_t)n * 128)); CK(cudaMemset(b, 2, (size_t)n * 128));
    Timer t;
    for (int bps : {4, 8}) {
        const int grid = g_sms * bps;
        for (int i = 0; i < 6; ++i) { gather<8><<<grid, 256>>>(a, b, n, i); uint4* s = a; a = b; b = s; }
        CK(cudaDeviceSynchronize());
        const int reps = 40;
        t.start();
        for (int i = 0; i < reps; ++i) { gather<8><<<grid, 256>>>(a, b, n, 100 + i); uint4* s = a; a = b; b = s; }
        const double us = t.stop_us(reps);
        printf("A gather 2r+1w rows=2^%d (%.0f MiB/buffer) bps=%d : %8.2f us/pass  %6.2f ns/1k children  requested %7.1f GB/s\n",
               31 - __builtin_clz(n), n * 128.0 / 1048576.0, bps, us, 1e6 * us / n, 3.0 * n * 128 / us / 1e3);
    }
    cudaFree(a); cudaFree(b);
}

// ---- B: read-only row gathers --------------------------------------------------------------------
template <int UNROLL>
__global__ void __launch_bounds__(256) read_rows(const uint4* __restrict__ cur, uint32_t n, uint32_t total, uint32_t salt, uint4* sink)
{
    const uint32_t lane = threadIdx.x & 31u, gl = lane & 7u, grp = lane >> 3;
    const uint32_t gwarp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, nwarps = (gridDim.x * blockDim.x) >> 5;
    uint4 acc = make_uint4(0, 0, 0, 0);
    for (uint64_t base = (uint64_t)gwarp * 4 * UNROLL; base < total; base += (uint64_t)nwarps * 4 * UNROLL) {
        uint4 x[UNROLL];
#pragma unroll
        for (int u = 0; u < UNROLL; ++u) {
            const uint32_t r = mix((uint32_t)base + u * 4 + grp + salt) % n;
            x[u] = __ldg(cur + (size_t)r * 8 + gl);
        }
#pragma unroll
        for (int u = 0; u < UNROLL; ++u) { acc.x ^= x[u].x; acc.y ^= x[u].y; acc.z ^= x[u].z; acc.w ^= x[u].w; }
    }
    if (acc.x == 0x12345678u && acc.y == 0x9abcdef0u) sink[threadIdx.x] = acc;  // never true: keeps the loads alive
}

static void run_read_rows(uint32_t n, uint32_t total)
{
    uint4 *a, *sink;
    CK(cudaMalloc(&a, (size_t)n * 128)); CK(cudaMalloc(&sink, 4096));
    CK(cudaMemset(a, 1, (size_t)n * 128));
    Timer t;
    for (int bps : {4, 8}) {
        const int grid = g_sms * bps;
        for (int i = 0; i < 4; ++i) read_rows<16><<<grid, 256>>>(a, n, total, i, sink);
        CK(cudaDeviceSynchronize());
        const int reps = 20;
        t.start();
        for (int i = 0; i < reps; ++i) read_rows<16><<<grid, 256>>>(a, n, total, 100 + i, sink);
        const double us = t.stop_us(reps);
        printf("B read-only random 128 B rows, buffer 2^%d rows (%.0f MiB), %u rows read, bps=%d : %8.2f us  %7.1f GB/s L2->SM\n",
               31 - __builtin_clz(n), n * 128.0 / 1048576.0, total, bps, us, (double)total * 128 / us / 1e3);
    }
    cudaFree(a); cudaFree(sink);
}

// ---- C: random 4-byte gathers ------------------------------------------------------------------
template <int UNROLL>
__global__ void __launch_bounds__(256) read_words(const float* __restrict__ s, uint32_t n, uint32_t total, uint32_t salt, float* sink)
{
    const uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x, nthreads = gridDim.x * blockDim.x;
    float acc = 0.f;
    for (uint64_t base = (uint64_t)tid * UNROLL; base < total; base += (uint64_t)nthreads * UNROLL) {
        float x[UNROLL];
#pragma unroll
        for (int u = 0; u < UNROLL; ++u) x[u] = __ldg(s + mix((uint32_t)base + u + salt) % n);
#pragma unroll
        for (int u = 0; u < UNROLL; ++u) acc += x[u];
    }
    if (acc == 123456.789f) sink[threadIdx.x] = acc;
}

static void run_read_words(uint32_t n, uint32_t total)
{
    float *a, *sink;
    CK(cudaMalloc(&a, (size_t)n * 4)); CK(cudaMalloc(&sink, 4096));
    CK(cudaMemset(a, 0, (size_t)n * 4));
    Timer t;
    const int grid = g_sms * 8;
    for (int i = 0; i < 4; ++i) read_words<8><<<grid, 256>>>(a, n, total, i, sink);
    CK(cudaDeviceSynchronize());
    const int reps = 20;
    t.start();
    for (int i = 0; i < reps; ++i) read_words<8><<<grid, 256>>>(a, n, total, 100 + i, sink);
    const double us = t.stop_us(reps);
    printf("C random 4-byte gathers from %.1f MiB, %u gathers : %8.2f us  %6.2f Ggathers/s  = %7.1f GB/s of 32 B sectors\n", n * 4.0 / 1048576.0, total, us,
           total / us / 1e3, (double)total * 32 / us / 1e3);
    cudaFree(a); cudaFree(sink);
}

// ---- D: streaming row writes ---------------------------------------------------------------------
__global__ void __launch_bounds__(256) write_rows(uint4* __restrict__ out, uint32_t n, uint32_t salt)
{
    const uint64_t tid = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x, nthreads = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t i = tid; i < (uint64_t)n * 8; i += nthreads) out[i] = make_uint4(salt, (uint32_t)i, salt, salt);
}

static void run_write_rows(uint32_t n)
{
    uint4* a;
    CK(cudaMalloc(&a, (size_t)n * 128));
    Timer t;
    const int grid = g_sms * 8;
    for (int i = 0; i < 4; ++i) write_rows<<<grid, 256>>>(a, n, i);
    CK(cudaDeviceSynchronize());
    const int reps = 20;
    t.start();
    for (int i = 0; i < reps; ++i) write_rows<<<grid, 256>>>(a, n, 100 + i);
    const double us = t.stop_us(reps);
    printf("D streaming writes, %.0f MiB : %8.2f us  %7.1f GB/s\n", n * 128.0 / 1048576.0, us, (double)n * 128 / us / 1e3);
    cudaFree(a);
}

// ---- E: staged gather with TMA tile::gather4 ------------------------------------------------------
// MODE 0: one cp.async.bulk per row (round 1), MODE 1: 16-byte cp.async, MODE 2: tile::gather4
struct StageErr { unsigned int timeouts; };

template <int WARPS, int MODE>
__global__ void __launch_bounds__(WARPS * 32) staged(const uint4* __restrict__ cur, uint4* __restrict__ nxt, uint32_t n, uint32_t salt,
                                                     const __grid_constant__ CUtensorMap tmap, StageErr* err)
{
    constexpr int ROWB = 128, NBLK = 8, G = 8, CPT = 4;
    constexpr int STAGE = 32 * 2 * ROWB;
    extern __shared__ __align__(128) unsigned char smem[];
    __shared__ __align__(8) uint64_t bars[WARPS][2];
    const uint32_t lane = threadIdx.x & 31u, w = threadIdx.x >> 5;
    unsigned char* my = smem + (size_t)w * 2 * STAGE;
    const uint32_t gwarp = blockIdx.x * WARPS + w, nwarps = gridDim.x * WARPS;
    if (MODE != 1 && lane == 0) {
        for (int s = 0; s < 2; ++s) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(sa(&bars[w][s])));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncwarp();
    const uint32_t iters = (n / 32 + nwarps - 1 - gwarp) / nwarps;  // n multiple of 32
    bool dead = false;
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
        } else if (MODE == 2) {
            // lane l < 16 fetches row slots 4l .. 4l+3 = both parents of children 2l and 2l+1
            if (lane == 0) asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(sa(&bars[w][it & 1])), "r"(STAGE) : "memory");
            __syncwarp();
            if (lane < 16) {
                const uint32_t c0 = base + 2u * lane, c1 = c0 + 1u;
                const int32_t r0 = (int32_t)(mix(c0 * 2u + salt) % n), r1 = (int32_t)(mix(c0 * 2u + 1u + salt) % n);
                const int32_t r2 = (int32_t)(mix(c1 * 2u + salt) % n), r3 = (int32_t)(mix(c1 * 2u + 1u + salt) % n);
                asm volatile(
                    "cp.async.bulk.tensor.2d.shared::cluster.global.tile::gather4.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4, %5, %6}], [%7];" ::"r"(
                        sa(st + lane * 4 * ROWB)),
                    "l"(&tmap), "r"(0), "r"(r0), "r"(r1), "r"(r2), "r"(r3), "r"(sa(&bars[w][it & 1]))
                    : "memory");
            }
        } else {
            for (int r0 = 0; r0 < 64; r0 += 4) {
                const uint32_t r = r0 + lane / NBLK;
                const uint32_t child = base + (r >> 1);
                const uint32_t src = mix(child * 2u + (r & 1u) + salt) % n;
                asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(sa(st + r * ROWB + (lane % NBLK) * 16)),
                             "l"(cur + (size_t)src * NBLK + (lane % NBLK)) : "memory");
            }
            asm volatile("cp.async.commit_group;" ::: "memory");
        }
    };
    if (iters) issue(0);
    for (uint32_t it = 0; it < iters && !dead; ++it) {
        if (it + 1 < iters) {
            if (MODE != 1) asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            issue(it + 1);
        }
        if (MODE != 1) {
            uint32_t done = 0;
            for (uint32_t spin = 0; spin < (1u << 22) && !done; ++spin)  // bounded: a wrong tensor map must not hang the GPU
                asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                             : "=r"(done) : "r"(sa(&bars[w][it & 1])), "r"((it >> 1) & 1u) : "memory");
            if (!done) {
                if (lane == 0) atomicAdd(&err->timeouts, 1u);
                dead = true;
                break;
            }
        } else {
            if (it + 1 < iters) asm volatile("cp.async.wait_group 1;" ::: "memory");
            else asm volatile("cp.async.wait_group 0;" ::: "memory");
            __syncwarp();
        }
        const uint32_t base = (gwarp + it * nwarps) * 32u;
        const unsigned char* st = my + (it & 1) * STAGE;
        const uint32_t gl = lane % G, grp = lane / G;
#pragma unroll
        for (int tile = 0; tile < G; ++tile) {
            const uint32_t c = tile * CPT + grp;
            const uint4 x = *reinterpret_cast<const uint4*>(st + (c * 2 + 0) * ROWB + gl * 16);
            const uint4 y = *reinterpret_cast<const uint4*>(st + (c * 2 + 1) * ROWB + gl * 16);
            nxt[(size_t)(base + c) * NBLK + gl] = make_uint4(x.x ^ y.x, x.y ^ y.y, x.z ^ y.z, x.w ^ y.w);
        }
        __syncwarp();
    }
}

typedef CUresult (*encode_fn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                              const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static encode_fn get_encode()
{
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess) return nullptr;
    return (encode_fn)fn;
}

static uint64_t checksum(const uint4* dev, size_t bytes)
{
    uint32_t* h = (uint32_t*)malloc(bytes);
    CK(cudaMemcpy(h, dev, bytes, cudaMemcpyDeviceToHost));
    uint64_t s = 0xcbf29ce484222325ull;
    for (size_t i = 0; i < bytes / 4; ++i) { s ^= h[i]; s *= 0x100000001b3ull; }
    free(h);
    return s;
}

__global__ void fill_rows(uint32_t* p, size_t words) { for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < words; i += (size_t)gridDim.x * blockDim.x) p[i] = mix((uint32_t)i); }

template <int WARPS, int MODE>
static void run_staged(uint32_t n, int bps, int box_rows, uint64_t* sum_out)
{
    uint4 *a, *b;
    StageErr* err;
    CK(cudaMalloc(&a, (size_t)n * 128)); CK(cudaMalloc(&b, (size_t)n * 128)); CK(cudaMalloc(&err, sizeof(StageErr)));
    fill_rows<<<1024, 256>>>((uint32_t*)a, (size_t)n * 32);
    CK(cudaMemset(b, 0, (size_t)n * 128)); CK(cudaMemset(err, 0, sizeof(StageErr)));
    CUtensorMap tmap;
    memset(&tmap, 0, sizeof(tmap));
    if (MODE == 2) {
        encode_fn enc = get_encode();
        if (!enc) { printf("E gather4: cuTensorMapEncodeTiled entry point not available\n"); return; }
        cuuint64_t gdim[2] = { 32, n };
        cuuint64_t gstride[1] = { 128 };
        cuuint32_t box[2] = { 32, (cuuint32_t)box_rows };
        cuuint32_t estr[2] = { 1, 1 };
        CUresult r = enc(&tmap, CU_TENSOR_MAP_DATA_TYPE_UINT32, 2, a, gdim, gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                         CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r != CUDA_SUCCESS) { printf("E gather4 box_rows=%d: cuTensorMapEncodeTiled failed (%d)\n", box_rows, (int)r); return; }
    }
    const size_t smem = (size_t)WARPS * 2 * 32 * 2 * 128;
    CK(cudaFuncSetAttribute(staged<WARPS, MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const int grid = g_sms * bps;
    // correctness pass first (salt 7, a -> b)
    staged<WARPS, MODE><<<grid, WARPS * 32, smem>>>(a, b, n, 7, tmap, err);
    cudaError_t e = cudaDeviceSynchronize();
    StageErr herr;
    if (e != cudaSuccess) { printf("E mode %d box_rows=%d: kernel failed: %s\n", MODE, box_rows, cudaGetErrorString(e)); exit(1); }
    CK(cudaMemcpy(&herr, err, sizeof(herr), cudaMemcpyDeviceToHost));
    const uint64_t sum = checksum(b, (size_t)n * 128);
    if (sum_out) *sum_out = sum;
    if (herr.timeouts) { printf("E mode %d box_rows=%d: %u warps timed out waiting for the copies (tensor map not accepted)\n", MODE, box_rows, herr.timeouts); return; }
    Timer t;
    for (int i = 0; i < 4; ++i) staged<WARPS, MODE><<<grid, WARPS * 32, smem>>>(a, b, n, i, tmap, err);
    CK(cudaDeviceSynchronize());
    const int reps = 20;
    t.start();
    for (int i = 0; i < reps; ++i) staged<WARPS, MODE><<<grid, WARPS * 32, smem>>>(a, b, n, 100 + i, tmap, err);
    const double us = t.stop_us(reps);
    const char* names[3] = { "TMA bulk per row", "cp.async 16 B", "TMA tile::gather4" };
    printf("E staged %-18s warps/block=%d blocks/SM=%d box_rows=%d : %8.2f us/pass  requested %7.1f GB/s  checksum %016llx\n", names[MODE], WARPS, bps, box_rows, us,
           3.0 * n * 128 / us / 1e3, (unsigned long long)sum);
    cudaFree(a); cudaFree(b); cudaFree(err);
}

int main(int argc, char** argv)
{
    CK(cudaDeviceGetAttribute(&g_sms, cudaDevAttrMultiProcessorCount, 0));
    const char* which = argc > 1 ? argv[1] : "ABCDE";
    if (strchr(which, 'A'))
        for (int lg : {17, 18, 19, 20, 21, 22}) run_gather(1u << lg);
    if (strchr(which, 'B')) {
        run_read_rows(1u << 16, 1u << 22);
        run_read_rows(1u << 18, 1u << 22);
        run_read_rows(1u << 20, 1u << 22);
    }
    if (strchr(which, 'C')) {
        run_read_words(1u << 20, 1u << 24);
        run_read_words(1u << 22, 1u << 24);
    }
    if (strchr(which, 'D')) {
        run_write_rows(1u << 18);
        run_write_rows(1u << 20);
        run_write_rows(1u << 22);
    }
    if (strchr(which, 'E')) {
        const uint32_t n = 1u << 20;
        uint64_t ref = 0, got = 0;
        run_staged<4, 1>(n, 2, 0, &ref);
        run_staged<8, 1>(n, 1, 0, nullptr);
        run_staged<4, 0>(n, 3, 0, &got);
        printf("E   bulk-per-row result %s the cp.async result\n", got == ref ? "matches" : "DIFFERS from");
        for (int box_rows : {1, 4}) {
            got = 0;
            run_staged<4, 2>(n, 2, box_rows, &got);
            printf("E   gather4 box_rows=%d result %s the cp.async result\n", box_rows, got == ref ? "matches" : "DIFFERS from");
            if (got == ref) {
                run_staged<4, 2>(n, 3, box_rows, nullptr);
                run_staged<8, 2>(n, 1, box_rows, nullptr);
                run_staged<12, 2>(n, 1, box_rows, nullptr);
                break;
            }
        }
    }
    return 0;
}
