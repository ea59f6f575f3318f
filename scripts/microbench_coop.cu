// What does a cooperative launch cost behind another kernel, against a plain launch of the same
// kernel?  And what does the library's hand-rolled grid barrier cost per round?
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o scripts/microbench_coop.bin scripts/microbench_coop.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at line %d\n", cudaGetErrorString(e), __LINE__); return 1; } } while (0)

__global__ void busy(float* out, int iters)
{
    float v = threadIdx.x;
    for (int i = 0; i < iters; ++i) v = v * 1.0001f + 0.5f;
    if (v == 123.f) out[0] = v;
}

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

__global__ void __launch_bounds__(256) rounds(unsigned long long* ctr, unsigned long long base, int nrounds)
{
    for (int r = 0; r < nrounds; ++r) grid_barrier(ctr, base + (unsigned long long)(r + 1) * gridDim.x);
}

int main()
{
    int sms = 0;
    CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
    float* out; unsigned long long* ctr;
    CK(cudaMalloc(&out, 4096)); CK(cudaMalloc(&ctr, 8)); CK(cudaMemset(ctr, 0, 8));
    cudaStream_t s; CK(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    unsigned long long used = 0;
    for (int grid : {10, 148, 296, 512}) for (int nrounds : {0, 1, 4}) for (int coop = 0; coop < 2; ++coop) {
        float total = 0.f; const int reps = 50;
        for (int i = 0; i < reps + 5; ++i) {
            busy<<<sms * 4, 256, 0, s>>>(out, 2000);
            cudaEventRecord(e0, s);
            unsigned long long base = used; int nr = nrounds;
            if (coop) {
                void* args[] = { &ctr, &base, &nr };
                CK(cudaLaunchCooperativeKernel((const void*)rounds, dim3(grid), dim3(256), args, 0, s));
            } else {
                rounds<<<grid, 256, 0, s>>>(ctr, base, nr);
            }
            used += (unsigned long long)nrounds * grid;
            cudaEventRecord(e1, s);
            CK(cudaStreamSynchronize(s));
            float ms; cudaEventElapsedTime(&ms, e0, e1);
            if (i >= 5) total += ms;
        }
        printf("grid %3d x 256, %d barrier round(s), %s launch behind a busy kernel: %7.2f us\n", grid, nrounds, coop ? "cooperative" : "plain      ", 1e3 * total / reps);
    }
    return 0;
}
