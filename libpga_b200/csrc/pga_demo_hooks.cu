// pga_demo_hooks.cu -- the hooks of the reference's three demo programs, written
// afresh as device functions inside the shared library, so that bindings which
// cannot compile CUDA (ctypes, cgo, JNI) can still drive the hook path.
//
// They mirror, in behaviour: test/test.cu:24-30 (sum of genes),
// test2/test.cu:25-36 (bounded knapsack), test3/test.cu:24-64 (tour length with a
// duplicate penalty, and a duplicate-avoiding crossover).  A program that links
// libpga.a defines its own hooks exactly as with the reference; these are only
// conveniences and bench workloads.
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "../../include/pga_b200.h"

#define DEMO_MAX_CITIES 110

__constant__ float demo_item_value[6] = { 75.f, 150.f, 250.f, 35.f, 10.f, 100.f };
__constant__ float demo_item_weight[6] = { 7.f, 8.f, 6.f, 4.f, 3.f, 9.f };
__constant__ float demo_city_matrix[DEMO_MAX_CITIES][DEMO_MAX_CITIES];

__device__ float demo_sum_obj(gene* g, unsigned len)
{
    float total = 0.f;
    for (unsigned k = 0; k < len; ++k) total += g[k];
    return total;
}

__device__ float demo_knapsack_obj(gene* g, unsigned len)
{
    const float capacity = 10.0f;
    float value = 0.f, weight = 0.f;
    for (unsigned k = 0; k < len; ++k) {
        const int copies = (int)(g[k] * 2);  // up to MAX_ITEM_COUNT = 2 of each
        value += demo_item_value[k] * copies;
        weight += demo_item_weight[k] * copies;
    }
    return weight <= capacity ? value : capacity - weight;
}

__device__ float demo_tsp_obj(gene* g, unsigned len)
{
    float cost = 0.f;
    for (unsigned k = 1; k < len; ++k) cost += demo_city_matrix[(int)(g[k - 1] * len)][(int)(g[k] * len)];
    for (unsigned k = 0; k < len; ++k) {
        const int here = (int)(g[k] * len);
        for (unsigned m = 0; m < len; ++m)
            if (k != m && here == (int)(g[m] * len)) cost += 10000;
    }
    return -1 * cost;
}

__device__ void demo_tsp_crossover(gene* p1, gene* p2, gene* child, float* rnd, unsigned len)
{
    int used[DEMO_MAX_CITIES] = {};
    for (unsigned k = 0; k < len; ++k) {
        const int a = (int)(p1[k] * len), b = (int)(p2[k] * len);
        if (!used[a]) { child[k] = p1[k]; used[a] = 1; }
        else if (!used[b]) { child[k] = p2[k]; used[b] = 1; }
        else child[k] = rnd[k];
    }
}

// Test hooks with no counterpart in the reference's demos.  A crossover that writes
// only the even genes: the odd ones keep what the child row held before the call, which
// the reference defines as the stale next_gen row (src/pga.cu:302).
__device__ void demo_even_crossover(gene* p1, gene* p2, gene* child, float* rnd, unsigned len)
{
    for (unsigned k = 0; k < len; k += 2) child[k] = rnd[k] > 0.5f ? p1[k] : p2[k];
}

// A mutation hook that reads its whole rand row: every gene moves towards the draw with
// probability 1/8.
__device__ void demo_blend_mutate(gene* g, float* rnd, unsigned len)
{
    for (unsigned k = 0; k < len; ++k)
        if (rnd[k] < 0.125f) g[k] = 0.5f * (g[k] + rnd[(k + 1) % len]);
}

__device__ obj_f demo_sum_obj_ptr = demo_sum_obj;
__device__ crossover_f demo_even_crossover_ptr = demo_even_crossover;
__device__ mutate_f demo_blend_mutate_ptr = demo_blend_mutate;
__device__ obj_f demo_knapsack_obj_ptr = demo_knapsack_obj;
__device__ obj_f demo_tsp_obj_ptr = demo_tsp_obj;
__device__ crossover_f demo_tsp_crossover_ptr = demo_tsp_crossover;

static void check(cudaError_t e, int line)
{
    if (e != cudaSuccess) {
        fprintf(stderr, "GPUassert: %s %s %d\n", cudaGetErrorString(e), __FILE__, line);
        exit((int)e);
    }
}

extern "C" void* pgax_demo_hook(const char* name)
{
    void* fn = NULL;
    if (!name) return NULL;
    if (!strcmp(name, "sum_obj")) check(cudaMemcpyFromSymbol(&fn, demo_sum_obj_ptr, sizeof(fn)), __LINE__);
    else if (!strcmp(name, "knapsack_obj")) check(cudaMemcpyFromSymbol(&fn, demo_knapsack_obj_ptr, sizeof(fn)), __LINE__);
    else if (!strcmp(name, "tsp_obj")) check(cudaMemcpyFromSymbol(&fn, demo_tsp_obj_ptr, sizeof(fn)), __LINE__);
    else if (!strcmp(name, "tsp_crossover")) check(cudaMemcpyFromSymbol(&fn, demo_tsp_crossover_ptr, sizeof(fn)), __LINE__);
    else if (!strcmp(name, "even_crossover")) check(cudaMemcpyFromSymbol(&fn, demo_even_crossover_ptr, sizeof(fn)), __LINE__);
    else if (!strcmp(name, "blend_mutate")) check(cudaMemcpyFromSymbol(&fn, demo_blend_mutate_ptr, sizeof(fn)), __LINE__);
    return fn;
}

// test3 reads the matrix into a [110][110] host array and copies only the first
// city_count^2 floats of it to the device symbol (test3/test.cu:70-79); kept as is.
extern "C" int pgax_demo_tsp_load(const float* matrix, int city_count)
{
    if (!matrix || city_count < 1 || city_count > DEMO_MAX_CITIES) return PGAX_EINVAL;
    static float host[DEMO_MAX_CITIES][DEMO_MAX_CITIES];
    static float zero[DEMO_MAX_CITIES][DEMO_MAX_CITIES];
    memset(host, 0, sizeof(host));
    for (int r = 0; r < city_count; ++r)
        for (int c = 0; c < city_count; ++c) host[r][c] = matrix[r * city_count + c];
    check(cudaMemcpyToSymbol(demo_city_matrix, zero, sizeof(zero)), __LINE__);
    check(cudaMemcpyToSymbol(demo_city_matrix, host, sizeof(float) * (size_t)(city_count * city_count)), __LINE__);
    return 0;
}
