// ref_time.cu -- wall-clock time of pga_run on the reference's own test/ workload
// (/root/reference/test/test.cu:37-45: 40000 genomes of 100 genes, the demo's sum
// objective), written against include/pga.h only, so that the SAME object links against
//   * the unmodified reference library rebuilt for sm_100a (oracle/_ref/libpga_ref.a + cuRAND)
//     -> oracle/_ref/time/ref1
//   * this repository's libpga.a                                -> oracle/_ref/time/ours1
// and the two are timed on the same GPU.  The objective is the reference demo's own
// device function (test/test.cu, main renamed at compile time, reached through its
// `ofunction` pointer as test/test.cu:39-41 does).
//
// TEST / BENCH INFRASTRUCTURE ONLY (recipe: oracle/build_ref.sh); bench.py reports it as
// `reference_gpu` beside the test1 workload.  The reference is seeded by time(NULL)
// (src/pga.cu:154), so only the time is compared here; results are compared bit for bit by
// oracle/ref_pin.cu.
//
// usage: ref1|ours1 [--size N] [--gens G] [--reps R]
#include <pga.h>

#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>

#define CK(x)                                                                                     \
    do {                                                                                          \
        cudaError_t e_ = (x);                                                                     \
        if (e_ != cudaSuccess) {                                                                  \
            fprintf(stderr, "ref_time: %s at line %d\n", cudaGetErrorString(e_), __LINE__);       \
            return 2;                                                                             \
        }                                                                                         \
    } while (0)

extern __device__ obj_f ofunction;  // /root/reference/test/test.cu:33

int main(int argc, char** argv)
{
    unsigned long size = 40000;
    unsigned gens = 100, reps = 5;
    const unsigned len = 100;  // GENOME_LENGTH, test/test.cu:22
    for (int i = 1; i + 1 < argc; i += 2) {
        if (!strcmp(argv[i], "--size")) size = strtoul(argv[i + 1], NULL, 10);
        else if (!strcmp(argv[i], "--gens")) gens = (unsigned)strtoul(argv[i + 1], NULL, 10);
        else if (!strcmp(argv[i], "--reps")) reps = (unsigned)strtoul(argv[i + 1], NULL, 10);
    }

    pga_t* p = pga_init();
    if (!p) return 1;
    population_t* pop = pga_create_population(p, size, len, RANDOM_POPULATION);
    if (!pop) return 1;
    void* func = NULL;
    CK(cudaMemcpyFromSymbol(&func, ofunction, sizeof(obj_f)));
    pga_set_objective_function(p, (obj_f)func);

    pga_run(p, 10);  // warm-up: module load, first launches
    CK(cudaDeviceSynchronize());

    double best_s = 1e30, sum_s = 0.0;
    for (unsigned r = 0; r < reps; ++r) {
        CK(cudaDeviceSynchronize());
        const auto t0 = std::chrono::steady_clock::now();
        pga_run(p, gens);
        CK(cudaDeviceSynchronize());
        const auto t1 = std::chrono::steady_clock::now();
        const double s = std::chrono::duration<double>(t1 - t0).count();
        best_s = s < best_s ? s : best_s;
        sum_s += s;
    }
    const double mean_s = sum_s / reps;
    printf("{\"size\": %lu, \"genome_len\": %u, \"generations\": %u, \"reps\": %u, \"seconds_mean\": %.6f, \"seconds_min\": %.6f, "
           "\"us_per_generation\": %.3f, \"evals_per_s\": %.6g}\n",
           size, len, gens, reps, mean_s, best_s, 1e6 * mean_s / gens, (double)size * gens / mean_s);
    pga_deinit(p);
    return 0;
}
