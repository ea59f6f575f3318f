// ref_pin.cu -- lock-step comparison of the UNMODIFIED reference library with the
// CPU oracle, generation by generation, bit for bit.
//
// TEST INFRASTRUCTURE ONLY (recipe: oracle/build_ref.sh; outputs oracle/_ref/pin/pinN).
//
// One binary per reference demo program (-DPIN_CASE=1|2|3).  Each is device-linked from
//   * this file,
//   * the reference's own test*/test.cu, compiled where it lies with -Dmain=... so
//     that its __device__ hooks (objf / crossover, reached through the `ofunction` /
//     `cfunction` pointer variables, and test3's __constant__ city_matrix) are the
//     reference's own device code,
//   * oracle/_ref/libpga_ref.a = /root/reference/src/pga.cu compiled as it lies,
//   * oracle/curand_shim.c instead of -lcurand, which serves the oracle's Philox rand
//     rows through the four cuRAND calls of src/pga.cu:100, :153-154, :167,
//   * oracle/pga_oracle.c.
// The reference then draws exactly the uniforms the oracle draws, and every kernel of
// src/pga.cu (:81-86 init, :250-262 evaluate, :280-317 selection + crossover,
// :333-347 mutate) and pga_get_best (:218-236) is compared with the oracle's
// restatement: genomes and scores must be bit-identical after every step.
//
// The population's device buffers are reached through a mirror of the reference's
// struct population (src/pga.cu:37-46).
//
// Output: one "GEN ..." line per generation with FNV-1a digests (the GPU test replays
// the same workload on this repository's library and compares the digests), the
// reference's own "%f" line from pga_get_best between markers, and "PIN OK" / exit 0
// when everything matched; the first mismatch is printed and the exit code is 1.
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <pga.h>  // the reference's include/pga.h (-I/root/reference/include)

#include "pga_oracle.h"

#ifndef PIN_CASE
#error "compile with -DPIN_CASE=1, 2 or 3"
#endif

extern "C" void pin_shim_configure(uint64_t seed, unsigned genome_len);
extern "C" void pin_shim_reset(void);
extern "C" unsigned long long pin_shim_fills(void);

// defined by the reference's test*/test.cu
extern __device__ obj_f ofunction;
#if PIN_CASE == 3
extern __device__ crossover_f cfunction;
#define MAX_CITY_COUNT 110
extern __constant__ float city_matrix[MAX_CITY_COUNT][MAX_CITY_COUNT];
#endif

// src/pga.cu:37-46
struct ref_population {
    unsigned long size;
    unsigned genome_len;
    gene* _g;
    gene* current_gen;
    gene* next_gen;
    gene* _g2;
    float* score;
    float* rand;
};

#define CK(x)                                                                              \
    do {                                                                                   \
        cudaError_t e_ = (x);                                                              \
        if (e_ != cudaSuccess) {                                                           \
            fprintf(stderr, "ref_pin: %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); \
            exit(2);                                                                       \
        }                                                                                  \
    } while (0)

static int failures = 0;

static void compare(const char* what, unsigned gen, const void* ref_dev, const void* oracle_host, size_t words, unsigned row_words)
{
    uint32_t* h = (uint32_t*)malloc(words * 4 + 4);
    CK(cudaMemcpy(h, ref_dev, words * 4, cudaMemcpyDeviceToHost));
    const uint32_t* o = (const uint32_t*)oracle_host;
    size_t bad = 0, first = 0;
    for (size_t i = 0; i < words; ++i)
        if (h[i] != o[i]) {
            if (!bad) first = i;
            ++bad;
        }
    if (bad) {
        float a, b;
        memcpy(&a, &h[first], 4);
        memcpy(&b, &o[first], 4);
        printf("MISMATCH %s generation %u: %zu of %zu words differ; first at row %zu word %zu: reference %.9g (0x%08x) oracle %.9g (0x%08x)\n",
               what, gen, bad, words, first / row_words, first % row_words, a, h[first], b, o[first]);
        failures += 1;
    }
    free(h);
}

static uint64_t digest_dev(const void* dev, size_t bytes)
{
    void* h = malloc(bytes + 1);
    CK(cudaMemcpy(h, dev, bytes, cudaMemcpyDeviceToHost));
    uint64_t d = opga_digest(h, bytes);
    free(h);
    return d;
}

struct pin_case {
    unsigned long size;
    unsigned len;
    unsigned lockstep_gens, run_gens;
    const char* obj_name;
    const char* xov_name;
};

static void set_hooks(pga_t* p, opga_t* o, const pin_case& c)
{
    void* func;
    CK(cudaMemcpyFromSymbol(&func, ofunction, sizeof(obj_f)));  // as test/test.cu:39-41
    pga_set_objective_function(p, (obj_f)func);
    opga_set_objective_function(o, (opga_obj_f)opga_demo_hook(c.obj_name));
#if PIN_CASE == 3
    CK(cudaMemcpyFromSymbol(&func, cfunction, sizeof(crossover_f)));  // test3/test.cu:89-90
    pga_set_crossover_function(p, (crossover_f)func);
    opga_set_crossover_function(o, (opga_crossover_f)opga_demo_hook(c.xov_name));
#endif
}

int main(int argc, char** argv)
{
    uint64_t seed = 42;
    const char* instance = NULL;
    for (int i = 1; i < argc; ++i) {
        if (!strcmp(argv[i], "--seed") && i + 1 < argc) seed = strtoull(argv[++i], NULL, 0);
        else if (!strcmp(argv[i], "--instance") && i + 1 < argc) instance = argv[++i];
    }
#if PIN_CASE == 1
    pin_case c = { 40000, 100, 100, 100, "sum_obj", NULL };  // test/test.cu:37, :43
#elif PIN_CASE == 2
    pin_case c = { 100, 6, 5, 5, "knapsack_obj", NULL };  // test2/test.cu:43, :49
#else
    pin_case c = { 1000, 100, 60, 1000, "tsp_obj", "tsp_crossover" };  // test3/test.cu:85, :93
    {
        // the instance, loaded the way test3/test.cu:69-79 loads it (including the
        // city_count^2-float prefix copy of a [110][110] array)
        if (!instance) { fprintf(stderr, "ref_pin: case 3 needs --instance <gen output>\n"); return 2; }
        FILE* f = fopen(instance, "r");
        if (!f) { perror(instance); return 2; }
        static float host_city_matrix[MAX_CITY_COUNT][MAX_CITY_COUNT];
        int city_count = 0;
        if (fscanf(f, "%d", &city_count) != 1 || city_count > MAX_CITY_COUNT) return 2;
        float* flat = (float*)malloc(sizeof(float) * city_count * city_count);
        for (int i = 0; i < city_count; ++i)
            for (int j = 0; j < city_count; ++j) {
                if (fscanf(f, "%f", &host_city_matrix[i][j]) != 1) return 2;
                flat[i * city_count + j] = host_city_matrix[i][j];
            }
        fclose(f);
        CK(cudaMemcpyToSymbol(city_matrix, &host_city_matrix, sizeof(float) * (city_count * city_count)));
        opga_demo_tsp_load(flat, city_count);
        free(flat);
        c.len = (unsigned)city_count;
    }
#endif
    for (int i = 1; i < argc; ++i) {
        if (!strcmp(argv[i], "--size") && i + 1 < argc) c.size = strtoul(argv[++i], NULL, 0);
        else if (!strcmp(argv[i], "--gens") && i + 1 < argc) c.lockstep_gens = c.run_gens = (unsigned)strtoul(argv[++i], NULL, 0);
    }
    const size_t words = (size_t)c.size * c.len;
    printf("CASE %d size %lu len %u seed %llu lockstep %u run %u\n", PIN_CASE, c.size, c.len, (unsigned long long)seed,
           c.lockstep_gens, c.run_gens);

    // ---- part A: the stepwise API, compared after every step ------------------------
    {
        pin_shim_reset();
        pin_shim_configure(seed, c.len);
        pga_t* p = pga_init();                                                      // src/pga.cu:148
        population_t* pop = pga_create_population(p, c.size, c.len, RANDOM_POPULATION);  // :177, fills rand (epoch 0) and genes
        if (!p || !pop) { fprintf(stderr, "ref_pin: reference pga_init/pga_create_population failed\n"); return 2; }
        ref_population* rp = (ref_population*)pop;
        opga_t* o = opga_init(seed);
        opga_set_threads(o, 8);
        opga_set_quiet(o, 1);
        opga_pop_t* op = opga_create_population(o, c.size, c.len, OPGA_GENES_F32);
        set_hooks(p, o, c);

        compare("initial population (src/pga.cu:81-86)", 0, rp->current_gen, opga_pop_current(op), words, c.len);
        for (unsigned g = 1; g <= c.lockstep_gens && !failures; ++g) {
            pga_fill_random_values(p, pop);  // src/pga.cu:103-105
            opga_fill_random_values(o, op);
            pga_evaluate(p, pop);  // :264-270
            opga_evaluate(o, op);
            compare("scores after pga_evaluate (src/pga.cu:250-262)", g, rp->score, opga_pop_scores(op), c.size, 1);
            pga_crossover(p, pop, TOURNAMENT);  // :319-325
            opga_crossover(o, op);
            compare("next_gen after pga_crossover (src/pga.cu:280-317)", g, rp->next_gen, opga_pop_next(op), words, c.len);
            pga_mutate(p, pop);  // :349-354
            opga_mutate(o, op);
            compare("next_gen after pga_mutate (src/pga.cu:333-347)", g, rp->next_gen, opga_pop_next(op), words, c.len);
            pga_swap_generations(p, pop);  // :362-366
            opga_swap_generations(o, op);
            printf("GEN %u genes %016llx scores_before %016llx\n", g, (unsigned long long)digest_dev(rp->current_gen, words * 4),
                   (unsigned long long)digest_dev(rp->score, c.size * 4));
        }
        pga_evaluate(p, pop);
        opga_evaluate(o, op);
        compare("final scores", c.lockstep_gens + 1, rp->score, opga_pop_scores(op), c.size, 1);
        printf("STEPWISE fills %llu genes %016llx scores %016llx\n", pin_shim_fills(),
               (unsigned long long)digest_dev(rp->current_gen, words * 4), (unsigned long long)digest_dev(rp->score, c.size * 4));
        pga_deinit(p);
        opga_deinit(o);
    }

    // ---- part B: pga_run and pga_get_best -------------------------------------------------
    {
        pin_shim_reset();
        pin_shim_configure(seed, c.len);
        pga_t* p = pga_init();
        population_t* pop = pga_create_population(p, c.size, c.len, RANDOM_POPULATION);
        ref_population* rp = (ref_population*)pop;
        opga_t* o = opga_init(seed);
        opga_set_threads(o, 8);
        opga_set_quiet(o, 1);
        opga_pop_t* op = opga_create_population(o, c.size, c.len, OPGA_GENES_F32);
        set_hooks(p, o, c);
        pga_run(p, c.run_gens);  // src/pga.cu:376-391
        opga_run(o, c.run_gens);
        compare("current_gen after pga_run", c.run_gens, rp->current_gen, opga_pop_current(op), words, c.len);
        compare("scores after pga_run", c.run_gens, rp->score, opga_pop_scores(op), c.size, 1);
        printf("RUN fills %llu genes %016llx scores %016llx\n", pin_shim_fills(),
               (unsigned long long)digest_dev(rp->current_gen, words * 4), (unsigned long long)digest_dev(rp->score, c.size * 4));
        // pga_get_best prints the best score itself (src/pga.cu:230)
        printf("REF_BEST_BEGIN\n");
        fflush(stdout);
        gene* best = pga_get_best(p, pop);
        fflush(stdout);
        printf("REF_BEST_END\n");
        float oscore = 0;
        long oid = opga_best_index(o, op, &oscore);
        printf("ORACLE_BEST %f\n", oscore);
        const float* orow = (const float*)opga_pop_current(op) + (size_t)oid * c.len;
        if (memcmp(best, orow, sizeof(float) * c.len) != 0) {
            printf("MISMATCH best genome: reference row differs from oracle row %ld\n", oid);
            failures += 1;
        }
        free(best);
        pga_deinit(p);
        opga_deinit(o);
    }

    if (failures) {
        printf("PIN FAILED (%d)\n", failures);
        return 1;
    }
    printf("PIN OK\n");
    return 0;
}
