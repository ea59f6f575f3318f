/*
 * pga_oracle.h -- CPU oracle for the libpga generation hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is part of the product: only
 * tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * legs may load it, and only as the checker or the timed CPU baseline.
 *
 * It is a serial (optionally OpenMP) C restatement of the reference's pga.h
 * semantics (/root/reference/src/pga.cu), with the reference's time-seeded cuRAND
 * XORWOW stream replaced by the counter-based Philox4x32-10 draw layout the
 * north star mandates.  Parity status: the reference holds NO golden vectors or
 * assertions for this path, so the oracle is pinned by
 *   - the published Philox4x32-10 known-answer vectors (tests/test_oracle_cpu.py),
 *   - digests of its own trajectories committed under tests/golden/,
 *   - the derivable optima of the three reference demo programs.
 * The stubs the reference never implemented (migration, islands, top lists) are
 * "parity unpinned": the oracle is their only specification.
 */
#ifndef PGA_ORACLE_H
#define PGA_ORACLE_H

#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct opga opga_t;
typedef struct opga_pop opga_pop_t;

typedef float (*opga_obj_f)(float *, unsigned);
typedef void (*opga_mutate_f)(float *, float *, unsigned);
typedef void (*opga_crossover_f)(float *, float *, float *, float *, unsigned);

#define OPGA_MAX_POPULATIONS 10 /* include/pga.h:44 */

enum { OPGA_GENES_F32 = 0, OPGA_GENES_BITS = 1 };
enum {
	OPGA_OBJ_HOOK = 0, /* user callback, serial per individual */
	OPGA_OBJ_SUM = 1,
	OPGA_OBJ_RASTRIGIN = 2,
	OPGA_OBJ_SPHERE = 3,
	OPGA_OBJ_ONEMAX = 4
};
enum { OPGA_TOPO_RING = 0, OPGA_TOPO_RANDOM = 1 };

/* draw streams (counter word 3, high byte) */
enum { OPGA_STREAM_ROW = 0, OPGA_STREAM_MASK = 1, OPGA_STREAM_INITBITS = 2, OPGA_STREAM_MIGRATE = 3 };

/* ---- RNG ---- */
void opga_philox4x32_10(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4]);
uint32_t opga_draw_word(uint64_t seed, uint32_t island, uint32_t epoch, uint32_t individual,
			uint32_t stream, uint32_t w);
float opga_draw_uniform(uint64_t seed, uint32_t island, uint32_t epoch, uint32_t individual, uint32_t j);
void opga_rand_row(uint64_t seed, uint32_t island, uint32_t epoch, uint32_t individual, uint32_t len, float *out);

/* ---- numerics shared with the device (same operation order) ---- */
float opga_cos2pi(float x);
float opga_builtin_objective_f32(int which, const float *g, unsigned len);
float opga_builtin_objective_bits(int which, const uint32_t *w, unsigned len_bits);

/* ---- pga.h mirror ---- */
opga_t *opga_init(uint64_t seed);
void opga_deinit(opga_t *);
void opga_set_threads(opga_t *, int nthreads); /* 1 = serial (default) */
void opga_set_island_base(opga_t *, uint32_t first_global_island);
void opga_set_quiet(opga_t *, int quiet);
opga_pop_t *opga_create_population(opga_t *, unsigned long size, unsigned genome_len, int encoding);
opga_pop_t *opga_create_population_on_island(opga_t *, unsigned long size, unsigned genome_len, int encoding,
					     uint32_t island);
void opga_set_objective_function(opga_t *, opga_obj_f);
void opga_set_builtin_objective(opga_t *, int which);
void opga_set_mutate_function(opga_t *, opga_mutate_f);
void opga_set_crossover_function(opga_t *, opga_crossover_f);
float *opga_get_best(opga_t *, opga_pop_t *);
long opga_best_index(opga_t *, opga_pop_t *, float *score_out);
float **opga_get_best_top(opga_t *, opga_pop_t *, unsigned length);
float *opga_get_best_all(opga_t *);
float **opga_get_best_top_all(opga_t *, unsigned length);
void opga_evaluate(opga_t *, opga_pop_t *);
void opga_evaluate_all(opga_t *);
void opga_crossover(opga_t *, opga_pop_t *);
void opga_crossover_all(opga_t *);
void opga_mutate(opga_t *, opga_pop_t *);
void opga_mutate_all(opga_t *);
void opga_swap_generations(opga_t *, opga_pop_t *);
void opga_fill_random_values(opga_t *, opga_pop_t *);
void opga_set_tournament(opga_t *, unsigned k);                      /* candidates per parent (2 = the reference) */
void opga_set_target(opga_t *, float value, unsigned check_every);   /* include/pga.h:141, :147 */
unsigned opga_generations_run(opga_t *);
void opga_run(opga_t *, unsigned n);
void opga_migrate(opga_t *, float pct);
void opga_migrate_between(opga_t *, opga_pop_t *, opga_pop_t *, float pct);
void opga_set_topology(opga_t *, int topo);
void opga_set_migration_delay(opga_t *, unsigned delay);
void opga_run_islands(opga_t *, unsigned n, unsigned m, float pct);

/* ---- raw access for the parity harness ---- */
unsigned long opga_pop_size(opga_pop_t *);
unsigned opga_pop_genome_len(opga_pop_t *);
unsigned opga_pop_row_words(opga_pop_t *); /* 32-bit words per stored row (no padding) */
uint32_t opga_pop_epoch(opga_pop_t *);
void *opga_pop_current(opga_pop_t *);
void *opga_pop_next(opga_pop_t *);
float *opga_pop_scores(opga_pop_t *);
int32_t *opga_pop_selection(opga_pop_t *); /* [2*size]: p1,p2 of the last crossover */
void opga_pop_set_current(opga_pop_t *, const void *rows);
void opga_pop_set_scores(opga_pop_t *, const float *scores);
void opga_pop_set_epoch(opga_pop_t *, uint32_t epoch);

/* total order used by every ranking: higher score first, then lower index;
 * NaN ranks below everything.  k best / k worst, both listed by ascending index. */
void opga_select_top(const float *score, unsigned long n, unsigned long k, int worst, uint32_t *out_idx);
unsigned long opga_migration_count(unsigned long size, float pct);
void opga_random_permutation(uint64_t seed, uint32_t round, unsigned count, uint32_t *out);

/* ---- the reference demo programs' hooks, restated ---- */
float opga_demo_sum_obj(float *g, unsigned len);       /* test/test.cu:24-30 */
float opga_demo_knapsack_obj(float *g, unsigned len);  /* test2/test.cu:25-36 */
float opga_demo_tsp_obj(float *g, unsigned len);       /* test3/test.cu:26-46 */
void opga_demo_tsp_crossover(float *p1, float *p2, float *c, float *rnd, unsigned len); /* test3/test.cu:48-64 */
void opga_demo_tsp_load(const float *matrix, int city_count); /* test3/test.cu:70-79 */
void *opga_demo_hook(const char *name);

uint64_t opga_digest(const void *data, size_t bytes); /* FNV-1a 64 */

#ifdef __cplusplus
}
#endif
#endif
