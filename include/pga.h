/*
 * pga.h -- the libpga C API, served by the B200-native engine in this repository.
 *
 * This header is the drop-in boundary.  Every type, enum and prototype below has
 * the same name, argument order and meaning as pbalcer/libpga's public header
 * (reference: include/pga.h:26-150), so programs written against that library --
 * its own test/test.cu, test2/test.cu and test3/test.cu included -- compile
 * against this file unchanged and link with `-L<repo> -lpga`.
 *
 * Behaviour that the reference leaves open or unfinished is documented per
 * function.  Knobs that libpga never had (seed, packed-bit genomes, built-in
 * objectives, island topology, host-buffer transfers, multi-GPU set-up) live in
 * pga_b200.h so that this file stays frozen.
 *
 * Conventions (reference: src/pga.cu:25-33, :218-236):
 *   - higher score is better everywhere;
 *   - any CUDA failure prints "GPUassert: <msg> <file> <line>" to stderr and exits;
 *   - calls block until their work is done (the engine is asynchronous inside).
 */
#ifndef PGA_H
#define PGA_H

#ifdef __cplusplus
extern "C" {
#endif

/* opaque handles: one solver, and one (sub)population / island inside it */
typedef struct pga pga_t;
typedef struct population population_t;

/* a gene is one float; integer problems decode it, e.g. (int)(g * K) */
typedef float gene;

enum population_type {
	RANDOM_POPULATION, /* genes start as uniform draws */
	MAX_POPULATION_TYPE
};

enum crossover_selection_type {
	TOURNAMENT, /* size-2 tournament; the only scheme the reference has */
	MAX_SELECTION_TYPE
};

#define MAX_POPULATIONS 10

/*
 * User hooks.  Each value is a DEVICE code address: declare a __device__
 * function-pointer variable initialised with the __device__ function and fetch
 * it with cudaMemcpyFromSymbol (reference: include/pga.h:65-67, test/test.cu:32-41).
 *   obj_f(genome, len)                       -> fitness of one individual
 *   mutate_f(genome, rand_row, len)          -> mutates genome in place
 *   crossover_f(p1, p2, child, rand_row, len)-> writes child
 * genome/p1/p2/child are contiguous float[len]; rand_row is that individual's
 * float[len] of uniform draws in [0,1) for the current generation.
 */
typedef float (*obj_f)(gene *, unsigned);
typedef void (*mutate_f)(gene *, float *, unsigned);
typedef void (*crossover_f)(gene *, gene *, gene *, float *, unsigned);

/* Creates a solver on the current CUDA device.  NULL on allocation failure.
 * Seed: $PGA_SEED if set, otherwise the wall clock (reference: src/pga.cu:148-164). */
pga_t *pga_init();

/* Frees every population of the solver and the solver itself (src/pga.cu:166-175). */
void pga_deinit(pga_t *);

/* Adds a population of `size` individuals with `genome_len` genes each.
 * NULL when MAX_POPULATIONS are already present, genome_len < 4, or memory runs
 * out (src/pga.cu:177-204).  The handle is owned by the solver. */
population_t *pga_create_population(pga_t *, unsigned long size, unsigned genome_len, enum population_type type);

/* Objective hook; required before evaluate/run unless a built-in objective is
 * selected through pga_b200.h (src/pga.cu:206-208). */
void pga_set_objective_function(pga_t *, obj_f);

/* Mutation hook; NULL restores the default: with probability 0.01 one random
 * gene is replaced by a uniform draw (include/pga.h:74-78, src/pga.cu:127-133). */
void pga_set_mutate_function(pga_t *, mutate_f);

/* Crossover hook; NULL restores the default uniform crossover
 * (include/pga.h:80-85, src/pga.cu:135-143). */
void pga_set_crossover_function(pga_t *, crossover_f);

/* Best individual of one population: prints its score as "%f\n" on stdout and
 * returns a malloc'ed float[genome_len] the caller frees.  The first maximum
 * wins ties (src/pga.cu:218-236). */
gene *pga_get_best(pga_t *, population_t *);
/* The `length` best individuals, best first: a malloc'ed array of malloc'ed rows,
 * all owned by the caller (a stub in the reference, src/pga.cu:238-240). */
gene **pga_get_best_top(pga_t *, population_t *, unsigned length);
/* Best individual over all populations; prints like pga_get_best (stub: src/pga.cu:242-244). */
gene *pga_get_best_all(pga_t *);
/* The `length` best over all populations, best first (stub: src/pga.cu:246-248). */
gene **pga_get_best_top_all(pga_t *, unsigned length);

/* Scores the current generation with the objective (src/pga.cu:250-276). */
void pga_evaluate(pga_t *, population_t *);
void pga_evaluate_all(pga_t *);

/* Builds the next generation: child i = crossover of two tournament winners
 * (src/pga.cu:278-331).  The selection-type argument is accepted and, as in the
 * reference, only TOURNAMENT exists. */
void pga_crossover(pga_t *, population_t *, enum crossover_selection_type);
void pga_crossover_all(pga_t *, enum crossover_selection_type);

/* One exchange between populations: the top `pct` percent of every population
 * moves to a partner chosen by the island topology (random by default, as the
 * reference words it: include/pga.h:108-111; stub at src/pga.cu:368-370). */
void pga_migrate(pga_t *, float pct);
/* Copies the top `pct` percent of `from` over the worst individuals of `to`
 * (include/pga.h:112-115; stub at src/pga.cu:372-374). */
void pga_migrate_between(pga_t *, population_t *from, population_t *to, float pct);

/* Applies the mutation operator to the next generation (src/pga.cu:333-360). */
void pga_mutate(pga_t *, population_t *);
void pga_mutate_all(pga_t *);

/* Makes the next generation current (src/pga.cu:362-366). */
void pga_swap_generations(pga_t *, population_t *);

/* Advances the population to a fresh set of random draws; every step between
 * two calls sees the same draws (src/pga.cu:99-105). */
void pga_fill_random_values(pga_t *, population_t *);

/* n generations of {new draws, evaluate, crossover, mutate, swap} on the FIRST
 * population, then a final evaluate so scores match the population on return
 * (src/pga.cu:376-391).  Here the whole generation is one fused kernel. */
void pga_run(pga_t *, unsigned n);

/* n generations on every population; every m generations the top `pct` percent
 * migrate (include/pga.h:145-150; stub at src/pga.cu:393-395). */
void pga_run_islands(pga_t *, unsigned n, unsigned m, float pct);

#ifdef __cplusplus
}
#endif

#endif
