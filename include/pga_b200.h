/*
 * pga_b200.h -- extension surface of the B200-native libpga engine.
 *
 * Everything libpga's frozen header (pga.h) cannot express: the seed, packed-bit
 * genomes, built-in objectives that run without indirect calls, island topology
 * and delay, host-buffer transfers, multi-GPU set-up and timing.  All entry points
 * are C-ABI (plain pointers and sizes); a binding in any language needs nothing
 * else.  The reference has no counterpart for most of them; where one exists the
 * comment cites it (paths relative to the reference checkout).
 *
 * Status returns: 0 = success, negative = PGAX_E*.  CUDA/NCCL failures follow
 * libpga's convention instead: "GPUassert: ..." on stderr and exit
 * (reference: src/pga.cu:25-33).
 */
#ifndef PGA_B200_H
#define PGA_B200_H

#include <stddef.h>
#include <stdint.h>

#include "pga.h"

#ifdef __cplusplus
extern "C" {
#endif

#define PGAX_OK 0
#define PGAX_EINVAL (-1)   /* bad argument or state */
#define PGAX_ERANGE (-2)   /* size beyond a documented limit */
#define PGAX_ENOSYS (-3)   /* combination not supported (e.g. hooks on packed bits) */
#define PGAX_ENCCL (-4)    /* NCCL library not loadable */
#define PGAX_ECUDA (-5)    /* a CUDA call failed (only from entry points that say so) */
#define PGAX_EIO (-6)      /* a file could not be read or written */

enum pgax_gene_encoding {
	PGAX_GENES_F32 = 0, /* libpga's `gene`: one float per gene */
	PGAX_GENES_BITS = 1 /* one bit per gene, 32 per word, LSB first */
};

enum pgax_objective {
	PGAX_OBJ_HOOK = 0,      /* the obj_f set through pga_set_objective_function */
	PGAX_OBJ_SUM = 1,       /* sum of genes */
	PGAX_OBJ_RASTRIGIN = 2, /* -(10 L + sum(x^2 - 10 cos 2 pi x)), x = (g - 0.5) * 10.24 */
	PGAX_OBJ_SPHERE = 3,    /* -sum(x^2), same x */
	PGAX_OBJ_ONEMAX = 4     /* number of set bits (PGAX_GENES_BITS only) */
};

enum pgax_topology {
	PGAX_TOPO_RANDOM = 1, /* a fresh random permutation per exchange (include/pga.h:148) */
	PGAX_TOPO_RING = 0    /* island i -> (i+1) mod K */
};

/* ---- configuration ------------------------------------------------------ */

/* Seed of the counter-based generator.  Replaces the reference's
 * curandSetPseudoRandomGeneratorSeed(time(NULL)) (src/pga.cu:153-154).
 * Affects populations created afterwards and all later draws. */
int pgax_set_seed(pga_t *, uint64_t seed);
uint64_t pgax_get_seed(pga_t *);

/* pga_create_population with an explicit gene encoding (src/pga.cu:177-204).
 * genome_len counts genes (bits for PGAX_GENES_BITS). */
population_t *pgax_create_population(pga_t *, unsigned long size, unsigned genome_len,
				     enum pgax_gene_encoding enc);

/* Selects a built-in objective (inlined into the fused kernel) or, with
 * PGAX_OBJ_HOOK, returns to the user hook. */
int pgax_set_builtin_objective(pga_t *, enum pgax_objective which);

/* ---- selection schemes -----------------------------------------------------
 * The reference declares enum crossover_selection_type as "pretty much just a
 * placeholder" with the single member TOURNAMENT, notes "XXX add different types of
 * selections" (include/pga.h:36-42, :103) and ignores the argument of pga_crossover
 * (src/pga.cu:319, :329).  pga.h stays frozen, so the further schemes are values of the
 * same enum type defined here:
 *   TOURNAMENT              the reference's tournament of TOURNAMENT_POPULATION = 2
 *                           candidates (src/pga.cu:278-292)
 *   PGAX_TOURNAMENT_OF(k)   the same rule among k candidates, 1 <= k <= PGAX_MAX_TOURNAMENT:
 *                           parent 1 from draws 0..k-1 of the child's rand row, parent 2
 *                           from draws k..2k-1; a later candidate wins only if strictly
 *                           better.  k = 1 is uniform selection (no pressure).
 * pga_crossover / pga_crossover_all honour their type argument (an unknown value ends the
 * program with a message, it is not ignored); pga_run and pga_run_islands use the scheme
 * set here (default TOURNAMENT, as src/pga.cu:384 passes). */
#define PGAX_SELECTION_BASE 0x100u
#define PGAX_MAX_TOURNAMENT 16u
#define PGAX_TOURNAMENT_OF(k) ((enum crossover_selection_type)(PGAX_SELECTION_BASE + (unsigned)(k)))
#define PGAX_SELECT_UNIFORM PGAX_TOURNAMENT_OF(1)
int pgax_set_selection(pga_t *, enum crossover_selection_type);

/* ---- early stop ----------------------------------------------------------------
 * include/pga.h:141 and :147 promise that pga_run / pga_run_islands run "until
 * n-generations or obj_func(best_genome) == value"; no signature carries the value and
 * the reference has no such code.  pgax_set_target(value, check_every) arms it: the run
 * ends after the first generation g, g a multiple of check_every, in which the best
 * score of the population (of any island, on any rank) is >= value.  The check costs one
 * 4-byte copy that overlaps the next generation; that next generation is dropped when
 * the check succeeds.  check_every = 0 disarms.  pgax_generations_run tells how many
 * generations the last pga_run / pga_run_islands executed. */
int pgax_set_target(pga_t *, float value, unsigned check_every);
unsigned pgax_generations_run(pga_t *);

/* ---- dump / restore --------------------------------------------------------------
 * The reference has no checkpointing.  pgax_save writes the solver's seed, island
 * schedule state and, per population, shape, epoch, current generation and scores to
 * `path`; pgax_restore reads them back into a solver that is empty (the populations
 * are created) or holds populations of the same shapes.  A run resumed from the file
 * continues exactly as the uninterrupted run would.  Hooks are device addresses of the
 * running process and are not saved: set them again after a restore.  next_gen is not
 * saved (it only matters to a crossover hook that leaves genes unwritten). */
int pgax_save(pga_t *, const char *path);
int pgax_restore(pga_t *, const char *path);

/* Suppresses the "%f\n" line pga_get_best prints (src/pga.cu:230). */
int pgax_set_quiet(pga_t *, int quiet);

/* Runs the solver's kernels on a caller-owned cudaStream_t (passed as void*);
 * NULL returns to the solver's own stream. */
int pgax_set_stream(pga_t *, void *cuda_stream);

/* Records the two tournament winners of every child during the next crossover /
 * fused generations (for parity checks); off by default. */
int pgax_capture_selection(pga_t *, population_t *, int enable);

/* ---- host-buffer transfers (H2D / D2H inside the call) ------------------ */

/* Row layout on the host: [size][row_words] 32-bit words, row_words =
 * genome_len (f32) or ceil(genome_len/32) (bits); no padding. */
population_t *pgax_population(pga_t *, unsigned index); /* the index-th population, NULL past the end */
int pgax_encoding(population_t *);                       /* enum pgax_gene_encoding */
unsigned pgax_row_words(population_t *);
unsigned long pgax_population_size(population_t *);
unsigned pgax_genome_len(population_t *);
uint32_t pgax_epoch(population_t *);
int pgax_set_epoch(population_t *, uint32_t epoch);
uint32_t pgax_island_id(population_t *);

int pgax_upload_genomes(pga_t *, population_t *, const void *host_rows, size_t bytes);
int pgax_download_genomes(pga_t *, population_t *, void *host_rows, size_t bytes);
int pgax_download_next_genomes(pga_t *, population_t *, void *host_rows, size_t bytes);
int pgax_upload_scores(pga_t *, population_t *, const float *host_scores, size_t count);
int pgax_download_scores(pga_t *, population_t *, float *host_scores, size_t count);
int pgax_download_selection(pga_t *, population_t *, int32_t *host_pairs, size_t count_pairs);

/* pga_get_best without the print and without the row: index and score of the
 * first maximum (src/pga.cu:221-229). */
int pgax_best(pga_t *, population_t *, unsigned long *index, float *score);
/* Packed-bit populations: the best row as stored (row_words words). */
int pgax_get_best_words(pga_t *, population_t *, uint32_t *row_out);

/* ---- ranking and migration ---------------------------------------------- */

/* Indices of the k best (worst = 0) or k worst (worst = 1) individuals under the
 * engine's total order (higher score first, then lower index; NaN last), listed
 * in ascending index order.  host_indices receives min(k, size) entries. */
int pgax_select_top(pga_t *, population_t *, unsigned long k, int worst, uint32_t *host_indices);

/* No counterpart in the reference.  pgax_select_top for callers that cannot let
 * the library allocate: a process whose GPU memory belongs entirely to its own
 * pool, or a loop that cannot take the device-wide stall of cudaMalloc/cudaFree.
 * It needs no solver and no population: `dev_scores` (n floats) and `dev_indices`
 * (room for min(k, n) indices) are DEVICE pointers supplied by the caller, and the
 * working memory is a fixed 72 KB array inside the library image, held by every
 * GPU that loads the library until the process exits.  Same result and order as
 * pgax_select_top.  Rules:
 *   - n may not exceed PGAX_NOALLOC_MAX_INDIVIDUALS; larger inputs return
 *     PGAX_ERANGE with a message on stderr and launch nothing;
 *   - the work is queued on `cuda_stream` (a cudaStream_t, NULL = default stream)
 *     and the call returns without waiting; the indices are valid once the stream
 *     has drained;
 *   - the FIRST call on each device also looks up the array's device address
 *     (cudaGetSymbolAddress) with that device current: make that first call
 *     before any stream capture begins;
 *   - all calls on a device share the one array: calls whose kernels could overlap
 *     on the GPU (two streams, two threads) race on it and must be serialised by
 *     the caller.  It never calls cudaMalloc, cudaFree or any allocating routine.
 * Returns PGAX_OK, PGAX_EINVAL (NULL pointer), PGAX_ERANGE or PGAX_ECUDA. */
#define PGAX_NOALLOC_MAX_INDIVIDUALS (1u << 24)
int pgax_select_top_noalloc(const float *dev_scores, unsigned long n, unsigned long k, int worst,
			    uint32_t *dev_indices, void *cuda_stream);

int pgax_set_topology(pga_t *, enum pgax_topology);
/* Generations between choosing emigrants and inserting them at the destination
 * in pga_run_islands: 1 (default) overlaps the transfer with one generation of
 * compute, 0 is synchronous. */
int pgax_set_migration_delay(pga_t *, unsigned generations);
/* Number of individuals "top pct percent" means: floor(size * pct / 100). */
unsigned long pgax_migration_count(unsigned long size, float pct);

/* ---- one island group per GPU: process-per-GPU over NCCL ----------------- */

/* Rank 0 calls pgax_dist_unique_id, ships the 128 bytes to every rank by any
 * means (torch.distributed, MPI, a file), then every rank calls pgax_dist_init.
 * Afterwards local population s of rank r is island s*world + r of a ring (or
 * random topology) over world*K islands, K = populations per rank.  Every rank
 * must hold the same number of populations, of the same size and genome length.
 * Create populations AFTER pgax_dist_init / pgax_dist_set_layout so that their
 * initial draws use the global island number.  NCCL is dlopen'ed on first use;
 * programs that never call these link without it (reference link line:
 * test/Makefile:3). */
#define PGAX_UNIQUE_ID_BYTES 128
int pgax_dist_unique_id(void *out_bytes);
int pgax_dist_init(pga_t *, int rank, int world, const void *unique_id_bytes);
/* Host-only island bookkeeping without a communicator (tests, planning). */
int pgax_dist_set_layout(pga_t *, int rank, int world);
int pgax_dist_rank(pga_t *);
int pgax_dist_world(pga_t *);

/* The exchange schedule, as a pure host function (no GPU, no solver): who sends
 * to whom in exchange number `round`.  Global island g lives on rank g % world as
 * local population g / world.  Transfers are listed in ascending source-island
 * order; every rank posts its sends and receives in that order.  Returns the
 * number of transfers written (self-transfers are skipped) or PGAX_E*. */
#define PGAX_MAX_GLOBAL_ISLANDS (MAX_POPULATIONS * 64)
typedef struct pgax_transfer {
	uint32_t src_island, dst_island;
	int32_t src_rank, dst_rank;
	uint32_t src_local, dst_local;
} pgax_transfer;
int pgax_plan_exchange(uint64_t seed, uint32_t round, enum pgax_topology topology, int world,
		       unsigned islands_per_rank, pgax_transfer *out, unsigned capacity);

/* ---- timing --------------------------------------------------------------- */

/* Device time in ms of the generation loop of the last pga_run/pga_run_islands,
 * from CUDA events on the solver's stream. */
float pgax_last_run_ms(pga_t *);
/* Kernels launched by this solver so far. */
unsigned long long pgax_kernel_launches(pga_t *);
/* Bytes this solver has copied from the device to the host so far (results, scores,
 * rows, index lists).  pga_get_best copies all N scores in the reference
 * (src/pga.cu:219-220); here it is a 16-byte key and one row. */
unsigned long long pgax_bytes_to_host(pga_t *);

/* ---- the reference's demo hooks, compiled into the shared library --------- */

/* Device addresses of re-written versions of the hooks in the reference's demo
 * programs, for bindings that cannot compile CUDA themselves:
 * "sum_obj" (test/test.cu:24-30), "knapsack_obj" (test2/test.cu:28-36),
 * "tsp_obj" (test3/test.cu:26-46), "tsp_crossover" (test3/test.cu:48-64).
 * NULL for unknown names. */
void *pgax_demo_hook(const char *name);
/* Uploads a city_count x city_count matrix the way test3 does (test3/test.cu:70-79). */
int pgax_demo_tsp_load(const float *matrix, int city_count);

const char *pgax_version(void);

#ifdef __cplusplus
}
#endif
#endif
