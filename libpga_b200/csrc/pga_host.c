/*
 * pga_host.c -- the host side of the B200-native libpga engine, in plain C.
 *
 * Implements every function of include/pga.h (reference: src/pga.cu:148-395) and
 * the extension surface of include/pga_b200.h on top of the thin device C-ABI in
 * pga_dev.h.  This file holds no CUDA: handles, validation, hook bookkeeping, the
 * generation loop and the migration schedule live here; buffers, kernels, streams
 * and NCCL live in pga_dev.cu.
 *
 * There is no CPU fallback.  If the device layer cannot reach a GPU the first
 * CUDA call prints "GPUassert: ..." and exits, like the reference.
 */
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

#include "../../include/pga.h"
#include "../../include/pga_b200.h"
#include "pga_dev.h"

struct population { /* reference: src/pga.cu:37-46 */
	unsigned long size;
	unsigned genome_len;
	int encoding;
	uint32_t island; /* global island id: counter word of every draw */
	uint32_t epoch;  /* number of pga_fill_random_values calls so far */
	int scores_valid; /* score[] describes current_gen */
	int best_in_kernel; /* the last scoring launch also reduced the best */
	pgadev_pop *dev;
};

struct pga { /* reference: src/pga.cu:48-56 */
	unsigned p_count;
	population_t *populations[MAX_POPULATIONS];
	obj_f objective;
	mutate_f mutate;       /* NULL = default operator */
	crossover_f crossover; /* NULL = default operator */
	int builtin;           /* enum pgax_objective */
	uint64_t seed;
	int quiet;
	int topology;
	unsigned delay;
	uint32_t migrate_round;
	int rank, world; /* island layout across processes */
	int have_comm;
	unsigned reserve_sms;   /* blocks held back in the next launches (an exchange is in flight beside them) */
	unsigned reserve_conf;  /* $PGA_RESERVE_SMS, default 8 */
	float last_run_ms;
	unsigned tournament;      /* candidates per parent used by pga_run / pga_run_islands; 2 = TOURNAMENT */
	int target_armed;         /* pgax_set_target: stop when the best score reaches target_value */
	float target_value;
	unsigned target_every;
	unsigned generations_run; /* by the last pga_run / pga_run_islands */
	pgadev_ctx *dev;
};

/* an exchange in flight: chosen at generation g, lands at g + delay */
struct exchange {
	int active;
	unsigned land_gen;
	unsigned start_gen;
	uint64_t k_in[MAX_POPULATIONS];         /* immigrants per local island */
};


/* ------------------------------------------------------------------------- */

pga_t *pga_init() /* src/pga.cu:148-164 */
{
	pga_t *p = (pga_t *)calloc(1, sizeof(*p));
	if (p == NULL)
		return NULL;
	const char *env = getenv("PGA_SEED");
	p->seed = env && *env ? strtoull(env, NULL, 0) : (uint64_t)time(NULL); /* :154 */
	p->builtin = PGAX_OBJ_HOOK;
	p->topology = PGAX_TOPO_RANDOM; /* "random migrations", include/pga.h:148 */
	p->delay = 1;
	p->tournament = 2; /* TOURNAMENT_POPULATION, src/pga.cu:278 */
	env = getenv("PGA_RESERVE_SMS");
	p->reserve_conf = env && *env ? (unsigned)strtoul(env, NULL, 0) : 8u;
	p->rank = 0;
	p->world = 1;
	p->dev = pgadev_ctx_create();
	if (p->dev == NULL) {
		free(p);
		return NULL;
	}
	return p;
}

void pga_deinit(pga_t *p) /* src/pga.cu:166-175 */
{
	if (p == NULL)
		return;
	pgadev_ctx_sync(p->dev);
	for (unsigned i = 0; i < p->p_count; ++i) {
		pgadev_pop_destroy(p->dev, p->populations[i]->dev);
		free(p->populations[i]);
	}
	pgadev_ctx_destroy(p->dev);
	free(p);
}

population_t *pgax_create_population(pga_t *p, unsigned long size, unsigned genome_len,
				     enum pgax_gene_encoding enc)
{
	if (p->p_count == MAX_POPULATIONS) /* src/pga.cu:180 */
		return NULL;
	if (genome_len < 4) /* src/pga.cu:184: selection and mutation read draws 0..3 */
		return NULL;
	if (enc != PGAX_GENES_F32 && enc != PGAX_GENES_BITS)
		return NULL;
	if (size >= 0xFFFFFFFFul) /* indices travel as 32-bit words */
		return NULL;
	population_t *pop = (population_t *)calloc(1, sizeof(*pop));
	if (pop == NULL)
		return NULL;
	pop->size = size;
	pop->genome_len = genome_len;
	pop->encoding = enc;
	pop->dev = pgadev_pop_create(p->dev, size, genome_len, enc);
	if (pop->dev == NULL) {
		free(pop);
		return NULL;
	}
	/* global island id (counter word of every draw): local index * world + rank */
	pop->island = p->p_count * (uint32_t)p->world + (uint32_t)p->rank;
	p->populations[p->p_count++] = pop;
	pgadev_pop_init_random(p->dev, pop->dev, p->seed, pop->island); /* :116-117, :81-86 */
	return pop;
}

population_t *pga_create_population(pga_t *p, unsigned long size, unsigned genome_len, enum population_type type)
{
	if ((int)type < 0 || type >= MAX_POPULATION_TYPE) /* :95-97 has one generator */
		return NULL;
	return pgax_create_population(p, size, genome_len, PGAX_GENES_F32);
}

void pga_set_objective_function(pga_t *p, obj_f f) /* src/pga.cu:206-208 */
{
	p->objective = f;
	p->builtin = PGAX_OBJ_HOOK;
	for (unsigned i = 0; i < p->p_count; ++i)
		p->populations[i]->scores_valid = 0;
}

void pga_set_mutate_function(pga_t *p, mutate_f f) { p->mutate = f; }          /* src/pga.cu:210-212 */
void pga_set_crossover_function(pga_t *p, crossover_f f) { p->crossover = f; } /* src/pga.cu:214-216 */

/* ------------------------------------------------------------------------- */

static void die(const char *what)
{
	fprintf(stderr, "libpga: %s\n", what);
	exit(2);
}

static void fill_step(pga_t *p, population_t *pop, unsigned phases, pgadev_step *st)
{
	st->phases = phases;
	st->seed = p->seed;
	st->island = pop->island;
	st->epoch = pop->epoch;
	st->objective = p->builtin;
	st->obj_hook = (void *)p->objective;
	st->mutate_hook = (void *)p->mutate;
	st->crossover_hook = (void *)p->crossover;
	st->reserve_sms = p->reserve_sms;
	st->tournament = p->tournament;
	if ((phases & PGADEV_PH_EVALUATE) && p->builtin == PGAX_OBJ_HOOK && p->objective == NULL)
		die("no objective function set (pga_set_objective_function)");
}

static void launch(pga_t *p, population_t *pop, unsigned phases)
{
	pgadev_step st;
	fill_step(p, pop, phases, &st);
	int rc = pgadev_launch(p->dev, pop->dev, &st);
	if (rc == PGAX_ENOSYS)
		die("this combination of gene encoding, objective and hooks is not supported");
	if (rc == PGAX_ERANGE)
		die("genome too long for the hook path (rows are staged in shared memory)");
	if (rc != 0)
		die("launch rejected");
}

static void evaluate_async(pga_t *p, population_t *pop)
{
	launch(p, pop, PGADEV_PH_EVALUATE);
	pop->scores_valid = 1;
	pop->best_in_kernel = 1;
}

/* one whole generation in one kernel: draws, selection, crossover, mutation,
 * scoring of the children, best-of-population (body of src/pga.cu:381-387 plus
 * the evaluate that follows it) */
static void generation_async(pga_t *p, population_t *pop)
{
	if (!pop->scores_valid)
		evaluate_async(p, pop);
	pop->epoch += 1;
	launch(p, pop, PGADEV_PH_FUSED);
	pgadev_swap_genes(pop->dev);
	pgadev_swap_scores(pop->dev);
	pop->scores_valid = 1;
	pop->best_in_kernel = 1;
}

void pga_evaluate(pga_t *p, population_t *pop) /* src/pga.cu:264-270 */
{
	evaluate_async(p, pop);
	pgadev_ctx_sync(p->dev);
}

void pga_evaluate_all(pga_t *p) /* src/pga.cu:272-276 */
{
	for (unsigned i = 0; i < p->p_count; ++i)
		evaluate_async(p, p->populations[i]);
	pgadev_ctx_sync(p->dev);
}

/* Candidates per parent a selection type stands for, 0 if it is not a selection type.
 * The reference declares the enum as "pretty much just a placeholder" and drops the
 * argument (include/pga.h:36-42, :103 "XXX add different types of selections";
 * src/pga.cu:319, :329).  Here TOURNAMENT is the reference's size-2 tournament and
 * PGAX_TOURNAMENT_OF(k) a tournament among k candidates (k = 1: uniform selection). */
static unsigned selection_size(enum crossover_selection_type type)
{
	if (type == TOURNAMENT)
		return 2;
	unsigned v = (unsigned)type;
	if (v > PGAX_SELECTION_BASE && v <= PGAX_SELECTION_BASE + PGAX_MAX_TOURNAMENT)
		return v - PGAX_SELECTION_BASE;
	return 0;
}

static void crossover_with(pga_t *p, population_t *pop, enum crossover_selection_type type)
{
	unsigned k = selection_size(type), keep = p->tournament;
	if (k == 0)
		die("pga_crossover: unknown selection type");
	p->tournament = k;
	launch(p, pop, PGADEV_PH_CROSSOVER);
	p->tournament = keep;
}

void pga_crossover(pga_t *p, population_t *pop, enum crossover_selection_type type) /* src/pga.cu:319-325 */
{
	crossover_with(p, pop, type);
	pgadev_ctx_sync(p->dev);
}

void pga_crossover_all(pga_t *p, enum crossover_selection_type type) /* src/pga.cu:327-331 */
{
	for (unsigned i = 0; i < p->p_count; ++i)
		crossover_with(p, p->populations[i], type);
	pgadev_ctx_sync(p->dev);
}

int pgax_set_selection(pga_t *p, enum crossover_selection_type type)
{
	unsigned k = selection_size(type);
	if (k == 0)
		return PGAX_EINVAL;
	p->tournament = k;
	return 0;
}

void pga_mutate(pga_t *p, population_t *pop) /* src/pga.cu:349-354 */
{
	launch(p, pop, PGADEV_PH_MUTATE);
	pgadev_ctx_sync(p->dev);
}

void pga_mutate_all(pga_t *p) /* src/pga.cu:356-360 */
{
	for (unsigned i = 0; i < p->p_count; ++i)
		launch(p, p->populations[i], PGADEV_PH_MUTATE);
	pgadev_ctx_sync(p->dev);
}

void pga_swap_generations(pga_t *p, population_t *pop) /* src/pga.cu:362-366 */
{
	(void)p;
	pgadev_swap_genes(pop->dev);
	pop->scores_valid = 0; /* score[] still describes the old generation */
	pop->best_in_kernel = 0;
}

void pga_fill_random_values(pga_t *p, population_t *pop) /* src/pga.cu:103-105 */
{
	(void)p;
	pop->epoch += 1; /* counter-based draws: a new epoch IS a new rand buffer */
}

/* ---- early stop -----------------------------------------------------------------
 * include/pga.h:141 and :147 say pga_run / pga_run_islands run "until n-generations or
 * obj_func(best_genome) == value"; no signature carries the value and the reference has
 * no such code.  pgax_set_target arms it: after every generation g that is a multiple of
 * `every`, the best score of the population(s) -- already reduced inside the generation
 * kernel -- is copied to the host (8 bytes) while generation g+1 runs; if it had reached
 * the target, generation g+1 is dropped (the two buffers are swapped back) and the run
 * ends with generation g current.  Nothing is launched or copied when no target is set. */
static int target_check_due(pga_t *p, unsigned g, unsigned n)
{
	return p->target_armed && p->target_every && g % p->target_every == 0 && g < n;
}

static void target_check_begin(pga_t *p, unsigned count)
{
	pgadev_pop *pops[MAX_POPULATIONS];
	int reduced[MAX_POPULATIONS];
	for (unsigned s = 0; s < count; ++s) {
		pops[s] = p->populations[s]->dev;
		reduced[s] = p->populations[s]->best_in_kernel || p->populations[s]->size == 0;
		p->populations[s]->best_in_kernel = 1;
	}
	pgadev_check_begin(p->dev, pops, reduced, count, p->world > 1 && p->have_comm);
}

static int target_reached(pga_t *p)
{
	float best = pgadev_check_wait(p->dev);
	return best >= p->target_value;
}

/* drop the generation that was launched speculatively: the previous one is current again */
static void undo_generation(population_t *pop)
{
	pgadev_swap_genes(pop->dev);
	pgadev_swap_scores(pop->dev);
	pop->epoch -= 1;
	pop->scores_valid = 1;
	pop->best_in_kernel = 0; /* the best slot of the dropped launch is stale: reduce again when asked */
}

void pga_run(pga_t *p, unsigned n) /* src/pga.cu:376-391 */
{
	if (p->p_count == 0)
		return;
	population_t *pop = p->populations[0];
	if (!pop->scores_valid)
		evaluate_async(p, pop);
	pgadev_timer_start(p->dev);
	unsigned done = n;
	int pending = 0;
	for (unsigned g = 1; g <= n; ++g) {
		generation_async(p, pop);
		/* the check of generation g-1 was read back while generation g ran */
		if (pending && target_reached(p)) {
			undo_generation(pop);
			done = g - 1;
			break;
		}
		pending = 0;
		if (target_check_due(p, g, n)) {
			target_check_begin(p, 1); /* pga_run works on populations[0] only (src/pga.cu:382-386) */
			pending = 1;
		}
	}
	p->generations_run = done;
	pgadev_timer_stop(p->dev);
	p->last_run_ms = pgadev_timer_ms(p->dev);
	pgadev_ctx_sync(p->dev);
}

/* ---- best-individual queries ---------------------------------------------- */

static gene *row_to_genes(pga_t *p, population_t *pop, uint64_t index)
{
	unsigned rw = pgadev_pop_row_words(pop->dev);
	gene *out = (gene *)malloc(sizeof(gene) * pop->genome_len);
	if (out == NULL)
		return NULL;
	if (pop->encoding == PGAX_GENES_BITS) {
		uint32_t *w = (uint32_t *)malloc(sizeof(uint32_t) * rw);
		pgadev_download_row(p->dev, pop->dev, index, w);
		for (unsigned j = 0; j < pop->genome_len; ++j)
			out[j] = ((w[j >> 5] >> (j & 31)) & 1u) ? 1.0f : 0.0f;
		free(w);
	} else {
		pgadev_download_row(p->dev, pop->dev, index, out);
	}
	return out;
}

static int best_of(pga_t *p, population_t *pop, uint64_t *index, float *score)
{
	if (pop->size == 0)
		return PGAX_EINVAL;
	if (!pop->scores_valid)
		evaluate_async(p, pop);
	pgadev_best(p->dev, pop->dev, pop->best_in_kernel, index, score);
	return 0;
}

gene *pga_get_best(pga_t *p, population_t *pop) /* src/pga.cu:218-236 */
{
	uint64_t index;
	float score;
	if (best_of(p, pop, &index, &score) != 0)
		return NULL;
	if (!p->quiet) {
		printf("%f\n", score); /* src/pga.cu:230 */
		fflush(stdout);
	}
	return row_to_genes(p, pop, index);
}

int pgax_best(pga_t *p, population_t *pop, unsigned long *index, float *score)
{
	uint64_t i;
	float s;
	int rc = best_of(p, pop, &i, &s);
	if (rc != 0)
		return rc;
	if (index) *index = (unsigned long)i;
	if (score) *score = s;
	return 0;
}

int pgax_get_best_words(pga_t *p, population_t *pop, uint32_t *row_out)
{
	uint64_t i;
	float s;
	int rc = best_of(p, pop, &i, &s);
	if (rc != 0)
		return rc;
	pgadev_download_row(p->dev, pop->dev, i, row_out);
	return 0;
}

/* total order shared with the device: higher score first, then lower index, NaN last */
static uint32_t order_key(float s)
{
	uint32_t b;
	if (s != s)
		return 0u;
	s = s + 0.0f;
	memcpy(&b, &s, 4);
	return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}

struct ranked { uint32_t key, pop, idx; };

static int ranked_cmp(const void *a, const void *b)
{
	const struct ranked *x = (const struct ranked *)a, *y = (const struct ranked *)b;
	if (x->key != y->key) return x->key < y->key ? 1 : -1;
	if (x->pop != y->pop) return x->pop < y->pop ? -1 : 1;
	if (x->idx != y->idx) return x->idx < y->idx ? -1 : 1;
	return 0;
}

/* the `length` best of one population as (key, pop, idx), best first */
static unsigned top_of(pga_t *p, unsigned pi, unsigned length, struct ranked *out)
{
	population_t *pop = p->populations[pi];
	if (length > pop->size)
		length = (unsigned)pop->size;
	if (length == 0)
		return 0;
	if (!pop->scores_valid)
		evaluate_async(p, pop);
	uint32_t *idx = (uint32_t *)malloc(sizeof(uint32_t) * length);
	float *top = (float *)malloc(sizeof(float) * length);
	pgadev_select_top(p->dev, pop->dev, length, 0, 0);
	pgadev_download_top(p->dev, pop->dev, 0, length, idx);
	pgadev_download_top_scores(p->dev, pop->dev, 0, length, top); /* `length` scores, not the whole array */
	for (unsigned j = 0; j < length; ++j) {
		out[j].key = order_key(top[j]);
		out[j].pop = pi;
		out[j].idx = idx[j];
	}
	free(top);
	free(idx);
	qsort(out, length, sizeof(*out), ranked_cmp);
	return length;
}

gene **pga_get_best_top(pga_t *p, population_t *pop, unsigned length) /* stub: src/pga.cu:238-240 */
{
	unsigned pi = 0;
	while (pi < p->p_count && p->populations[pi] != pop)
		++pi;
	if (pi == p->p_count)
		return NULL;
	struct ranked *r = (struct ranked *)malloc(sizeof(*r) * (length ? length : 1));
	unsigned n = top_of(p, pi, length, r);
	gene **out = (gene **)calloc(length ? length : 1, sizeof(gene *));
	for (unsigned j = 0; j < n; ++j)
		out[j] = row_to_genes(p, pop, r[j].idx);
	free(r);
	return out;
}

gene *pga_get_best_all(pga_t *p) /* stub: src/pga.cu:242-244 */
{
	int have = 0;
	unsigned best_pop = 0;
	uint64_t best_idx = 0;
	float best = 0;
	for (unsigned i = 0; i < p->p_count; ++i) {
		uint64_t idx;
		float s;
		if (best_of(p, p->populations[i], &idx, &s) != 0)
			continue;
		if (!have || best < s) { /* ties stay with the earlier population */
			have = 1; best = s; best_idx = idx; best_pop = i;
		}
	}
	if (!have)
		return NULL;
	if (!p->quiet) {
		printf("%f\n", best);
		fflush(stdout);
	}
	return row_to_genes(p, p->populations[best_pop], best_idx);
}

gene **pga_get_best_top_all(pga_t *p, unsigned length) /* stub: src/pga.cu:246-248 */
{
	struct ranked *r = (struct ranked *)malloc(sizeof(*r) * ((size_t)length * p->p_count + 1));
	unsigned n = 0;
	for (unsigned i = 0; i < p->p_count; ++i)
		n += top_of(p, i, length, r + n);
	qsort(r, n, sizeof(*r), ranked_cmp);
	gene **out = (gene **)calloc(length ? length : 1, sizeof(gene *));
	for (unsigned j = 0; j < length && j < n; ++j)
		out[j] = row_to_genes(p, p->populations[r[j].pop], r[j].idx);
	free(r);
	return out;
}

/* ---- migration -------------------------------------------------------------- */

unsigned long pgax_migration_count(unsigned long size, float pct)
{
	if (!(pct > 0.0f))
		return 0;
	double k = floor((double)size * (double)pct / 100.0);
	if (k > (double)size)
		k = (double)size;
	return (unsigned long)k;
}

/* Philox words for the host-side permutation: same generator, same layout as the
 * device draws (stream 3 = MIGRATE). */
static void philox4x32_10(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4])
{
	uint32_t c[4] = { ctr[0], ctr[1], ctr[2], ctr[3] }, k0 = key[0], k1 = key[1];
	for (int r = 0; r < 10; ++r) {
		uint64_t a = (uint64_t)0xD2511F53u * c[0], b = (uint64_t)0xCD9E8D57u * c[2];
		uint32_t n0 = (uint32_t)(b >> 32) ^ c[1] ^ k0, n2 = (uint32_t)(a >> 32) ^ c[3] ^ k1;
		c[1] = (uint32_t)b; c[3] = (uint32_t)a; c[0] = n0; c[2] = n2;
		k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
	}
	memcpy(out, c, sizeof(c));
}

/* destination of every global island for exchange number `round`; pure */
static void destinations(uint64_t seed, uint32_t round, int topology, unsigned total, uint32_t *dest)
{
	for (unsigned i = 0; i < total; ++i)
		dest[i] = i;
	if (topology == PGAX_TOPO_RANDOM) {
		uint32_t key[2] = { (uint32_t)seed, (uint32_t)(seed >> 32) };
		for (unsigned i = total; i-- > 1;) { /* Fisher-Yates */
			uint32_t ctr[4] = { i >> 2, 0, round, 3u << 24 }, w[4];
			philox4x32_10(ctr, key, w);
			unsigned j = w[i & 3] % (i + 1);
			uint32_t t = dest[i]; dest[i] = dest[j]; dest[j] = t;
		}
	} else {
		for (unsigned i = 0; i < total; ++i)
			dest[i] = (i + 1) % total;
	}
}

int pgax_plan_exchange(uint64_t seed, uint32_t round, enum pgax_topology topology, int world,
		       unsigned islands_per_rank, pgax_transfer *out, unsigned capacity)
{
	if (world < 1 || out == NULL)
		return PGAX_EINVAL;
	unsigned total = islands_per_rank * (unsigned)world;
	if (total > PGAX_MAX_GLOBAL_ISLANDS)
		return PGAX_ERANGE;
	uint32_t dest[PGAX_MAX_GLOBAL_ISLANDS];
	destinations(seed, round, topology, total, dest);
	unsigned count = 0;
	for (unsigned g = 0; g < total; ++g) { /* ascending source order */
		if (dest[g] == g)
			continue;
		if (count == capacity)
			return PGAX_ERANGE;
		out[count].src_island = g;
		out[count].dst_island = dest[g];
		out[count].src_rank = (int32_t)(g % (unsigned)world);
		out[count].dst_rank = (int32_t)(dest[g] % (unsigned)world);
		out[count].src_local = g / (unsigned)world;
		out[count].dst_local = dest[g] / (unsigned)world;
		++count;
	}
	return (int)count;
}

/* choose emigrants on every local island and put the transfers on the side stream */
static void exchange_start(pga_t *p, struct exchange *ex, float pct)
{
	unsigned K = p->p_count;
	pgax_transfer plan[PGAX_MAX_GLOBAL_ISLANDS];
	int count = pgax_plan_exchange(p->seed, p->migrate_round, (enum pgax_topology)p->topology, p->world, K, plan,
				       PGAX_MAX_GLOBAL_ISLANDS);
	if (count < 0)
		die("too many islands");
	p->migrate_round += 1;
	memset(ex->k_in, 0, sizeof(ex->k_in));

	/* boxes first: allocation must stay outside the NCCL group.  An island's inbox takes
	 * what its source sends, min(count(source), own size), which exceeds its own count
	 * when the source is the larger island: size both boxes for the largest local sender */
	uint64_t k_max = 0;
	for (unsigned s = 0; s < K; ++s) {
		uint64_t k = pgax_migration_count(p->populations[s]->size, pct);
		if (k > k_max) k_max = k;
	}
	for (unsigned s = 0; s < K; ++s) {
		uint64_t cap = p->populations[s]->size < k_max ? p->populations[s]->size : k_max;
		pgadev_reserve_boxes(p->dev, p->populations[s]->dev, cap);
	}

	/* Emigrants are ranked on the main stream, right behind the generation that produced
	 * them: the single cooperative launch takes ~15 us with the whole GPU to itself, while
	 * beside the next generation it would either wait for that kernel to drain (all of its
	 * blocks must be resident at once) or crawl on the few SMs held back for side work.
	 * Packing and the transfer only read the finished generation and go to the side
	 * stream, where they overlap the next generation. */
	uint64_t k_out[MAX_POPULATIONS];
	memset(k_out, 0, sizeof(k_out));
	for (int t = 0; t < count; ++t) {
		if (plan[t].src_rank != p->rank)
			continue;
		population_t *src = p->populations[plan[t].src_local];
		uint64_t k = pgax_migration_count(src->size, pct);
		if (plan[t].dst_rank == p->rank) { /* local destination: its shape is known */
			population_t *dst = p->populations[plan[t].dst_local];
			if (k > dst->size) k = dst->size;
			if (pgadev_pop_row_words(dst->dev) != pgadev_pop_row_words(src->dev)) k = 0;
		}
		k_out[plan[t].src_local] = k;
		if (k)
			pgadev_select_top(p->dev, src->dev, k, 0, 0);
	}
	pgadev_trace_mark(p->dev, "emigrants ranked");
	pgadev_side_after_main(p->dev);
	pgadev_route_to_side(p->dev, 1);
	for (unsigned s = 0; s < K; ++s)
		if (k_out[s])
			pgadev_pack_emigrants(p->dev, p->populations[s]->dev, k_out[s]);
	pgadev_route_to_side(p->dev, 0);
	pgadev_mark_packed(p->dev);

	/* the plan lists transfers in ascending source order on every rank, so the
	 * messages between one pair of ranks are posted in the same order on both */
	int use_nccl = p->world > 1;
	if (use_nccl) {
		if (!p->have_comm)
			die("pgax_dist_init has not been called on this solver");
		pgadev_nccl_group_begin(p->dev);
	}
	for (int t = 0; t < count; ++t) {
		int src_here = plan[t].src_rank == p->rank, dst_here = plan[t].dst_rank == p->rank;
		if (src_here && dst_here) {
			uint64_t k = k_out[plan[t].src_local];
			if (k)
				pgadev_transfer_local(p->dev, p->populations[plan[t].src_local]->dev,
						      p->populations[plan[t].dst_local]->dev, k);
			ex->k_in[plan[t].dst_local] = k;
		} else if (src_here) {
			uint64_t k = k_out[plan[t].src_local];
			if (k)
				pgadev_nccl_send_outbox(p->dev, p->populations[plan[t].src_local]->dev, k, plan[t].dst_rank);
		} else if (dst_here) {
			/* islands have the same shape on every rank (see pga_b200.h) */
			population_t *dst = p->populations[plan[t].dst_local];
			uint64_t k = pgax_migration_count(dst->size, pct);
			if (k)
				pgadev_nccl_recv_inbox(p->dev, dst->dev, k, plan[t].src_rank);
			ex->k_in[plan[t].dst_local] = k;
		}
	}
	if (use_nccl)
		pgadev_nccl_group_end(p->dev);
	ex->active = 1;
}

/* immigrants replace the worst of whatever generation is current now */
static void exchange_finish(pga_t *p, struct exchange *ex)
{
	pgadev_main_after_side(p->dev);
	pgadev_trace_mark(p->dev, "transfer waited for");
	for (unsigned s = 0; s < p->p_count; ++s) {
		uint64_t k = ex->k_in[s];
		if (!k)
			continue;
		population_t *dst = p->populations[s];
		pgadev_select_top(p->dev, dst->dev, k, 1, 1);
		pgadev_trace_mark(p->dev, "victims ranked");
		pgadev_unpack_immigrants(p->dev, dst->dev, k);
		dst->best_in_kernel = 0;
	}
	pgadev_trace_mark(p->dev, "immigrants unpacked");
	ex->active = 0;
}

static void ensure_scores(pga_t *p)
{
	for (unsigned s = 0; s < p->p_count; ++s)
		if (!p->populations[s]->scores_valid)
			evaluate_async(p, p->populations[s]);
}

void pga_migrate(pga_t *p, float pct) /* stub: src/pga.cu:368-370; include/pga.h:108-111 */
{
	if (p->p_count * (unsigned)p->world < 2)
		return;
	struct exchange ex;
	memset(&ex, 0, sizeof(ex));
	ensure_scores(p);
	exchange_start(p, &ex, pct);
	exchange_finish(p, &ex);
	pgadev_ctx_sync(p->dev);
}

void pga_migrate_between(pga_t *p, population_t *from, population_t *to, float pct) /* stub: src/pga.cu:372-374 */
{
	if (from == to || pgadev_pop_row_words(from->dev) != pgadev_pop_row_words(to->dev))
		return;
	if (!from->scores_valid) evaluate_async(p, from);
	if (!to->scores_valid) evaluate_async(p, to);
	uint64_t k = pgax_migration_count(from->size, pct);
	if (k > to->size) k = to->size;
	if (!k)
		return;
	pgadev_reserve_boxes(p->dev, to->dev, k);
	pgadev_select_top(p->dev, from->dev, k, 0, 0);
	pgadev_pack_emigrants(p->dev, from->dev, k);
	pgadev_side_after_main(p->dev);
	pgadev_transfer_local(p->dev, from->dev, to->dev, k);
	pgadev_main_after_side(p->dev);
	pgadev_select_top(p->dev, to->dev, k, 1, 1);
	pgadev_unpack_immigrants(p->dev, to->dev, k);
	to->best_in_kernel = 0;
	pgadev_ctx_sync(p->dev);
}

void pga_run_islands(pga_t *p, unsigned n, unsigned m, float pct) /* stub: src/pga.cu:393-395; include/pga.h:145-150 */
{
	if (p->p_count == 0)
		return;
	struct exchange ex;
	memset(&ex, 0, sizeof(ex));
	unsigned total = p->p_count * (unsigned)p->world;
	ensure_scores(p);
	pgadev_timer_start(p->dev);
	unsigned done = n;
	int pending = 0;
	for (unsigned g = 1; g <= n; ++g) {
		/* generation start+2 overwrites the buffers the emigrants are packed from */
		if (ex.active && g == ex.start_gen + 2)
			pgadev_main_wait_packed(p->dev);
		/* while an exchange is being ranked, packed and sent on the side stream the
		 * generation kernel leaves a few SMs free for it */
		p->reserve_sms = ex.active ? p->reserve_conf : 0;
		pgadev_trace_mark(p->dev, ex.active ? "generation (beside)" : "generation");
		for (unsigned s = 0; s < p->p_count; ++s)
			generation_async(p, p->populations[s]);
		p->reserve_sms = 0;
		pgadev_trace_mark(p->dev, "generation done");
		if (pending && target_reached(p)) { /* include/pga.h:147; see pga_run */
			for (unsigned s = 0; s < p->p_count; ++s)
				undo_generation(p->populations[s]);
			done = g - 1;
			ex.active = 0; /* an exchange still in flight never lands */
			break;
		}
		pending = 0;
		if (ex.active && ex.land_gen == g)
			exchange_finish(p, &ex);
		if (m && g % m == 0 && g + p->delay <= n && total > 1 && !ex.active) {
			exchange_start(p, &ex, pct);
			ex.start_gen = g;
			ex.land_gen = g + p->delay;
			if (p->delay == 0)
				exchange_finish(p, &ex);
		}
		if (target_check_due(p, g, n)) {
			target_check_begin(p, p->p_count);
			pending = 1;
		}
	}
	p->generations_run = done;
	pgadev_timer_stop(p->dev);
	p->last_run_ms = pgadev_timer_ms(p->dev);
	pgadev_ctx_sync(p->dev);
	pgadev_trace_report(p->dev);
}

/* ---- extension surface -------------------------------------------------------- */

int pgax_set_seed(pga_t *p, uint64_t seed) { p->seed = seed; return 0; }
uint64_t pgax_get_seed(pga_t *p) { return p->seed; }
int pgax_set_quiet(pga_t *p, int q) { p->quiet = q; return 0; }
int pgax_set_stream(pga_t *p, void *s) { pgadev_ctx_sync(p->dev); pgadev_ctx_set_stream(p->dev, s); return 0; }

int pgax_set_builtin_objective(pga_t *p, enum pgax_objective which)
{
	if ((int)which < PGAX_OBJ_HOOK || which > PGAX_OBJ_ONEMAX)
		return PGAX_EINVAL;
	p->builtin = which;
	for (unsigned i = 0; i < p->p_count; ++i)
		p->populations[i]->scores_valid = 0;
	return 0;
}

int pgax_capture_selection(pga_t *p, population_t *pop, int enable)
{
	pgadev_pop_capture_selection(p->dev, pop->dev, enable);
	return 0;
}

population_t *pgax_population(pga_t *p, unsigned index) { return index < p->p_count ? p->populations[index] : NULL; }
int pgax_encoding(population_t *pop) { return pop->encoding; }
unsigned pgax_row_words(population_t *pop) { return pgadev_pop_row_words(pop->dev); }
unsigned long pgax_population_size(population_t *pop) { return pop->size; }
unsigned pgax_genome_len(population_t *pop) { return pop->genome_len; }
uint32_t pgax_epoch(population_t *pop) { return pop->epoch; }
int pgax_set_epoch(population_t *pop, uint32_t e) { pop->epoch = e; return 0; }
uint32_t pgax_island_id(population_t *pop) { return pop->island; }

static size_t rows_bytes(population_t *pop)
{
	return (size_t)pop->size * pgadev_pop_row_words(pop->dev) * 4u;
}

int pgax_upload_genomes(pga_t *p, population_t *pop, const void *host, size_t bytes)
{
	if (host == NULL || bytes != rows_bytes(pop))
		return PGAX_EINVAL;
	pgadev_upload_rows(p->dev, pop->dev, host, 0);
	pop->scores_valid = 0;
	pop->best_in_kernel = 0;
	return 0;
}

int pgax_download_genomes(pga_t *p, population_t *pop, void *host, size_t bytes)
{
	if (host == NULL || bytes != rows_bytes(pop))
		return PGAX_EINVAL;
	pgadev_download_rows(p->dev, pop->dev, host, 0);
	return 0;
}

int pgax_download_next_genomes(pga_t *p, population_t *pop, void *host, size_t bytes)
{
	if (host == NULL || bytes != rows_bytes(pop))
		return PGAX_EINVAL;
	pgadev_download_rows(p->dev, pop->dev, host, 1);
	return 0;
}

int pgax_upload_scores(pga_t *p, population_t *pop, const float *host, size_t count)
{
	if (host == NULL || count != pop->size)
		return PGAX_EINVAL;
	pgadev_upload_scores(p->dev, pop->dev, host);
	pop->scores_valid = 1;
	pop->best_in_kernel = 0;
	return 0;
}

int pgax_download_scores(pga_t *p, population_t *pop, float *host, size_t count)
{
	if (host == NULL || count != pop->size)
		return PGAX_EINVAL;
	pgadev_download_scores(p->dev, pop->dev, host);
	return 0;
}

int pgax_download_selection(pga_t *p, population_t *pop, int32_t *host_pairs, size_t count_pairs)
{
	if (host_pairs == NULL || count_pairs != pop->size)
		return PGAX_EINVAL;
	return pgadev_download_selection(p->dev, pop->dev, host_pairs);
}

int pgax_select_top(pga_t *p, population_t *pop, unsigned long k, int worst, uint32_t *host_indices)
{
	if (host_indices == NULL)
		return PGAX_EINVAL;
	if (k > pop->size)
		k = pop->size;
	if (k == 0)
		return 0;
	if (!pop->scores_valid)
		evaluate_async(p, pop);
	pgadev_select_top(p->dev, pop->dev, k, worst ? 1 : 0, worst ? 1 : 0);
	pgadev_download_top(p->dev, pop->dev, worst ? 1 : 0, k, host_indices);
	return 0;
}

int pgax_set_target(pga_t *p, float value, unsigned check_every)
{
	p->target_armed = check_every != 0;
	p->target_value = value;
	p->target_every = check_every;
	return 0;
}

unsigned pgax_generations_run(pga_t *p) { return p->generations_run; }

/* ---- dump / restore (the reference has no checkpointing) ---- */
#define PGAX_FILE_MAGIC "PGAB200"
#define PGAX_FILE_VERSION 1u

struct file_header {
	char magic[8];
	uint32_t version, p_count;
	uint64_t seed;
	uint32_t migrate_round, delay;
	int32_t topology, builtin;
	uint32_t tournament, pad;
};

struct file_pop {
	uint64_t size;
	uint32_t genome_len, row_words, island, epoch;
	int32_t encoding, scores_valid;
};

int pgax_save(pga_t *p, const char *path)
{
	if (path == NULL)
		return PGAX_EINVAL;
	pgadev_ctx_sync(p->dev);
	FILE *f = fopen(path, "wb");
	if (f == NULL)
		return PGAX_EIO;
	struct file_header h;
	memset(&h, 0, sizeof(h));
	memcpy(h.magic, PGAX_FILE_MAGIC, 8);
	h.version = PGAX_FILE_VERSION;
	h.p_count = p->p_count;
	h.seed = p->seed;
	h.migrate_round = p->migrate_round;
	h.delay = p->delay;
	h.topology = p->topology;
	h.builtin = p->builtin;
	h.tournament = p->tournament;
	int ok = fwrite(&h, sizeof(h), 1, f) == 1;
	for (unsigned i = 0; ok && i < p->p_count; ++i) {
		population_t *pop = p->populations[i];
		struct file_pop fp;
		memset(&fp, 0, sizeof(fp));
		fp.size = pop->size;
		fp.genome_len = pop->genome_len;
		fp.row_words = pgadev_pop_row_words(pop->dev);
		fp.island = pop->island;
		fp.epoch = pop->epoch;
		fp.encoding = pop->encoding;
		fp.scores_valid = pop->scores_valid;
		ok = fwrite(&fp, sizeof(fp), 1, f) == 1;
		size_t bytes = rows_bytes(pop);
		void *rows = malloc(bytes ? bytes : 1);
		float *scores = (float *)malloc(sizeof(float) * (pop->size ? pop->size : 1));
		if (rows == NULL || scores == NULL) {
			free(rows); free(scores);
			ok = 0;
			break;
		}
		if (pop->size) {
			pgadev_download_rows(p->dev, pop->dev, rows, 0);
			pgadev_download_scores(p->dev, pop->dev, scores);
		}
		ok = ok && fwrite(rows, 1, bytes, f) == bytes && fwrite(scores, sizeof(float), pop->size, f) == pop->size;
		free(rows);
		free(scores);
	}
	if (fclose(f) != 0)
		ok = 0;
	return ok ? 0 : PGAX_EIO;
}

int pgax_restore(pga_t *p, const char *path)
{
	if (path == NULL)
		return PGAX_EINVAL;
	FILE *f = fopen(path, "rb");
	if (f == NULL)
		return PGAX_EIO;
	struct file_header h;
	int rc = PGAX_EIO;
	if (fread(&h, sizeof(h), 1, f) != 1 || memcmp(h.magic, PGAX_FILE_MAGIC, 8) != 0 || h.version != PGAX_FILE_VERSION ||
	    h.p_count > MAX_POPULATIONS)
		goto out;
	if (p->p_count != 0 && p->p_count != h.p_count) {
		rc = PGAX_EINVAL;
		goto out;
	}
	int create = p->p_count == 0;
	pgadev_ctx_sync(p->dev);
	p->seed = h.seed;
	p->migrate_round = h.migrate_round;
	p->delay = h.delay;
	p->topology = h.topology;
	p->builtin = h.builtin;
	p->tournament = h.tournament ? h.tournament : 2;
	for (unsigned i = 0; i < h.p_count; ++i) {
		struct file_pop fp;
		if (fread(&fp, sizeof(fp), 1, f) != 1)
			goto out;
		population_t *pop;
		if (create) {
			pop = pgax_create_population(p, (unsigned long)fp.size, fp.genome_len, (enum pgax_gene_encoding)fp.encoding);
			if (pop == NULL) {
				rc = PGAX_EINVAL;
				goto out;
			}
		} else {
			pop = p->populations[i];
			if (pop->size != fp.size || pop->genome_len != fp.genome_len || pop->encoding != fp.encoding) {
				rc = PGAX_EINVAL;
				goto out;
			}
		}
		if (pgadev_pop_row_words(pop->dev) != fp.row_words) {
			rc = PGAX_EINVAL;
			goto out;
		}
		pop->island = fp.island;
		pop->epoch = fp.epoch;
		size_t bytes = rows_bytes(pop);
		void *rows = malloc(bytes ? bytes : 1);
		float *scores = (float *)malloc(sizeof(float) * (pop->size ? pop->size : 1));
		int ok = rows != NULL && scores != NULL && fread(rows, 1, bytes, f) == bytes &&
			 fread(scores, sizeof(float), pop->size, f) == pop->size;
		if (ok && pop->size) {
			pgadev_upload_rows(p->dev, pop->dev, rows, 0);
			pgadev_upload_scores(p->dev, pop->dev, scores);
		}
		free(rows);
		free(scores);
		if (!ok)
			goto out;
		pop->scores_valid = fp.scores_valid;
		pop->best_in_kernel = 0;
	}
	rc = 0;
out:
	fclose(f);
	return rc;
}

int pgax_set_topology(pga_t *p, enum pgax_topology t)
{
	if (t != PGAX_TOPO_RING && t != PGAX_TOPO_RANDOM)
		return PGAX_EINVAL;
	p->topology = t;
	return 0;
}

int pgax_set_migration_delay(pga_t *p, unsigned g) { p->delay = g; return 0; }

int pgax_dist_unique_id(void *out) { return pgadev_nccl_unique_id(out); }

int pgax_dist_set_layout(pga_t *p, int rank, int world)
{
	if (world < 1 || rank < 0 || rank >= world)
		return PGAX_EINVAL;
	p->rank = rank;
	p->world = world;
	/* re-number islands that already exist; their genomes keep the draws they were
	 * created with, so create populations AFTER this call for rank-distinct islands */
	for (unsigned i = 0; i < p->p_count; ++i)
		p->populations[i]->island = i * (uint32_t)world + (uint32_t)rank;
	return 0;
}

int pgax_dist_init(pga_t *p, int rank, int world, const void *id)
{
	int rc = pgax_dist_set_layout(p, rank, world);
	if (rc != 0)
		return rc;
	rc = pgadev_nccl_init(p->dev, rank, world, id);
	if (rc != 0)
		return rc;
	p->have_comm = 1;
	return 0;
}

int pgax_dist_rank(pga_t *p) { return p->rank; }
int pgax_dist_world(pga_t *p) { return p->world; }

float pgax_last_run_ms(pga_t *p) { return p->last_run_ms; }
unsigned long long pgax_kernel_launches(pga_t *p) { return pgadev_ctx_launches(p->dev); }
unsigned long long pgax_bytes_to_host(pga_t *p) { return pgadev_ctx_d2h_bytes(p->dev); }

const char *pgax_version(void) { return "libpga-b200 0.1 (sm_100a)"; }
