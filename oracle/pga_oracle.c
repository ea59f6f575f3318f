/*
 * pga_oracle.c -- CPU oracle for the libpga generation hot path.
 *
 * TEST INFRASTRUCTURE ONLY (see pga_oracle.h).  Build: oracle/Makefile
 * (gcc -O2 -ffp-contract=off -fopenmp).  With one thread this is the serial C
 * oracle; with more threads the loops over individuals run under OpenMP and give
 * identical bits, because every random draw is a pure function of
 * (seed, island, epoch, individual, position).
 *
 * Each function names the reference lines it restates; paths are relative to
 * /root/reference.  Where the reference is a stub the header comment in
 * include/pga.h is the only source and the behaviour is fixed HERE.
 */
#include "pga_oracle.h"

#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

/* ------------------------------------------------------------------------- */
/* Philox4x32-10 (Salmon et al., SC'11; same constants and round function as  */
/* curand_philox4x32_x.h in the CUDA toolkit).  Replaces the cuRAND XORWOW    */
/* generator of src/pga.cu:35,99-101,153-154.                                 */
/* ------------------------------------------------------------------------- */
void opga_philox4x32_10(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4])
{
	uint32_t c0 = ctr[0], c1 = ctr[1], c2 = ctr[2], c3 = ctr[3];
	uint32_t k0 = key[0], k1 = key[1];
	for (int r = 0; r < 10; ++r) {
		uint64_t p0 = (uint64_t)0xD2511F53u * c0;
		uint64_t p1 = (uint64_t)0xCD9E8D57u * c2;
		uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0;
		uint32_t n1 = (uint32_t)p1;
		uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1;
		uint32_t n3 = (uint32_t)p0;
		c0 = n0; c1 = n1; c2 = n2; c3 = n3;
		k0 += 0x9E3779B9u;
		k1 += 0xBB67AE85u;
	}
	out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

/*
 * Draw layout.  word w of stream s for (island, epoch, individual):
 *   philox(ctr = {w/4, individual, epoch, s<<24 | island}, key = seed)[w%4]
 * "epoch" counts pga_fill_random_values calls on the population (src/pga.cu:103);
 * epoch 0 is the fill done inside pga_create_population (src/pga.cu:116).
 */
uint32_t opga_draw_word(uint64_t seed, uint32_t island, uint32_t epoch, uint32_t individual,
			uint32_t stream, uint32_t w)
{
	uint32_t ctr[4] = { w >> 2, individual, epoch, (stream << 24) | (island & 0xFFFFFFu) };
	uint32_t key[2] = { (uint32_t)seed, (uint32_t)(seed >> 32) };
	uint32_t out[4];
	opga_philox4x32_10(ctr, key, out);
	return out[w & 3];
}

/* uniform in [0,1): top 24 bits.  cuRAND's range is (0,1] (curand.h), which lets
 * the reference index one past the end (src/pga.cu:129,284); [0,1) cannot. */
static inline float word_to_uniform(uint32_t x)
{
	return (float)(x >> 8) * 0x1.0p-24f;
}

float opga_draw_uniform(uint64_t seed, uint32_t island, uint32_t epoch, uint32_t individual, uint32_t j)
{
	return word_to_uniform(opga_draw_word(seed, island, epoch, individual, OPGA_STREAM_ROW, j));
}

/* the four words of Philox block b of a stream */
static void draw_block_words(uint64_t seed, uint32_t island, uint32_t epoch, uint32_t individual, uint32_t stream,
			     uint32_t b, uint32_t out[4])
{
	uint32_t key[2] = { (uint32_t)seed, (uint32_t)(seed >> 32) };
	uint32_t ctr[4] = { b, individual, epoch, (stream << 24) | (island & 0xFFFFFFu) };
	opga_philox4x32_10(ctr, key, out);
}

/* the four uniforms of Philox block b of an individual's rand row */
static void rand_block(uint64_t seed, uint32_t island, uint32_t epoch, uint32_t individual, uint32_t b, float out[4])
{
	uint32_t key[2] = { (uint32_t)seed, (uint32_t)(seed >> 32) };
	uint32_t ctr[4] = { b, individual, epoch, ((uint32_t)OPGA_STREAM_ROW << 24) | (island & 0xFFFFFFu) };
	uint32_t w[4];
	opga_philox4x32_10(ctr, key, w);
	for (int e = 0; e < 4; ++e)
		out[e] = word_to_uniform(w[e]);
}

/* one individual's "rand row": the slice rand + index*genome_len of src/pga.cu:298 */
void opga_rand_row(uint64_t seed, uint32_t island, uint32_t epoch, uint32_t individual, uint32_t len, float *out)
{
	uint32_t key[2] = { (uint32_t)seed, (uint32_t)(seed >> 32) };
	for (uint32_t b = 0; b * 4 < len; ++b) {
		uint32_t ctr[4] = { b, individual, epoch, ((uint32_t)OPGA_STREAM_ROW << 24) | (island & 0xFFFFFFu) };
		uint32_t w[4];
		opga_philox4x32_10(ctr, key, w);
		for (uint32_t e = 0; e < 4 && b * 4 + e < len; ++e)
			out[b * 4 + e] = word_to_uniform(w[e]);
	}
}

/* ------------------------------------------------------------------------- */
/* Built-in objectives (not in the reference; BASELINE.json configs 2 and 3).  */
/* The operation order below IS the specification: the device evaluates the    */
/* same expression tree, so float scores agree bit for bit.                    */
/* ------------------------------------------------------------------------- */
float opga_cos2pi(float x)
{
	float k = rintf(x);
	float t = x - k; /* exact, |t| <= 0.5 */
	float a = fabsf(t);
	int flip = a > 0.25f;
	float r = flip ? 0.5f - a : a; /* [0, 0.25] */
	float s = r * r;
	float p = 0x1.f9d38ap+2f;
	p = fmaf(p, s, -0x1.a6d1f2p+4f);
	p = fmaf(p, s, 0x1.e1f506p+5f);
	p = fmaf(p, s, -0x1.55d3c8p+6f);
	p = fmaf(p, s, 0x1.03c1f0p+6f);
	p = fmaf(p, s, -0x1.3bd3ccp+4f);
	p = fmaf(p, s, 1.0f);
	return flip ? -p : p;
}

static inline float obj_accumulate(int which, float acc, float g)
{
	if (which == OPGA_OBJ_SUM)
		return acc + g;
	float x = (g - 0.5f) * 10.24f;
	if (which == OPGA_OBJ_SPHERE)
		return fmaf(x, x, acc);
	/* Rastrigin term x^2 - 10 cos(2 pi x) */
	float c = opga_cos2pi(x);
	float term = fmaf(x, x, -(10.0f * c));
	return acc + term;
}

/* Canonical reduction shape, a function of len only: G = min(32, pow2ceil(len/4))
 * virtual lanes; lane v owns the 4-gene blocks v, v+G, ...; then an xor butterfly. */
float opga_builtin_objective_f32(int which, const float *g, unsigned len)
{
	unsigned nblk = (len + 3) / 4, G = 1;
	while (G < nblk && G < 32)
		G <<= 1;
	float part[32], tmp[32];
	for (unsigned v = 0; v < G; ++v) {
		float acc = 0.0f;
		for (unsigned b = v; b < nblk; b += G)
			for (unsigned e = 0; e < 4; ++e)
				if (4 * b + e < len)
					acc = obj_accumulate(which, acc, g[4 * b + e]);
		part[v] = acc;
	}
	for (unsigned off = G >> 1; off >= 1; off >>= 1) {
		for (unsigned v = 0; v < G; ++v)
			tmp[v] = part[v] + part[v ^ off];
		memcpy(part, tmp, G * sizeof(float));
	}
	float total = part[0];
	switch (which) {
	case OPGA_OBJ_SUM: return total;
	case OPGA_OBJ_SPHERE: return -total;
	default: return -((float)(10u * len) + total);
	}
}

float opga_builtin_objective_bits(int which, const uint32_t *w, unsigned len_bits)
{
	(void)which; /* OneMax: number of set bits, exact in f32 below 2^24 */
	unsigned words = (len_bits + 31) / 32, n = 0;
	for (unsigned i = 0; i < words; ++i)
		n += (unsigned)__builtin_popcount(w[i]);
	return (float)n;
}

/* ------------------------------------------------------------------------- */
/* state: struct population / struct pga of src/pga.cu:37-56                   */
/* ------------------------------------------------------------------------- */
struct opga_pop {
	unsigned long size;
	unsigned genome_len;
	int encoding;
	unsigned row_words; /* 32-bit words per row */
	uint32_t island;    /* global island id (counter word 3) */
	uint32_t epoch;
	uint32_t *cur, *nxt; /* f32 or packed bits, row-major [size][row_words] */
	float *score;
	int32_t *sel;
	int scores_valid;
};

struct pending_migration {
	int active;
	unsigned land_gen;
	unsigned long k[OPGA_MAX_POPULATIONS];
	unsigned dest[OPGA_MAX_POPULATIONS];
	uint32_t *rows[OPGA_MAX_POPULATIONS];
	float *scores[OPGA_MAX_POPULATIONS];
};

struct opga {
	unsigned p_count;
	opga_pop_t *populations[OPGA_MAX_POPULATIONS];
	opga_obj_f objective;
	int builtin;
	opga_mutate_f mutate;       /* NULL = default (include/pga.h:76) */
	opga_crossover_f crossover; /* NULL = default (include/pga.h:83) */
	uint64_t seed;
	uint32_t island_base;
	int threads;
	int quiet;
	int topology;
	unsigned delay;
	uint32_t migrate_round;
	unsigned tournament; /* candidates per parent; 2 = TOURNAMENT_POPULATION, src/pga.cu:278 */
	int target_armed;    /* include/pga.h:141, :147: "... or obj_func(best_genome) == value" */
	float target_value;
	unsigned target_every;
	unsigned generations_run;
};

opga_t *opga_init(uint64_t seed) /* src/pga.cu:148-164 */
{
	opga_t *p = calloc(1, sizeof(*p));
	if (!p)
		return NULL;
	p->seed = seed;
	p->threads = 1;
	p->delay = 1;
	p->tournament = 2;
	p->topology = OPGA_TOPO_RANDOM; /* "random migrations", include/pga.h:148 */
	return p;
}

void opga_deinit(opga_t *p) /* src/pga.cu:166-175 */
{
	for (unsigned i = 0; i < p->p_count; ++i) {
		opga_pop_t *q = p->populations[i];
		free(q->cur); free(q->nxt); free(q->score); free(q->sel); free(q);
	}
	free(p);
}

void opga_set_threads(opga_t *p, int n) { p->threads = n < 1 ? 1 : n; }
void opga_set_island_base(opga_t *p, uint32_t b) { p->island_base = b; }
void opga_set_quiet(opga_t *p, int q) { p->quiet = q; }
void opga_set_topology(opga_t *p, int t) { p->topology = t; }
void opga_set_migration_delay(opga_t *p, unsigned d) { p->delay = d; }
void opga_set_objective_function(opga_t *p, opga_obj_f f) { p->objective = f; p->builtin = OPGA_OBJ_HOOK; } /* src/pga.cu:206 */
void opga_set_builtin_objective(opga_t *p, int w) { p->builtin = w; }
void opga_set_mutate_function(opga_t *p, opga_mutate_f f) { p->mutate = f; }          /* src/pga.cu:210 */
void opga_set_crossover_function(opga_t *p, opga_crossover_f f) { p->crossover = f; } /* src/pga.cu:214 */

static uint32_t tail_mask(unsigned len_bits)
{
	unsigned r = len_bits & 31u;
	return r ? ((1u << r) - 1u) : 0xFFFFFFFFu;
}

/* src/pga.cu:177-204 (validation), :107-118 (fill), :81-86 (genes := rand row) */
opga_pop_t *opga_create_population(opga_t *p, unsigned long size, unsigned genome_len, int encoding)
{
	return opga_create_population_on_island(p, size, genome_len, encoding, p->island_base + p->p_count);
}

/* same, with an explicit global island number (multi-process layouts, where the
 * local populations of a rank are islands rank, rank + world, ...) */
opga_pop_t *opga_create_population_on_island(opga_t *p, unsigned long size, unsigned genome_len, int encoding,
					     uint32_t island)
{
	if (p->p_count == OPGA_MAX_POPULATIONS)
		return NULL;
	if (genome_len < 4)
		return NULL;
	opga_pop_t *q = calloc(1, sizeof(*q));
	if (!q)
		return NULL;
	q->size = size;
	q->genome_len = genome_len;
	q->encoding = encoding;
	q->row_words = encoding == OPGA_GENES_BITS ? (genome_len + 31) / 32 : genome_len;
	q->island = island;
	q->epoch = 0;
	size_t words = (size_t)size * q->row_words;
	q->cur = calloc(words ? words : 1, 4);
	q->nxt = calloc(words ? words : 1, 4);
	q->score = calloc(size ? size : 1, sizeof(float));
	q->sel = calloc(size ? 2 * size : 1, sizeof(int32_t));
	p->populations[p->p_count++] = q;

#pragma omp parallel for num_threads(p->threads) schedule(static)
	for (long i = 0; i < (long)size; ++i) {
		uint32_t *row = q->cur + (size_t)i * q->row_words;
		if (encoding == OPGA_GENES_BITS) {
			for (unsigned w = 0; w < q->row_words; ++w)
				row[w] = opga_draw_word(p->seed, q->island, 0, (uint32_t)i, OPGA_STREAM_INITBITS, w);
			row[q->row_words - 1] &= tail_mask(genome_len);
		} else {
			opga_rand_row(p->seed, q->island, 0, (uint32_t)i, genome_len, (float *)row);
		}
	}
	return q;
}

void opga_fill_random_values(opga_t *p, opga_pop_t *q) /* src/pga.cu:99-105 */
{
	(void)p;
	q->epoch += 1;
}

/* src/pga.cu:250-270 */
void opga_evaluate(opga_t *p, opga_pop_t *q)
{
#pragma omp parallel for num_threads(p->threads) schedule(static)
	for (long i = 0; i < (long)q->size; ++i) {
		uint32_t *row = q->cur + (size_t)i * q->row_words;
		if (q->encoding == OPGA_GENES_BITS)
			q->score[i] = opga_builtin_objective_bits(p->builtin, row, q->genome_len);
		else if (p->builtin != OPGA_OBJ_HOOK)
			q->score[i] = opga_builtin_objective_f32(p->builtin, (float *)row, q->genome_len);
		else
			q->score[i] = p->objective((float *)row, q->genome_len);
	}
	q->scores_valid = 1;
}

void opga_evaluate_all(opga_t *p) /* src/pga.cu:272-276 */
{
	for (unsigned i = 0; i < p->p_count; ++i)
		opga_evaluate(p, p->populations[i]);
}

static inline long draw_index(float u, unsigned long n)
{
	long v = (long)(u * (float)n); /* src/pga.cu:284, :129 */
	return v < (long)n ? v : (long)n - 1;
}

/* src/pga.cu:280-292: size-2 tournament, the later candidate wins only if strictly better */
static inline long tournament(const float *score, float ua, float ub, unsigned long n)
{
	long a = draw_index(ua, n), b = draw_index(ub, n);
	return score[a] < score[b] ? b : a;
}

/* src/pga.cu:280-292 with TOURNAMENT_POPULATION = K: candidates from draws first..first+K-1
 * of the individual's rand row; a later candidate wins only if strictly better */
static long tournament_of(const opga_t *p, const opga_pop_t *q, uint32_t individual, unsigned first, unsigned K)
{
	long best = -1;
	for (unsigned i = 0; i < K; ++i) {
		float u = opga_draw_uniform(p->seed, q->island, q->epoch, individual, first + i);
		long c = draw_index(u, q->size);
		if (best == -1 || q->score[best] < q->score[c])
			best = c;
	}
	return best;
}

void opga_set_tournament(opga_t *p, unsigned k) { p->tournament = k ? k : 2; }

/* src/pga.cu:294-325 with the default operator of :135-143 */
void opga_crossover(opga_t *p, opga_pop_t *q)
{
#pragma omp parallel for num_threads(p->threads) schedule(static)
	for (long i = 0; i < (long)q->size; ++i) {
		float u[4];
		opga_rand_row(p->seed, q->island, q->epoch, (uint32_t)i, 4, u);
		long a, b;
		if (p->tournament == 2) {
			a = tournament(q->score, u[0], u[1], q->size); /* src/pga.cu:306 */
			b = tournament(q->score, u[2], u[3], q->size); /* src/pga.cu:307 */
		} else {
			a = tournament_of(p, q, (uint32_t)i, 0, p->tournament);
			b = tournament_of(p, q, (uint32_t)i, p->tournament, p->tournament);
		}
		q->sel[2 * i] = (int32_t)a;
		q->sel[2 * i + 1] = (int32_t)b;
		const uint32_t *p1 = q->cur + (size_t)a * q->row_words;
		const uint32_t *p2 = q->cur + (size_t)b * q->row_words;
		uint32_t *c = q->nxt + (size_t)i * q->row_words;
		if (q->encoding == OPGA_GENES_BITS) {
			uint32_t m[4];
			for (unsigned w = 0; w < q->row_words; ++w) {
				if ((w & 3) == 0) /* one Philox block yields four mask words */
					draw_block_words(p->seed, q->island, q->epoch, (uint32_t)i, OPGA_STREAM_MASK, w >> 2, m);
				c[w] = (p1[w] & m[w & 3]) | (p2[w] & ~m[w & 3]);
			}
		} else if (p->crossover) {
			float *rnd = malloc(sizeof(float) * q->genome_len);
			opga_rand_row(p->seed, q->island, q->epoch, (uint32_t)i, q->genome_len, rnd);
			p->crossover((float *)p1, (float *)p2, (float *)c, rnd, q->genome_len);
			free(rnd);
		} else {
			float rnd[4];
			for (unsigned j = 0; j < q->genome_len; ++j) {
				if ((j & 3) == 0)
					rand_block(p->seed, q->island, q->epoch, (uint32_t)i, j >> 2, rnd);
				c[j] = rnd[j & 3] > 0.5f ? p1[j] : p2[j]; /* src/pga.cu:137-141 */
			}
		}
	}
}

void opga_crossover_all(opga_t *p) /* src/pga.cu:327-331 */
{
	for (unsigned i = 0; i < p->p_count; ++i)
		opga_crossover(p, p->populations[i]);
}

/* src/pga.cu:333-354 with the default operator of :127-133 */
void opga_mutate(opga_t *p, opga_pop_t *q)
{
	const float chance = 0.01f;
#pragma omp parallel for num_threads(p->threads) schedule(static)
	for (long i = 0; i < (long)q->size; ++i) {
		uint32_t *c = q->nxt + (size_t)i * q->row_words;
		if (q->encoding == OPGA_GENES_F32 && p->mutate) {
			float *rnd = malloc(sizeof(float) * q->genome_len);
			opga_rand_row(p->seed, q->island, q->epoch, (uint32_t)i, q->genome_len, rnd);
			p->mutate((float *)c, rnd, q->genome_len);
			free(rnd);
			continue;
		}
		float u[4];
		opga_rand_row(p->seed, q->island, q->epoch, (uint32_t)i, 4, u);
		long idx = draw_index(u[0], q->genome_len);
		if (u[1] <= chance) {
			if (q->encoding == OPGA_GENES_BITS) {
				uint32_t bit = 1u << (idx & 31);
				if (u[2] > 0.5f)
					c[idx >> 5] |= bit;
				else
					c[idx >> 5] &= ~bit;
			} else {
				((float *)c)[idx] = u[2];
			}
		}
	}
}

void opga_mutate_all(opga_t *p) /* src/pga.cu:356-360 */
{
	for (unsigned i = 0; i < p->p_count; ++i)
		opga_mutate(p, p->populations[i]);
}

void opga_swap_generations(opga_t *p, opga_pop_t *q) /* src/pga.cu:362-366 */
{
	(void)p;
	uint32_t *t = q->nxt;
	q->nxt = q->cur;
	q->cur = t;
	q->scores_valid = 0;
}

static void one_generation(opga_t *p, opga_pop_t *q) /* body of src/pga.cu:381-387 */
{
	opga_fill_random_values(p, q);
	if (!q->scores_valid) /* pure objective: re-scoring an unchanged generation is a no-op */
		opga_evaluate(p, q);
	opga_crossover(p, q);
	opga_mutate(p, q);
	opga_swap_generations(p, q);
}

/* include/pga.h:141, :147: run "until n-generations or obj_func(best_genome) == value".  No
 * signature carries the value and the reference has no such code; fixed here: after every
 * generation g that is a multiple of `every` (g < n) the run ends if the best score of any
 * population is >= value.  every = 0 disarms. */
void opga_set_target(opga_t *p, float value, unsigned every)
{
	p->target_armed = every != 0;
	p->target_value = value;
	p->target_every = every;
}

unsigned opga_generations_run(opga_t *p) { return p->generations_run; }

static int target_reached(opga_t *p, unsigned g, unsigned n)
{
	if (!p->target_armed || g % p->target_every != 0 || g >= n)
		return 0;
	for (unsigned s = 0; s < p->p_count; ++s) {
		opga_pop_t *q = p->populations[s];
		if (q->size == 0)
			continue;
		if (!q->scores_valid)
			opga_evaluate(p, q);
		float best;
		opga_best_index(p, q, &best);
		if (best >= p->target_value)
			return 1;
	}
	return 0;
}

void opga_run(opga_t *p, unsigned n) /* src/pga.cu:376-391 */
{
	if (p->p_count == 0)
		return;
	/* pga_run works on populations[0] only (src/pga.cu:382-386); so does its stop test */
	unsigned keep = p->p_count, g;
	p->generations_run = n;
	for (g = 1; g <= n; ++g) {
		one_generation(p, p->populations[0]);
		p->p_count = 1;
		int stop = target_reached(p, g, n);
		p->p_count = keep;
		if (stop) {
			p->generations_run = g;
			break;
		}
	}
	opga_evaluate(p, p->populations[0]);
}

/* src/pga.cu:218-236: first maximum wins; a NaN at index 0 is never displaced */
long opga_best_index(opga_t *p, opga_pop_t *q, float *score_out)
{
	(void)p;
	float best = 0;
	long id = -1;
	for (long i = 0; i < (long)q->size; ++i)
		if (id == -1 || best < q->score[i]) {
			best = q->score[i];
			id = i;
		}
	if (score_out)
		*score_out = best;
	return id;
}

static float *row_as_floats(opga_pop_t *q, long id)
{
	float *out = malloc(sizeof(float) * q->genome_len);
	const uint32_t *row = q->cur + (size_t)id * q->row_words;
	if (q->encoding == OPGA_GENES_BITS)
		for (unsigned j = 0; j < q->genome_len; ++j)
			out[j] = (row[j >> 5] >> (j & 31)) & 1u ? 1.0f : 0.0f;
	else
		memcpy(out, row, sizeof(float) * q->genome_len);
	return out;
}

float *opga_get_best(opga_t *p, opga_pop_t *q)
{
	float best;
	long id = opga_best_index(p, q, &best);
	if (id < 0)
		return NULL;
	if (!p->quiet)
		printf("%f\n", best); /* src/pga.cu:230 */
	return row_as_floats(q, id);
}

/* ---- ranking: one total order for every "top" / "worst" query ------------ */
static inline uint32_t order_key(float s)
{
	if (s != s)
		return 0u; /* NaN below everything */
	s = s + 0.0f;      /* -0 -> +0 */
	uint32_t b;
	memcpy(&b, &s, 4);
	return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}

struct ranked { uint64_t key; };

static int cmp_desc(const void *a, const void *b)
{
	uint64_t x = ((const struct ranked *)a)->key, y = ((const struct ranked *)b)->key;
	return x < y ? 1 : (x > y ? -1 : 0);
}

static int cmp_u32(const void *a, const void *b)
{
	uint32_t x = *(const uint32_t *)a, y = *(const uint32_t *)b;
	return x < y ? -1 : (x > y ? 1 : 0);
}

/* rank key: (order_key(score) << 32) | ~index  -> higher score first, then lower index */
static struct ranked *rank_all(const float *score, unsigned long n)
{
	struct ranked *r = malloc(sizeof(*r) * (n ? n : 1));
	for (unsigned long i = 0; i < n; ++i)
		r[i].key = ((uint64_t)order_key(score[i]) << 32) | (uint32_t)~(uint32_t)i;
	qsort(r, n, sizeof(*r), cmp_desc);
	return r;
}

void opga_select_top(const float *score, unsigned long n, unsigned long k, int worst, uint32_t *out_idx)
{
	if (k > n)
		k = n;
	struct ranked *r = rank_all(score, n);
	for (unsigned long j = 0; j < k; ++j) {
		unsigned long pos = worst ? n - 1 - j : j;
		out_idx[j] = ~(uint32_t)r[pos].key;
	}
	free(r);
	qsort(out_idx, k, sizeof(uint32_t), cmp_u32);
}

/* include/pga.h:91 (stub at src/pga.cu:238-240): `length` best rows, best first */
float **opga_get_best_top(opga_t *p, opga_pop_t *q, unsigned length)
{
	(void)p;
	if (length > q->size)
		length = (unsigned)q->size;
	struct ranked *r = rank_all(q->score, q->size);
	float **out = malloc(sizeof(float *) * (length ? length : 1));
	for (unsigned j = 0; j < length; ++j)
		out[j] = row_as_floats(q, (long)(~(uint32_t)r[j].key));
	free(r);
	return out;
}

/* include/pga.h:92 (stub at src/pga.cu:242-244): best over all populations,
 * ties to the earlier population */
float *opga_get_best_all(opga_t *p)
{
	long best_id = -1;
	unsigned best_pop = 0;
	float best = 0;
	for (unsigned i = 0; i < p->p_count; ++i) {
		float s;
		long id = opga_best_index(p, p->populations[i], &s);
		if (id >= 0 && (best_id == -1 || best < s)) {
			best = s; best_id = id; best_pop = i;
		}
	}
	if (best_id < 0)
		return NULL;
	if (!p->quiet)
		printf("%f\n", best);
	return row_as_floats(p->populations[best_pop], best_id);
}

/* include/pga.h:93 (stub at src/pga.cu:246-248): `length` best rows over all
 * populations, best first; ties to the earlier population, then lower index */
float **opga_get_best_top_all(opga_t *p, unsigned length)
{
	unsigned long total = 0;
	for (unsigned i = 0; i < p->p_count; ++i)
		total += p->populations[i]->size;
	if (length > total)
		length = (unsigned)total;
	/* 96-bit order: score key, then ~population, then ~index */
	struct triple { uint32_t k; uint32_t pop; uint32_t idx; } *all = malloc(sizeof(*all) * (total ? total : 1));
	unsigned long t = 0;
	for (unsigned i = 0; i < p->p_count; ++i)
		for (unsigned long j = 0; j < p->populations[i]->size; ++j) {
			all[t].k = order_key(p->populations[i]->score[j]);
			all[t].pop = i;
			all[t].idx = (uint32_t)j;
			++t;
		}
	float **out = malloc(sizeof(float *) * (length ? length : 1));
	/* selection by repeated scan: length is small in practice */
	char *taken = calloc(total ? total : 1, 1);
	for (unsigned r = 0; r < length; ++r) {
		long bi = -1;
		for (unsigned long j = 0; j < total; ++j)
			if (!taken[j] && (bi < 0 || all[j].k > all[bi].k))
				bi = (long)j;
		taken[bi] = 1;
		out[r] = row_as_floats(p->populations[all[bi].pop], all[bi].idx);
	}
	free(taken);
	free(all);
	return out;
}

/* ---- migration: include/pga.h:108-115, :145-150; stubs at src/pga.cu:368-374, :393-395.
 * Parity unpinned by the reference; fixed here:
 *   k = floor(size * pct / 100) ("top %pct"), emigrants = k best of the source,
 *   victims = k worst of the destination, both in ascending index order,
 *   immigrant j overwrites victim j (row and score). */
unsigned long opga_migration_count(unsigned long size, float pct)
{
	if (!(pct > 0.0f))
		return 0;
	double k = floor((double)size * (double)pct / 100.0);
	if (k > (double)size)
		k = (double)size;
	return (unsigned long)k;
}

static void snapshot_emigrants(opga_pop_t *src, unsigned long k, uint32_t **rows, float **scores)
{
	uint32_t *idx = malloc(sizeof(uint32_t) * (k ? k : 1));
	opga_select_top(src->score, src->size, k, 0, idx);
	*rows = malloc((size_t)(k ? k : 1) * src->row_words * 4);
	*scores = malloc(sizeof(float) * (k ? k : 1));
	for (unsigned long j = 0; j < k; ++j) {
		memcpy(*rows + (size_t)j * src->row_words, src->cur + (size_t)idx[j] * src->row_words, src->row_words * 4);
		(*scores)[j] = src->score[idx[j]];
	}
	free(idx);
}

static void insert_immigrants(opga_pop_t *dst, unsigned long k, const uint32_t *rows, const float *scores)
{
	if (k > dst->size)
		k = dst->size;
	uint32_t *idx = malloc(sizeof(uint32_t) * (k ? k : 1));
	opga_select_top(dst->score, dst->size, k, 1, idx);
	for (unsigned long j = 0; j < k; ++j) {
		memcpy(dst->cur + (size_t)idx[j] * dst->row_words, rows + (size_t)j * dst->row_words, dst->row_words * 4);
		dst->score[idx[j]] = scores[j];
	}
	free(idx);
}

void opga_migrate_between(opga_t *p, opga_pop_t *from, opga_pop_t *to, float pct)
{
	if (from == to || from->row_words != to->row_words)
		return;
	if (!from->scores_valid) opga_evaluate(p, from);
	if (!to->scores_valid) opga_evaluate(p, to);
	unsigned long k = opga_migration_count(from->size, pct);
	if (k > to->size) k = to->size;
	if (!k)
		return;
	uint32_t *rows; float *scores;
	snapshot_emigrants(from, k, &rows, &scores);
	insert_immigrants(to, k, rows, scores);
	free(rows); free(scores);
}

/* Fisher-Yates over stream MIGRATE: word i of (island 0, epoch = round, individual 0) */
void opga_random_permutation(uint64_t seed, uint32_t round, unsigned count, uint32_t *out)
{
	for (unsigned i = 0; i < count; ++i)
		out[i] = i;
	for (unsigned i = count; i-- > 1;) {
		uint32_t w = opga_draw_word(seed, 0, round, 0, OPGA_STREAM_MIGRATE, i);
		unsigned j = w % (i + 1);
		uint32_t t = out[i]; out[i] = out[j]; out[j] = t;
	}
}

static void destinations(opga_t *p, unsigned *dest)
{
	unsigned K = p->p_count;
	if (p->topology == OPGA_TOPO_RANDOM) {
		uint32_t perm[OPGA_MAX_POPULATIONS];
		opga_random_permutation(p->seed, p->migrate_round, K, perm);
		for (unsigned s = 0; s < K; ++s) dest[s] = perm[s];
	} else {
		for (unsigned s = 0; s < K; ++s) dest[s] = (s + 1) % K;
	}
	p->migrate_round += 1;
}

static void start_exchange(opga_t *p, struct pending_migration *pm, float pct)
{
	destinations(p, pm->dest);
	for (unsigned s = 0; s < p->p_count; ++s) {
		opga_pop_t *src = p->populations[s], *dst = p->populations[pm->dest[s]];
		unsigned long k = opga_migration_count(src->size, pct);
		if (k > dst->size) k = dst->size;
		if (pm->dest[s] == s || src->row_words != dst->row_words) k = 0;
		pm->k[s] = k;
		snapshot_emigrants(src, k, &pm->rows[s], &pm->scores[s]);
	}
	pm->active = 1;
}

static void finish_exchange(opga_t *p, struct pending_migration *pm)
{
	for (unsigned s = 0; s < p->p_count; ++s) {
		if (pm->k[s])
			insert_immigrants(p->populations[pm->dest[s]], pm->k[s], pm->rows[s], pm->scores[s]);
		free(pm->rows[s]); free(pm->scores[s]);
	}
	pm->active = 0;
}

/* one synchronous exchange over the configured topology */
void opga_migrate(opga_t *p, float pct)
{
	if (p->p_count < 2)
		return;
	for (unsigned s = 0; s < p->p_count; ++s)
		if (!p->populations[s]->scores_valid)
			opga_evaluate(p, p->populations[s]);
	struct pending_migration pm;
	memset(&pm, 0, sizeof(pm));
	start_exchange(p, &pm, pct);
	finish_exchange(p, &pm);
}

/* n generations on every island; every m generations the k best of each island
 * leave for its neighbour and land `delay` generations later (delay 1 = the
 * transfer overlaps one generation of compute; 0 = synchronous). */
void opga_run_islands(opga_t *p, unsigned n, unsigned m, float pct)
{
	if (p->p_count == 0)
		return;
	struct pending_migration pm;
	memset(&pm, 0, sizeof(pm));
	for (unsigned s = 0; s < p->p_count; ++s)
		if (!p->populations[s]->scores_valid)
			opga_evaluate(p, p->populations[s]);
	for (unsigned g = 1; g <= n; ++g) {
		for (unsigned s = 0; s < p->p_count; ++s) {
			one_generation(p, p->populations[s]);
			opga_evaluate(p, p->populations[s]);
		}
		if (pm.active && pm.land_gen == g)
			finish_exchange(p, &pm);
		if (m && g % m == 0 && g + p->delay <= n && p->p_count > 1 && !pm.active) {
			start_exchange(p, &pm, pct);
			pm.land_gen = g + p->delay;
			if (p->delay == 0)
				finish_exchange(p, &pm);
		}
		if (target_reached(p, g, n)) { /* include/pga.h:147 */
			p->generations_run = g;
			if (pm.active) /* an exchange in flight never lands */
				for (unsigned s = 0; s < p->p_count; ++s) {
					free(pm.rows[s]); free(pm.scores[s]);
				}
			return;
		}
	}
	p->generations_run = n;
}

/* ---- raw access ----------------------------------------------------------- */
unsigned long opga_pop_size(opga_pop_t *q) { return q->size; }
unsigned opga_pop_genome_len(opga_pop_t *q) { return q->genome_len; }
unsigned opga_pop_row_words(opga_pop_t *q) { return q->row_words; }
uint32_t opga_pop_epoch(opga_pop_t *q) { return q->epoch; }
void *opga_pop_current(opga_pop_t *q) { return q->cur; }
void *opga_pop_next(opga_pop_t *q) { return q->nxt; }
float *opga_pop_scores(opga_pop_t *q) { return q->score; }
int32_t *opga_pop_selection(opga_pop_t *q) { return q->sel; }
void opga_pop_set_current(opga_pop_t *q, const void *rows)
{
	memcpy(q->cur, rows, (size_t)q->size * q->row_words * 4);
	q->scores_valid = 0;
}
void opga_pop_set_scores(opga_pop_t *q, const float *s)
{
	memcpy(q->score, s, q->size * sizeof(float));
	q->scores_valid = 1;
}
void opga_pop_set_epoch(opga_pop_t *q, uint32_t e) { q->epoch = e; }

/* ---- reference demo hooks, restated as host C ------------------------------ */
float opga_demo_sum_obj(float *g, unsigned len) /* test/test.cu:24-30 */
{
	float s = 0;
	for (unsigned i = 0; i < len; ++i)
		s += g[i];
	return s;
}

static const float knap_value[6] = { 75, 150, 250, 35, 10, 100 };  /* test2/test.cu:25 */
static const float knap_weight[6] = { 7, 8, 6, 4, 3, 9 };          /* test2/test.cu:26 */

float opga_demo_knapsack_obj(float *g, unsigned len) /* test2/test.cu:28-36 */
{
	float s = 0, w = 0;
	for (unsigned i = 0; i < len; ++i) {
		int count = (int)(g[i] * 2);
		s += knap_value[i] * count;
		w += knap_weight[i] * count;
	}
	return w <= 10.0 ? s : (float)(10.0 - w);
}

#define TSP_MAX 110
static float tsp_matrix[TSP_MAX][TSP_MAX]; /* test3/test.cu:24 */

/* test3/test.cu:70-79: the host fills a [110][110] array row by row, then copies
 * only the first city_count^2 floats of it to the device symbol -- so for 100
 * cities rows 0..89 arrive whole, row 90 up to column 99, rows 91..99 stay zero. */
void opga_demo_tsp_load(const float *matrix, int city_count)
{
	static float host[TSP_MAX][TSP_MAX];
	memset(host, 0, sizeof(host));
	memset(tsp_matrix, 0, sizeof(tsp_matrix));
	for (int i = 0; i < city_count; ++i)
		for (int j = 0; j < city_count; ++j)
			host[i][j] = matrix[i * city_count + j];
	memcpy(tsp_matrix, host, sizeof(float) * (size_t)(city_count * city_count));
}

float opga_demo_tsp_obj(float *g, unsigned len) /* test3/test.cu:26-46 */
{
	float length = 0.f;
	for (unsigned i = 1; i < len; ++i) {
		int a = (int)(g[i - 1] * len), b = (int)(g[i] * len);
		length += tsp_matrix[a][b];
	}
	for (unsigned i = 0; i < len; ++i) {
		int a = (int)(g[i] * len);
		for (unsigned j = 0; j < len; ++j) {
			int b = (int)(g[j] * len);
			if (i != j && a == b)
				length += 10000;
		}
	}
	return -1 * length;
}

void opga_demo_tsp_crossover(float *p1, float *p2, float *c, float *rnd, unsigned len) /* test3/test.cu:48-64 */
{
	int seen[TSP_MAX] = { 0 };
	for (unsigned i = 0; i < len; ++i) {
		int v1 = (int)(p1[i] * len), v2 = (int)(p2[i] * len);
		if (seen[v1] == 0) {
			c[i] = p1[i];
			seen[v1] += 1;
		} else if (seen[v2] == 0) {
			c[i] = p2[i];
			seen[v2] += 1;
		} else {
			c[i] = rnd[i];
		}
	}
}

/* test hooks without a counterpart in the reference's demos (see pga_demo_hooks.cu) */
void opga_demo_even_crossover(float *p1, float *p2, float *c, float *rnd, unsigned len)
{
	for (unsigned k = 0; k < len; k += 2)
		c[k] = rnd[k] > 0.5f ? p1[k] : p2[k];
}

void opga_demo_blend_mutate(float *g, float *rnd, unsigned len)
{
	for (unsigned k = 0; k < len; ++k)
		if (rnd[k] < 0.125f)
			g[k] = 0.5f * (g[k] + rnd[(k + 1) % len]);
}

void *opga_demo_hook(const char *name)
{
	if (!strcmp(name, "even_crossover")) return (void *)opga_demo_even_crossover;
	if (!strcmp(name, "blend_mutate")) return (void *)opga_demo_blend_mutate;
	if (!strcmp(name, "sum_obj")) return (void *)opga_demo_sum_obj;
	if (!strcmp(name, "knapsack_obj")) return (void *)opga_demo_knapsack_obj;
	if (!strcmp(name, "tsp_obj")) return (void *)opga_demo_tsp_obj;
	if (!strcmp(name, "tsp_crossover")) return (void *)opga_demo_tsp_crossover;
	return NULL;
}

uint64_t opga_digest(const void *data, size_t bytes)
{
	const unsigned char *d = data;
	uint64_t h = 0xcbf29ce484222325ull;
	for (size_t i = 0; i < bytes; ++i) {
		h ^= d[i];
		h *= 0x100000001b3ull;
	}
	return h;
}
