/*
 * onemax.c -- the engine driven from a plain C program: no CUDA in the caller.
 *
 *   gcc -O2 -Iinclude examples/onemax.c -o onemax -Llibpga_b200 -lpga_b200 -Wl,-rpath,$PWD/libpga_b200
 *   ./onemax [population] [bits] [generations] [seed]
 *
 * Uses libpga's own entry points (include/pga.h; reference: include/pga.h:53-143) plus the
 * extension calls that select packed-bit genomes and the built-in OneMax objective.
 * Prints the best score the way pga_get_best does ("%f\n", reference: src/pga.cu:230)
 * and a summary line.
 */
#include <stdio.h>
#include <stdlib.h>

#include "pga.h"
#include "pga_b200.h"

int main(int argc, char **argv)
{
	unsigned long size = argc > 1 ? strtoul(argv[1], NULL, 0) : 1ul << 16;
	unsigned bits = argc > 2 ? (unsigned)strtoul(argv[2], NULL, 0) : 1024;
	unsigned generations = argc > 3 ? (unsigned)strtoul(argv[3], NULL, 0) : 50;
	unsigned long long seed = argc > 4 ? strtoull(argv[4], NULL, 0) : 42;

	pga_t *p = pga_init();
	if (!p)
		return 1;
	pgax_set_seed(p, seed);
	population_t *pop = pgax_create_population(p, size, bits, PGAX_GENES_BITS);
	if (!pop)
		return 2;
	pgax_set_builtin_objective(p, PGAX_OBJ_ONEMAX);

	pga_run(p, generations);

	gene *best = pga_get_best(p, pop); /* prints the score; one float (0/1) per bit */
	unsigned ones = 0;
	for (unsigned i = 0; i < bits; ++i)
		ones += best[i] != 0.0f;
	unsigned long index;
	float score;
	pgax_best(p, pop, &index, &score);
	printf("best individual %lu: %u of %u bits set, score %.1f, %.3f ms on the device, %llu kernel launches\n", index, ones,
	       bits, score, pgax_last_run_ms(p), pgax_kernel_launches(p));
	free(best);
	pga_deinit(p);
	return ones == (unsigned)score ? 0 : 3;
}
