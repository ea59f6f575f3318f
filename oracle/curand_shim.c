/*
 * curand_shim.c -- link-time stand-in for the four cuRAND entry points the
 * reference calls, so that the UNMODIFIED reference library can be run with the
 * oracle's random draws and compared with it bit for bit.
 *
 * TEST INFRASTRUCTURE ONLY.  Linked into oracle/_ref/pin/* in place of -lcurand
 * (recipe: oracle/build_ref.sh); never part of libpga.a or libpga_b200.so.
 *
 * The reference's only source of non-determinism is cuRAND
 * (/root/reference/src/pga.cu):
 *   :153  curandCreateGenerator(&randGen, CURAND_RNG_PSEUDO_DEFAULT)
 *   :154  curandSetPseudoRandomGeneratorSeed(randGen, time(NULL))
 *   :100  curandGenerateUniform(randGen, p->rand, p->size*p->genome_len)
 *   :167  curandDestroyGenerator(randGen)
 * Here curandGenerateUniform(gen, ptr, n) fills the device buffer `ptr` with the
 * oracle's rand rows: element [i*L + j] = R(seed, island, epoch, i, j), computed on
 * the host by opga_rand_row (oracle/pga_oracle.c) and copied to the device.
 *   island = order in which distinct buffers were first seen (the reference fills a
 *            population's buffer once inside pga_create_population, :116)
 *   epoch  = number of earlier fills of that buffer (the first fill is epoch 0)
 *   L      = genome length, which the cuRAND API does not carry: the harness states
 *            it (and the seed) with pin_shim_configure() before creating a population;
 *            the time(NULL) seed of :154 is ignored.
 */
#include <cuda_runtime_api.h>
#include <curand.h>
#include <stdio.h>
#include <stdlib.h>

#include "pga_oracle.h"

#define SHIM_MAX_BUFFERS 16

static struct {
	uint64_t seed;
	unsigned genome_len;
	unsigned count;
	struct { float *ptr; unsigned genome_len; uint32_t epoch; } buf[SHIM_MAX_BUFFERS];
	unsigned long long fills;
} shim;

/* harness-side controls (not cuRAND symbols) */
void pin_shim_configure(uint64_t seed, unsigned genome_len)
{
	shim.seed = seed;
	shim.genome_len = genome_len;
}

void pin_shim_reset(void)
{
	shim.count = 0;
	shim.fills = 0;
}

unsigned long long pin_shim_fills(void) { return shim.fills; }

curandStatus_t CURANDAPI curandCreateGenerator(curandGenerator_t *generator, curandRngType_t rng_type)
{
	(void)rng_type;
	*generator = (curandGenerator_t)&shim;
	return CURAND_STATUS_SUCCESS;
}

curandStatus_t CURANDAPI curandSetPseudoRandomGeneratorSeed(curandGenerator_t generator, unsigned long long seed)
{
	(void)generator;
	(void)seed; /* time(NULL) in the reference; the pinned seed comes from pin_shim_configure */
	return CURAND_STATUS_SUCCESS;
}

curandStatus_t CURANDAPI curandDestroyGenerator(curandGenerator_t generator)
{
	(void)generator;
	return CURAND_STATUS_SUCCESS;
}

curandStatus_t CURANDAPI curandGenerateUniform(curandGenerator_t generator, float *outputPtr, size_t num)
{
	(void)generator;
	unsigned slot = 0;
	while (slot < shim.count && shim.buf[slot].ptr != outputPtr)
		++slot;
	if (slot == shim.count) {
		if (shim.count == SHIM_MAX_BUFFERS || shim.genome_len == 0) {
			fprintf(stderr, "curand_shim: not configured / too many buffers\n");
			exit(3);
		}
		shim.buf[slot].ptr = outputPtr;
		shim.buf[slot].genome_len = shim.genome_len;
		shim.buf[slot].epoch = 0;
		shim.count += 1;
	}
	const unsigned L = shim.buf[slot].genome_len;
	if (num % L != 0) {
		fprintf(stderr, "curand_shim: %zu floats is not a whole number of %u-gene rows\n", num, L);
		exit(3);
	}
	const size_t rows = num / L;
	float *host = (float *)malloc(sizeof(float) * (num ? num : 1));
	const uint32_t epoch = shim.buf[slot].epoch;
#pragma omp parallel for schedule(static)
	for (long i = 0; i < (long)rows; ++i)
		opga_rand_row(shim.seed, slot, epoch, (uint32_t)i, L, host + (size_t)i * L);
	cudaError_t e = cudaMemcpy(outputPtr, host, sizeof(float) * num, cudaMemcpyHostToDevice);
	free(host);
	if (e != cudaSuccess) {
		fprintf(stderr, "curand_shim: cudaMemcpy failed: %s\n", cudaGetErrorString(e));
		exit(3);
	}
	shim.buf[slot].epoch += 1;
	shim.fills += 1;
	return CURAND_STATUS_SUCCESS;
}
