/*
 * pga_dev.h -- the thin C-ABI between the plain-C host (pga_host.c) and the
 * sm_100a device layer (pga_dev.cu).  Internal; not installed.
 *
 * The host owns handles, validation, hook bookkeeping, the generation and
 * migration schedules.  The device layer owns every CUDA object: buffers,
 * streams, events, kernels, the NCCL communicator.  CUDA/NCCL errors never come
 * back as codes: the device layer prints "GPUassert: ..." and exits, as libpga
 * does (reference: src/pga.cu:25-33).
 */
#ifndef PGA_DEV_H
#define PGA_DEV_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct pgadev_ctx pgadev_ctx; /* per solver: device, streams, events, NCCL */
typedef struct pgadev_pop pgadev_pop; /* per population: device buffers */

/* phases of the generation kernel; the fused generation is all three */
#define PGADEV_PH_CROSSOVER 1u /* select two parents, write child to next_gen */
#define PGADEV_PH_MUTATE 2u    /* mutate the row of next_gen */
#define PGADEV_PH_EVALUATE 4u  /* score the row, reduce best-of-population */
#define PGADEV_PH_FUSED 7u

/* everything one launch needs besides the buffers */
typedef struct pgadev_step {
	uint32_t phases;
	uint64_t seed;
	uint32_t island;
	uint32_t epoch;
	int objective;   /* enum pgax_objective */
	void *obj_hook;  /* device code addresses, NULL = default / built-in */
	void *mutate_hook;
	void *crossover_hook;
	uint32_t reserve_sms; /* leave this many SMs' worth of blocks unlaunched, for work on the side stream */
	uint32_t tournament;  /* candidates per parent in the selection; 2 = the reference's TOURNAMENT */
} pgadev_step;

pgadev_ctx *pgadev_ctx_create(void);
void pgadev_ctx_destroy(pgadev_ctx *);
void pgadev_ctx_set_stream(pgadev_ctx *, void *stream);
void pgadev_ctx_sync(pgadev_ctx *);
unsigned long long pgadev_ctx_launches(pgadev_ctx *);
unsigned long long pgadev_ctx_d2h_bytes(pgadev_ctx *);
/* event pair on the main stream around a region */
void pgadev_timer_start(pgadev_ctx *);
void pgadev_timer_stop(pgadev_ctx *);
float pgadev_timer_ms(pgadev_ctx *); /* syncs on the stop event */

/* $PGA_TRACE=1: named marks on the main stream; the report prints the mean gap per pair of marks */
void pgadev_trace_mark(pgadev_ctx *, const char *label);
void pgadev_trace_report(pgadev_ctx *);

pgadev_pop *pgadev_pop_create(pgadev_ctx *, uint64_t size, uint32_t genome_len, int encoding);
void pgadev_pop_destroy(pgadev_ctx *, pgadev_pop *);
uint32_t pgadev_pop_row_words(pgadev_pop *);
void pgadev_pop_init_random(pgadev_ctx *, pgadev_pop *, uint64_t seed, uint32_t island);
void pgadev_pop_capture_selection(pgadev_ctx *, pgadev_pop *, int enable);

/* Launches one pass on the main stream (asynchronous).
 * EVALUATE alone reads current_gen and writes score[]; CROSSOVER reads current_gen
 * + score[] and writes next_gen; MUTATE rewrites next_gen; FUSED does all three in
 * one kernel and writes the children's scores to the shadow score array. */
int pgadev_launch(pgadev_ctx *, pgadev_pop *, const pgadev_step *);
void pgadev_swap_genes(pgadev_pop *);  /* next_gen <-> current_gen */
void pgadev_swap_scores(pgadev_pop *); /* shadow scores become score[] */

/* first maximum of score[] (reference tie rule), from the in-kernel reduction when
 * `reduced_in_kernel`, else from a reduction pass.  Blocking. */
void pgadev_best(pgadev_ctx *, pgadev_pop *, int reduced_in_kernel, uint64_t *index, float *score);

/* Early-stop check.  begin: the maximum best score over `count` local populations (and over
 * all ranks when all_ranks), queued on the main stream behind the launches so far, 4 bytes
 * to pinned host memory.  reduced_in_kernel[i] says whether population i's best slot is
 * current; if not a reduction pass runs first.  wait: blocks until that copy landed. */
void pgadev_check_begin(pgadev_ctx *, pgadev_pop **pops, const int *reduced_in_kernel, unsigned count, int all_ranks);
float pgadev_check_wait(pgadev_ctx *);

/* blocking copies between host rows (unpadded) and device rows (padded) */
void pgadev_upload_rows(pgadev_ctx *, pgadev_pop *, const void *host, int next_gen);
void pgadev_download_rows(pgadev_ctx *, pgadev_pop *, void *host, int next_gen);
void pgadev_download_row(pgadev_ctx *, pgadev_pop *, uint64_t index, void *host_words);
void pgadev_upload_scores(pgadev_ctx *, pgadev_pop *, const float *host);
void pgadev_download_scores(pgadev_ctx *, pgadev_pop *, float *host);
int pgadev_download_selection(pgadev_ctx *, pgadev_pop *, int32_t *host_pairs);

/* ---- ranking / migration -------------------------------------------------- */
/* k best (worst=0) or k worst (worst=1) of score[], ascending index order, into the
 * population's device index list `slot` (0 = emigrants, 1 = victims).  Async. */
void pgadev_select_top(pgadev_ctx *, pgadev_pop *, uint64_t k, int worst, int slot);
void pgadev_download_top(pgadev_ctx *, pgadev_pop *, int slot, uint64_t k, uint32_t *host_idx);
/* the scores of those k individuals, gathered on the device (k floats cross the bus) */
void pgadev_download_top_scores(pgadev_ctx *, pgadev_pop *, int slot, uint64_t k, float *host_scores);
/* size the outbox/inbox for k rows (allocates; call outside NCCL groups) */
void pgadev_reserve_boxes(pgadev_ctx *, pgadev_pop *, uint64_t k);
/* gather the k emigrants (rows + scores) listed in slot 0 into the outbox */
void pgadev_pack_emigrants(pgadev_ctx *, pgadev_pop *, uint64_t k);
/* outbox(src) -> inbox(dst) on the same device, on the side stream */
void pgadev_transfer_local(pgadev_ctx *, pgadev_pop *src, pgadev_pop *dst, uint64_t k);
/* overwrite the k victims listed in slot 1 with the inbox (rows + scores) */
void pgadev_unpack_immigrants(pgadev_ctx *, pgadev_pop *, uint64_t k);
/* route the following select_top / pack_emigrants calls to the side stream (1) or
 * back to the main stream (0) */
void pgadev_route_to_side(pgadev_ctx *, int on);
/* record "emigrants packed" on the side stream / make the main stream wait for it */
void pgadev_mark_packed(pgadev_ctx *);
void pgadev_main_wait_packed(pgadev_ctx *);
/* side stream waits for the main stream / main stream waits for the side stream */
void pgadev_side_after_main(pgadev_ctx *);
void pgadev_main_after_side(pgadev_ctx *);

/* ---- NCCL (dlopen'ed) ------------------------------------------------------ */
int pgadev_nccl_unique_id(void *out128);
int pgadev_nccl_init(pgadev_ctx *, int rank, int world, const void *id128);
void pgadev_nccl_group_begin(pgadev_ctx *);
void pgadev_nccl_group_end(pgadev_ctx *);
/* on the side stream, inside a group */
void pgadev_nccl_send_outbox(pgadev_ctx *, pgadev_pop *src, uint64_t k, int peer);
void pgadev_nccl_recv_inbox(pgadev_ctx *, pgadev_pop *dst, uint64_t k, int peer);

#ifdef __cplusplus
}
#endif
#endif
