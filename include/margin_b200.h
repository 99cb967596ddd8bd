/*
 * margin_b200.h — C ABI of the B200-native pair-HMM hot path of MarginPolish.
 *
 * This is the drop-in boundary.  Everything is plain C: POD
 * structs, pointers and sizes.  No torch / sonLib / margin types appear here.
 * The reference-side binding (how `pairwiseAligner.h`'s stList/SymbolString
 * entry points are routed onto these calls) is shown in INTEGRATION.md and
 * implemented in compat/pairwiseAligner_b200.c.
 *
 * Reference interfaces replaced (all paths relative to the margin tree):
 *   margin_b200_posteriors_batch  <- getAlignedPairsWithIndelsUsingAnchors  inc/pairwiseAligner.h:81-84
 *                                    getAlignedPairsUsingAnchors            inc/pairwiseAligner.h:77-79
 *                                    (the per-read loop of poa_realign,     impl/poa.c:682-713)
 *   margin_b200_forward_batch     <- computeForwardProbability              inc/pairwiseAligner.h:54-57
 *                                    (the allele x read loop,               impl/bubbleGraph.c:1045-1073)
 *   margin_b200_model             <- StateMachine3 + (Rle)NucleotideEmissions + RepeatSubMatrix
 *                                    impl/stateMachine.c:507-519, :716-720; inc/stateMachine.h:77-82
 *   margin_b200_params            <- PairwiseAlignmentParameters            inc/pairwiseAligner.h:28-39
 *   margin_b200_bands             <- band_constructDynamic / band_construct impl/pairwiseAligner.c:120-226
 */
#ifndef MARGIN_B200_H_
#define MARGIN_B200_H_

#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MARGIN_B200_MAX_RL 51          /* MAXIMUM_REPEAT_LENGTH, inc/margin.h:133 */
#define MARGIN_B200_PROB_1 10000000    /* PAIR_ALIGNMENT_PROB_1, inc/pairwiseAligner.h:26 */

/* Transition slots; value = natural log of the transition probability.
 * Names follow StateMachine3 (impl/stateMachine.c:507-519). */
enum {
    MB_T_MATCH_CONTINUE = 0,   /* M -> M */
    MB_T_MATCH_FROM_GAP_X = 1, /* X -> M */
    MB_T_MATCH_FROM_GAP_Y = 2, /* Y -> M */
    MB_T_GAP_OPEN_X = 3,       /* M -> X */
    MB_T_GAP_EXTEND_X = 4,     /* X -> X */
    MB_T_GAP_SWITCH_TO_X = 5,  /* Y -> X */
    MB_T_GAP_OPEN_Y = 6,       /* M -> Y */
    MB_T_GAP_EXTEND_Y = 7,     /* Y -> Y */
    MB_T_GAP_SWITCH_TO_Y = 8   /* X -> Y */
};

/* One flattened state machine (one strand).  `rle` selects the emission
 * function: 0 = NucleotideEmissions (impl/stateMachine.c:363-383), symbols are
 * compared whole (anything >= 4 is ambiguous); 1 = RleNucleotideEmissions
 * (impl/stateMachine.c:722-738): base = symbol & 255, run length = symbol >> 8,
 * match emission = match[xb*4+yb] + 2.3025 * rsm[tbl(xb)][rlX][rlY] where
 * tbl(b) = b for b<4 else 0 and `rsm` is ALREADY resolved for the strand, i.e.
 * rsm[b] holds logProbabilities[(strand ? b : 3-b)] of impl/repeatSubMatrix.c:28-31,
 * laid out [base][underlying = ref run length][observed = read run length]. */
typedef struct margin_b200_model {
    double t[9];
    double match[16];
    double gap_x[4];
    double gap_y[4];
    int32_t rle;
    int32_t max_rl; /* row stride of rsm (51); run lengths must be < max_rl */
    double rsm[4 * MARGIN_B200_MAX_RL * MARGIN_B200_MAX_RL];
} margin_b200_model;

/* PairwiseAlignmentParameters (inc/pairwiseAligner.h:28-39), hot-path fields only. */
typedef struct margin_b200_params {
    double threshold;
    int64_t min_diags_between_traceback;
    int64_t traceback_diagonals;
    int64_t diagonal_expansion;
    int64_t split_matrix_bigger_than_this;
    int32_t dynamic_anchor_expansion;
    int32_t reserved;
} margin_b200_params;

/* A batch of independent alignments (one per read of a chunk, or per allele x
 * read substring).  Sequences are concatenated Symbol (uint16) arrays;
 * alignment i uses sym_x[off_x[i] .. off_x[i+1]) (the reference, "X") and
 * sym_y[off_y[i] .. off_y[i+1]) (the read, "Y").  Anchors are (x, y, expansion)
 * int32 triples, 0-based sequence coordinates, strictly increasing in x and y,
 * expansion even (impl/pairwiseAligner.c:143-159); alignment i owns triples
 * off_a[i] .. off_a[i+1).  model_slot[i] picks the state machine (0 = forward-
 * strand read, 1 = reverse-strand read; impl/poa.c:646-649). */
typedef struct margin_b200_batch {
    int64_t n;
    const uint16_t *sym_x;
    const int64_t *off_x;
    const uint16_t *sym_y;
    const int64_t *off_y;
    const int32_t *anchors;
    const int64_t *off_a;
    const uint8_t *model_slot;
    const uint8_t *ragged_left;
    const uint8_t *ragged_right;
    /* Optional compact form of the anchors (NULL = use `anchors`): one 32-bit word per anchor,
     *   bits 0-11  x - x_prev   (x_prev = -1 before an alignment's first anchor, so the field is >= 1)
     *   bits 12-23 y - y_prev   (likewise)
     *   bits 24-31 expansion / 2
     * with the same off_a.  A host program that walks the reference's stList of stIntTuple anchors anyway
     * (compat/pairwiseAligner_b200.c) emits this instead of triples and moves a third of the bytes over PCIe; the
     * device expands it back to triples.  A batch with a step >= 4096 or an expansion >= 512 must use `anchors`
     * (the margin_b200_pack_anchor helper below returns 0 for such an anchor). */
    const uint32_t *anchors_delta;
    /* Optional still more compact forms (NULL = not used); both exist because the host link, not the GPU, bounds a box
     * of eight GPUs.
     * packed_x / packed_y: the symbols as one byte each, min(base, 4) * 51 + run length (run length 0 without repeat
     * counts), same offsets as sym_x / sym_y, which may then be NULL.  The host checks run lengths < 51 while packing
     * (margin_b200_pack_symbol returns 255 for a symbol that does not fit; such a batch must send sym_x / sym_y).
     * anchors_runs / off_r: the anchors as runs, one 32-bit word per run (alignment i owns words off_r[i] .. off_r[i+1]):
     *   bits 0-7   run: this many FURTHER anchors follow at step (+1, +1) with the same expansion
     *   bits 8-17  x - x_prev of the run's first anchor (x_prev = -1 before an alignment's first anchor), 1..1023
     *   bits 18-27 y - y_prev, likewise
     *   bits 28-31 expansion / 2 (expansion even, 0..30)
     * off_a still counts anchors.  Takes precedence over anchors_delta and anchors; `anchors` should still be given
     * (the rare host-side paths — large-gap splitting, the message for a bad anchor — read the triples). */
    const uint8_t *packed_x;
    const uint8_t *packed_y;
    const uint32_t *anchors_runs;
    const int64_t *off_r;
} margin_b200_batch;

/* The packed byte of a Symbol (inc/stateMachine.h:15: base | run length << 8 with repeat counts, else the base), or 255
 * when it does not fit (run length >= 51). */
static inline uint8_t margin_b200_pack_symbol(uint16_t symbol, int rle) {
    const unsigned b = rle ? (symbol & 255u) : symbol, r = rle ? (unsigned) (symbol >> 8) : 0u;
    if (r >= (unsigned) MARGIN_B200_MAX_RL) return 255;
    return (uint8_t) ((b > 4u ? 4u : b) * MARGIN_B200_MAX_RL + r);
}

/* The first word of a run (run = 0), or 0 when the anchor does not fit (dx, dy in 1..1023, expansion even in 0..30);
 * a following anchor at step (+1, +1) with the same expansion extends the run by adding 1 to the word, up to 255 times. */
static inline uint32_t margin_b200_pack_anchor_run(int64_t dx, int64_t dy, int64_t expansion) {
    if (dx < 1 || dx > 1023 || dy < 1 || dy > 1023 || expansion < 0 || expansion > 30 || (expansion & 1)) return 0;
    return ((uint32_t) dx << 8) | ((uint32_t) dy << 18) | ((uint32_t) (expansion / 2) << 28);
}

/* The compact anchor word, or 0 when the anchor does not fit (dx, dy in 1..4095, expansion even in 0..510). */
static inline uint32_t margin_b200_pack_anchor(int64_t dx, int64_t dy, int64_t expansion) {
    if (dx < 1 || dx > 4095 || dy < 1 || dy > 4095 || expansion < 0 || expansion > 510 || (expansion & 1)) return 0;
    return (uint32_t) dx | ((uint32_t) dy << 12) | ((uint32_t) (expansion / 2) << 24);
}

/* One posterior tuple, same meaning as the reference's stIntTuple (weight, x, y)
 * (impl/pairwiseAligner.c:598-608): weight = floor(p * 1e7), x/y 0-based, gap
 * lists may hold -1. */
typedef struct margin_b200_pair {
    int32_t w, x, y;
} margin_b200_pair;

/* Tuple formats of a result (margin_b200_set_pair_format).  PACKED64 carries the same tuple in one 64-bit word,
 *   bits 0-23 weight, bits 24-43 x + 1, bits 44-63 y + 1,
 * which moves 8 instead of 12 bytes per tuple to the host; it needs every sequence of the batch to be shorter than
 * 2^20 - 1 symbols, otherwise the call falls back to TRIPLES and says so in margin_b200_result.pair_format. */
enum { MARGIN_B200_PAIRS_TRIPLES = 0, MARGIN_B200_PAIRS_PACKED64 = 1, MARGIN_B200_PAIRS_DELTA32 = 2 };

/* DELTA32: one 32-bit record per tuple — bits 0-23 weight, bits 24-31 the step from the previous tuple of the same list
 * of the same alignment: (anti-diagonal step 0..15) << 4 | (x step + 8), i.e. x += (code & 15) - 8 and
 * y += (code >> 4) - ((code & 15) - 8); code 0xFF takes (x, y) from the list's escape array (margin_b200_result.escapes,
 * consumed in order) instead.  The first tuple of a list is always an escape.  Half the bytes of PACKED64 over the host
 * link; no limit on sequence lengths.  The inline decoder below walks a list with a cursor. */
typedef struct margin_b200_xy {
    int32_t x, y;
} margin_b200_xy;

typedef struct margin_b200_delta_cursor {
    const uint32_t *rec;       /* next record */
    const margin_b200_xy *esc; /* next escape */
    int32_t x, y;              /* coordinates of the previous tuple */
} margin_b200_delta_cursor;

static inline margin_b200_pair margin_b200_delta_next(margin_b200_delta_cursor *c) {
    const uint32_t r = *c->rec++;
    const uint32_t code = r >> 24;
    margin_b200_pair p;
    if (code == 0xFFu) {
        c->x = c->esc->x;
        c->y = c->esc->y;
        c->esc++;
    } else {
        const int32_t dx = (int32_t) (code & 15u) - 8;
        c->x += dx;
        c->y += (int32_t) (code >> 4) - dx;
    }
    p.w = (int32_t) (r & 0xffffffu);
    p.x = c->x;
    p.y = c->y;
    return p;
}

static inline margin_b200_pair margin_b200_unpack_pair(uint64_t v) {
    margin_b200_pair p;
    p.w = (int32_t) (v & 0xffffffu);
    p.x = (int32_t) ((v >> 24) & 0xfffffu) - 1;
    p.y = (int32_t) (v >> 44) - 1;
    return p;
}

/* Result of a posterior batch.  Memory is pinned host memory owned by the
 * context and valid until the next call on the same context.  List l of
 * alignment i is pairs[l][off[l][i] .. off[l][i+1]), in exactly the order the
 * reference's stList would have (impl/pairwiseAligner.c:1104-1123).
 * l: 0 = match, 1 = gapX (delete), 2 = gapY (insert); lists 1,2 are empty when
 * want_indels == 0. */
typedef struct margin_b200_result {
    int64_t n;
    const margin_b200_pair *pairs[3];
    const int64_t *off[3];
    int64_t band_cells;      /* lattice cells inside the bands of this batch */
    int64_t diagonals;       /* anti-diagonals swept */
    int64_t jobs;            /* device jobs (alignments after large-gap splitting) */
    int64_t segments;        /* traceback segments */
    float gpu_ms;            /* device time of this call (CUDA events) */
    int32_t kernel_launches; /* kernels launched by this call */
    int64_t tiles;           /* independent forward/backward work units the segments were cut into */
    int32_t pair_format;     /* MARGIN_B200_PAIRS_*: with PACKED64, pairs[l] points at uint64_t words */
    int32_t reserved;
    /* where the band cells went: lane-group tile kernels (the rest), the general tile path for tiles with a diagonal
     * wider than 32 cells, or a sequential sweep of the whole job (degenerate matrices, MARGIN_B200_MODE=sequential,
     * jobs re-run after a failed seam) */
    int64_t wide_cells;
    int64_t sequential_cells;
    /* seam guard: tile edges whose independently computed copies of F / B were compared, and how many of them differed
     * by more than the tolerance (their jobs were re-run sequentially inside the call, so results are unaffected) */
    int32_t guard_seams;
    int32_t guard_failures;
    /* MARGIN_B200_PAIRS_DELTA32 only (NULL otherwise): list l of alignment i consumes escapes[l] from
     * escape_off[l][i] on; pairs[l] then points at uint32_t records */
    const margin_b200_xy *escapes[3];
    const int64_t *escape_off[3];
} margin_b200_result;

typedef struct margin_b200_ctx margin_b200_ctx;

/* All functions return 0 on success, non-zero on error; the message is
 * available from margin_b200_last_error().  There is no CPU fallback: a call
 * without a usable CUDA device fails. */
int margin_b200_device_count(void); /* usable CUDA devices (0 when there is none) */
int margin_b200_create(margin_b200_ctx **ctx, int device);
/* Optional: initialises the CUDA runtime, the primary context of every device and this library's kernel images ahead
 * of the first margin_b200_create (safe from any thread, e.g. a helper thread started while the host program still
 * parses its inputs: `margin polish` reads parameters and BAM records for a second or more before poa_realign
 * (impl/polish.c:476 -> impl/poa.c:1955) needs the aligner).  Returns the device count. */
int margin_b200_prewarm(void);
/* The same for devices 0 .. n-1 only (a primary context costs 0.5-0.7 s per B200 and the driver creates them one after
 * the other: a host program that will drive two of eight GPUs should not wait for eight). */
int margin_b200_prewarm_devices(int n);
void margin_b200_destroy(margin_b200_ctx *ctx);
/* Page-locked host memory for the arrays a batch points at; optional: any host memory works, page-locked memory
 * makes the host->device copies asynchronous and several times faster).  NULL on failure (margin_b200_last_error). */
void *margin_b200_pinned_alloc(size_t bytes);
void margin_b200_pinned_free(void *p);
const char *margin_b200_last_error(void);
int margin_b200_version(void);

int margin_b200_set_model(margin_b200_ctx *ctx, int slot, const margin_b200_model *model);
int margin_b200_set_params(margin_b200_ctx *ctx, const margin_b200_params *params);
/* CUDA events around every kernel launch, read back with margin_b200_last_kernel_ms (default off: measurement
 * only, two driver calls per launch). */
int margin_b200_set_kernel_timing(margin_b200_ctx *ctx, int on);
/* Arithmetic of the lane-group tile kernels (default MARGIN_B200_F64, or the environment's MARGIN_B200_PRECISION).
 *   F64  the reference's own arithmetic (impl/pairwiseAligner.c is double throughout): every add and multiply in the
 *        reference's order; tuple weights equal the reference's up to one unit of 1e-7.
 *   F32  cell values in FP32, renormalised every 32 diagonals, same piece-wise cubic logAdd: posteriors within 1e-4 of
 *        the reference's (BASELINE.json north_star), about twice the throughput.  Tiles wider than 32 cells, sequential
 *        sweeps and margin_b200_forward_batch stay FP64 in both modes. */
enum { MARGIN_B200_F64 = 0, MARGIN_B200_F32 = 1 };
int margin_b200_set_precision(margin_b200_ctx *ctx, int precision);
int margin_b200_get_precision(margin_b200_ctx *ctx);
/* Tuple format of the results of this context (default MARGIN_B200_PAIRS_TRIPLES). */
int margin_b200_set_pair_format(margin_b200_ctx *ctx, int format);

/* Banded forward/backward/posterior for every alignment of the batch. */
int margin_b200_posteriors_batch(margin_b200_ctx *ctx, const margin_b200_batch *batch, int want_indels,
                                 margin_b200_result *out);

/* The same call in three steps, so that the kernels can be timed with inputs
 * already resident in HBM: upload copies the batch host->device;
 * posteriors_resident runs every kernel on the uploaded batch and leaves the
 * result lists in HBM (out->pairs / out->off are NULL, the counters are set);
 * download copies the lists device->host and fills out->pairs / out->off.
 * margin_b200_posteriors_batch == upload + resident + download. */
int margin_b200_upload_batch(margin_b200_ctx *ctx, const margin_b200_batch *batch);
int margin_b200_posteriors_resident(margin_b200_ctx *ctx, int want_indels, margin_b200_result *out);
int margin_b200_download_result(margin_b200_ctx *ctx, margin_b200_result *out);

/* Forward-only log-likelihood per alignment (computeForwardProbability):
 * out_log_prob[i] is the natural-log total probability (double, the reference
 * truncates to float at the call site, impl/bubbleGraph.c:1067). */
int margin_b200_forward_batch(margin_b200_ctx *ctx, const margin_b200_batch *batch, double *out_log_prob,
                              margin_b200_result *stats);
/* The forward-only likelihoods of SMALL matrices with no allocation at all.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight
 * serving loops, harnesses that forbid it).  Same kernels and same results, bit for bit, as margin_b200_forward_batch;
 * the scratch (packed symbols, job table, work lists) is a fixed 16 MiB device array inside the library, held by every
 * GPU that loads it until exit.  Limits per call: at most MARGIN_B200_NOALLOC_MAX_ALIGNMENTS alignments,
 * MARGIN_B200_NOALLOC_MAX_SYMBOLS symbols per side, no anchors, no ragged ends, and every matrix must have a shorter
 * side of at most 31 symbols (the bubble-graph case); anything beyond is rejected with a clear error, nothing is
 * truncated.  All pointers except ctx are DEVICE pointers owned by the caller (offsets as in margin_b200_batch,
 * total_x = off_x[n], total_y = off_y[n]; d_model_slot may be NULL = slot 0; d_out_log_prob holds n doubles).  The call
 * runs on the context's stream and returns when the results are in d_out_log_prob.  It calls no cudaMalloc / cudaFree
 * or any allocating routine; the models must have been set before (margin_b200_set_model allocates, once).  All calls
 * on a device share the scratch array, which keeps the last call's contents (what must start at zero is cleared by
 * every call): calls that overlap on the GPU race, so issue them one at a time per device. */
#define MARGIN_B200_NOALLOC_MAX_ALIGNMENTS 65536
#define MARGIN_B200_NOALLOC_MAX_SYMBOLS (4 << 20)
int margin_b200_forward_tiny_noalloc(margin_b200_ctx *ctx, int64_t n, const uint16_t *d_sym_x, const int64_t *d_off_x,
                                     int64_t total_x, const uint16_t *d_sym_y, const int64_t *d_off_y, int64_t total_y,
                                     const uint8_t *d_model_slot, double *d_out_log_prob);
/* Kernel-only step of the above on a batch staged by margin_b200_upload_batch. */
int margin_b200_forward_resident(margin_b200_ctx *ctx, margin_b200_result *stats);

/* Device time of the last *_resident / *_batch call split by kernel class (CUDA events recorded on the
 * context's stream around every launch): ms[k] = summed duration, launches[k] = number of launches.
 * Classes, in order: band, schedule, forward, backward, scan, gather, offsets, final, guard. */
#define MARGIN_B200_KERNEL_CLASSES 9
int margin_b200_last_kernel_ms(margin_b200_ctx *ctx, float *ms, int32_t *launches, int n_class);

/* Band geometry only (band_constructDynamic / band_construct): for alignment i
 * and anti-diagonal d in [0, lX+lY], xmy_l / xmy_r at off_d[i] + d, where
 * off_d is the exclusive prefix sum of (lX+lY+1).  Buffers are caller-owned
 * host memory. */
int margin_b200_bands(margin_b200_ctx *ctx, const margin_b200_batch *batch, int32_t *xmy_l, int32_t *xmy_r);

#ifdef __cplusplus
}
#endif
#endif
