/*
 * pairwiseAligner_b200.c — the reference-side binding of libmargin_b200.so.
 *
 * This file is compiled AGAINST margin's own headers (inc/margin.h, sonLib) and linked INTO margin in place of the
 * bodies of seven functions; everything else of impl/pairwiseAligner.c and impl/poa.c stays the reference's code:
 *
 *   getAlignedPairsWithIndelsUsingAnchors   inc/pairwiseAligner.h:81-84   (impl/pairwiseAligner.c:1144-1171)
 *   getAlignedPairsUsingAnchors             inc/pairwiseAligner.h:77-79   (impl/pairwiseAligner.c:1125-1142)
 *   getAlignedPairsWithIndels               inc/pairwiseAligner.h:68-72   (impl/pairwiseAligner.c:1183-1191)
 *   getAlignedPairs                         inc/pairwiseAligner.h:62-66   (impl/pairwiseAligner.c:1173-1181)
 *   poa_realign                             inc/margin.h:684              (impl/poa.c:668-716), one batched call per chunk
 *   computeForwardProbability               inc/pairwiseAligner.h:54-57   (impl/pairwiseAligner.c:849-905)
 *   poa_polish                              inc/margin.h:820              (impl/poa.c:1852-1873), one batched forward call
 *                                                                          per bubble graph (impl/bubbleGraph.c:1067 unchanged)
 *   poa_estimateRepeatCountsUsingBayesianModel  inc/margin.h              (impl/poa.c:1715-1727), flat arrays, same sums
 *
 * Same signatures, same ownership (callee-allocated stLists of individually allocated stIntTuples with
 * stIntTuple_destruct as destructor), same list order, same error behaviour (invalid band ->
 * PAIRWISE_ALIGNMENT_EXCEPTION_ID exception; anything else st_errAbort).  Thread model as in the reference: called
 * concurrently from OpenMP threads (polish.c:476); every thread owns one library context on device
 * (thread number % device count).  All computation happens in libmargin_b200.so on the GPU; there is no CPU fallback.
 *
 * See INTEGRATION.md for how it is linked (oracle/Makefile builds oracle/_ref/margin_b200cli this way for the
 * end-to-end parity test).
 */
#include <omp.h>
#include <pthread.h>

#include "margin.h"
#include "margin_b200.h"

/* Private layouts of impl/stateMachine.c (:507-519, :716-720), re-declared to flatten a StateMachine
 * (the reference offers no accessor). */
typedef struct {
    StateMachine model;
    double TRANSITION_MATCH_CONTINUE;
    double TRANSITION_MATCH_FROM_GAP_X;
    double TRANSITION_MATCH_FROM_GAP_Y;
    double TRANSITION_GAP_OPEN_X;
    double TRANSITION_GAP_OPEN_Y;
    double TRANSITION_GAP_EXTEND_X;
    double TRANSITION_GAP_EXTEND_Y;
    double TRANSITION_GAP_SWITCH_TO_X;
    double TRANSITION_GAP_SWITCH_TO_Y;
} B200StateMachine3;

typedef struct {
    NucleotideEmissions ne;
    RepeatSubMatrix *repeatSubMatrix;
    bool strand;
} B200RleNucleotideEmissions;

struct B200Deferred;

typedef struct {
    margin_b200_ctx *ctx;
    StateMachine *sm[2];                 /* machines currently loaded in slots 0 / 1 */
    PairwiseAlignmentParameters params;  /* parameters currently loaded */
    int haveParams;
    margin_b200_model scratch;
    margin_b200_model loaded[2];   /* copies of what slots 0 / 1 hold (only read by the MARGIN_B200_DUMP hook) */
    margin_b200_params loadedParams;
    int dynamicAnchors;            /* whether b200_batch_add reads the anchors' third field (p->dynamicAnchorExpansion) */
    struct B200Deferred *deferred; /* non-NULL while poa_polish collects forward jobs (see the last section) */
    struct B200Deferred *replay;   /* non-NULL while a record / replay wrapper hands the results back in order */
} B200Thread;

static __thread B200Thread tls;

static void b200_fail(void) {
    const char *msg = margin_b200_last_error();
    if (strstr(msg, "PAIRWISE_ALIGNMENT_EXCEPTION") != NULL) {
        stThrowNew(PAIRWISE_ALIGNMENT_EXCEPTION_ID, "%s", msg);
    }
    st_errAbort("margin_b200: %s\n", msg);
}

/*
 * CUDA start-up (runtime, primary context, kernel images) costs about a second; `margin polish` spends at least that
 * long parsing parameters and BAM records before the first poa_realign, so a helper thread started at load time pays
 * it beside them.  MARGIN_B200_PREWARM=0 turns it off.  The runtime serialises its own initialisation: a worker that
 * reaches margin_b200_create first simply waits for the helper.
 */
static pthread_t b200_prewarmThread;
static int b200_prewarmStarted = 0;

static void *b200_prewarm_main(void *arg) {
    (void) arg;
    /* device 0 only: how many devices the run will drive is known when the OpenMP team exists (b200_device_of_thread);
     * the first thread mapped to another device creates that context itself, beside the others' work */
    margin_b200_prewarm_devices(1);
    return NULL;
}

/*
 * OpenMP thread t -> device.  The aligner is a few per cent of a chunk's wall, so one B200 keeps
 * well over 16 host threads fed, while every further device costs its primary context (0.5-0.7 s, created one after the
 * other by the driver), its own scratch pool and its own cold block caches: `margin polish -t 32` on the 2 Mb contig
 * took 14 s with one device, 14.7 s with four and 26 s with eight when every visible device was used.  So a device is
 * added per MARGIN_B200_THREADS_PER_GPU threads of the team (default 16), up to the visible devices.
 */
static int b200_device_of_thread(int devices) {
    static int perGpu = 0;
    if (perGpu == 0) {
        const char *e = getenv("MARGIN_B200_THREADS_PER_GPU");
        perGpu = e != NULL && atoi(e) > 0 ? atoi(e) : 16;
    }
    int team = omp_get_num_threads(), used = (team + perGpu - 1) / perGpu;
    if (used > devices) used = devices;
    if (used < 1) used = 1;
    return omp_get_thread_num() % used;
}

/*
 * A CUDA process pays for every device it can SEE when the driver initialises (measured: context set-up of the first
 * chunks 0.5 s with one visible B200, 2.4 s with four, about 5 s with eight), whether or not it ever uses them.  The
 * thread count of a margin run is on its command line (-t / --threads, polish.c:150, phase.c), so before CUDA starts the
 * visible devices are cut down to the ones b200_device_of_thread will use: the first ceil(threads / 16) of
 * CUDA_VISIBLE_DEVICES (or of 0, 1, 2, ... when it is unset).  MARGIN_B200_KEEP_VISIBLE_DEVICES=1 leaves the list alone.
 */
static void b200_limit_visible_devices(void) {
    const char *keep = getenv("MARGIN_B200_KEEP_VISIBLE_DEVICES");
    if (keep != NULL && keep[0] != '0') return;
    FILE *fh = fopen("/proc/self/cmdline", "rb");
    if (fh == NULL) return;
    static char buf[1 << 16];
    size_t len = fread(buf, 1, sizeof(buf) - 2, fh);
    fclose(fh);
    buf[len] = buf[len + 1] = '\0';
    long threads = 1; /* margin's default */
    for (char *a = buf; a < buf + len; a += strlen(a) + 1) {
        const char *value = NULL;
        if (strcmp(a, "-t") == 0 || strcmp(a, "--threads") == 0) value = a + strlen(a) + 1;
        else if (strncmp(a, "--threads=", 10) == 0) value = a + 10;
        else if (a[0] == '-' && a[1] == 't' && a[2] >= '0' && a[2] <= '9') value = a + 2;
        if (value != NULL && value < buf + len && atol(value) > 0) threads = atol(value);
    }
    const char *per = getenv("MARGIN_B200_THREADS_PER_GPU");
    long perGpu = per != NULL && atol(per) > 0 ? atol(per) : 16, wanted = (threads + perGpu - 1) / perGpu;
    if (wanted < 1) wanted = 1;
    if (wanted > 64) return;
    const char *visible = getenv("CUDA_VISIBLE_DEVICES");
    char list[1024];
    size_t at = 0;
    if (visible == NULL) {
        for (long d = 0; d < wanted && at + 8 < sizeof(list); d++) at += (size_t) snprintf(list + at, sizeof(list) - at, d ? ",%ld" : "%ld", d);
    } else {
        long seen = 0;
        for (const char *c = visible; *c != '\0' && at + 1 < sizeof(list); c++) {
            if (*c == ',' && ++seen >= wanted) break;
            list[at++] = *c;
        }
    }
    list[at] = '\0';
    setenv("CUDA_VISIBLE_DEVICES", list, 1);
}

__attribute__((constructor)) static void b200_prewarm_start(void) {
    b200_limit_visible_devices();
    const char *e = getenv("MARGIN_B200_PREWARM");
    if (e != NULL && e[0] == '0') return;
    if (pthread_create(&b200_prewarmThread, NULL, b200_prewarm_main, NULL) == 0) b200_prewarmStarted = 1;
}

__attribute__((destructor)) static void b200_prewarm_finish(void) {
    if (b200_prewarmStarted) pthread_join(b200_prewarmThread, NULL);
    b200_prewarmStarted = 0;
}

static margin_b200_ctx *b200_ctx(void) {
    if (tls.ctx == NULL) {
        int n = margin_b200_device_count();
        if (n <= 0) st_errAbort("margin_b200: no CUDA device; this build of margin has no CPU aligner\n");
        if (margin_b200_create(&tls.ctx, b200_device_of_thread(n)) != 0) b200_fail();
        /* 4-byte step records over PCIe (half of the 8-byte words, a third of the triples: the device->host link is what
         * bounds a box of eight GPUs); b200_list() / b200_pairs() decode them while they walk the lists anyway */
        if (margin_b200_set_pair_format(tls.ctx, MARGIN_B200_PAIRS_DELTA32) != 0) b200_fail();
    }
    return tls.ctx;
}

/* RLE emissions are recognised by probing: only getRleNucleotideMatchProb looks at the run-length byte. */
static int b200_is_rle(StateMachine *sM) {
    Emissions *e = sM->emissions;
    return e->emission(e, (Symbol) ((1 << 8) | 0), (Symbol) ((1 << 8) | 0)) !=
           e->emission(e, (Symbol) ((2 << 8) | 0), (Symbol) ((1 << 8) | 0));
}

/*
 * Layout self-check.  The two structs above are private to impl/stateMachine.c; if the reference
 * ever changes them the flattened model would silently be wrong.  So every freshly flattened model is compared with
 * what the reference's own function pointers say: stateMachine3_cellCalculate (impl/stateMachine.c:562-586) is run
 * with a recording doTransition callback, which yields the nine transitions and the three emissions it pairs them
 * with, and emissions->emission / gapEmissionX / gapEmissionY are sampled over all base pairs and a spread of run
 * lengths.  Values must agree bit for bit (the device forms match + 2.3025 * rsm with the same two operations).
 */
typedef struct {
    double eP[3][3], tP[3][3];
    int seen[3][3];
} B200Probe;

static void b200_probe_transition(double *fromCells, double *toCells, int64_t from, int64_t to, double eP, double tP,
                                  void *extraArgs) {
    (void) fromCells;
    (void) toCells;
    B200Probe *pr = extraArgs;
    if (from < 0 || from > 2 || to < 0 || to > 2) st_errAbort("margin_b200: state machine with unexpected state numbers\n");
    pr->eP[from][to] = eP;
    pr->tP[from][to] = tP;
    pr->seen[from][to] = 1;
}

static double b200_flat_emission(const margin_b200_model *m, Symbol cX, Symbol cY) {
    unsigned xb = m->rle ? (cX & 255u) : cX, yb = m->rle ? (cY & 255u) : cY;
    double sub = (xb >= 4 || yb >= 4) ? -2.772588722 : m->match[xb * 4 + yb]; /* impl/stateMachine.c:378-383 */
    if (!m->rle) return sub;
    unsigned tb = xb >= 4 ? 0 : xb;
    return sub + 2.3025 * m->rsm[((size_t) tb * MARGIN_B200_MAX_RL + (cX >> 8)) * MARGIN_B200_MAX_RL + (cY >> 8)];
}

static int b200_same_bits(double a, double b) { return memcmp(&a, &b, sizeof(double)) == 0; }

static void b200_check_flattening(StateMachine *sM, const margin_b200_model *m) {
    if (sM->matchState != 0 || sM->gapXState != 1 || sM->gapYState != 2 || sM->stateNumber != 3)
        st_errAbort("margin_b200: state machine is not the 3-state (match, gapX, gapY) machine\n");
    B200Probe pr;
    memset(&pr, 0, sizeof(pr));
    double cells[4][3];
    memset(cells, 0, sizeof(cells));
    Symbol cX = (Symbol) ((m->rle ? 3 << 8 : 0) | 1), cY = (Symbol) ((m->rle ? 2 << 8 : 0) | 2);
    sM->cellCalculate(sM, cells[0], cells[1], cells[2], cells[3], cX, cY, b200_probe_transition, &pr);
    /* [from][to] in (match, gapX, gapY) order against the MB_T_* slots */
    static const int slotOf[3][3] = {{MB_T_MATCH_CONTINUE, MB_T_GAP_OPEN_X, MB_T_GAP_OPEN_Y},
                                     {MB_T_MATCH_FROM_GAP_X, MB_T_GAP_EXTEND_X, MB_T_GAP_SWITCH_TO_Y},
                                     {MB_T_MATCH_FROM_GAP_Y, MB_T_GAP_SWITCH_TO_X, MB_T_GAP_EXTEND_Y}};
    Emissions *e = sM->emissions;
    double want[3] = {e->emission(e, cX, cY), e->gapEmissionX(e, cX), e->gapEmissionY(e, cY)};
    for (int from = 0; from < 3; from++) {
        for (int to = 0; to < 3; to++) {
            if (!pr.seen[from][to] || !b200_same_bits(pr.tP[from][to], m->t[slotOf[from][to]]) ||
                !b200_same_bits(pr.eP[from][to], want[to]))
                st_errAbort("margin_b200: StateMachine3 layout check failed for transition %d -> %d "
                            "(private structs of impl/stateMachine.c changed?)\n", from, to);
        }
    }
    static const int runLengths[6] = {0, 1, 2, 7, 23, 50};
    for (int xb = 0; xb < 5; xb++) {
        if (xb < 4 && (!b200_same_bits(e->gapEmissionX(e, (Symbol) xb), m->gap_x[xb]) ||
                       !b200_same_bits(e->gapEmissionY(e, (Symbol) xb), m->gap_y[xb])))
            st_errAbort("margin_b200: emission layout check failed for the gap emissions of base %d\n", xb);
        for (int yb = 0; yb < 5; yb++) {
            for (int i = 0; i < (m->rle ? 6 : 1); i++) {
                for (int j = 0; j < (m->rle ? 6 : 1); j++) {
                    Symbol sx = (Symbol) ((m->rle ? runLengths[i] << 8 : 0) | xb), sy = (Symbol) ((m->rle ? runLengths[j] << 8 : 0) | yb);
                    if (!b200_same_bits(e->emission(e, sx, sy), b200_flat_emission(m, sx, sy)))
                        st_errAbort("margin_b200: emission layout check failed for symbols %d / %d "
                                    "(private structs of impl/stateMachine.c changed?)\n", (int) sx, (int) sy);
                }
            }
        }
    }
}

static void b200_load_machine(int slot, StateMachine *sM) {
    if (sM->type != threeState && sM->type != threeStateAsymmetric) st_errAbort("margin_b200: unsupported state machine\n");
    B200StateMachine3 *s = (B200StateMachine3 *) sM;
    margin_b200_model *m = &tls.scratch;
    memset(m, 0, sizeof(*m));
    m->t[MB_T_MATCH_CONTINUE] = s->TRANSITION_MATCH_CONTINUE;
    m->t[MB_T_MATCH_FROM_GAP_X] = s->TRANSITION_MATCH_FROM_GAP_X;
    m->t[MB_T_MATCH_FROM_GAP_Y] = s->TRANSITION_MATCH_FROM_GAP_Y;
    m->t[MB_T_GAP_OPEN_X] = s->TRANSITION_GAP_OPEN_X;
    m->t[MB_T_GAP_EXTEND_X] = s->TRANSITION_GAP_EXTEND_X;
    m->t[MB_T_GAP_SWITCH_TO_X] = s->TRANSITION_GAP_SWITCH_TO_X;
    m->t[MB_T_GAP_OPEN_Y] = s->TRANSITION_GAP_OPEN_Y;
    m->t[MB_T_GAP_EXTEND_Y] = s->TRANSITION_GAP_EXTEND_Y;
    m->t[MB_T_GAP_SWITCH_TO_Y] = s->TRANSITION_GAP_SWITCH_TO_Y;
    NucleotideEmissions *ne = (NucleotideEmissions *) sM->emissions;
    memcpy(m->match, ne->EMISSION_MATCH_PROBS, sizeof(m->match));
    memcpy(m->gap_x, ne->EMISSION_GAP_X_PROBS, sizeof(m->gap_x));
    memcpy(m->gap_y, ne->EMISSION_GAP_Y_PROBS, sizeof(m->gap_y));
    m->rle = b200_is_rle(sM);
    m->max_rl = MARGIN_B200_MAX_RL;
    if (m->rle) {
        B200RleNucleotideEmissions *re = (B200RleNucleotideEmissions *) sM->emissions;
        RepeatSubMatrix *r = re->repeatSubMatrix;
        if (r->maximumRepeatLength != MARGIN_B200_MAX_RL) st_errAbort("margin_b200: repeat matrix must be 51 x 51\n");
        int64_t sq = r->maximumRepeatLength * r->maximumRepeatLength;
        for (int64_t b = 0; b < 4; b++) { /* repeatSubMatrix_setLogProb index, impl/repeatSubMatrix.c:28-31 */
            memcpy(m->rsm + b * sq, r->logProbabilities + (re->strand ? b : 3 - b) * sq, sizeof(double) * sq);
        }
    }
    /* The slot is keyed by CONTENT, not by the StateMachine pointer: a machine freed and rebuilt at the same address
     * (the reference's test suites do that in loops) or modified in place must not keep a stale model on the device. */
    if (tls.sm[slot] != NULL && memcmp(m, &tls.loaded[slot], sizeof(*m)) == 0) {
        tls.sm[slot] = sM;
        return;
    }
    b200_check_flattening(sM, m);
    if (margin_b200_set_model(b200_ctx(), slot, m) != 0) b200_fail();
    tls.sm[slot] = sM;
    tls.loaded[slot] = *m;
}

static void b200_load_params(PairwiseAlignmentParameters *p) {
    tls.dynamicAnchors = p->dynamicAnchorExpansion;
    if (tls.haveParams && tls.params.threshold == p->threshold &&
        tls.params.minDiagsBetweenTraceBack == p->minDiagsBetweenTraceBack &&
        tls.params.traceBackDiagonals == p->traceBackDiagonals && tls.params.diagonalExpansion == p->diagonalExpansion &&
        tls.params.splitMatrixBiggerThanThis == p->splitMatrixBiggerThanThis &&
        tls.params.dynamicAnchorExpansion == p->dynamicAnchorExpansion)
        return;
    margin_b200_params q;
    memset(&q, 0, sizeof(q));
    q.threshold = p->threshold;
    q.min_diags_between_traceback = p->minDiagsBetweenTraceBack;
    q.traceback_diagonals = p->traceBackDiagonals;
    q.diagonal_expansion = p->diagonalExpansion;
    q.split_matrix_bigger_than_this = p->splitMatrixBiggerThanThis;
    q.dynamic_anchor_expansion = p->dynamicAnchorExpansion;
    if (margin_b200_set_params(b200_ctx(), &q) != 0) b200_fail();
    tls.loadedParams = q;
    tls.params = *p;
    tls.haveParams = 1;
}

/* ------------------------------------------------------------------------------------------------ batches */

typedef struct {
    int64_t n, capN, capX, capY, capA;
    uint16_t *symX, *symY;
    int32_t *anchors;
    uint32_t *anchorsDelta; /* the compact words of margin_b200_batch.anchors_delta, kept beside the triples */
    int compactOk;          /* every anchor so far fits a compact word */
    uint8_t *packedX, *packedY; /* one byte per symbol (margin_b200_pack_symbol): what travels when packedOk */
    int packedOk;
    uint32_t *runs;         /* the anchors as runs (margin_b200_batch.anchors_runs): what travels when runsOk */
    int64_t nRuns;
    int runsOk;
    int64_t *offX, *offY, *offA, *offR;
    uint8_t *slot, *raggedL, *raggedR;
} B200Batch;

/* One retired set of bulk arrays per thread, so that a batch (of one, often) does not page-lock memory on every call. */
static __thread struct {
    uint16_t *symX, *symY;
    int32_t *anchors;
    uint32_t *anchorsDelta, *runs;
    uint8_t *packedX, *packedY;
    int64_t capX, capY, capA;
    int full;
} b200_spare;

static B200Batch *b200_batch_construct(int64_t n) {
    B200Batch *b = st_calloc(1, sizeof(B200Batch));
    b->compactOk = b->packedOk = b->runsOk = 1;
    if (b200_spare.full) {
        b->packedX = b200_spare.packedX;
        b->packedY = b200_spare.packedY;
        b->runs = b200_spare.runs;
        b->symX = b200_spare.symX;
        b->symY = b200_spare.symY;
        b->anchors = b200_spare.anchors;
        b->anchorsDelta = b200_spare.anchorsDelta;
        b->capX = b200_spare.capX;
        b->capY = b200_spare.capY;
        b->capA = b200_spare.capA;
        b200_spare.full = 0;
    }
    b->capN = n < 1 ? 1 : n;
    b->offX = st_calloc(b->capN + 1, sizeof(int64_t));
    b->offY = st_calloc(b->capN + 1, sizeof(int64_t));
    b->offA = st_calloc(b->capN + 1, sizeof(int64_t));
    b->offR = st_calloc(b->capN + 1, sizeof(int64_t));
    b->slot = st_calloc(b->capN + 1, 1);
    b->raggedL = st_calloc(b->capN + 1, 1);
    b->raggedR = st_calloc(b->capN + 1, 1);
    return b;
}

/* Room for one more alignment (the deferred forward queue does not know its size in advance). */
static void b200_batch_grow(B200Batch *b) {
    if (b->n < b->capN) return;
    b->capN = b->capN * 2 + 1024;
    b->offX = st_realloc(b->offX, (b->capN + 1) * sizeof(int64_t));
    b->offY = st_realloc(b->offY, (b->capN + 1) * sizeof(int64_t));
    b->offA = st_realloc(b->offA, (b->capN + 1) * sizeof(int64_t));
    b->offR = st_realloc(b->offR, (b->capN + 1) * sizeof(int64_t));
    b->slot = st_realloc(b->slot, b->capN + 1);
    b->raggedL = st_realloc(b->raggedL, b->capN + 1);
    b->raggedR = st_realloc(b->raggedR, b->capN + 1);
}

/* The arrays that cross PCIe (symbols, compact anchor words) live in page-locked memory (margin_b200_pinned_alloc):
 * the copies to the device are then asynchronous DMA instead of the driver's staged pageable path. */
static void *b200_pinned_grow(void *old, int64_t oldBytes, int64_t newBytes) {
    void *p = margin_b200_pinned_alloc((size_t) newBytes);
    if (p == NULL) b200_fail();
    if (old != NULL) {
        memcpy(p, old, (size_t) oldBytes);
        margin_b200_pinned_free(old);
    }
    return p;
}

static void b200_batch_destruct(B200Batch *b) {
    if (!b200_spare.full) {
        b200_spare.symX = b->symX;
        b200_spare.symY = b->symY;
        b200_spare.anchors = b->anchors;
        b200_spare.anchorsDelta = b->anchorsDelta;
        b200_spare.runs = b->runs;
        b200_spare.packedX = b->packedX;
        b200_spare.packedY = b->packedY;
        b200_spare.capX = b->capX;
        b200_spare.capY = b->capY;
        b200_spare.capA = b->capA;
        b200_spare.full = 1;
    } else {
        free(b->symX);
        free(b->symY);
        free(b->anchors);
        free(b->anchorsDelta);
        margin_b200_pinned_free(b->runs);
        margin_b200_pinned_free(b->packedX);
        margin_b200_pinned_free(b->packedY);
    }
    free(b->offX);
    free(b->offY);
    free(b->offA);
    free(b->offR);
    free(b->slot);
    free(b->raggedL);
    free(b->raggedR);
    free(b);
}

/* Appends one alignment; anchor x coordinates are shifted by -shiftX (the crop of impl/poa.c:637-639). */
static void **b200_list_array(stList *list); /* defined with the augment code below */

/* Room for nX / nY symbols and nA anchors in all (the filled parts are kept).  poa_realign knows the totals of a chunk
 * before it adds the first read, so that the page-locked arrays are taken once instead of doubling their way up. */
static void b200_batch_reserve(B200Batch *b, int64_t nX, int64_t nY, int64_t nA) {
    int64_t oX = b->offX[b->n], oY = b->offY[b->n], oA = b->offA[b->n];
    if (nX > b->capX) {
        b->symX = st_realloc(b->symX, nX * sizeof(uint16_t)); /* fallback forms stay pageable: they travel only when packing fails */
        b->packedX = b200_pinned_grow(b->packedX, oX, nX);
        b->capX = nX;
    }
    if (nY > b->capY) {
        b->symY = st_realloc(b->symY, nY * sizeof(uint16_t));
        b->packedY = b200_pinned_grow(b->packedY, oY, nY);
        b->capY = nY;
    }
    if (nA > b->capA) {
        b->anchors = st_realloc(b->anchors, nA * 3 * sizeof(int32_t));
        b->anchorsDelta = st_realloc(b->anchorsDelta, nA * sizeof(uint32_t));
        b->runs = b200_pinned_grow(b->runs, b->nRuns * sizeof(uint32_t), nA * sizeof(uint32_t));
        b->capA = nA;
    }
}

static void b200_batch_add(B200Batch *b, SymbolString sX, SymbolString sY, stList *anchorPairs, int64_t shiftX, int slot,
                           bool raggedL, bool raggedR, int rle) {
    b200_batch_grow(b);
    int64_t i = b->n++;
    int64_t nA = anchorPairs == NULL ? 0 : stList_length(anchorPairs);
    int64_t oX = b->offX[i], oY = b->offY[i], oA = b->offA[i];
    b->n--; /* b200_batch_reserve looks at the filled part */
    if (oX + (int64_t) sX.length > b->capX || oY + (int64_t) sY.length > b->capY || oA + nA > b->capA)
        b200_batch_reserve(b, oX + (int64_t) sX.length > b->capX ? (oX + sX.length) * 2 + 1024 : b->capX,
                           oY + (int64_t) sY.length > b->capY ? (oY + sY.length) * 2 + 1024 : b->capY,
                           oA + nA > b->capA ? (oA + nA) * 2 + 1024 : b->capA);
    b->n++;
    memcpy(b->symX + oX, sX.sequence, sX.length * sizeof(uint16_t)); /* Symbol is uint16_t, inc/stateMachine.h:15 */
    memcpy(b->symY + oY, sY.sequence, sY.length * sizeof(uint16_t));
    /* one byte per symbol for the host link (a symbol that does not fit keeps the whole batch on the uint16 arrays, where
     * the library reports it) */
    for (int64_t k = 0; k < sX.length && b->packedOk; k++)
        if ((b->packedX[oX + k] = margin_b200_pack_symbol(sX.sequence[k], rle)) == 255) b->packedOk = 0;
    for (int64_t k = 0; k < sY.length && b->packedOk; k++)
        if ((b->packedY[oY + k] = margin_b200_pack_symbol(sY.sequence[k], rle)) == 255) b->packedOk = 0;
    {
        /* an stIntTuple is an int64_t array with its length in front (sonLibTuples.c:32-39; poa.c:601-606 writes into it the
         * same way), the list's backing array is read directly (b200_list_array): 4.7 M anchors per chunk */
        void **tuples = nA > 0 ? b200_list_array(anchorPairs) : NULL;
        int32_t *triples = b->anchors + 3 * oA;
        uint32_t *steps = b->anchorsDelta + oA, *runs = b->runs;
        int64_t px = -1, py = -1, pe = -1, nRuns = b->nRuns;
        const int dynamic = tls.dynamicAnchors;
        for (int64_t k = 0; k < nA; k++) {
            const int64_t *a = tuples[k];
            /* anchors may come as pairs (x, y): band_construct then uses p->diagonalExpansion and never reads a third field */
            const int64_t x = (int32_t) (a[1] - shiftX), y = (int32_t) a[2], e = (dynamic && a[0] > 2) ? (int32_t) a[3] : 0;
            triples[3 * k] = (int32_t) x;
            triples[3 * k + 1] = (int32_t) y;
            triples[3 * k + 2] = (int32_t) e;
            const int64_t dx = x - px, dy = y - py;
            /* the same anchor as a 4-byte step from its predecessor ... */
            const uint32_t w = margin_b200_pack_anchor(dx, dy, e);
            if (w == 0) b->compactOk = 0;
            steps[k] = w;
            /* ... and as runs: BAM-derived anchors are one per match column, so most steps are (+1, +1) with the same expansion */
            if (k > 0 && dx == 1 && dy == 1 && e == pe && (runs[nRuns - 1] & 255u) < 255u && runs[nRuns - 1] != 0) {
                runs[nRuns - 1]++;
            } else {
                const uint32_t r = margin_b200_pack_anchor_run(dx, dy, e);
                if (r == 0) b->runsOk = 0;
                runs[nRuns++] = r;
            }
            px = x;
            py = y;
            pe = e;
        }
        b->nRuns = nRuns;
    }
    b->offR[i + 1] = b->nRuns;
    b->offX[i + 1] = oX + sX.length;
    b->offY[i + 1] = oY + sY.length;
    b->offA[i + 1] = oA + nA;
    b->slot[i] = (uint8_t) slot;
    b->raggedL[i] = raggedL;
    b->raggedR[i] = raggedR;
}

static const uint16_t b200_noSymbols[1] = {0};
static const int32_t b200_noAnchors[3] = {0, 0, 0};

static void b200_batch_describe(const B200Batch *b, margin_b200_batch *mb) {
    memset(mb, 0, sizeof(*mb));
    mb->n = b->n;
    mb->sym_x = b->symX ? b->symX : b200_noSymbols;
    mb->off_x = b->offX;
    mb->sym_y = b->symY ? b->symY : b200_noSymbols;
    mb->off_y = b->offY;
    mb->anchors = b->anchors ? b->anchors : b200_noAnchors;
    if (b->compactOk && b->anchorsDelta) mb->anchors_delta = b->anchorsDelta; /* takes precedence over the triples */
    if (b->runsOk && b->nRuns > 0) { /* ... and the runs over both */
        mb->anchors_runs = b->runs;
        mb->off_r = b->offR;
    }
    if (b->packedOk && b->packedX && b->packedY) {
        mb->packed_x = b->packedX;
        mb->packed_y = b->packedY;
    }
    mb->off_a = b->offA;
    mb->model_slot = b->slot;
    mb->ragged_left = b->raggedL;
    mb->ragged_right = b->raggedR;
}

/* Debug hook: MARGIN_B200_DUMP=<prefix> writes every posterior batch (inputs, models, parameters) to
 * <prefix>.<thread>.<n>.bin so that scripts/replay_dump.py can replay it against the oracle outside margin. */
static void b200_batch_dump(const B200Batch *b, int wantIndels) {
    static __thread int counter = 0;
    const char *prefix = getenv("MARGIN_B200_DUMP");
    if (prefix == NULL || prefix[0] == 0) return;
    char *name = stString_print("%s.%d.%d.bin", prefix, omp_get_thread_num(), counter++);
    FILE *fh = fopen(name, "wb");
    if (fh == NULL) st_errAbort("margin_b200: cannot write %s\n", name);
    int64_t tX = b->offX[b->n], tY = b->offY[b->n], tA = b->offA[b->n];
    int64_t head[8] = {0x42323030, b->n, wantIndels, tX, tY, tA, sizeof(margin_b200_model), sizeof(margin_b200_params)};
    fwrite(head, sizeof(head), 1, fh);
    fwrite(&tls.loadedParams, sizeof(margin_b200_params), 1, fh);
    fwrite(tls.loaded, sizeof(margin_b200_model), 2, fh);
    fwrite(b->offX, sizeof(int64_t), b->n + 1, fh);
    fwrite(b->offY, sizeof(int64_t), b->n + 1, fh);
    fwrite(b->offA, sizeof(int64_t), b->n + 1, fh);
    fwrite(b->slot, 1, b->n, fh);
    fwrite(b->raggedL, 1, b->n, fh);
    fwrite(b->raggedR, 1, b->n, fh);
    if (tX) fwrite(b->symX, sizeof(uint16_t), tX, fh);
    if (tY) fwrite(b->symY, sizeof(uint16_t), tY, fh);
    if (tA) fwrite(b->anchors, 3 * sizeof(int32_t), tA, fh);
    fclose(fh);
    free(name);
}

static void b200_batch_run(B200Batch *b, int wantIndels, margin_b200_result *res) {
    margin_b200_batch mb;
    b200_batch_describe(b, &mb);
    b200_batch_dump(b, wantIndels);
    if (margin_b200_posteriors_batch(b200_ctx(), &mb, wantIndels, res) != 0) b200_fail();
}

/* List l of alignment i as an stList of stIntTuple(weight, x + shiftX, y), the reference's own representation. */
static stList *b200_list(const margin_b200_result *res, int l, int64_t i, int64_t shiftX) {
    stList *out = stList_construct3(0, (void (*)(void *)) stIntTuple_destruct);
    if (res->pair_format == MARGIN_B200_PAIRS_DELTA32) {
        margin_b200_delta_cursor cur = {(const uint32_t *) res->pairs[l] + res->off[l][i], res->escapes[l] + res->escape_off[l][i], 0, 0};
        for (int64_t k = res->off[l][i]; k < res->off[l][i + 1]; k++) {
            margin_b200_pair q = margin_b200_delta_next(&cur);
            stList_append(out, stIntTuple_construct3(q.w, (int64_t) q.x + shiftX, q.y));
        }
        return out;
    }
    if (res->pair_format == MARGIN_B200_PAIRS_PACKED64) {
        const uint64_t *v = (const uint64_t *) res->pairs[l];
        for (int64_t k = res->off[l][i]; k < res->off[l][i + 1]; k++) {
            margin_b200_pair q = margin_b200_unpack_pair(v[k]);
            stList_append(out, stIntTuple_construct3(q.w, (int64_t) q.x + shiftX, q.y));
        }
        return out;
    }
    const margin_b200_pair *p = res->pairs[l];
    for (int64_t k = res->off[l][i]; k < res->off[l][i + 1]; k++) {
        stList_append(out, stIntTuple_construct3(p[k].w, (int64_t) p[k].x + shiftX, p[k].y));
    }
    return out;
}

/* ------------------------------------------------------------------- inc/pairwiseAligner.h entry points */

void getAlignedPairsWithIndelsUsingAnchors(StateMachine *sM, SymbolString sX, SymbolString sY, stList *anchorPairs,
                                           PairwiseAlignmentParameters *p, stList **alignedPairs, stList **gapXPairs,
                                           stList **gapYPairs, bool alignmentHasRaggedLeftEnd,
                                           bool alignmentHasRaggedRightEnd) {
    b200_load_machine(0, sM);
    b200_load_params(p);
    B200Batch *b = b200_batch_construct(1);
    b200_batch_add(b, sX, sY, anchorPairs, 0, 0, alignmentHasRaggedLeftEnd, alignmentHasRaggedRightEnd, tls.loaded[0].rle);
    margin_b200_result res;
    b200_batch_run(b, 1, &res);
    *alignedPairs = b200_list(&res, 0, 0, 0);
    *gapXPairs = b200_list(&res, 1, 0, 0);
    *gapYPairs = b200_list(&res, 2, 0, 0);
    b200_batch_destruct(b);
}

stList *getAlignedPairsUsingAnchors(StateMachine *sM, SymbolString sX, SymbolString sY, stList *anchorPairs,
                                    PairwiseAlignmentParameters *p, bool alignmentHasRaggedLeftEnd,
                                    bool alignmentHasRaggedRightEnd) {
    b200_load_machine(0, sM);
    b200_load_params(p);
    B200Batch *b = b200_batch_construct(1);
    b200_batch_add(b, sX, sY, anchorPairs, 0, 0, alignmentHasRaggedLeftEnd, alignmentHasRaggedRightEnd, tls.loaded[0].rle);
    margin_b200_result res;
    b200_batch_run(b, 0, &res);
    stList *alignedPairs = b200_list(&res, 0, 0, 0);
    b200_batch_destruct(b);
    return alignedPairs;
}

stList *getAlignedPairs(StateMachine *sM, SymbolString sX, SymbolString sY, PairwiseAlignmentParameters *p,
                        bool alignmentHasRaggedLeftEnd, bool alignmentHasRaggedRightEnd) {
    return getAlignedPairsUsingAnchors(sM, sX, sY, NULL, p, alignmentHasRaggedLeftEnd, alignmentHasRaggedRightEnd);
}

void getAlignedPairsWithIndels(StateMachine *sM, SymbolString sX, SymbolString sY, PairwiseAlignmentParameters *p,
                               stList **alignedPairs, stList **gapXPairs, stList **gapYPairs,
                               bool alignmentHasRaggedLeftEnd, bool alignmentHasRaggedRightEnd) {
    getAlignedPairsWithIndelsUsingAnchors(sM, sX, sY, NULL, p, alignedPairs, gapXPairs, gapYPairs,
                                          alignmentHasRaggedLeftEnd, alignmentHasRaggedRightEnd);
}

/* ------------------------------------------------------- impl/poa.c:317-543 on the library's flat tuple arrays */

/*
 * poa_augment for one read, reading the (weight, x, y) records of a margin_b200_result directly (SURVEY.md §8 f N1).
 * The reference's version wants three stLists of individually allocated stIntTuples, builds an AVL set of the match
 * tuples and sorts the gap lists through their pointers; with the aligner on the GPU that was half of the wall time
 * of `margin polish`.  Same graph updates in the same order (base and repeat-count weights, observations, complete
 * inserts and deletes with their left shift), hence the same floating-point sums:
 *   - the match set is an open-addressing table of packed coordinates (isMatch, impl/poa.c:168-174);
 *   - gap records are sorted in place by (x, y) / (y, x) (cmpAlignedPairsByCoordinates / ...Inverted, :137-166);
 *   - a maximal run of gap records is scanned once per start k with a running minimum for the weight (:389-446).
 * poaBaseObservation_construct, poaInsert_construct, poaDelete_construct and getMaxCommonSuffixLength are exported by
 * impl/poa.c but not declared in inc/margin.h.
 */
PoaBaseObservation *poaBaseObservation_construct(int64_t readNo, int64_t offset, double weight);
PoaInsert *poaInsert_construct(RleString *insert, double weight, bool strand);
PoaDelete *poaDelete_construct(int64_t length, double weight, bool strand);
int64_t getMaxCommonSuffixLength(RleString *str1, int64_t length1, RleString *str2, bool compareRepeatCounts);

typedef struct {
    int64_t x, y, w;
} B200Pair;

typedef struct {
    B200Pair *pairs[3];
    int64_t cap[3];
    uint64_t *table; /* 0 = empty slot, else packed coordinates + 1 */
    uint64_t tableSize;
    uint64_t *keys[2]; /* radix sort buffers of b200_sort_pairs */
    int64_t capKeys;
    RleString sub; /* the candidate insert / deleted substring (b200_substring), buffers reused across candidates */
    uint64_t capSub;
} B200AugmentScratch;

static __thread B200AugmentScratch augmentScratch;
static __thread double augmentPhase[4]; /* match loop, match set, inserts (sort + enumeration), deletes */

/* List l of alignment i as flat records, x shifted back by the crop. */
static int64_t b200_pairs(const margin_b200_result *res, int l, int64_t i, int64_t shiftX) {
    B200AugmentScratch *sc = &augmentScratch;
    int64_t first = res->off[l][i], n = res->off[l][i + 1] - first;
    if (n > sc->cap[l]) {
        sc->cap[l] = n * 2 + 1024;
        sc->pairs[l] = st_realloc(sc->pairs[l], sc->cap[l] * sizeof(B200Pair));
    }
    B200Pair *out = sc->pairs[l];
    if (res->pair_format == MARGIN_B200_PAIRS_DELTA32) {
        margin_b200_delta_cursor cur = {(const uint32_t *) res->pairs[l] + first, res->escapes[l] + res->escape_off[l][i], 0, 0};
        for (int64_t k = 0; k < n; k++) {
            margin_b200_pair q = margin_b200_delta_next(&cur);
            out[k].x = (int64_t) q.x + shiftX;
            out[k].y = q.y;
            out[k].w = q.w;
        }
    } else if (res->pair_format == MARGIN_B200_PAIRS_PACKED64) {
        const uint64_t *v = (const uint64_t *) res->pairs[l] + first;
        for (int64_t k = 0; k < n; k++) {
            margin_b200_pair q = margin_b200_unpack_pair(v[k]);
            out[k].x = (int64_t) q.x + shiftX;
            out[k].y = q.y;
            out[k].w = q.w;
        }
    } else {
        const margin_b200_pair *q = res->pairs[l] + first;
        for (int64_t k = 0; k < n; k++) {
            out[k].x = (int64_t) q[k].x + shiftX;
            out[k].y = q[k].y;
            out[k].w = q[k].w;
        }
    }
    return n;
}

static inline uint64_t b200_coordinate_key(int64_t x, int64_t y) {
    return (((uint64_t) x << 32) | (uint64_t) y) + 1;
}

static inline uint64_t b200_table_slot(uint64_t key, uint64_t mask) {
    return ((key * 0x9E3779B97F4A7C15ULL) >> 20) & mask;
}

static void b200_match_set_fill(const B200Pair *matches, int64_t n) {
    B200AugmentScratch *sc = &augmentScratch;
    uint64_t size = 1024;
    while (size < (uint64_t) n * 2 + 2) size *= 2;
    if (size > sc->tableSize) {
        free(sc->table);
        sc->table = st_malloc(size * sizeof(uint64_t));
    }
    sc->tableSize = size;
    memset(sc->table, 0, size * sizeof(uint64_t));
    for (int64_t k = 0; k < n; k++) {
        uint64_t key = b200_coordinate_key(matches[k].x, matches[k].y), at = b200_table_slot(key, size - 1);
        while (sc->table[at] != 0) at = (at + 1) & (size - 1);
        sc->table[at] = key;
    }
}

static inline bool b200_is_match(int64_t x, int64_t y) {
    const B200AugmentScratch *sc = &augmentScratch;
    if (x < 0 || y < 0) return 0;
    uint64_t key = b200_coordinate_key(x, y), mask = sc->tableSize - 1, at = b200_table_slot(key, mask);
    while (sc->table[at] != 0) {
        if (sc->table[at] == key) return 1;
        at = (at + 1) & mask;
    }
    return 0;
}

static int b200_cmp_xy(const void *a, const void *b) {
    const B200Pair *p = a, *q = b;
    if (p->x != q->x) return p->x < q->x ? -1 : 1;
    return p->y < q->y ? -1 : (p->y > q->y ? 1 : 0);
}

static int b200_cmp_yx(const void *a, const void *b) {
    const B200Pair *p = a, *q = b;
    if (p->y != q->y) return p->y < q->y ? -1 : 1;
    return p->x < q->x ? -1 : (p->x > q->x ? 1 : 0);
}

/*
 * Sorts gap records by (x, y) (byYX = 0: cmpAlignedPairsByCoordinates, impl/poa.c:137-151) or by (y, x) (byYX = 1:
 * ...Inverted, :153-166).  Coordinates are >= -1 and weights < 2^24, so a record packs into one 64-bit key
 * (major + 1) << 44 | (minor + 1) << 24 | weight, which an LSD radix sort of four 10-bit digits orders in O(n); records
 * are unique per cell, so the order equals the comparison sort's.  Falls back to qsort for short lists and for
 * coordinates beyond 2^20.
 */
#define B200_RADIX_BITS 10
static void b200_sort_pairs(B200Pair *p, int64_t n, int byYX) {
    B200AugmentScratch *sc = &augmentScratch;
    int ok = n >= 256;
    for (int64_t i = 0; ok && i < n; i++) {
        ok = p[i].x >= -1 && p[i].y >= -1 && p[i].x < (1 << 20) - 1 && p[i].y < (1 << 20) - 1 && p[i].w >= 0 &&
             p[i].w < (1 << 24);
    }
    if (!ok) {
        qsort(p, n, sizeof(B200Pair), byYX ? b200_cmp_yx : b200_cmp_xy);
        return;
    }
    if (n > sc->capKeys) {
        sc->capKeys = n * 2;
        sc->keys[0] = st_realloc(sc->keys[0], sc->capKeys * sizeof(uint64_t));
        sc->keys[1] = st_realloc(sc->keys[1], sc->capKeys * sizeof(uint64_t));
    }
    uint64_t *a = sc->keys[0], *b = sc->keys[1];
    for (int64_t i = 0; i < n; i++) {
        uint64_t major = (uint64_t) ((byYX ? p[i].y : p[i].x) + 1), minor = (uint64_t) ((byYX ? p[i].x : p[i].y) + 1);
        a[i] = (major << 44) | (minor << 24) | (uint64_t) p[i].w;
    }
    for (int shift = 24; shift < 64; shift += B200_RADIX_BITS) {
        int64_t count[1 << B200_RADIX_BITS];
        memset(count, 0, sizeof(count));
        for (int64_t i = 0; i < n; i++) count[(a[i] >> shift) & ((1 << B200_RADIX_BITS) - 1)]++;
        if (count[(a[0] >> shift) & ((1 << B200_RADIX_BITS) - 1)] == n) continue; /* all keys agree on this digit */
        int64_t at = 0;
        for (int d = 0; d < (1 << B200_RADIX_BITS); d++) {
            int64_t c = count[d];
            count[d] = at;
            at += c;
        }
        for (int64_t i = 0; i < n; i++) b[count[(a[i] >> shift) & ((1 << B200_RADIX_BITS) - 1)]++] = a[i];
        uint64_t *t = a;
        a = b;
        b = t;
    }
    for (int64_t i = 0; i < n; i++) {
        int64_t major = (int64_t) (a[i] >> 44) - 1, minor = (int64_t) ((a[i] >> 24) & 0xfffff) - 1;
        p[i].x = byYX ? minor : major;
        p[i].y = byYX ? major : minor;
        p[i].w = (int64_t) (a[i] & 0xffffff);
    }
}

/* rleString_copySubstring (impl/rle.c:82-102) into the thread's scratch RleString: no allocation per candidate. */
static RleString *b200_substring(const RleString *from, int64_t start, int64_t length) {
    B200AugmentScratch *sc = &augmentScratch;
    if ((uint64_t) length + 1 > sc->capSub) {
        sc->capSub = (uint64_t) length * 2 + 64;
        sc->sub.rleString = st_realloc(sc->sub.rleString, sc->capSub);
        sc->sub.repeatCounts = st_realloc(sc->sub.repeatCounts, sc->capSub * sizeof(uint64_t));
    }
    RleString *sub = &sc->sub;
    sub->length = (uint64_t) length;
    sub->nonRleLength = 0;
    memcpy(sub->rleString, from->rleString + start, (size_t) length);
    sub->rleString[length] = '\0';
    for (int64_t i = 0; i < length; i++) {
        sub->repeatCounts[i] = from->repeatCounts[start + i];
        sub->nonRleLength += sub->repeatCounts[i];
    }
    return sub;
}

/* The backing array of an stList (its first field, externalTools/sonLib/C/impl/sonLibListPrivate.h:11-16), checked once
 * against stList_get: the per-tuple loops below would otherwise spend a tenth of a chunk in that accessor. */
static void **b200_list_array(stList *list) {
    static int checked = 0;
    if (!checked) {
        stList *probe = stList_construct();
        stList_append(probe, &checked);
        stList_append(probe, &probe);
        void **a = *(void ***) probe;
        if (a[0] != stList_get(probe, 0) || a[1] != stList_get(probe, 1)) st_errAbort("margin_b200: stList layout check failed\n");
        stList_destruct(probe);
        checked = 1;
    }
    return *(void ***) list;
}

/*
 * Observation slabs.  poa_augment allocates one 24-byte PoaBaseObservation per tuple (impl/poa.c:21-29,
 * :338, :203, :232): nine million callocs per sweep of a 100 kb chunk and as many frees when the graph dies, more time
 * than everything else the host does for the chunk.  A graph built by this file's poa_realign takes its observations from
 * slabs instead: the match observations of all reads from ONE block laid out node by node (counted first, so every
 * node's list is created at its final size and filled in place), insert / delete observations from a bump allocator.  The
 * lists that hold slab observations get a NULL element destructor; the slabs hang off a registry keyed by the Poa and are
 * released by poa_destruct below, which replaces the reference's (weakened, oracle/Makefile) and otherwise does what
 * impl/poa.c:128-132 does.  Order and content of every list are those of the reference's read-by-read appends.
 */
typedef struct B200Slab {
    struct B200Slab *next;
    int64_t bytes;
} B200Slab;

typedef struct {
    Poa *poa;
    B200Slab *slabs;
} B200SlabOwner;

static pthread_mutex_t b200_slabLock = PTHREAD_MUTEX_INITIALIZER;
static B200SlabOwner **b200_slabOwners; /* live graphs that own slabs (at most a few per thread) */
static int64_t b200_slabOwnersN, b200_slabOwnersCap;

static __thread struct {
    B200SlabOwner *owner;         /* NULL: observations come from poaBaseObservation_construct */
    char *at, *end;               /* the bump allocator's current slab */
    PoaBaseObservation **nextObs; /* per node: where its next match observation goes ... */
    void ***nextSlot;             /* ... and the slot of its list that will point at it */
    int64_t capNodes;
} slabState;

#define B200_SLAB_BYTES (4 << 20)

static void *b200_slab_new(B200SlabOwner *owner, int64_t bytes) {
    B200Slab *slab = st_malloc(sizeof(B200Slab) + (size_t) bytes);
    slab->bytes = bytes;
    slab->next = owner->slabs; /* only the thread that builds the graph touches the chain until poa_destruct */
    owner->slabs = slab;
    return slab + 1;
}

static B200SlabOwner *b200_slab_owner(Poa *poa) {
    B200SlabOwner *owner = st_calloc(1, sizeof(B200SlabOwner));
    owner->poa = poa;
    pthread_mutex_lock(&b200_slabLock);
    if (b200_slabOwnersN == b200_slabOwnersCap) {
        b200_slabOwnersCap = b200_slabOwnersCap * 2 + 64;
        b200_slabOwners = st_realloc(b200_slabOwners, b200_slabOwnersCap * sizeof(B200SlabOwner *));
    }
    b200_slabOwners[b200_slabOwnersN++] = owner;
    pthread_mutex_unlock(&b200_slabLock);
    return owner;
}

static inline void *b200_slab_alloc(int64_t bytes) { /* 8-byte aligned, not zeroed */
    bytes = (bytes + 7) & ~(int64_t) 7;
    if (slabState.at + bytes > slabState.end) {
        int64_t size = bytes > B200_SLAB_BYTES ? bytes : B200_SLAB_BYTES;
        slabState.at = b200_slab_new(slabState.owner, size);
        slabState.end = slabState.at + size;
    }
    void *out = slabState.at;
    slabState.at += bytes;
    return out;
}

static inline PoaBaseObservation *b200_observation(int64_t readNo, int64_t offset, double weight) {
    if (slabState.owner == NULL) return poaBaseObservation_construct(readNo, offset, weight);
    PoaBaseObservation *obs = b200_slab_alloc(sizeof(PoaBaseObservation));
    obs->readNo = readNo;
    obs->offset = offset;
    obs->weight = weight;
    return obs;
}

void poa_destruct(Poa *poa) {
    B200Slab *slabs = NULL;
    pthread_mutex_lock(&b200_slabLock);
    for (int64_t i = 0; i < b200_slabOwnersN; i++) {
        if (b200_slabOwners[i]->poa != poa) continue;
        slabs = b200_slabOwners[i]->slabs;
        free(b200_slabOwners[i]);
        b200_slabOwners[i] = b200_slabOwners[--b200_slabOwnersN];
        break;
    }
    pthread_mutex_unlock(&b200_slabLock);
    stList_destruct(poa->nodes); /* lists of slab observations carry no element destructor */
    rleString_destruct(poa->refString);
    free(poa);
    while (slabs != NULL) {
        B200Slab *next = slabs->next;
        free(slabs);
        slabs = next;
    }
}

/*
 * Where an insert / delete of a node already exists.  poa_augment (impl/poa.c:180-232) scans the node's list for every
 * candidate (rleString_eq per entry): with a few hundred reads the lists grow to dozens of low-weight entries and the
 * scans were a quarter of the augment.  During one poa_realign the graph starts without inserts and deletes, so an
 * open-addressing table filled by this file's own appends knows every entry: key = (node position, delete length), or
 * (node position, hash of the insert) verified with rleString_eq.  The lists themselves and their order are untouched.
 */
typedef struct {
    uint64_t key; /* 0 = empty */
    const void *owner; /* the PoaNode whose list holds the entry */
    void *entry;
} B200IndelSlot;

typedef struct {
    B200IndelSlot *slots;
    uint64_t size, used;
    int active; /* 0: the graph may hold entries this table does not know (test hook): scan the lists */
} B200IndelTable;

static __thread B200IndelTable indelTable;

static void b200_indel_table_begin(int fresh) {
    B200IndelTable *t = &indelTable;
    if (t->slots == NULL) {
        t->size = 1 << 16;
        t->slots = st_calloc(t->size, sizeof(B200IndelSlot));
    } else if (t->used > 0) {
        memset(t->slots, 0, t->size * sizeof(B200IndelSlot));
    }
    t->used = 0;
    t->active = fresh;
}

static inline uint64_t b200_mix(uint64_t h) {
    h ^= h >> 33;
    h *= 0xff51afd7ed558ccdULL;
    h ^= h >> 33;
    h *= 0xc4ceb9fe1a85ec53ULL;
    h ^= h >> 33;
    return h | 1; /* never 0 */
}

static void b200_indel_table_put(uint64_t key, const void *owner, void *entry) {
    B200IndelTable *t = &indelTable;
    if ((t->used + 1) * 2 > t->size) { /* grow and re-insert */
        B200IndelSlot *old = t->slots;
        uint64_t oldSize = t->size;
        t->size *= 4;
        t->slots = st_calloc(t->size, sizeof(B200IndelSlot));
        for (uint64_t i = 0; i < oldSize; i++) {
            if (old[i].key == 0) continue;
            uint64_t at = old[i].key & (t->size - 1);
            while (t->slots[at].key != 0) at = (at + 1) & (t->size - 1);
            t->slots[at] = old[i];
        }
        free(old);
    }
    uint64_t at = key & (t->size - 1);
    while (t->slots[at].key != 0) at = (at + 1) & (t->size - 1);
    t->slots[at].key = key;
    t->slots[at].owner = owner;
    t->slots[at].entry = entry;
    t->used++;
}

static inline uint64_t b200_insert_key(const PoaNode *node, const RleString *insert) {
    uint64_t h = (uint64_t) (uintptr_t) node * 0x9E3779B97F4A7C15ULL + insert->length;
    for (uint64_t i = 0; i < insert->length; i++)
        h = (h ^ ((uint64_t) (unsigned char) insert->rleString[i] | (insert->repeatCounts[i] << 8))) * 0x100000001B3ULL;
    return b200_mix(h);
}

static void b200_node_add_insert(PoaNode *node, RleString *insert, double weight, bool strand, PoaBaseObservation *obs) {
    PoaInsert *hit = NULL;
    uint64_t key = 0;
    if (indelTable.active) {
        const B200IndelTable *t = &indelTable;
        key = b200_insert_key(node, insert);
        for (uint64_t at = key & (t->size - 1); t->slots[at].key != 0 && hit == NULL; at = (at + 1) & (t->size - 1)) {
            if (t->slots[at].key != key || t->slots[at].owner != node) continue;
            PoaInsert *candidate = t->slots[at].entry;
            if (rleString_eq(candidate->insert, insert)) hit = candidate;
        }
    } else {
        for (int64_t m = 0; m < stList_length(node->inserts) && hit == NULL; m++) {
            PoaInsert *candidate = stList_get(node->inserts, m);
            if (rleString_eq(candidate->insert, insert)) hit = candidate;
        }
    }
    if (hit == NULL) {
        hit = poaInsert_construct(rleString_copy(insert), 0, FALSE);
        if (slabState.owner != NULL) stList_setDestructor(hit->observations, NULL);
        stList_append(node->inserts, hit);
        if (indelTable.active) b200_indel_table_put(key, node, hit);
    }
    if (strand) hit->weightForwardStrand += weight;
    else hit->weightReverseStrand += weight;
    stList_append(hit->observations, obs);
}

static void b200_node_add_delete(PoaNode *node, int64_t length, double weight, bool strand, PoaBaseObservation *obs) {
    PoaDelete *hit = NULL;
    uint64_t key = 0;
    if (indelTable.active) {
        const B200IndelTable *t = &indelTable;
        key = b200_mix(((uint64_t) (uintptr_t) node * 0x9E3779B97F4A7C15ULL) ^ ((uint64_t) length * 0xD6E8FEB86659FD93ULL));
        for (uint64_t at = key & (t->size - 1); t->slots[at].key != 0 && hit == NULL; at = (at + 1) & (t->size - 1)) {
            if (t->slots[at].key != key || t->slots[at].owner != node) continue;
            PoaDelete *candidate = t->slots[at].entry;
            if (candidate->length == length) hit = candidate;
        }
    } else {
        for (int64_t m = 0; m < stList_length(node->deletes) && hit == NULL; m++) {
            PoaDelete *candidate = stList_get(node->deletes, m);
            if (candidate->length == length) hit = candidate;
        }
    }
    if (hit == NULL) {
        hit = poaDelete_construct(length, 0, FALSE);
        if (slabState.owner != NULL) stList_setDestructor(hit->observations, NULL);
        stList_append(node->deletes, hit);
        if (indelTable.active) b200_indel_table_put(key, node, hit);
    }
    if (strand) hit->weightForwardStrand += weight;
    else hit->weightReverseStrand += weight;
    stList_append(hit->observations, obs);
}

/*
 * Inserts and deletes, node by node.  poa_augment looks every candidate up in its node's list the
 * moment it is found (impl/poa.c:180-232); read after read that touches a different node, entry, observation list and
 * string each time, a handful of cache misses per candidate and most of what is left of the augment.  With slabs on, the
 * read loop only RECORDS candidates (node, read, offset, weight, strand, the shifted and rotated string); after the last
 * read b200_candidates_commit orders them by node with a stable counting sort and walks each node once: its few distinct
 * strings / lengths are matched in a small local table, every entry's observation list is created at its final size and
 * the observations come from the slab.  The sort is stable and the recording order is the reference's processing order,
 * so every list (inserts of a node, deletes of a node, observations of an entry) has the reference's order; weights are
 * sums of integers below 2^24 and exact in any order.
 */
typedef struct {
    int32_t node, readNo, offset, weight;
    int64_t at;      /* inserts: where the string starts in B200Candidates.bases / counts; deletes: the length */
    uint64_t hash;   /* inserts: of the string; deletes: unused */
    int32_t length;  /* inserts: symbols of the string */
    int32_t strand;
} B200Candidate;

typedef struct {
    B200Candidate *items, *sorted;
    int64_t n, cap, capSorted;
    char *bases;
    uint64_t *counts;
    int64_t nSymbols, capSymbols;
} B200Candidates;

static __thread B200Candidates candidates[2]; /* 0 = inserts, 1 = deletes */
static __thread int64_t candidateStats[4];    /* candidates and distinct entries of the last commit, per kind */

static inline B200Candidate *b200_candidate_push(B200Candidates *c, int64_t node, int64_t readNo, int64_t offset, double weight,
                                                 bool strand) {
    if (c->n == c->cap) {
        c->cap = c->cap * 2 + (1 << 16);
        c->items = st_realloc(c->items, c->cap * sizeof(B200Candidate));
    }
    B200Candidate *it = &c->items[c->n++];
    it->node = (int32_t) node;
    it->readNo = (int32_t) readNo;
    it->offset = (int32_t) offset;
    it->weight = (int32_t) weight;
    it->strand = strand;
    return it;
}

static inline void b200_candidate_insert(int64_t node, const RleString *insert, double weight, bool strand, int64_t readNo,
                                         int64_t offset) {
    B200Candidates *c = &candidates[0];
    B200Candidate *it = b200_candidate_push(c, node, readNo, offset, weight, strand);
    if (c->nSymbols + (int64_t) insert->length > c->capSymbols) {
        c->capSymbols = (c->capSymbols + (int64_t) insert->length) * 2 + (1 << 16);
        c->bases = st_realloc(c->bases, c->capSymbols);
        c->counts = st_realloc(c->counts, c->capSymbols * sizeof(uint64_t));
    }
    it->at = c->nSymbols;
    it->length = (int32_t) insert->length;
    uint64_t h = insert->length;
    for (uint64_t i = 0; i < insert->length; i++) {
        c->bases[c->nSymbols + i] = insert->rleString[i];
        c->counts[c->nSymbols + i] = insert->repeatCounts[i];
        h = (h ^ ((uint64_t) (unsigned char) insert->rleString[i] | (insert->repeatCounts[i] << 8))) * 0x100000001B3ULL;
    }
    it->hash = h;
    c->nSymbols += insert->length;
}

static inline void b200_candidate_delete(int64_t node, int64_t length, double weight, bool strand, int64_t readNo, int64_t offset) {
    B200Candidate *it = b200_candidate_push(&candidates[1], node, readNo, offset, weight, strand);
    it->at = length;
    it->hash = 0;
    it->length = 0;
}

static void b200_candidates_sort(B200Candidates *c, int64_t nNodes) {
    if (c->n > c->capSorted) {
        c->capSorted = c->cap;
        free(c->sorted);
        c->sorted = st_malloc(c->capSorted * sizeof(B200Candidate));
    }
    int64_t *first = st_calloc(nNodes + 1, sizeof(int64_t));
    for (int64_t k = 0; k < c->n; k++) first[c->items[k].node + 1]++;
    for (int64_t j = 0; j < nNodes; j++) first[j + 1] += first[j];
    for (int64_t k = 0; k < c->n; k++) c->sorted[first[c->items[k].node]++] = c->items[k];
    free(first);
}

/* Inserts and deletes of a slab graph live in the slab themselves (struct, RleString, its two arrays): only their
 * observation lists are separate allocations (so that a list can still be appended to), and releasing those is all
 * the element destructor of node->inserts / node->deletes has left to do (cf. poaInsert_destruct, impl/poa.c:49-53). */
static void b200_slab_insert_destruct(void *insert) { stList_destruct(((PoaInsert *) insert)->observations); }
static void b200_slab_delete_destruct(void *delete) { stList_destruct(((PoaDelete *) delete)->observations); }

static PoaInsert *b200_slab_insert(const char *bases, const uint64_t *counts, int64_t length) {
    PoaInsert *insert = b200_slab_alloc(sizeof(PoaInsert));
    RleString *string = b200_slab_alloc(sizeof(RleString));
    string->rleString = b200_slab_alloc(length + 1);
    string->repeatCounts = b200_slab_alloc(length * (int64_t) sizeof(uint64_t));
    memcpy(string->rleString, bases, length);
    string->rleString[length] = '\0';
    string->length = (uint64_t) length;
    string->nonRleLength = 0;
    for (int64_t i = 0; i < length; i++) string->nonRleLength += (string->repeatCounts[i] = counts[i]);
    insert->insert = string;
    insert->weightForwardStrand = insert->weightReverseStrand = 0.0;
    insert->observations = NULL;
    return insert;
}

static PoaDelete *b200_slab_delete(int64_t length) {
    PoaDelete *delete = b200_slab_alloc(sizeof(PoaDelete));
    delete->length = length;
    delete->weightForwardStrand = delete->weightReverseStrand = 0.0;
    delete->observations = NULL;
    return delete;
}

static void b200_candidates_commit(Poa *poa) {
    int64_t nNodes = stList_length(poa->nodes);
    void **nodes = b200_list_array(poa->nodes);
    int64_t capLocal = 64, *entryOf = NULL, capEntryOf = 0;
    struct {
        uint64_t key;
        void *entry;
        int64_t count, filled;
        void **slots;
    } *local = st_malloc(capLocal * sizeof(*local));
    for (int kind = 0; kind < 2; kind++) {
        B200Candidates *c = &candidates[kind];
        b200_candidates_sort(c, nNodes);
        candidateStats[2 * kind] = c->n;
        candidateStats[2 * kind + 1] = 0;
        for (int64_t i = 0, j; i < c->n; i = j) {
            PoaNode *node = nodes[c->sorted[i].node];
            for (j = i + 1; j < c->n && c->sorted[j].node == c->sorted[i].node; j++);
            if (j - i > capEntryOf) {
                capEntryOf = (j - i) * 2 + 64;
                entryOf = st_realloc(entryOf, capEntryOf * sizeof(int64_t));
            }
            int64_t nLocal = 0;
            for (int64_t k = i; k < j; k++) { /* which entry of the node each candidate belongs to, new entries in order */
                const B200Candidate *it = &c->sorted[k];
                uint64_t key = kind == 0 ? it->hash : (uint64_t) it->at;
                int64_t e = 0;
                for (; e < nLocal; e++) {
                    if (local[e].key != key) continue;
                    if (kind == 1) break;
                    const RleString *have = ((PoaInsert *) local[e].entry)->insert;
                    if ((int64_t) have->length == it->length && memcmp(have->rleString, c->bases + it->at, it->length) == 0 &&
                        memcmp(have->repeatCounts, c->counts + it->at, it->length * sizeof(uint64_t)) == 0) break;
                }
                if (e == nLocal) {
                    if (nLocal == capLocal) {
                        capLocal *= 2;
                        local = st_realloc(local, capLocal * sizeof(*local));
                    }
                    local[e].key = key;
                    local[e].count = 0;
                    local[e].filled = 0;
                    local[e].entry = kind == 0 ? (void *) b200_slab_insert(c->bases + it->at, c->counts + it->at, it->length)
                                               : (void *) b200_slab_delete(it->at);
                    nLocal++;
                }
                local[e].count++;
                entryOf[k - i] = e;
            }
            candidateStats[2 * kind + 1] += nLocal;
            /* the node's list and every entry's observation list at their final sizes */
            stList **entries = kind == 0 ? &node->inserts : &node->deletes;
            if (stList_length(*entries) != 0) st_errAbort("margin_b200: the graph is not fresh\n");
            stList_destruct(*entries);
            *entries = stList_construct3(nLocal, kind == 0 ? b200_slab_insert_destruct : b200_slab_delete_destruct);
            void **entrySlots = b200_list_array(*entries);
            for (int64_t e = 0; e < nLocal; e++) {
                stList *observations = stList_construct3(local[e].count, NULL);
                if (kind == 0) ((PoaInsert *) local[e].entry)->observations = observations;
                else ((PoaDelete *) local[e].entry)->observations = observations;
                local[e].slots = b200_list_array(observations);
                entrySlots[e] = local[e].entry;
            }
            for (int64_t k = i; k < j; k++) {
                const B200Candidate *it = &c->sorted[k];
                int64_t e = entryOf[k - i];
                double *forward = kind == 0 ? &((PoaInsert *) local[e].entry)->weightForwardStrand
                                            : &((PoaDelete *) local[e].entry)->weightForwardStrand;
                double *reverse = kind == 0 ? &((PoaInsert *) local[e].entry)->weightReverseStrand
                                            : &((PoaDelete *) local[e].entry)->weightReverseStrand;
                *(it->strand ? forward : reverse) += it->weight;
                local[e].slots[local[e].filled++] = b200_observation(it->readNo, it->offset, it->weight);
            }
        }
        c->n = 0;
        c->nSymbols = 0;
    }
    free(local);
    free(entryOf);
}

/*
 * Counts the match tuples of ALL reads per node, takes one block for their observations and gives every node of the
 * (fresh) graph an observation list of its final length; b200_poa_augment then writes observation k of a node straight
 * into its place.  Reads are still walked in order, so each list holds what the reference's appends would (impl/poa.c:338).
 */
static void b200_match_lists_presize(Poa *poa, const margin_b200_result *res, int64_t n, const int64_t *shiftX) {
    int64_t nNodes = stList_length(poa->nodes), total = 0;
    if (nNodes + 1 > slabState.capNodes) {
        slabState.capNodes = nNodes + 1;
        slabState.nextObs = st_realloc(slabState.nextObs, slabState.capNodes * sizeof(PoaBaseObservation *));
        slabState.nextSlot = st_realloc(slabState.nextSlot, slabState.capNodes * sizeof(void **));
    }
    int64_t *counts = st_calloc(nNodes + 1, sizeof(int64_t));
    for (int64_t i = 0; i < n; i++) {
        int64_t m = b200_pairs(res, 0, i, shiftX[i]);
        const B200Pair *matches = augmentScratch.pairs[0];
        for (int64_t k = 0; k < m; k++) {
            if (matches[k].x + 1 < 1 || matches[k].x + 1 >= nNodes) st_errAbort("margin_b200: a match beyond the graph\n");
            counts[matches[k].x + 1]++;
        }
        total += m;
    }
    slabState.owner = b200_slab_owner(poa);
    slabState.at = slabState.end = NULL;
    PoaBaseObservation *block = b200_slab_new(slabState.owner, (total + 1) * (int64_t) sizeof(PoaBaseObservation));
    void **nodes = b200_list_array(poa->nodes);
    for (int64_t j = 0; j < nNodes; j++) {
        PoaNode *node = nodes[j];
        if (stList_length(node->observations) != 0) st_errAbort("margin_b200: the graph is not fresh\n");
        stList_destruct(node->observations);
        node->observations = stList_construct3(counts[j], NULL);
        slabState.nextObs[j] = block;
        slabState.nextSlot[j] = b200_list_array(node->observations);
        block += counts[j];
    }
    free(counts);
}

static void b200_poa_augment(Poa *poa, RleString *read, bool readStrand, int64_t readNo, const margin_b200_result *res,
                             int64_t aln, int64_t shiftX, PolishParams *polishParams) {
    B200AugmentScratch *sc = &augmentScratch;
    bool compareRL = polishParams->poaConstructCompareRepeatCounts;
    int64_t nMatch = b200_pairs(res, 0, aln, shiftX);
    int64_t nDelete = b200_pairs(res, 1, aln, shiftX); /* gapX: reference base without read base */
    int64_t nInsert = b200_pairs(res, 2, aln, shiftX); /* gapY */
    B200Pair *matches = sc->pairs[0], *deletes = sc->pairs[1], *inserts = sc->pairs[2];

    /* match weights and observations, in list order (impl/poa.c:324-341) */
    double tPhase0 = omp_get_wtime();
    void **nodes = b200_list_array(poa->nodes);
    for (int64_t k = 0; k < nMatch; k++) {
        PoaNode *node = nodes[matches[k].x + 1];
        int64_t y = matches[k].y;
        node->baseWeights[poa->alphabet->convertCharToSymbol(read->rleString[y])] += matches[k].w;
        int64_t rc = read->repeatCounts[y] < poa->maxRepeatCount ? read->repeatCounts[y] : poa->maxRepeatCount - 1;
        node->repeatCountWeights[rc] += matches[k].w;
        if (slabState.owner != NULL) { /* the node's list has its final size already (b200_match_lists_presize) */
            PoaBaseObservation *obs = slabState.nextObs[matches[k].x + 1]++;
            obs->readNo = readNo;
            obs->offset = y;
            obs->weight = matches[k].w;
            *slabState.nextSlot[matches[k].x + 1]++ = obs;
        } else {
            stList_append(node->observations, poaBaseObservation_construct(readNo, y, matches[k].w));
        }
    }
    double tPhase1 = omp_get_wtime();
    b200_match_set_fill(matches, nMatch);
    double tPhase2 = omp_get_wtime();

    /* complete inserts: same reference coordinate, consecutive read coordinates, flanked by matches or the ends */
    b200_sort_pairs(inserts, nInsert, 0);
    for (int64_t i = 0, j; i < nInsert; i = j) {
        int64_t x = inserts[i].x, y0 = inserts[i].y;
        for (j = i + 1; j < nInsert && inserts[j].x == x && inserts[j].y == y0 + (j - i); j++);
        for (int64_t k = i; k < j; k++) {
            int64_t yk = y0 + (k - i);
            if (yk - 1 > -1 && !b200_is_match(x, yk - 1)) continue;
            double weight = UINT_MAX;
            for (int64_t l = k; l < j; l++) {
                weight = weight < inserts[l].w ? weight : inserts[l].w;
                int64_t yl = y0 + (l - i);
                if (yl + 1 < read->length && !b200_is_match(x + 1, yl + 1)) continue;
                RleString *insert = b200_substring(read, yk, l + 1 - k);
                int64_t position = getShift(poa->refString, x + 1, insert, compareRL);
                int64_t suffix = getMaxCommonSuffixLength(poa->refString, position, insert, compareRL);
                if (suffix > 0) {
                    rleString_rotateString(insert, suffix, polishParams->useRunLengthEncoding);
                    position -= suffix;
                }
                if (slabState.owner != NULL) b200_candidate_insert(position, insert, weight, readStrand, readNo, yk);
                else b200_node_add_insert(nodes[position], insert, weight, readStrand, b200_observation(readNo, yk, weight));
            }
        }
    }

    /* complete deletes: same read coordinate, consecutive reference coordinates */
    double tPhase3 = omp_get_wtime();
    b200_sort_pairs(deletes, nDelete, 1);
    int64_t refLength = poa->refString->length;
    for (int64_t i = 0, j; i < nDelete; i = j) {
        int64_t x0 = deletes[i].x, y = deletes[i].y;
        for (j = i + 1; j < nDelete && deletes[j].y == y && deletes[j].x == x0 + (j - i); j++);
        for (int64_t k = i; k < j; k++) {
            int64_t xk = x0 + (k - i);
            if (xk - 1 > -1 && !b200_is_match(xk - 1, y)) continue;
            double weight = UINT_MAX;
            for (int64_t l = k; l < j; l++) {
                weight = weight < deletes[l].w ? weight : deletes[l].w;
                int64_t xl = x0 + (l - i);
                if (xl + 1 < refLength && !b200_is_match(xl + 1, y + 1)) continue;
                int64_t length = l - k + 1;
                RleString *deleted = b200_substring(poa->refString, xk, length);
                int64_t position = getShift(poa->refString, xk, deleted, compareRL);
                position -= getMaxCommonSuffixLength(poa->refString, position, deleted, compareRL);
                if (slabState.owner != NULL) b200_candidate_delete(position, length, weight, readStrand, readNo, y);
                else b200_node_add_delete(nodes[position], length, weight, readStrand, b200_observation(readNo, y, weight));
            }
        }
    }
    double tPhase4 = omp_get_wtime();
    augmentPhase[0] += tPhase1 - tPhase0;
    augmentPhase[1] += tPhase2 - tPhase1;
    augmentPhase[2] += tPhase3 - tPhase2;
    augmentPhase[3] += tPhase4 - tPhase3;
}

/*
 * Test hook (oracle/b200_tests.c): b200_poa_augment on the reference's own list representation, so that
 * tests/polisherTest.c:163-244 (hand-made tuples, golden POA) can be replayed against it.  Lists are
 * (weight, x, y) stIntTuples: matches, inserts (gapY), deletes (gapX), as poa_augment takes them (impl/poa.c:317).
 */
void margin_b200_binding_augment_lists(Poa *poa, RleString *read, bool readStrand, int64_t readNo, stList *matches,
                                       stList *inserts, stList *deletes, PolishParams *polishParams) {
    stList *lists[3] = {matches, deletes, inserts}; /* result order: match, gapX = delete, gapY = insert */
    margin_b200_result res;
    memset(&res, 0, sizeof(res));
    margin_b200_pair *pairs[3];
    int64_t off[3][2];
    for (int l = 0; l < 3; l++) {
        int64_t n = stList_length(lists[l]);
        pairs[l] = st_calloc(n + 1, sizeof(margin_b200_pair));
        for (int64_t k = 0; k < n; k++) {
            stIntTuple *t = stList_get(lists[l], k);
            pairs[l][k].w = (int32_t) stIntTuple_get(t, 0);
            pairs[l][k].x = (int32_t) stIntTuple_get(t, 1);
            pairs[l][k].y = (int32_t) stIntTuple_get(t, 2);
        }
        off[l][0] = 0;
        off[l][1] = n;
        res.pairs[l] = pairs[l];
        res.off[l] = off[l];
    }
    res.n = 1;
    res.pair_format = MARGIN_B200_PAIRS_TRIPLES;
    b200_indel_table_begin(0); /* the graph may already hold inserts and deletes: scan the lists as the reference does */
    b200_poa_augment(poa, read, readStrand, readNo, &res, 0, 0, polishParams);
    for (int l = 0; l < 3; l++) free(pairs[l]);
}

/* ---------------------------------------------------- impl/poa.c:1678-1727 + impl/repeatSubMatrix.c:65-151, flat */

/*
 * poa_estimateRepeatCountsUsingBayesianModel (SURVEY.md §8 f N4: the per-node run-length argmax, 5 % of a chunk once the
 * aligner is on the GPU).  The reference walks every node's observation list once for the range of observed run lengths
 * and once more per candidate length, through stList_get, the read list and repeatSubMatrix_getLogProb's index
 * arithmetic each time.  Here: one pass that resolves every observation to (table row, weight), then one pass that
 * updates all candidates per observation.  Every candidate's sum still receives its terms in list order, one multiply and
 * one add each (this file is compiled with the reference's own flags), so the log-probabilities, their maximum and the
 * consensus are bit for bit the reference's.
 */
typedef struct {
    const double *row; /* logProbabilities + strand-resolved base block + observed run length */
    double weight;
} B200RlTerm;

static __thread B200RlTerm *rlTerms;
static __thread int64_t rlTermsCap;

static int64_t b200_ml_repeat_count(Poa *poa, PoaNode *node, void **reads, RepeatSubMatrix *rsm) {
    int64_t M = rsm->maximumRepeatLength;
    Symbol base = poa->alphabet->convertCharToSymbol(node->base);
    if (base >= rsm->alphabet->alphabetSize - 1) { /* repeatSubMatrix_setLogProb (impl/repeatSubMatrix.c:16-27): N counts as A */
        if (base > rsm->alphabet->alphabetSize - 1) st_errAbort("margin_b200: base > 'Nn' (%d) not supported for repeat estimation!\n", (int) base);
        base = 0;
    }
    const double *blockF = rsm->logProbabilities + (int64_t) base * M * M, *blockR = rsm->logProbabilities + (int64_t) (3 - base) * M * M;
    int64_t n = stList_length(node->observations);
    void **obs = b200_list_array(node->observations);
    if (n > rlTermsCap) {
        rlTermsCap = n * 2 + 64;
        rlTerms = st_realloc(rlTerms, rlTermsCap * sizeof(B200RlTerm));
    }
    int64_t minRL = M, maxRL = 0;
    char *maxRLReadId = NULL;
    for (int64_t k = 0; k < n; k++) {
        PoaBaseObservation *o = obs[k];
        BamChunkRead *read = reads[o->readNo];
        int64_t observed = read->rleRead->repeatCounts[o->offset];
        if (observed < minRL) minRL = observed;
        if (observed > maxRL) {
            maxRL = observed;
            maxRLReadId = read->readName;
        }
        if (observed >= M) observed = M - 1; /* "Be robust to over-long repeat count observations" */
        rlTerms[k].row = (read->forwardStrand ? blockF : blockR) + observed;
        rlTerms[k].weight = o->weight;
    }
    if (maxRL >= M) {
        char *logIdentifier = getLogIdentifier();
        st_logInfo(" %s Got overlong repeat observation(s), max: %" PRIi64 " (%s), ignoring this and cutting off overlong repeat counts to max\n",
                   logIdentifier, maxRL, maxRLReadId);
        maxRL = M - 1;
        free(logIdentifier);
    }
    if (minRL == M) return 0; /* no valid observations: a repeat length of 0 (the caller turns it into 1) */
    if (minRL > maxRL) st_errAbort("margin_b200: every repeat count observation of a node is over-long\n");
    double logProb[MARGIN_B200_MAX_RL + 1];
    if (M > MARGIN_B200_MAX_RL) st_errAbort("margin_b200: repeat matrix larger than %d\n", MARGIN_B200_MAX_RL);
    int64_t range = maxRL - minRL + 1;
    for (int64_t u = 0; u < range; u++) logProb[u] = LOG_ONE;
    for (int64_t k = 0; k < n; k++) {
        const double *row = rlTerms[k].row + minRL * M; /* [underlying][observed]: one row of M entries per candidate */
        double w = rlTerms[k].weight;
        for (int64_t u = 0; u < range; u++) logProb[u] += row[u * M] * w;
    }
    int64_t best = 0;
    double bestValue = logProb[0] / PAIR_ALIGNMENT_PROB_1;
    for (int64_t u = 1; u < range; u++) { /* getMax (impl/repeatSubMatrix.c:153-167): the first of equal maxima */
        double v = logProb[u] / PAIR_ALIGNMENT_PROB_1;
        if (v > bestValue) {
            bestValue = v;
            best = u;
        }
    }
    return best + minRL;
}

/* getRunLengthMode (impl/poa.c:1678-1699), for parameter sets without a repeat matrix: the run length that first reaches
 * the highest count among the observations of the node's own base. */
static int64_t b200_run_length_mode(Poa *poa, PoaNode *node, void **reads) {
    Symbol base = poa->alphabet->convertCharToSymbol(node->base);
    int64_t n = stList_length(node->observations), nDistinct = 0, maxCount = 0, maxRL = 0;
    void **obs = b200_list_array(node->observations);
    int64_t values[64], counts[64];
    for (int64_t k = 0; k < n; k++) {
        PoaBaseObservation *o = obs[k];
        RleString *rleString = ((BamChunkRead *) reads[o->readNo])->rleRead;
        if (poa->alphabet->convertCharToSymbol(rleString->rleString[o->offset]) != base) continue;
        int64_t rl = rleString->repeatCounts[o->offset], at = 0;
        while (at < nDistinct && values[at] != rl) at++;
        if (at == nDistinct) {
            if (nDistinct == 64) st_errAbort("margin_b200: more than 64 distinct run lengths at one node\n");
            values[nDistinct] = rl;
            counts[nDistinct++] = 0;
        }
        if (++counts[at] > maxCount) {
            maxCount = counts[at];
            maxRL = rl;
        }
    }
    return maxRL;
}

void poa_estimateRepeatCountsUsingBayesianModel(Poa *poa, stList *bamChunkReads, RepeatSubMatrix *repeatSubMatrix) {
    double tStart = omp_get_wtime();
    poa->refString->nonRleLength = 0;
    int64_t nNodes = stList_length(poa->nodes);
    void **nodes = b200_list_array(poa->nodes), **reads = b200_list_array(bamChunkReads);
    for (int64_t i = 1; i < nNodes; i++) {
        PoaNode *node = nodes[i];
        int64_t rc = repeatSubMatrix == NULL ? b200_run_length_mode(poa, node, reads)
                                             : b200_ml_repeat_count(poa, node, reads, repeatSubMatrix);
        if (rc == 0) rc = 1; /* Prevent zero length estimates */
        poa->refString->repeatCounts[i - 1] = rc;
        node->repeatCount = rc;
        poa->refString->nonRleLength += rc;
    }
    st_logInfo(" margin_b200 repeat counts: %" PRIi64 " nodes in %.3f s\n", nNodes - 1, omp_get_wtime() - tStart);
}

/* ----------------------------------------------------------------------------- impl/poa.c:668-716, batched */

/*
 * poa_realign with ONE aligner call per chunk instead of one per read: pass 1 prepares every read exactly as
 * getAlignedPairsWithIndelsCroppingReference does (impl/poa.c:612-666: crop the reference to the anchored span,
 * shift the anchors, build both SymbolStrings, pick the strand's state machine), the batch runs on the GPU, pass 2
 * applies each read's three lists to the graph in read order (b200_poa_augment above: matches, inserts <- gapY,
 * deletes <- gapX, x shifted back by the crop).
 */
static int b200_referenceAugment = -1; /* -1 = not decided yet: MARGIN_B200_REFERENCE_AUGMENT=1 turns it on */

static int b200_reference_augment(void) {
    if (b200_referenceAugment < 0) {
        const char *e = getenv("MARGIN_B200_REFERENCE_AUGMENT");
        b200_referenceAugment = e != NULL && e[0] == '1';
    }
    return b200_referenceAugment;
}

/* A/B hook of the test suite (oracle/b200_tests.c): 1 = apply the lists with the reference's own poa_augment
 * (impl/poa.c:317-543) on stLists, 0 = with b200_poa_augment on the flat arrays. */
void margin_b200_binding_set_reference_augment(int on) { b200_referenceAugment = on != 0; }

static int b200_observationSlabs = -1; /* MARGIN_B200_OBSERVATION_SLABS=0: one calloc per observation, as the reference */

static int b200_observation_slabs(void) {
    if (b200_observationSlabs < 0) {
        const char *e = getenv("MARGIN_B200_OBSERVATION_SLABS");
        b200_observationSlabs = e == NULL || atoi(e) != 0;
    }
    return b200_observationSlabs;
}

void margin_b200_binding_set_observation_slabs(int on) { b200_observationSlabs = on != 0; }

Poa *poa_realign(stList *bamChunkReads, stList *anchorAlignments, RleString *reference, PolishParams *polishParams) {
    uint64_t maximumRepeatLength = 2; /* MRL is exclusive */
    if (polishParams->useRunLengthEncoding) {
        maximumRepeatLength = polishParams->repeatSubMatrix != NULL ? polishParams->repeatSubMatrix->maximumRepeatLength
                                                                    : MAXIMUM_REPEAT_LENGTH;
    }
    double tSetup = omp_get_wtime();
    Poa *poa = poa_getReferenceGraph(reference, polishParams->alphabet, maximumRepeatLength);
    int64_t n = stList_length(bamChunkReads);
    if (n == 0) return poa;

    b200_load_machine(0, polishParams->stateMachineForForwardStrandRead);
    b200_load_machine(1, polishParams->stateMachineForReverseStrandRead);
    b200_load_params(polishParams->p);

    double tStart = omp_get_wtime(), tLists = 0.0, tAugment = 0.0, tFree = 0.0;
    memset(augmentPhase, 0, sizeof(augmentPhase));
    B200Batch *b = b200_batch_construct(n);
    {
        /* totals of the chunk: read symbols exactly, reference symbols from the reads' spans (bounded by the anchors'
         * ends plus the read's overhang, or the whole reference without anchors), anchors exactly */
        int64_t totalX = 0, totalY = 0, totalA = 0;
        for (int64_t i = 0; i < n; i++) {
            BamChunkRead *chunkRead = stList_get(bamChunkReads, i);
            stList *anchorPairs = anchorAlignments == NULL ? NULL : stList_get(anchorAlignments, i);
            int64_t nA = anchorPairs == NULL ? 0 : stList_length(anchorPairs), span = reference->length;
            if (nA > 0) {
                stIntTuple *fPair = stList_get(anchorPairs, 0), *lPair = stList_peek(anchorPairs);
                span = stIntTuple_get(lPair, 0) - stIntTuple_get(fPair, 0) + 1 + chunkRead->rleRead->length -
                       (stIntTuple_get(lPair, 1) - stIntTuple_get(fPair, 1));
                if (span > (int64_t) reference->length) span = reference->length;
            }
            totalX += span;
            totalY += chunkRead->rleRead->length;
            totalA += nA;
        }
        b200_batch_reserve(b, totalX + 1024, totalY + 1024, totalA + 1024);
    }
    int64_t *firstRefPositions = st_calloc(n, sizeof(int64_t));
    for (int64_t i = 0; i < n; i++) {
        BamChunkRead *chunkRead = stList_get(bamChunkReads, i);
        RleString *read = chunkRead->rleRead;
        stList *anchorPairs = anchorAlignments == NULL ? NULL : stList_get(anchorAlignments, i);
        int64_t firstRefPosition = 0, endRefPosition = reference->length;
        uint64_t maxRL = poa->maxRepeatCount - 1; /* the anchor-less branch, impl/poa.c:688-691 */
        if (anchorAlignments != NULL) {
            if (stList_length(anchorPairs) > 0) { /* impl/poa.c:621-633 */
                stIntTuple *fPair = stList_get(anchorPairs, 0);
                firstRefPosition = stIntTuple_get(fPair, 0) - stIntTuple_get(fPair, 1);
                firstRefPosition = firstRefPosition < 0 ? 0 : firstRefPosition;
                stIntTuple *lPair = stList_peek(anchorPairs);
                endRefPosition = 1 + stIntTuple_get(lPair, 0) + (read->length - stIntTuple_get(lPair, 1));
                endRefPosition = endRefPosition > reference->length ? reference->length : endRefPosition;
            }
            maxRL = polishParams->useRunLengthEncoding ? (uint64_t) polishParams->repeatSubMatrix->maximumRepeatLength : 2;
        }
        SymbolString sX = rleString_constructSymbolString(reference, firstRefPosition, endRefPosition - firstRefPosition,
                                                          polishParams->alphabet,
                                                          polishParams->useRepeatCountsInAlignment, maxRL);
        SymbolString sY = rleString_constructSymbolString(read, 0, read->length, polishParams->alphabet,
                                                          polishParams->useRepeatCountsInAlignment, maxRL);
        b200_batch_add(b, sX, sY, anchorPairs, firstRefPosition, chunkRead->forwardStrand ? 0 : 1, 0, 0, tls.loaded[0].rle);
        symbolString_destruct(sX);
        symbolString_destruct(sY);
        firstRefPositions[i] = firstRefPosition;
    }

    double tPrepared = omp_get_wtime();
    margin_b200_result res;
    b200_batch_run(b, 1, &res);
    double tAligned = omp_get_wtime();

    b200_indel_table_begin(1); /* the graph is fresh: every insert and delete from here on goes through the table */
    slabState.owner = NULL;
    if (!b200_reference_augment() && b200_observation_slabs()) {
        double t0 = omp_get_wtime();
        b200_match_lists_presize(poa, &res, n, firstRefPositions);
        tLists += omp_get_wtime() - t0;
    }
    for (int64_t i = 0; i < n; i++) {
        BamChunkRead *chunkRead = stList_get(bamChunkReads, i);
        double t0 = omp_get_wtime();
        if (b200_reference_augment()) { /* MARGIN_B200_REFERENCE_AUGMENT=1: the reference's poa_augment on stLists (A/B check) */
            stList *matches = b200_list(&res, 0, i, firstRefPositions[i]);
            stList *deletes = b200_list(&res, 1, i, firstRefPositions[i]); /* gapX: reference base without read base */
            stList *inserts = b200_list(&res, 2, i, firstRefPositions[i]); /* gapY */
            double t1 = omp_get_wtime();
            poa_augment(poa, chunkRead->rleRead, chunkRead->forwardStrand, i, matches, inserts, deletes, polishParams);
            double t2 = omp_get_wtime();
            stList_destruct(matches);
            stList_destruct(inserts);
            stList_destruct(deletes);
            tLists += t1 - t0;
            tAugment += t2 - t1;
            tFree += omp_get_wtime() - t2;
        } else {
            b200_poa_augment(poa, chunkRead->rleRead, chunkRead->forwardStrand, i, &res, i, firstRefPositions[i], polishParams);
            tAugment += omp_get_wtime() - t0;
        }
    }
    double tCommit = 0.0;
    if (slabState.owner != NULL) { /* the recorded inserts and deletes, node by node */
        double t0 = omp_get_wtime();
        b200_candidates_commit(poa);
        tCommit = omp_get_wtime() - t0;
        tAugment += tCommit;
    }
    slabState.owner = NULL;
    free(firstRefPositions);
    b200_batch_destruct(b);
    st_logInfo(" margin_b200 poa_realign: %" PRIi64 " reads, setup (graph, context, model) %.3f s, prepare %.3f s, aligner (GPU, whole batch) %.3f s, "
               "tuple lists / list sizing %.3f s, poa_augment %.3f s (matches %.3f, match set %.3f, inserts %.3f, deletes %.3f, "
               "node-ordered commit %.3f: %" PRIi64 " insert candidates -> %" PRIi64 " inserts, %" PRIi64 " -> %" PRIi64 " deletes), free %.3f s\n",
               n, tStart - tSetup, tPrepared - tStart, tAligned - tPrepared, tLists, tAugment, augmentPhase[0], augmentPhase[1],
               augmentPhase[2], augmentPhase[3], tCommit, candidateStats[0], candidateStats[1], candidateStats[2], candidateStats[3],
               tFree);
    return poa;
}

/* ------------------------------------------------- impl/pairwiseAligner.c:849-905 and impl/poa.c:1852-1873, batched */

/*
 * Forward-only likelihoods.  bubbleGraph_constructFromPoaAndVCF asks for one small unbanded
 * forward matrix per (allele, read substring) pair — 444 k calls on the bundled region, median 11 x 11 — and only
 * stores each result as a float in Bubble.alleleReadSupports (impl/bubbleGraph.c:1067), copying it for repeated
 * read substrings (:1058-1061); nothing reads the values before the graph is returned.  The wrapper of
 * bubbleGraph_constructFromPoaAndVCF below uses that: while the reference's body builds the graph,
 * computeForwardProbability queues the job and returns a quiet NaN whose payload is the job number (the double ->
 * float conversion and the copies keep the payload bits); after the graph is built all jobs run as ONE
 * margin_b200_forward_batch and every tagged float is replaced by its job's result.  bubbleGraph.c is compiled
 * unchanged.  Everywhere else computeForwardProbability is a batch of one: the VCF-driven constructors of `margin
 * phase` (impl/bubbleGraph.c:1126, :1338) could be wrapped the same way, bubbleGraph_partitionFilteredReads (:1500)
 * reads each value at once (:1663) and cannot.
 */
#define B200_TAG_LOW_MAX 0x3fffffU                    /* 22 payload bits below the quiet bit of a float NaN */
#define B200_MAX_DEFERRED (2 * (int64_t) B200_TAG_LOW_MAX) /* the sign bit doubles the range; low bits 0 = a real NaN */
#define B200_FLUSH_SYMBOLS (64 << 20)                 /* run the queue early beyond this many queued symbols */

typedef struct B200Deferred {
    B200Batch *queue;          /* jobs not yet run */
    int64_t done;              /* jobs already run; results[0 .. done) are final */
    double *results;
    int64_t capResults;
    StateMachine *machines[2]; /* the machines of the queued jobs (slot = index) */
    int rle[2];                /* ... and whether their symbols carry run lengths */
    PairwiseAlignmentParameters params;
    int haveParams;
    int warned;
    /* record / replay (b200_partition wrapper below): lengths of every recorded job, to check that the replay asks the
     * same questions in the same order */
    int64_t *lengths;
    int64_t capLengths;
    int64_t cursor;
} B200Deferred;

static double b200_tagged_nan(int64_t job) {
    uint64_t low = (uint64_t) (job % B200_TAG_LOW_MAX) + 1, sign = (uint64_t) (job / B200_TAG_LOW_MAX);
    uint64_t bits = (sign << 63) | 0x7ff8000000000000ULL | (low << 29); /* float: sign | 0x7fc00000 | low */
    double d;
    memcpy(&d, &bits, sizeof(d));
    return d;
}

/* Job number of a tagged float, or -1 when f is an ordinary value (including an untagged NaN). */
static int64_t b200_tag_of(float f) {
    uint32_t u;
    memcpy(&u, &f, sizeof(u));
    if ((u & 0x7fc00000U) != 0x7fc00000U || (u & B200_TAG_LOW_MAX) == 0) return -1;
    return (int64_t) (u >> 31) * B200_TAG_LOW_MAX + (int64_t) (u & B200_TAG_LOW_MAX) - 1;
}

static __thread double forwardDeviceMs; /* device time of the forward batches of this thread (CUDA events) */

static void b200_forward_run(B200Batch *b, double *out) {
    margin_b200_batch mb;
    margin_b200_result stats;
    b200_batch_describe(b, &mb);
    if (margin_b200_forward_batch(b200_ctx(), &mb, out, &stats) != 0) b200_fail();
    forwardDeviceMs += stats.gpu_ms;
}

/* Runs the queued jobs; their results land behind the ones already known. */
static void b200_deferred_flush(B200Deferred *d) {
    int64_t n = d->queue->n;
    if (n > 0) {
        if (d->done + n > d->capResults) {
            d->capResults = (d->done + n) * 2;
            d->results = st_realloc(d->results, d->capResults * sizeof(double));
        }
        for (int s = 0; s < 2; s++) {
            if (d->machines[s] != NULL) b200_load_machine(s, d->machines[s]);
        }
        b200_load_params(&d->params);
        b200_forward_run(d->queue, d->results + d->done);
        d->done += n;
    }
    b200_batch_destruct(d->queue);
    d->queue = b200_batch_construct(1024);
    d->machines[0] = d->machines[1] = NULL;
}

static int b200_same_params(const PairwiseAlignmentParameters *a, const PairwiseAlignmentParameters *b) {
    return a->threshold == b->threshold && a->minDiagsBetweenTraceBack == b->minDiagsBetweenTraceBack &&
           a->traceBackDiagonals == b->traceBackDiagonals && a->diagonalExpansion == b->diagonalExpansion &&
           a->splitMatrixBiggerThanThis == b->splitMatrixBiggerThanThis &&
           a->dynamicAnchorExpansion == b->dynamicAnchorExpansion;
}

double computeForwardProbability(SymbolString sX, SymbolString sY, stList *anchorPairs, PairwiseAlignmentParameters *p,
                                 StateMachine *sM, bool alignmentHasRaggedLeftEnd, bool alignmentHasRaggedRightEnd) {
    if (sX.length + sY.length == 0) return LOG_ONE; /* impl/pairwiseAligner.c:859-862 */
    if (tls.replay != NULL) { /* second pass of a record / replay wrapper: the answers are already known */
        B200Deferred *r = tls.replay;
        if (r->cursor >= r->done || r->lengths[2 * r->cursor] != sX.length || r->lengths[2 * r->cursor + 1] != sY.length)
            st_errAbort("margin_b200: replayed forward job %" PRIi64 " does not match the recorded one\n", r->cursor);
        return r->results[r->cursor++];
    }
    B200Deferred *d = tls.deferred;
    if (d != NULL && d->capLengths == 0 && d->done + d->queue->n >= B200_MAX_DEFERRED) { /* tags run out (not when recording) */
        if (!d->warned) st_logInfo("margin_b200: more than %" PRIi64 " forward jobs in one bubble graph; the rest run one by one\n",
                                   (int64_t) B200_MAX_DEFERRED);
        d->warned = 1;
        d = NULL;
    }
    if (d == NULL) { /* a batch of one */
        b200_load_machine(0, sM);
        b200_load_params(p);
        B200Batch *b = b200_batch_construct(1);
        b200_batch_add(b, sX, sY, anchorPairs, 0, 0, alignmentHasRaggedLeftEnd, alignmentHasRaggedRightEnd, tls.loaded[0].rle);
        double out = LOG_ZERO;
        b200_forward_run(b, &out);
        b200_batch_destruct(b);
        return out;
    }
    if (d->haveParams && !b200_same_params(&d->params, p)) b200_deferred_flush(d);
    d->params = *p;
    d->haveParams = 1;
    int slot = d->machines[0] == sM ? 0 : (d->machines[1] == sM ? 1 : -1);
    if (slot < 0) {
        if (d->machines[0] != NULL && d->machines[1] != NULL) b200_deferred_flush(d); /* a third machine */
        slot = d->machines[0] == NULL ? 0 : 1;
        d->machines[slot] = sM;
        d->rle[slot] = b200_is_rle(sM);
    }
    tls.dynamicAnchors = p->dynamicAnchorExpansion;
    int64_t job = d->done + d->queue->n;
    if (d->lengths != NULL || d->capLengths < 0) { /* recording for a replay */
        if (2 * (job + 1) > d->capLengths) {
            d->capLengths = d->capLengths < 0 ? 4096 : d->capLengths * 2;
            d->lengths = st_realloc(d->lengths, d->capLengths * sizeof(int64_t));
        }
        d->lengths[2 * job] = sX.length;
        d->lengths[2 * job + 1] = sY.length;
    }
    b200_batch_add(d->queue, sX, sY, anchorPairs, 0, slot, alignmentHasRaggedLeftEnd, alignmentHasRaggedRightEnd, d->rle[slot]);
    if (d->queue->offX[d->queue->n] + d->queue->offY[d->queue->n] > B200_FLUSH_SYMBOLS) b200_deferred_flush(d);
    return b200_tagged_nan(job);
}

/*
 * bubbleGraph_constructFromPoaAndVCF (impl/bubbleGraph.c:918-1124) with its forward matrices run as one batch.  Both
 * callers come through here: poa_polish below (haploid consensus, via the reference's bubbleGraph_constructFromPoa
 * arguments: no VCF, phasing = FALSE) and the diploid path of polish.c:676 (phasing = TRUE, once per bubble-finding
 * iteration) — without this a diploid run would issue every matrix as a GPU batch of one.  The reference's body is
 * linked under the name bubbleGraph_constructFromPoaAndVCF_reference (oracle/Makefile renames the symbol in
 * bubbleGraph.o; INTEGRATION.md shows the one-line source patch) and is called unchanged.
 */
BubbleGraph *bubbleGraph_constructFromPoaAndVCF_reference(Poa *poa, stList *bamChunkReads, stList *vcfEntries,
                                                          PolishParams *params, bool phasing);

static int b200_bubbleBatching = -1;

/* A/B hook of the test suite (oracle/b200_tests.c): 0 = every forward matrix as a GPU batch of one. */
void margin_b200_binding_set_bubble_batching(int on) { b200_bubbleBatching = on != 0; }

static int b200_batching_on(void) {
    if (b200_bubbleBatching < 0) { /* MARGIN_B200_BUBBLE_BATCH=0: every matrix as a batch of one (A/B check only) */
        const char *e = getenv("MARGIN_B200_BUBBLE_BATCH");
        b200_bubbleBatching = !(e != NULL && e[0] == '0');
    }
    return b200_bubbleBatching && tls.deferred == NULL && tls.replay == NULL; /* nested calls do not occur in the reference */
}

static void b200_deferred_begin(B200Deferred *d) {
    memset(d, 0, sizeof(*d));
    d->queue = b200_batch_construct(1024);
    tls.deferred = d;
    forwardDeviceMs = 0.0;
}

/* Runs what is queued and replaces every tagged float of the graph by its job's result. */
static void b200_deferred_patch(B200Deferred *d, BubbleGraph *bg) {
    tls.deferred = NULL;
    b200_deferred_flush(d);
    for (uint64_t i = 0; bg != NULL && i < bg->bubbleNo; i++) {
        Bubble *b = &bg->bubbles[i];
        for (uint64_t k = 0; k < b->alleleNo * b->readNo; k++) {
            int64_t job = b200_tag_of(b->alleleReadSupports[k]);
            if (job < 0) continue;
            if (job >= d->done) st_errAbort("margin_b200: forward job %" PRIi64 " of %" PRIi64 " was never queued\n", job, d->done);
            b->alleleReadSupports[k] = (float) d->results[job]; /* the conversion of impl/bubbleGraph.c:1067 */
        }
    }
}

static void b200_deferred_end(B200Deferred *d) {
    b200_batch_destruct(d->queue);
    free(d->results);
    free(d->lengths);
}

BubbleGraph *bubbleGraph_constructFromPoaAndVCF(Poa *poa, stList *bamChunkReads, stList *vcfEntries, PolishParams *params,
                                                bool phasing) {
    if (!b200_batching_on())
        return bubbleGraph_constructFromPoaAndVCF_reference(poa, bamChunkReads, vcfEntries, params, phasing);
    B200Deferred d;
    b200_deferred_begin(&d);
    double tStart = omp_get_wtime();
    BubbleGraph *bg = bubbleGraph_constructFromPoaAndVCF_reference(poa, bamChunkReads, vcfEntries, params, phasing);
    double tBuilt = omp_get_wtime();
    b200_deferred_patch(&d, bg);
    double tScored = omp_get_wtime();
    st_logInfo(" margin_b200 bubble graph%s: %" PRIi64 " forward matrices, graph %.3f s, forward batch %.3f s "
               "(%.1f ms on the device)\n", phasing ? " (phasing)" : "", d.done, tBuilt - tStart, tScored - tBuilt,
               forwardDeviceMs);
    b200_deferred_end(&d);
    return bg;
}

/*
 * The VCF-driven constructors of `margin phase` / `margin polish --vcf` (impl/bubbleGraph.c:1126, :1338) store their
 * forward likelihoods in Bubble.alleleReadSupports exactly like the constructor above (:1232, :1453) and return the
 * graph: same deferral.  Linked under *_reference names like it (oracle/Makefile).
 */
BubbleGraph *bubbleGraph_constructFromPoaAndVCFOnlyVCFAllele_reference(Poa *poa, stList *bamChunkReads,
                                                                       RleString *referenceSeqRLE, stList *vcfEntries,
                                                                       Params *params);
BubbleGraph *bubbleGraph_constructFromVCFAndBamChunkReadVcfEntrySubstrings_reference(stList *bamChunkReads,
                                                                                     stList *vcfEntries, Params *params,
                                                                                     stList **vcfEntriesToBubbleIdx);

BubbleGraph *bubbleGraph_constructFromPoaAndVCFOnlyVCFAllele(Poa *poa, stList *bamChunkReads, RleString *referenceSeqRLE,
                                                             stList *vcfEntries, Params *params) {
    if (!b200_batching_on())
        return bubbleGraph_constructFromPoaAndVCFOnlyVCFAllele_reference(poa, bamChunkReads, referenceSeqRLE, vcfEntries, params);
    B200Deferred d;
    b200_deferred_begin(&d);
    BubbleGraph *bg = bubbleGraph_constructFromPoaAndVCFOnlyVCFAllele_reference(poa, bamChunkReads, referenceSeqRLE, vcfEntries, params);
    b200_deferred_patch(&d, bg);
    st_logInfo(" margin_b200 bubble graph (VCF alleles): %" PRIi64 " forward matrices in one batch\n", d.done);
    b200_deferred_end(&d);
    return bg;
}

BubbleGraph *bubbleGraph_constructFromVCFAndBamChunkReadVcfEntrySubstrings(stList *bamChunkReads, stList *vcfEntries,
                                                                           Params *params, stList **vcfEntriesToBubbleIdx) {
    if (!b200_batching_on())
        return bubbleGraph_constructFromVCFAndBamChunkReadVcfEntrySubstrings_reference(bamChunkReads, vcfEntries, params,
                                                                                       vcfEntriesToBubbleIdx);
    B200Deferred d;
    b200_deferred_begin(&d);
    BubbleGraph *bg = bubbleGraph_constructFromVCFAndBamChunkReadVcfEntrySubstrings_reference(bamChunkReads, vcfEntries, params,
                                                                                              vcfEntriesToBubbleIdx);
    b200_deferred_patch(&d, bg);
    st_logInfo(" margin_b200 bubble graph (VCF read substrings): %" PRIi64 " forward matrices in one batch\n", d.done);
    b200_deferred_end(&d);
    return bg;
}

/*
 * bubbleGraph_partitionFilteredReads (impl/bubbleGraph.c:1500-1745; the diploid path, polish.c:821) consumes every
 * likelihood in the iteration that asks for it (:1657-1670), so tagging does not work.  But which matrices it asks for
 * does not depend on the answers: the wrapper runs the reference's body TWICE.  The first pass records (every
 * computeForwardProbability queues its job; outputs go to throw-away sets, no file), the queue runs as one batch, and
 * the second pass is the real one, with computeForwardProbability handing the recorded answers back in order (each
 * checked against the recorded matrix size).  The price is the function's host-side bookkeeping twice; what it buys is
 * one GPU batch instead of one launch + synchronisation per 11 x 11 matrix.
 */
BubbleGraph *bubbleGraph_partitionFilteredReads_reference(Poa *poa, stList *bamChunkReads, stGenomeFragment *gF,
                                                          BubbleGraph *bg, BamChunk *bamChunk,
                                                          uint64_t *reference_rleToNonRleCoordMap, stSet *hap1Reads,
                                                          stSet *hap2Reads, PolishParams *params, FILE *out,
                                                          char *logIdentifier);

BubbleGraph *bubbleGraph_partitionFilteredReads(Poa *poa, stList *bamChunkReads, stGenomeFragment *gF, BubbleGraph *bg,
                                                BamChunk *bamChunk, uint64_t *reference_rleToNonRleCoordMap,
                                                stSet *hap1Reads, stSet *hap2Reads, PolishParams *params, FILE *out,
                                                char *logIdentifier) {
    if (!b200_batching_on())
        return bubbleGraph_partitionFilteredReads_reference(poa, bamChunkReads, gF, bg, bamChunk, reference_rleToNonRleCoordMap,
                                                            hap1Reads, hap2Reads, params, out, logIdentifier);
    B200Deferred d;
    b200_deferred_begin(&d);
    d.capLengths = -1; /* record the matrix sizes too */
    stSet *scratch1 = stSet_construct(), *scratch2 = stSet_construct();
    char *dryRunId = stString_print("%s (margin_b200 dry run: queueing forward matrices)", logIdentifier);
    bubbleGraph_partitionFilteredReads_reference(poa, bamChunkReads, gF, bg, bamChunk, reference_rleToNonRleCoordMap, scratch1,
                                                 scratch2, params, NULL, dryRunId);
    free(dryRunId);
    stSet_destruct(scratch1);
    stSet_destruct(scratch2);
    b200_deferred_patch(&d, NULL);
    tls.replay = &d;
    d.cursor = 0;
    BubbleGraph *r = bubbleGraph_partitionFilteredReads_reference(poa, bamChunkReads, gF, bg, bamChunk,
                                                                  reference_rleToNonRleCoordMap, hap1Reads, hap2Reads, params,
                                                                  out, logIdentifier);
    tls.replay = NULL;
    if (d.cursor != d.done) st_errAbort("margin_b200: replay used %" PRIi64 " of %" PRIi64 " recorded forward jobs\n", d.cursor, d.done);
    st_logInfo(" margin_b200 partition of filtered reads: %" PRIi64 " forward matrices in one batch\n", d.done);
    b200_deferred_end(&d);
    return r;
}

/* poa_polish (impl/poa.c:1852-1873): the reference's body, its bubble graph built through the batching wrapper above. */
RleString *poa_polish(Poa *poa, stList *bamChunkReads, PolishParams *params, int64_t **poaToConsensusMap) {
    BubbleGraph *bg = bubbleGraph_constructFromPoaAndVCF(poa, bamChunkReads, NULL, params, FALSE);

    uint64_t *consensusPath = bubbleGraph_getConsensusPath(bg, params);
    RleString *newConsensusString = bubbleGraph_getConsensusString(bg, consensusPath, poaToConsensusMap, params);
    free(consensusPath);
    bubbleGraph_destruct(bg);
    return newConsensusString;
}
