/*
 * ref_harness.c — flat-array doorway into the UNMODIFIED reference aligner.
 *
 * TEST INFRASTRUCTURE ONLY (see oracle/pairhmm_oracle.c).  This file is linked
 * with the reference's own sources, compiled where they lie under
 * /root/reference by oracle/Makefile, into oracle/_ref/libmargin_ref.so.  It
 * contains no aligner logic of its own: it converts flat arrays to the
 * reference's stList / SymbolString / StateMachine objects, calls the
 * reference's public functions (inc/pairwiseAligner.h) and flattens the
 * results, so that tests can pin oracle/pairhmm_oracle.c and the CUDA path
 * against the real thing, and bench.py --impl reference can time it.
 */
#include "margin.h"
#include "../include/margin_b200.h"

/* Private layouts re-declared from the reference so a StateMachine can be
 * flattened.  impl/stateMachine.c:507-519 and :716-720. */
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
} RefStateMachine3;

typedef struct {
    NucleotideEmissions ne;
    RepeatSubMatrix *repeatSubMatrix;
    bool strand;
} RefRleNucleotideEmissions;

/* ---------------------------------------------------------------- params */

void *ref_params_load(const char *path) {
    return params_readParams((char *) path);
}

void *ref_params_state_machine(void *params, int slot) {
    PolishParams *pp = ((Params *) params)->polishParams;
    return slot == 0 ? pp->stateMachineForForwardStrandRead : pp->stateMachineForReverseStrandRead;
}

void ref_params_alignment(void *params, margin_b200_params *out, int *use_rle, int *use_repeat_counts, int *max_rl) {
    PolishParams *pp = ((Params *) params)->polishParams;
    out->threshold = pp->p->threshold;
    out->min_diags_between_traceback = pp->p->minDiagsBetweenTraceBack;
    out->traceback_diagonals = pp->p->traceBackDiagonals;
    out->diagonal_expansion = pp->p->diagonalExpansion;
    out->split_matrix_bigger_than_this = pp->p->splitMatrixBiggerThanThis;
    out->dynamic_anchor_expansion = pp->p->dynamicAnchorExpansion;
    out->reserved = 0;
    *use_rle = pp->useRunLengthEncoding;
    *use_repeat_counts = pp->useRepeatCountsInAlignment;
    *max_rl = pp->repeatSubMatrix ? (int) pp->repeatSubMatrix->maximumRepeatLength : 2;
}

void *ref_default_nucleotide_state_machine(void) {
    return stateMachine3_constructNucleotide(threeState);
}

/* Flatten a StateMachine3 (+ emissions) into the POD model.  `rle` tells which
 * emission struct hangs off it (the reference offers no way to ask). */
int ref_flatten_state_machine(void *sm, int rle, margin_b200_model *m) {
    RefStateMachine3 *s = (RefStateMachine3 *) sm;
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
    NucleotideEmissions *ne = (NucleotideEmissions *) s->model.emissions;
    memcpy(m->match, ne->EMISSION_MATCH_PROBS, sizeof(m->match));
    memcpy(m->gap_x, ne->EMISSION_GAP_X_PROBS, sizeof(m->gap_x));
    memcpy(m->gap_y, ne->EMISSION_GAP_Y_PROBS, sizeof(m->gap_y));
    m->rle = rle;
    m->max_rl = MARGIN_B200_MAX_RL;
    if (rle) {
        RefRleNucleotideEmissions *re = (RefRleNucleotideEmissions *) s->model.emissions;
        RepeatSubMatrix *r = re->repeatSubMatrix;
        if (r->maximumRepeatLength != MARGIN_B200_MAX_RL) return 1;
        int64_t sq = r->maximumRepeatLength * r->maximumRepeatLength;
        for (int64_t b = 0; b < 4; b++) {
            memcpy(m->rsm + b * sq, r->logProbabilities + (re->strand ? b : 3 - b) * sq, sizeof(double) * sq);
        }
    }
    return 0;
}

/* Raw probes of the reference's emission callbacks, to validate a flattened model. */
double ref_emission_match(void *sm, int cx, int cy) {
    StateMachine *s = (StateMachine *) sm;
    return s->emissions->emission(s->emissions, (Symbol) cx, (Symbol) cy);
}
double ref_emission_gap_x(void *sm, int cx) {
    StateMachine *s = (StateMachine *) sm;
    return s->emissions->gapEmissionX(s->emissions, (Symbol) cx);
}
double ref_emission_gap_y(void *sm, int cy) {
    StateMachine *s = (StateMachine *) sm;
    return s->emissions->gapEmissionY(s->emissions, (Symbol) cy);
}

/* ----------------------------------------------------------- conversions */

static PairwiseAlignmentParameters *to_ref_params(const margin_b200_params *p) {
    PairwiseAlignmentParameters *q = pairwiseAlignmentBandingParameters_construct();
    q->threshold = p->threshold;
    q->minDiagsBetweenTraceBack = p->min_diags_between_traceback;
    q->traceBackDiagonals = p->traceback_diagonals;
    q->diagonalExpansion = p->diagonal_expansion;
    q->splitMatrixBiggerThanThis = p->split_matrix_bigger_than_this;
    q->dynamicAnchorExpansion = p->dynamic_anchor_expansion;
    return q;
}

static stList *to_anchor_list(const int32_t *anchors, int64_t na) {
    stList *l = stList_construct3(0, (void (*)(void *)) stIntTuple_destruct);
    for (int64_t i = 0; i < na; i++) {
        stList_append(l, stIntTuple_construct3(anchors[3 * i], anchors[3 * i + 1], anchors[3 * i + 2]));
    }
    return l;
}

static SymbolString to_symbol_string(const uint16_t *s, int64_t l) {
    SymbolString ss;
    ss.alphabet = NULL;
    ss.sequence = (Symbol *) s;
    ss.length = l;
    return ss;
}

static margin_b200_pair *flatten_list(stList *l, int64_t *n) {
    *n = stList_length(l);
    margin_b200_pair *v = malloc(sizeof(margin_b200_pair) * (size_t) (*n > 0 ? *n : 1));
    for (int64_t i = 0; i < *n; i++) {
        stIntTuple *t = stList_get(l, i);
        v[i].w = (int32_t) stIntTuple_get(t, 0);
        v[i].x = (int32_t) stIntTuple_get(t, 1);
        v[i].y = (int32_t) stIntTuple_get(t, 2);
    }
    return v;
}

/* ------------------------------------------------------------- functions */

double ref_logadd(double x, double y) {
    return logAdd(x, y);
}

void ref_band(const int32_t *anchors, int64_t na, int64_t lx, int64_t ly, int dynamic, int64_t expansion, int64_t *kl,
              int64_t *kr) {
    stList *a = to_anchor_list(anchors, na);
    Band *band = dynamic ? band_constructDynamic(a, lx, ly) : band_construct(a, lx, ly, expansion);
    BandIterator *it = bandIterator_construct(band);
    for (int64_t d = 0; d <= lx + ly; d++) {
        Diagonal dg = bandIterator_getNext(it);
        kl[d] = diagonal_getMinXmy(dg);
        kr[d] = diagonal_getMaxXmy(dg);
    }
    bandIterator_destruct(it);
    band_destruct(band);
    stList_destruct(a);
}

int64_t ref_split_points(const int32_t *anchors, int64_t na, int64_t lx, int64_t ly, int64_t limit, int ragged_left,
                         int ragged_right, int64_t *sp) {
    stList *a = to_anchor_list(anchors, na);
    stList *s = getSplitPoints(a, lx, ly, limit, ragged_left, ragged_right);
    int64_t n = stList_length(s);
    for (int64_t i = 0; i < n; i++) {
        for (int j = 0; j < 4; j++) sp[4 * i + j] = stIntTuple_get(stList_get(s, i), j);
    }
    stList_destruct(s);
    stList_destruct(a);
    return n;
}

typedef struct {
    margin_b200_pair *pairs[3];
    int64_t count[3];
    int64_t cells, diagonals, tracebacks, max_width, blocks;
} ref_lists;

void ref_free_lists(ref_lists *l) {
    for (int i = 0; i < 3; i++) {
        free(l->pairs[i]);
        l->pairs[i] = NULL;
        l->count[i] = 0;
    }
}

int ref_posteriors(void *sm, const margin_b200_params *p, const uint16_t *sx, int64_t lx, const uint16_t *sy,
                   int64_t ly, const int32_t *anchors, int64_t na, int ragged_left, int ragged_right, int want_indels,
                   ref_lists *res) {
    memset(res, 0, sizeof(*res));
    PairwiseAlignmentParameters *q = to_ref_params(p);
    stList *a = to_anchor_list(anchors, na);
    SymbolString ssx = to_symbol_string(sx, lx), ssy = to_symbol_string(sy, ly);
    if (want_indels) {
        stList *l[3];
        getAlignedPairsWithIndelsUsingAnchors((StateMachine *) sm, ssx, ssy, a, q, &l[0], &l[1], &l[2], ragged_left,
                                              ragged_right);
        for (int i = 0; i < 3; i++) {
            res->pairs[i] = flatten_list(l[i], &res->count[i]);
            stList_destruct(l[i]);
        }
    } else {
        stList *l = getAlignedPairsUsingAnchors((StateMachine *) sm, ssx, ssy, a, q, ragged_left, ragged_right);
        res->pairs[0] = flatten_list(l, &res->count[0]);
        stList_destruct(l);
    }
    stList_destruct(a);
    pairwiseAlignmentBandingParameters_destruct(q);
    return 0;
}

double ref_forward(void *sm, const margin_b200_params *p, const uint16_t *sx, int64_t lx, const uint16_t *sy,
                   int64_t ly, const int32_t *anchors, int64_t na, int ragged_left, int ragged_right) {
    PairwiseAlignmentParameters *q = to_ref_params(p);
    stList *a = to_anchor_list(anchors, na);
    double r = computeForwardProbability(to_symbol_string(sx, lx), to_symbol_string(sy, ly), a, q, (StateMachine *) sm,
                                         ragged_left, ragged_right);
    stList_destruct(a);
    pairwiseAlignmentBandingParameters_destruct(q);
    return r;
}

/* Batch drivers (OpenMP over alignments, the reference's own threading unit is
 * coarser — one chunk per thread, polish.c:476 — but the work per alignment is
 * the same code).  sms[2] = forward / reverse strand machines.  Only the tuple
 * counts are kept: this is the timed reference arm of bench.py. */
int ref_posteriors_batch(void **sms, const margin_b200_params *p, const margin_b200_batch *b, int want_indels,
                         int64_t *counts3, int threads) {
    int64_t c0 = 0, c1 = 0, c2 = 0;
#pragma omp parallel for schedule(dynamic, 1) num_threads(threads > 0 ? threads : 1) reduction(+ : c0, c1, c2)
    for (int64_t i = 0; i < b->n; i++) {
        ref_lists r;
        ref_posteriors(sms[b->model_slot ? b->model_slot[i] : 0], p, b->sym_x + b->off_x[i],
                       b->off_x[i + 1] - b->off_x[i], b->sym_y + b->off_y[i], b->off_y[i + 1] - b->off_y[i],
                       b->anchors + 3 * b->off_a[i], b->off_a[i + 1] - b->off_a[i],
                       b->ragged_left ? b->ragged_left[i] : 0, b->ragged_right ? b->ragged_right[i] : 0, want_indels,
                       &r);
        c0 += r.count[0];
        c1 += r.count[1];
        c2 += r.count[2];
        ref_free_lists(&r);
    }
    counts3[0] = c0;
    counts3[1] = c1;
    counts3[2] = c2;
    return 0;
}

int ref_forward_batch(void **sms, const margin_b200_params *p, const margin_b200_batch *b, double *out, int threads) {
#pragma omp parallel for schedule(dynamic, 64) num_threads(threads > 0 ? threads : 1)
    for (int64_t i = 0; i < b->n; i++) {
        out[i] = ref_forward(sms[b->model_slot ? b->model_slot[i] : 0], p, b->sym_x + b->off_x[i],
                             b->off_x[i + 1] - b->off_x[i], b->sym_y + b->off_y[i], b->off_y[i + 1] - b->off_y[i],
                             b->anchors ? b->anchors + 3 * b->off_a[i] : NULL,
                             b->off_a ? b->off_a[i + 1] - b->off_a[i] : 0, b->ragged_left ? b->ragged_left[i] : 0,
                             b->ragged_right ? b->ragged_right[i] : 0);
    }
    return 0;
}

/* Build a reference StateMachine (+ emissions, + RepeatSubMatrix) from a flattened model, so the
 * reference can run where the params JSON is not available (the GPU box has no /root/reference).
 * The model's rsm is already strand-resolved, so the emissions are constructed with strand = 1
 * (identity mapping, impl/repeatSubMatrix.c:28).  Only reference constructors are used. */
void *ref_state_machine_from_model(const margin_b200_model *m) {
    StateMachine *sm = stateMachine3_construct(threeState, nucleotideEmissions_construct());
    RefStateMachine3 *s = (RefStateMachine3 *) sm;
    s->TRANSITION_MATCH_CONTINUE = m->t[MB_T_MATCH_CONTINUE];
    s->TRANSITION_MATCH_FROM_GAP_X = m->t[MB_T_MATCH_FROM_GAP_X];
    s->TRANSITION_MATCH_FROM_GAP_Y = m->t[MB_T_MATCH_FROM_GAP_Y];
    s->TRANSITION_GAP_OPEN_X = m->t[MB_T_GAP_OPEN_X];
    s->TRANSITION_GAP_EXTEND_X = m->t[MB_T_GAP_EXTEND_X];
    s->TRANSITION_GAP_SWITCH_TO_X = m->t[MB_T_GAP_SWITCH_TO_X];
    s->TRANSITION_GAP_OPEN_Y = m->t[MB_T_GAP_OPEN_Y];
    s->TRANSITION_GAP_EXTEND_Y = m->t[MB_T_GAP_EXTEND_Y];
    s->TRANSITION_GAP_SWITCH_TO_Y = m->t[MB_T_GAP_SWITCH_TO_Y];
    NucleotideEmissions *ne = (NucleotideEmissions *) sm->emissions;
    memcpy(ne->EMISSION_MATCH_PROBS, m->match, sizeof(m->match));
    memcpy(ne->EMISSION_GAP_X_PROBS, m->gap_x, sizeof(m->gap_x));
    memcpy(ne->EMISSION_GAP_Y_PROBS, m->gap_y, sizeof(m->gap_y));
    if (m->rle) {
        RepeatSubMatrix *r = repeatSubMatrix_constructEmpty(alphabet_constructNucleotide());
        int64_t sq = r->maximumRepeatLength * r->maximumRepeatLength;
        memcpy(r->logProbabilities, m->rsm, sizeof(double) * 4 * sq);
        sm->emissions = rleNucleotideEmissions_construct(sm->emissions, r, 1);
    }
    return sm;
}
