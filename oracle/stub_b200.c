/*
 * stub_b200.c — a CPU stand-in for libmargin_b200.so, built on the oracle (pairhmm_oracle.c).
 *
 * TEST INFRASTRUCTURE ONLY.  It exists so that the HOST side of the drop-in — compat/pairwiseAligner_b200.c: batch
 * assembly, list materialisation, b200_poa_augment, the deferred bubble-graph queue, the exception path — can be
 * exercised by the CPU test suite (`pytest -m "not gpu"`) in a container without a GPU: oracle/Makefile links it into
 * oracle/_ref/cutest_stub and oracle/_ref/margin_stubcli in place of the CUDA library.  Nothing under margin_b200/,
 * compat/ or bench.py's timed legs links or loads it; the product has no CPU path (margin_b200.cu fails without a
 * device).  It implements the subset of include/margin_b200.h the binding calls, with the same formats and errors.
 */
#include <math.h>
#include <omp.h>
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "../include/margin_b200.h"

typedef struct mb_oracle_lists {
    margin_b200_pair *pairs[3];
    int64_t count[3];
    int64_t cells, diagonals, tracebacks, max_width, blocks;
} mb_oracle_lists;

int mb_oracle_posteriors_batch(const margin_b200_model *models, const margin_b200_params *p, const margin_b200_batch *b,
                               int want_indels, mb_oracle_lists *res, int threads);
int mb_oracle_forward_batch(const margin_b200_model *models, const margin_b200_params *p, const margin_b200_batch *b,
                            double *out, int64_t *cells, int threads);
void mb_oracle_free_lists(mb_oracle_lists *l);

struct margin_b200_ctx {
    margin_b200_model models[2];
    int have_model[2];
    margin_b200_params params;
    int have_params;
    int pair_format;
    void *pairs[3];
    int64_t *off[3];
};

static __thread char g_err[1024];

static int fail(const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    return 1;
}

const char *margin_b200_last_error(void) { return g_err; }
int margin_b200_version(void) { return 100; }
int margin_b200_device_count(void) { return 1; }
int margin_b200_prewarm(void) { return 1; }
int margin_b200_prewarm_devices(int n) {
    (void) n;
    return 1;
}

void *margin_b200_pinned_alloc(size_t bytes) { return malloc(bytes ? bytes : 1); }
void margin_b200_pinned_free(void *p) { free(p); }
int margin_b200_set_precision(margin_b200_ctx *c, int precision) {
    (void) c;
    (void) precision;
    return 0;
}

int margin_b200_create(margin_b200_ctx **out, int device) {
    (void) device;
    *out = calloc(1, sizeof(margin_b200_ctx));
    return *out == NULL;
}

static void release(margin_b200_ctx *c) {
    for (int l = 0; l < 3; l++) {
        free(c->pairs[l]);
        free(c->off[l]);
        c->pairs[l] = NULL;
        c->off[l] = NULL;
    }
}

void margin_b200_destroy(margin_b200_ctx *c) {
    if (c == NULL) return;
    release(c);
    free(c);
}

int margin_b200_set_model(margin_b200_ctx *c, int slot, const margin_b200_model *m) {
    if (slot < 0 || slot > 1) return fail("margin_b200_set_model: slot must be 0 or 1");
    c->models[slot] = *m;
    c->have_model[slot] = 1;
    return 0;
}

int margin_b200_set_params(margin_b200_ctx *c, const margin_b200_params *p) {
    if (p->traceback_diagonals < 1) return fail("traceBackDiagonals must be >= 1");
    if (p->diagonal_expansion < 0 || p->diagonal_expansion % 2 != 0) return fail("diagonalExpansion must be even and >= 0");
    if (p->min_diags_between_traceback < 2) return fail("minDiagsBetweenTraceBack must be >= 2");
    if (p->traceback_diagonals + 1 >= p->min_diags_between_traceback)
        return fail("traceBackDiagonals + 1 must be < minDiagsBetweenTraceBack");
    c->params = *p;
    c->have_params = 1;
    return 0;
}

int margin_b200_set_kernel_timing(margin_b200_ctx *c, int on) {
    (void) c;
    (void) on;
    return 0;
}

int margin_b200_set_pair_format(margin_b200_ctx *c, int format) {
    c->pair_format = format;
    return 0;
}

static int check_batch(margin_b200_ctx *c, const margin_b200_batch *b) {
    if (!c->have_params) return fail("margin_b200_set_params has not been called");
    for (int64_t i = 0; i < b->n; i++) {
        int64_t lx = b->off_x[i + 1] - b->off_x[i], ly = b->off_y[i + 1] - b->off_y[i];
        int slot = b->model_slot ? b->model_slot[i] : 0;
        if (slot > 1 || !c->have_model[slot]) return fail("alignment %lld: model slot %d has not been set", (long long) i, slot);
        int64_t px = -1, py = -1;
        for (int64_t k = b->off_a[i]; k < b->off_a[i + 1]; k++) {
            int64_t x = b->anchors[3 * k], y = b->anchors[3 * k + 1], e = b->anchors[3 * k + 2];
            if (x <= px || y <= py || x < 0 || y < 0 || x >= lx || y >= ly || e < 0 || (e & 1))
                return fail("PAIRWISE_ALIGNMENT_EXCEPTION: alignment %lld: anchor %lld (%lld,%lld,%lld) is not strictly "
                            "increasing / in range / even", (long long) i, (long long) (k - b->off_a[i]), (long long) x,
                            (long long) y, (long long) e);
            px = x;
            py = y;
        }
        /* the compact forms the binding sends beside the plain arrays must expand to exactly those arrays */
        if (b->anchors_runs != NULL) {
            int64_t k = b->off_a[i];
            px = py = -1;
            for (int64_t r = b->off_r[i]; r < b->off_r[i + 1]; r++) {
                uint32_t w = b->anchors_runs[r];
                int64_t run = w & 255u, dx = (w >> 8) & 1023u, dy = (w >> 18) & 1023u, e = 2 * (int64_t) (w >> 28);
                for (int64_t t = 0; t <= run; t++, k++) {
                    px += t == 0 ? dx : 1;
                    py += t == 0 ? dy : 1;
                    if (k >= b->off_a[i + 1] || b->anchors[3 * k] != px || b->anchors[3 * k + 1] != py || b->anchors[3 * k + 2] != e)
                        return fail("alignment %lld: anchors_runs does not expand to the anchor triples at anchor %lld",
                                    (long long) i, (long long) (k - b->off_a[i]));
                }
            }
            if (k != b->off_a[i + 1]) return fail("alignment %lld: anchors_runs ends early", (long long) i);
        }
        if (b->packed_x != NULL && b->packed_y != NULL && b->sym_x != NULL && b->sym_y != NULL) {
            const int rle = c->models[slot].rle;
            for (int64_t k = b->off_x[i]; k < b->off_x[i + 1]; k++)
                if (b->packed_x[k] != margin_b200_pack_symbol(b->sym_x[k], rle))
                    return fail("alignment %lld: packed_x differs from sym_x at %lld", (long long) i, (long long) (k - b->off_x[i]));
            for (int64_t k = b->off_y[i]; k < b->off_y[i + 1]; k++)
                if (b->packed_y[k] != margin_b200_pack_symbol(b->sym_y[k], rle))
                    return fail("alignment %lld: packed_y differs from sym_y at %lld", (long long) i, (long long) (k - b->off_y[i]));
        }
    }
    return 0;
}

int margin_b200_posteriors_batch(margin_b200_ctx *c, const margin_b200_batch *b, int want_indels, margin_b200_result *out) {
    if (check_batch(c, b)) return 1;
    memset(out, 0, sizeof(*out));
    release(c);
    int64_t n = b->n;
    mb_oracle_lists *res = calloc(n > 0 ? n : 1, sizeof(mb_oracle_lists));
    mb_oracle_posteriors_batch(c->models, &c->params, b, want_indels, res, omp_get_num_procs());
    int packed = c->pair_format == MARGIN_B200_PAIRS_PACKED64;
    for (int64_t i = 0; i < n && packed; i++)
        if (b->off_x[i + 1] - b->off_x[i] >= (1 << 20) - 1 || b->off_y[i + 1] - b->off_y[i] >= (1 << 20) - 1) packed = 0;
    for (int l = 0; l < 3; l++) {
        c->off[l] = calloc(n + 1, sizeof(int64_t));
        for (int64_t i = 0; i < n; i++) c->off[l][i + 1] = c->off[l][i] + res[i].count[l];
        int64_t total = c->off[l][n];
        c->pairs[l] = malloc((total + 1) * (packed ? sizeof(uint64_t) : sizeof(margin_b200_pair)));
        for (int64_t i = 0; i < n; i++) {
            for (int64_t k = 0; k < res[i].count[l]; k++) {
                margin_b200_pair q = res[i].pairs[l][k];
                if (packed)
                    ((uint64_t *) c->pairs[l])[c->off[l][i] + k] =
                            (uint64_t) (uint32_t) q.w | ((uint64_t) (uint32_t) (q.x + 1) << 24) | ((uint64_t) (uint32_t) (q.y + 1) << 44);
                else
                    ((margin_b200_pair *) c->pairs[l])[c->off[l][i] + k] = q;
            }
        }
        out->pairs[l] = c->pairs[l];
        out->off[l] = c->off[l];
    }
    for (int64_t i = 0; i < n; i++) {
        out->band_cells += res[i].cells;
        out->diagonals += res[i].diagonals;
        out->segments += res[i].tracebacks;
        out->jobs += res[i].blocks;
        mb_oracle_free_lists(&res[i]);
    }
    free(res);
    out->n = n;
    out->pair_format = packed ? MARGIN_B200_PAIRS_PACKED64 : MARGIN_B200_PAIRS_TRIPLES;
    return 0;
}

int margin_b200_forward_batch(margin_b200_ctx *c, const margin_b200_batch *b, double *out_log_prob, margin_b200_result *stats) {
    if (check_batch(c, b)) return 1;
    int64_t cells = 0;
    mb_oracle_forward_batch(c->models, &c->params, b, out_log_prob, &cells, omp_get_num_procs());
    if (stats) {
        memset(stats, 0, sizeof(*stats));
        stats->n = b->n;
        stats->band_cells = cells;
    }
    return 0;
}
