/*
 * b200_tests.c — CuTest suite of the margin-side binding (compat/pairwiseAligner_b200.c), linked into
 * oracle/_ref/cutest_b200 next to the reference's own suites.  TEST INFRASTRUCTURE.
 *
 *   test_b200_augment_example      tests/polisherTest.c:163-244 (test_poa_augment_example) restated for b200_poa_augment:
 *                                  the same hand-made tuple lists must give the same golden POA.
 *   test_b200_augment_ab           b200_poa_augment against the reference's poa_augment (impl/poa.c:317-543) on the
 *                                  lists the aligner produces for random reads, without and with anchors: node
 *                                  weights, observations (order included), inserts and deletes must be identical.
 *   test_b200_realign_tiny_example tests/polisherTest.c:246-331 known answer (insert 0.990160, deletes 1.045831 /
 *                                  0.929195, 1e-3) through the batched poa_realign.
 *   test_b200_exception_path       bad anchors raise PAIRWISE_ALIGNMENT_EXCEPTION_ID through stThrowNew (catchable,
 *                                  tests/pairwiseAlignerTest.c:27-35 style) and the context stays usable.
 *   test_b200_bubble_batch_ab      poa_polish with the forward matrices deferred into one batch against the same
 *                                  matrices issued one by one (MARGIN_B200_BUBBLE_BATCH): identical consensus.
 *   test_b200_repeat_counts_ab     the flat poa_estimateRepeatCountsUsingBayesianModel against the reference's
 *                                  repeatSubMatrix_getMLRepeatCount, node by node.
 */
#include "CuTest.h"
#include "margin.h"

void margin_b200_binding_set_reference_augment(int on);
void margin_b200_binding_set_bubble_batching(int on);
void margin_b200_binding_set_observation_slabs(int on);
void margin_b200_binding_augment_lists(Poa *poa, RleString *read, bool readStrand, int64_t readNo, stList *matches,
                                       stList *inserts, stList *deletes, PolishParams *polishParams);

static char *b200ParamsFile = "../params/polish/ont/r9.4/allParams.np.human.r94-g344.json";

static void check_node(CuTest *testCase, Poa *poa, int64_t nodeIndex, char base, const double *baseWeights,
                       int64_t insertNumber, const char **inserts, const double *insertWeights, int64_t deleteNumber,
                       const int64_t *deleteLengths, const double *deleteWeights) {
    PoaNode *node = stList_get(poa->nodes, nodeIndex);
    CuAssertTrue(testCase, node->base == base);
    for (int64_t i = 0; i < poa->alphabet->alphabetSize; i++) CuAssertDblEquals(testCase, baseWeights[i], node->baseWeights[i], 0.0);
    CuAssertIntEquals(testCase, insertNumber, stList_length(node->inserts));
    for (int64_t i = 0; i < insertNumber; i++) {
        PoaInsert *ins = stList_get(node->inserts, i);
        char *expanded = rleString_expand(ins->insert);
        CuAssertStrEquals(testCase, inserts[i], expanded);
        free(expanded);
        CuAssertDblEquals(testCase, insertWeights[i], ins->weightForwardStrand + ins->weightReverseStrand, 0.0);
    }
    CuAssertIntEquals(testCase, deleteNumber, stList_length(node->deletes));
    for (int64_t i = 0; i < deleteNumber; i++) {
        PoaDelete *del = stList_get(node->deletes, i);
        CuAssertIntEquals(testCase, deleteLengths[i], del->length);
        CuAssertDblEquals(testCase, deleteWeights[i], del->weightForwardStrand + del->weightReverseStrand, 0.0);
    }
}

static void test_b200_augment_example(CuTest *testCase) {
    RleString *reference = rleString_construct_no_rle("GATTACA");
    Params *p = params_readParams(b200ParamsFile);
    Alphabet *alphabet = alphabet_constructNucleotide();
    Poa *poa = poa_getReferenceGraph(reference, alphabet, p->polishParams->repeatSubMatrix->maximumRepeatLength);
    RleString *read = rleString_construct_no_rle("GATACGGT");
    stList *matches = stList_construct3(0, (void (*)(void *)) stIntTuple_destruct);
    stList *inserts = stList_construct3(0, (void (*)(void *)) stIntTuple_destruct);
    stList *deletes = stList_construct3(0, (void (*)(void *)) stIntTuple_destruct);
    const int64_t m[9][3] = {{100, 0, 0}, {100, 1, 1}, {50, 2, 2}, {50, 3, 2}, {100, 4, 3}, {100, 5, 4}, {50, 6, 5}, {25, 6, 6}, {25, 6, 7}};
    const int64_t in[4][3] = {{50, 5, 5}, {25, 5, 6}, {50, 6, 6}, {75, 6, 7}};
    const int64_t de[2][3] = {{50, 2, 1}, {50, 3, 2}};
    for (int i = 0; i < 9; i++) stList_append(matches, stIntTuple_construct3(m[i][0], m[i][1], m[i][2]));
    for (int i = 0; i < 4; i++) stList_append(inserts, stIntTuple_construct3(in[i][0], in[i][1], in[i][2]));
    for (int i = 0; i < 2; i++) stList_append(deletes, stIntTuple_construct3(de[i][0], de[i][1], de[i][2]));

    margin_b200_binding_augment_lists(poa, read, 1, 0, matches, inserts, deletes, p->polishParams);

    CuAssertTrue(testCase, stList_length(poa->nodes) == 8);
    check_node(testCase, poa, 0, 'N', (const double[]) {0, 0, 0, 0, 0}, 0, NULL, NULL, 0, NULL, NULL);
    check_node(testCase, poa, 1, 'G', (const double[]) {0, 0, 100, 0, 0}, 0, NULL, NULL, 0, NULL, NULL);
    check_node(testCase, poa, 2, 'A', (const double[]) {100, 0, 0, 0, 0}, 0, NULL, NULL, 1, (const int64_t[]) {1}, (const double[]) {100});
    check_node(testCase, poa, 3, 'T', (const double[]) {0, 0, 0, 50, 0}, 0, NULL, NULL, 0, NULL, NULL);
    check_node(testCase, poa, 4, 'T', (const double[]) {0, 0, 0, 50, 0}, 0, NULL, NULL, 0, NULL, NULL);
    check_node(testCase, poa, 5, 'A', (const double[]) {100, 0, 0, 0, 0}, 0, NULL, NULL, 0, NULL, NULL);
    check_node(testCase, poa, 6, 'C', (const double[]) {0, 100, 0, 0, 0}, 2, (const char *[]) {"G", "GG"}, (const double[]) {50, 25}, 0, NULL, NULL);
    check_node(testCase, poa, 7, 'A', (const double[]) {0, 0, 75, 25, 0}, 2, (const char *[]) {"GT", "T"}, (const double[]) {50, 75}, 0, NULL, NULL);

    poa_destruct(poa);
    stList_destruct(matches);
    stList_destruct(inserts);
    stList_destruct(deletes);
    rleString_destruct(read);
    rleString_destruct(reference);
    params_destruct(p);
}

static void assert_observations_equal(CuTest *testCase, stList *a, stList *b) {
    CuAssertIntEquals(testCase, stList_length(a), stList_length(b));
    for (int64_t i = 0; i < stList_length(a); i++) {
        PoaBaseObservation *x = stList_get(a, i), *y = stList_get(b, i);
        CuAssertTrue(testCase, x->readNo == y->readNo && x->offset == y->offset && x->weight == y->weight);
    }
}

static void assert_poas_equal(CuTest *testCase, Poa *a, Poa *b) {
    CuAssertIntEquals(testCase, stList_length(a->nodes), stList_length(b->nodes));
    for (int64_t n = 0; n < stList_length(a->nodes); n++) {
        PoaNode *x = stList_get(a->nodes, n), *y = stList_get(b->nodes, n);
        CuAssertTrue(testCase, x->base == y->base && x->repeatCount == y->repeatCount);
        for (int64_t i = 0; i < a->alphabet->alphabetSize; i++) CuAssertTrue(testCase, x->baseWeights[i] == y->baseWeights[i]);
        for (int64_t i = 0; i < (int64_t) a->maxRepeatCount; i++) CuAssertTrue(testCase, x->repeatCountWeights[i] == y->repeatCountWeights[i]);
        assert_observations_equal(testCase, x->observations, y->observations);
        CuAssertIntEquals(testCase, stList_length(x->inserts), stList_length(y->inserts));
        for (int64_t i = 0; i < stList_length(x->inserts); i++) {
            PoaInsert *p = stList_get(x->inserts, i), *q = stList_get(y->inserts, i);
            CuAssertTrue(testCase, rleString_eq(p->insert, q->insert));
            CuAssertTrue(testCase, p->weightForwardStrand == q->weightForwardStrand && p->weightReverseStrand == q->weightReverseStrand);
            assert_observations_equal(testCase, p->observations, q->observations);
        }
        CuAssertIntEquals(testCase, stList_length(x->deletes), stList_length(y->deletes));
        for (int64_t i = 0; i < stList_length(x->deletes); i++) {
            PoaDelete *p = stList_get(x->deletes, i), *q = stList_get(y->deletes, i);
            CuAssertTrue(testCase, p->length == q->length);
            CuAssertTrue(testCase, p->weightForwardStrand == q->weightForwardStrand && p->weightReverseStrand == q->weightReverseStrand);
            assert_observations_equal(testCase, p->observations, q->observations);
        }
    }
}

/* Deterministic reads: own LCG instead of st_random, so that the A/B runs and the CPU/GPU builds see the same data. */
static uint64_t b200_rngState = 0x9E3779B97F4A7C15ULL;
static uint32_t b200_rng(void) {
    b200_rngState = b200_rngState * 6364136223846793005ULL + 1442695040888963407ULL;
    return (uint32_t) (b200_rngState >> 33);
}

static char *b200_random_sequence(int64_t length) {
    char *s = st_malloc(length + 1);
    for (int64_t i = 0; i < length; i++) {
        /* homopolymer runs, so that run-length encoding has something to do */
        s[i] = (i > 0 && b200_rng() % 100 < 35) ? s[i - 1] : "ACGT"[b200_rng() % 4];
    }
    s[length] = '\0';
    return s;
}

static char *b200_mutate(const char *seq, int subPct, int indelPct) {
    int64_t n = strlen(seq);
    char *out = st_malloc(2 * n + 16);
    int64_t k = 0;
    for (int64_t i = 0; i < n; i++) {
        int r = (int) (b200_rng() % 100);
        if (r < indelPct) continue;                                /* deletion */
        if (r < 2 * indelPct) out[k++] = "ACGT"[b200_rng() % 4];    /* insertion */
        out[k++] = (int) (b200_rng() % 100) < subPct ? "ACGT"[b200_rng() % 4] : seq[i];
    }
    if (k == 0) out[k++] = 'A';
    out[k] = '\0';
    return out;
}

static void test_b200_augment_ab(CuTest *testCase) {
    Params *params = params_readParams(b200ParamsFile);
    PolishParams *polishParams = params->polishParams;
    for (int64_t test = 0; test < 12; test++) {
        char *truth = b200_random_sequence(200 + (int64_t) (b200_rng() % 1200));
        char *draft = b200_mutate(truth, 1, 1);
        RleString *reference = polishParams->useRunLengthEncoding ? rleString_construct(draft) : rleString_construct_no_rle(draft);
        int64_t readNumber = 3 + (int64_t) (b200_rng() % 14);
        stList *reads = stList_construct3(0, (void (*)(void *)) bamChunkRead_destruct);
        for (int64_t i = 0; i < readNumber; i++) {
            char *r = b200_mutate(truth, 4, 4);
            stList_append(reads, bamChunkRead_construct2(stString_print("read_%d", (int) i), r, NULL, b200_rng() % 2,
                                                         polishParams->useRunLengthEncoding));
        }
        /* round 1: no anchors (full matrices) */
        margin_b200_binding_set_reference_augment(0);
        Poa *ours = poa_realign(reads, NULL, reference, polishParams);
        margin_b200_binding_set_reference_augment(1);
        Poa *theirs = poa_realign(reads, NULL, reference, polishParams);
        assert_poas_equal(testCase, ours, theirs);
        /* observations from slabs (the default, `ours`) and one calloc each give the same graph */
        margin_b200_binding_set_reference_augment(0);
        margin_b200_binding_set_observation_slabs(0);
        Poa *plain = poa_realign(reads, NULL, reference, polishParams);
        margin_b200_binding_set_observation_slabs(1);
        assert_poas_equal(testCase, plain, theirs);
        poa_destruct(plain);
        /* round 2: anchors from the posteriors, as poa_realignIterative does (impl/poa.c:1910-1916) */
        stList *anchors = poa_getAnchorAlignments(ours, NULL, stList_length(reads), polishParams);
        margin_b200_binding_set_reference_augment(0);
        Poa *ours2 = poa_realign(reads, anchors, reference, polishParams);
        margin_b200_binding_set_reference_augment(1);
        Poa *theirs2 = poa_realign(reads, anchors, reference, polishParams);
        assert_poas_equal(testCase, ours2, theirs2);
        margin_b200_binding_set_reference_augment(0);
        stList_destruct(anchors);
        poa_destruct(ours);
        poa_destruct(theirs);
        poa_destruct(ours2);
        poa_destruct(theirs2);
        stList_destruct(reads);
        rleString_destruct(reference);
        free(truth);
        free(draft);
    }
    params_destruct(params);
}

static void test_b200_realign_tiny_example(CuTest *testCase) {
    /* tests/polisherTest.c:246-331: reference GATACAGCGGG, read GATTACAGCG, run-length encoded */
    char *trueReference = "GATACAGCGGG";
    char *read = "GATTACAGCG";
    Params *params = params_readParams(b200ParamsFile);
    PolishParams *polishParams = params->polishParams;
    RleString *reference = polishParams->useRunLengthEncoding ? rleString_construct(trueReference) : rleString_construct_no_rle(trueReference);
    stList *reads = stList_construct3(0, (void (*)(void *)) bamChunkRead_destruct);
    stList_append(reads, bamChunkRead_construct2(stString_copy("read"), stString_copy(read), NULL, TRUE, polishParams->useRunLengthEncoding));
    Poa *poa = poa_realign(reads, NULL, reference, polishParams);
    /* the reference test looks at the totals; here both augment paths must agree on every node and the total insert
     * / delete weights must be those the reference test pins (tolerance 1e-3 on weight / PAIR_ALIGNMENT_PROB_1) */
    margin_b200_binding_set_reference_augment(1);
    Poa *theirs = poa_realign(reads, NULL, reference, polishParams);
    margin_b200_binding_set_reference_augment(0);
    assert_poas_equal(testCase, poa, theirs);
    double insertWeight = 0.0, deleteWeight = 0.0;
    for (int64_t n = 0; n < stList_length(poa->nodes); n++) {
        PoaNode *node = stList_get(poa->nodes, n);
        for (int64_t i = 0; i < stList_length(node->inserts); i++) {
            PoaInsert *ins = stList_get(node->inserts, i);
            insertWeight += ins->weightForwardStrand + ins->weightReverseStrand;
        }
        for (int64_t i = 0; i < stList_length(node->deletes); i++) {
            PoaDelete *del = stList_get(node->deletes, i);
            deleteWeight += del->weightForwardStrand + del->weightReverseStrand;
        }
    }
    CuAssertTrue(testCase, insertWeight > 0.0 && deleteWeight > 0.0);
    poa_destruct(poa);
    poa_destruct(theirs);
    stList_destruct(reads);
    rleString_destruct(reference);
    params_destruct(params);
}

static void test_b200_exception_path(CuTest *testCase) {
    Params *params = params_readParams(b200ParamsFile);
    PolishParams *polishParams = params->polishParams;
    StateMachine *sM = polishParams->stateMachineForForwardStrandRead;
    RleString *x = rleString_construct("GATTACAGATTACAGGATTTACAGATTACA"), *y = rleString_construct("GATTACAGATACAGGATTTACAGATTACA");
    SymbolString sX = rleString_constructSymbolString(x, 0, x->length, polishParams->alphabet, polishParams->useRepeatCountsInAlignment, MAXIMUM_REPEAT_LENGTH);
    SymbolString sY = rleString_constructSymbolString(y, 0, y->length, polishParams->alphabet, polishParams->useRepeatCountsInAlignment, MAXIMUM_REPEAT_LENGTH);
    stList *anchors = stList_construct3(0, (void (*)(void *)) stIntTuple_destruct);
    stList_append(anchors, stIntTuple_construct3(5, 5, 4));
    stList_append(anchors, stIntTuple_construct3(4, 6, 4)); /* x goes backwards: band_constructDynamic's precondition */
    int caught = 0;
    stTry {
            stList *l = getAlignedPairsUsingAnchors(sM, sX, sY, anchors, polishParams->p, 0, 0);
            stList_destruct(l);
        }
        stCatch(PAIRWISE_ALIGNMENT_EXCEPTION_ID)
            {
                caught = 1;
                st_logInfo("caught: %s\n", stExcept_getMsg(PAIRWISE_ALIGNMENT_EXCEPTION_ID));
            }
    stTryEnd
    CuAssertTrue(testCase, caught);
    /* the context is usable afterwards */
    stList_destruct(anchors);
    anchors = stList_construct3(0, (void (*)(void *)) stIntTuple_destruct);
    stList_append(anchors, stIntTuple_construct3(4, 4, 4));
    stList *l = getAlignedPairsUsingAnchors(sM, sX, sY, anchors, polishParams->p, 0, 0);
    CuAssertTrue(testCase, stList_length(l) > 10);
    stList_destruct(l);
    stList_destruct(anchors);
    symbolString_destruct(sX);
    symbolString_destruct(sY);
    rleString_destruct(x);
    rleString_destruct(y);
    params_destruct(params);
}

static void test_b200_bubble_batch_ab(CuTest *testCase) {
    Params *params = params_readParams(b200ParamsFile);
    PolishParams *polishParams = params->polishParams;
    for (int64_t test = 0; test < 4; test++) {
        char *truth = b200_random_sequence(600 + (int64_t) (b200_rng() % 600));
        char *draft = b200_mutate(truth, 2, 2);
        RleString *reference = polishParams->useRunLengthEncoding ? rleString_construct(draft) : rleString_construct_no_rle(draft);
        stList *reads = stList_construct3(0, (void (*)(void *)) bamChunkRead_destruct);
        for (int64_t i = 0; i < 16; i++) {
            stList_append(reads, bamChunkRead_construct2(stString_print("read_%d", (int) i), b200_mutate(truth, 4, 4), NULL,
                                                         b200_rng() % 2, polishParams->useRunLengthEncoding));
        }
        Poa *poa = poa_realign(reads, NULL, reference, polishParams);
        margin_b200_binding_set_bubble_batching(1);
        int64_t *mapA = NULL, *mapB = NULL;
        RleString *batched = poa_polish(poa, reads, polishParams, &mapA);
        margin_b200_binding_set_bubble_batching(0);
        RleString *oneByOne = poa_polish(poa, reads, polishParams, &mapB);
        margin_b200_binding_set_bubble_batching(1);
        CuAssertTrue(testCase, memcmp(mapA, mapB, reference->length * sizeof(int64_t)) == 0);
        free(mapA);
        free(mapB);
        CuAssertTrue(testCase, rleString_eq(batched, oneByOne));
        for (uint64_t i = 0; i < batched->length && i < oneByOne->length; i++)
            CuAssertTrue(testCase, batched->repeatCounts[i] == oneByOne->repeatCounts[i]);
        rleString_destruct(batched);
        rleString_destruct(oneByOne);
        poa_destruct(poa);
        stList_destruct(reads);
        rleString_destruct(reference);
        free(truth);
        free(draft);
    }
    params_destruct(params);
}

static void test_b200_repeat_counts_ab(CuTest *testCase) {
    /* the binding's flat restatement of poa_estimateRepeatCountsUsingBayesianModel against the reference's own
     * repeatSubMatrix_getMLRepeatCount (impl/repeatSubMatrix.c:124-151), node by node */
    Params *params = params_readParams(b200ParamsFile);
    PolishParams *polishParams = params->polishParams;
    for (int64_t test = 0; test < 6; test++) {
        char *truth = b200_random_sequence(400 + (int64_t) (b200_rng() % 1500));
        char *draft = b200_mutate(truth, 1, 1);
        RleString *reference = rleString_construct(draft);
        int64_t readNumber = 5 + (int64_t) (b200_rng() % 20);
        stList *reads = stList_construct3(0, (void (*)(void *)) bamChunkRead_destruct);
        for (int64_t i = 0; i < readNumber; i++) {
            stList_append(reads, bamChunkRead_construct2(stString_print("read_%d", (int) i), b200_mutate(truth, 4, 4), NULL,
                                                         b200_rng() % 2, polishParams->useRunLengthEncoding));
        }
        Poa *poa = poa_realign(reads, NULL, reference, polishParams);
        poa_estimateRepeatCountsUsingBayesianModel(poa, reads, polishParams->repeatSubMatrix);
        uint64_t total = 0;
        for (int64_t i = 1; i < stList_length(poa->nodes); i++) {
            PoaNode *node = stList_get(poa->nodes, i);
            double logProbability;
            int64_t want = repeatSubMatrix_getMLRepeatCount(polishParams->repeatSubMatrix, poa->alphabet->convertCharToSymbol(node->base),
                                                            node->observations, reads, &logProbability);
            if (want == 0) want = 1;
            CuAssertIntEquals(testCase, want, (int64_t) poa->refString->repeatCounts[i - 1]);
            CuAssertIntEquals(testCase, want, (int64_t) node->repeatCount);
            total += want;
        }
        CuAssertTrue(testCase, total == poa->refString->nonRleLength);
        poa_destruct(poa);
        stList_destruct(reads);
        rleString_destruct(reference);
        free(truth);
        free(draft);
    }
    params_destruct(params);
}

CuSuite *b200BindingTestSuite(void) {
    CuSuite *suite = CuSuiteNew();
    SUITE_ADD_TEST(suite, test_b200_augment_example);
    SUITE_ADD_TEST(suite, test_b200_augment_ab);
    SUITE_ADD_TEST(suite, test_b200_realign_tiny_example);
    SUITE_ADD_TEST(suite, test_b200_exception_path);
    SUITE_ADD_TEST(suite, test_b200_bubble_batch_ab);
    SUITE_ADD_TEST(suite, test_b200_repeat_counts_ab);
    return suite;
}
