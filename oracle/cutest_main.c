/*
 * cutest_main.c — runs the REFERENCE's own CuTest suites for the hot path (tests/pairwiseAlignerTest.c,
 * tests/polisherTest.c, tests/stitchingTest.c, compiled from /root/reference where they lie) plus the binding's own
 * A/B tests (b200_tests.c).  TEST INFRASTRUCTURE: oracle/Makefile links this main twice,
 *   _ref/cutest_ref    against the unmodified reference objects (CPU; what SURVEY.md §4 measured: OK (43 tests)),
 *   _ref/cutest_b200   against the same objects with compat/pairwiseAligner_b200.c linked in place of the seven
 *                      replaced functions, i.e. every getAlignedPairs* / poa_realign / computeForwardProbability the
 *                      suites call runs on the GPU through libmargin_b200.so.
 * The suites open ../params/... and ../tests/data/polishTestExamples/...: run from oracle/_ref/data/build.
 *
 *   cutest_b200 [logLevel] [suite ...]     suites: pairwise polisher stitching b200 (default: all)
 */
#include <string.h>

#include "CuTest.h"
#include "margin.h"

CuSuite *polisherTestSuite(void);
CuSuite *pairwiseAlignmentTestSuite(void);
CuSuite *stitchingTestSuite(void);
#ifdef WITH_B200_TESTS
CuSuite *b200BindingTestSuite(void);
#endif

static int wanted(int argc, char **argv, const char *name) {
    int any = 0;
    for (int i = 2; i < argc; i++) {
        any = 1;
        if (strcmp(argv[i], name) == 0) return 1;
    }
    return !any;
}

int main(int argc, char **argv) {
    if (argc >= 2) st_setLogLevelFromString(argv[1]);
    CuString *output = CuStringNew();
    CuSuite *suite = CuSuiteNew();
    if (wanted(argc, argv, "pairwise")) CuSuiteAddSuite(suite, pairwiseAlignmentTestSuite());
    if (wanted(argc, argv, "polisher")) CuSuiteAddSuite(suite, polisherTestSuite());
    if (wanted(argc, argv, "stitching")) CuSuiteAddSuite(suite, stitchingTestSuite());
#ifdef WITH_B200_TESTS
    if (wanted(argc, argv, "b200")) CuSuiteAddSuite(suite, b200BindingTestSuite());
#endif
    CuSuiteRun(suite);
    CuSuiteSummary(suite, output);
    CuSuiteDetails(suite, output);
    printf("%s\n", output->buffer);
    return suite->failCount > 0;
}
