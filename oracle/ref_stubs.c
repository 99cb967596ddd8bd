/*
 * ref_stubs.c — stand-ins for lp_solve, the one vendored dependency of the reference that
 * oracle/Makefile does not compile (65 k lines, only used by the read down-sampler,
 * impl/htsIntegration.c:898-1010, which runs when coverage exceeds maxDepth).  Any call aborts
 * loudly, so a parity run can never silently take a different path than the reference.
 * TEST INFRASTRUCTURE ONLY.
 */
#include <stdio.h>
#include <stdlib.h>

#define STUB(name)                                                                         \
    void *name() {                                                                         \
        fprintf(stderr, "oracle/_ref: lp_solve symbol %s is stubbed out (down-sampling)\n", #name); \
        abort();                                                                           \
        return NULL;                                                                       \
    }

STUB(make_lp)
STUB(set_add_rowmode)
STUB(add_constraint)
STUB(add_constraintex)
STUB(set_obj_fn)
STUB(set_maxim)
STUB(set_verbose)
STUB(solve)
STUB(get_variables)
STUB(delete_lp)
