/* ref_stubs.c -- link-time stand-ins for reference functions that the compiled
 * hot-path files mention but that belong to score / test families SURVEY.md
 * section 8 marks out of scope (conditional-Gaussian, permutation, NML,
 * predictive, custom callbacks, ...). Calling one is a harness bug, so each
 * raises error(). TEST INFRASTRUCTURE ONLY. */
#include "shim/minir.h"

#define STUB(name) void name(void) { Rf_error("oracle/_ref: " #name "() is out of scope and not compiled in"); }

STUB(R_qsort) STUB(R_qsort_int) STUB(R_qsort_int_I) STUB(revsort)
STUB(custom_score_function) STUB(custom_test_function)
STUB(em_irls_node) STUB(fitted_node_to_enum) 
STUB(predictive_loglik_cgnode)
STUB(predictive_loglik_dnode) STUB(predictive_loglik_gnode)

/* globals the reference keeps in src/globals.c:17 (not compiled here). */
double *regret_table = NULL;
