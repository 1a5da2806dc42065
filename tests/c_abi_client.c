/* A plain C client of include/bnl_b200.h -- what a patched bnlearn call site does
 * (INTEGRATION.md): register a data frame given as column pointers, run a batch
 * of CI tests and node scores, read the results. Built and run by
 * tests/test_gpu_c_client.py, which checks the printed numbers against the oracle.
 * usage: c_abi_client <nobs> <seed> */
#include <limits.h>
#include <stdio.h>
#include <stdlib.h>

#include "bnl_b200.h"

static unsigned long long s;
static unsigned next(void) { s = s * 6364136223846793005ULL + 1442695040888963407ULL; return (unsigned)(s >> 33); }

int main(int argc, char **argv) {
  long nobs = argc > 1 ? atol(argv[1]) : 1000;
  s = argc > 2 ? strtoull(argv[2], NULL, 10) : 1;
  int isnum[5] = {0, 0, 0, 0, 1};
  int *col[4];
  double *g = malloc(sizeof(double) * nobs);
  for (int c = 0; c < 4; c++) col[c] = malloc(sizeof(int) * nobs);
  for (long i = 0; i < nobs; i++) {
    col[0][i] = 1 + next() % 3;
    col[1][i] = 1 + (col[0][i] + next() % 2) % 2;
    col[2][i] = 1 + next() % 4;
    col[3][i] = (next() % 50 == 0) ? INT_MIN : 1 + (col[2][i] + next() % 3) % 3;
    g[i] = (double)(next() % 1000) / 1000.0;
  }
  const void *cols[5] = {col[0], col[1], col[2], col[3], g};
  int levels[5] = {3, 2, 4, 3, 0};

  bnl_ctx *ctx = NULL;
  bnl_data d = 0;
  if (bnl_init(&ctx, NULL, 0) != BNL_OK) { fprintf(stderr, "init: %s\n", bnl_last_error(ctx)); return 2; }
  if (bnl_data_register(ctx, 5, cols, isnum, levels, nobs, &d) != BNL_OK) {
    fprintf(stderr, "register: %s\n", bnl_last_error(ctx)); return 3;
  }
  int x[3] = {0, 2, 0}, y[3] = {1, 3, 3}, zoff[4] = {0, 0, 1, 3}, zidx[3] = {0, 1, 2};
  double st[3], df[3], pv[3], counter = 0;
  if (bnl_ci_discrete(ctx, d, BNL_MI, 3, x, y, zoff, zidx, st, df, pv) != BNL_OK) {
    fprintf(stderr, "ci: %s\n", bnl_last_error(ctx)); return 4;
  }
  for (int i = 0; i < 3; i++) printf("test %d %.17g %.17g %.17g\n", i, st[i], df[i], pv[i]);
  int tgt[2] = {1, 3}, poff[3] = {0, 1, 3}, pidx[3] = {0, 0, 2};
  double sc[2], np[2];
  if (bnl_score_discrete(ctx, d, BNL_BDE, 1.0, 2, tgt, poff, pidx, sc, np) != BNL_OK) {
    fprintf(stderr, "score: %s\n", bnl_last_error(ctx)); return 5;
  }
  for (int i = 0; i < 2; i++) printf("score %d %.17g\n", i, sc[i]);
  int xs[2] = {0, 2};
  double p2[2];
  if (bnl_indep_test(ctx, d, BNL_X2, 2, xs, 1, NULL, 0, p2, NULL, NULL, &counter) != BNL_OK) {
    fprintf(stderr, "indep: %s\n", bnl_last_error(ctx)); return 6;
  }
  printf("indep %.17g %.17g counter %.0f\n", p2[0], p2[1], counter);
  /* an error must come back as a code and a message, not a crash. */
  int bad = 9;
  if (bnl_ci_discrete(ctx, d, BNL_MI, 1, &bad, y, zoff, zidx, st, df, pv) != BNL_EINVAL) return 7;
  printf("error \"%s\"\n", bnl_last_error(ctx));
  bnl_data_release(ctx, d);
  bnl_shutdown(ctx);
  return 0;
}
