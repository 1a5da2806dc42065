/* bnl_b200.h -- C ABI of libbnl_b200.so, the B200 (sm_100a) engine for bnlearn's
 * structure-learning statistics (count tables, discrete/Gaussian node scores,
 * conditional-independence tests).
 *
 * The reference has no FFI for this path: it sits behind ordinary C functions
 * called from the .Call entry points. Each entry point
 * below names the reference function(s) whose per-candidate loop it replaces;
 * INTEGRATION.md shows the patched call sites. Paths are relative to the
 * reference's root (cran/bnlearn 5.2.1).
 *
 * Conventions
 *   - every function returns 0 on success or a BNL_E* code; the message is
 *     available from bnl_last_error(). Nothing longjmps or throws across the
 *     boundary (the reference's error() does, src/core/allocations.c:13-15).
 *   - all pointers are HOST pointers owned by the caller; outputs are
 *     pre-allocated by the caller. Data columns are only read during
 *     bnl_data_register(); afterwards the data set lives in device memory.
 *   - columns are identified by their 0-based position in the registered data
 *     set. Factor columns are int32, 1-based, NA = INT_MIN (R's encoding);
 *     numeric columns are double.
 *   - the ORDER of parents / conditioning variables is significant: it fixes
 *     the configuration index exactly as c_fast_config() does
 *     (src/core/sets.c:162-218: first variable varies fastest).
 *   - enum values are the reference's: test_e (src/tests/tests.h:8-36) and
 *     score_e (src/scores/scores.h:9-60).
 *   - a bnl_ctx is used by one host thread at a time (R is single-threaded).
 *   - there is no CPU fallback: without a CUDA device bnl_init() fails.
 */
#ifndef BNL_B200_H
#define BNL_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct bnl_ctx bnl_ctx;   /* opaque; one per process (or per R worker) */
typedef uint64_t bnl_data;        /* handle to a device-resident, packed data set */

enum {
  BNL_OK = 0,
  BNL_EINVAL = 1,       /* bad argument (index out of range, wrong column type, ...) */
  BNL_ECUDA = 2,        /* CUDA runtime error */
  BNL_ENOMEM = 3,       /* host or device allocation failed */
  BNL_EOVERFLOW = 4,    /* configuration count >= INT_MAX (src/core/sets.c:182-183) */
  BNL_EUNSUPPORTED = 5  /* valid in the reference, not handled by this engine yet */
};

/* test_e values handled (src/tests/tests.h:8-36). */
enum { BNL_MI = 1, BNL_MI_ADF = 2, BNL_X2 = 3, BNL_X2_ADF = 4, BNL_JT = 10, BNL_MI_CG = 30, BNL_MI_SH = 40,
       BNL_COR = 20, BNL_ZF = 21, BNL_MI_G = 22, BNL_MI_G_SH = 41,
       /* permutation, semiparametric and sequential permutation tests (tests.h:21-34) */
       BNL_MC_MI = 50, BNL_MC_X2 = 51, BNL_SP_MI = 52, BNL_SP_X2 = 53, BNL_MC_JT = 54,
       BNL_SMC_MI = 60, BNL_SMC_X2 = 61, BNL_SMC_JT = 62,
       BNL_MC_COR = 70, BNL_MC_MI_G = 71, BNL_MC_ZF = 72, BNL_SMC_COR = 80, BNL_SMC_MI_G = 81, BNL_SMC_ZF = 82 };
/* score_e values handled (src/scores/scores.h:9-60). */
enum { BNL_LOGLIK = 1, BNL_AIC = 3, BNL_BIC = 4, BNL_EBIC = 5, BNL_BDE = 6, BNL_BDS = 7, BNL_BDJ = 8, BNL_K2 = 9,
       BNL_MBDE = 10,
       BNL_FNML = 11, BNL_QNML = 12, BNL_NAL = 13, BNL_PNAL = 14,
       BNL_LOGLIK_G = 100, BNL_AIC_G = 102, BNL_BIC_G = 103, BNL_EBIC_G = 104, BNL_BGE = 105, BNL_NAL_G = 106,
       BNL_PNAL_G = 107,
       BNL_LOGLIK_CG = 200, BNL_AIC_CG = 202, BNL_BIC_CG = 203, BNL_EBIC_CG = 204, BNL_NAL_CG = 205,
       BNL_PNAL_CG = 206 };

/* ---- context ---------------------------------------------------------- */

/* Create a context on `ndev` CUDA devices (devices == NULL: the current device).
 * With ndev > 1 registered data is replicated on every device and each call's
 * units are split into contiguous blocks, one per device;
 * results do not depend on ndev. Replaces nothing in the reference (it has no
 * device); plays the role of R's parallel cluster set-up (R/utils-cluster.R:31-37). */
int bnl_init(bnl_ctx **ctx, const int *devices, int ndev);
void bnl_shutdown(bnl_ctx *ctx);
const char *bnl_last_error(const bnl_ctx *ctx);

/* Device-side timing of the most recent compute call, from CUDA events recorded
 * on the engine's own stream (max over devices): total_ms spans descriptor
 * upload .. result download; count_ms is the counting (or Gaussian solve)
 * kernel alone, epilogue_ms the statistic kernel alone; kernel_launches counts
 * this library's kernel launches in the call. Any pointer may be NULL. For
 * benchmarks; no counterpart in the reference. */
int bnl_last_call_profile(const bnl_ctx *ctx, double *total_ms, double *count_ms,
                          double *epilogue_ms, int *kernel_launches);

/* ---- data set ---------------------------------------------------------- */

/* Upload and pack a data frame. col[i] points to nobs int32 (is_numeric[i] == 0,
 * nlevels[i] levels) or nobs double (is_numeric[i] != 0). Factors are repacked
 * to uint8 (code - 1, NA -> 0xFF) when they have at most 127 levels and to
 * uint16 (NA -> 0xFFFF) up to 32767 levels; units that touch a uint16 column
 * are counted by the generic kernel.
 * Replaces the per-call column extraction of tabular_from_SEXP()
 * (src/core/data.table.c:143-225) and c_dataframe_column()
 * (src/minimal/data.frame.c:41). */
int bnl_data_register(bnl_ctx *ctx, int ncols, const void *const *col,
                      const int *is_numeric, const int *nlevels, int64_t nobs,
                      bnl_data *handle);
/* Hyper-parameters of the BGe score (BNL_BGE) for later calls on this data set:
 * the extra arguments iss.mu, iss.w and nu that per_node_score hands to
 * wishart_node (src/scores/per.node.score.c:250-266, src/scores/
 * wishart.posterior.c:108-160). iss_w <= 0 selects the front end's default
 * ncol(data) + 2, nu == NULL its default colMeans(data) (nu has one entry per
 * registered column; entries of factor columns are ignored). Uniform graph
 * prior. The k argument of bnl_score_gaussian is ignored for BNL_BGE. */
int bnl_set_bge_prior(bnl_ctx *ctx, bnl_data data, double iss_mu, double iss_w, const double *nu);

/* gamma of the extended BIC scores (BNL_EBIC, BNL_EBIC_G, BNL_EBIC_CG:
 * BIC - 4 gamma #parents log(ncol(data)), src/scores/per.node.score.c:116-134,
 * 155-173, 194-212) for later calls on this data set; 0.5 until set, the front
 * end's default. The k argument of the score calls is the BIC coefficient. */
int bnl_set_ebic_gamma(bnl_ctx *ctx, bnl_data data, double gamma);

/* A new data set made of rows[0..nrows) (0-based, repeats allowed) of a
 * registered one, gathered on the device: the bootstrap replicate
 * `data[resampling, , drop = FALSE]` of bootstrap.backend (R/bootstrap.R:19-36)
 * and arc.strength.boot (R/arc.strength.R:144-162) without the host-side copy and
 * the upload of a whole data frame per replicate. Only the row list crosses PCIe. */
int bnl_data_resample(bnl_ctx *ctx, bnl_data source, int64_t nrows, const int *rows,
                      bnl_data *handle);
/* A data set whose packed columns are filled on the device by the caller instead
 * of being uploaded: the way to replicate one data frame over the GPUs of a job
 * with one process per GPU ("NCCL broadcast from GPU 0") --
 * rank 0 registers the host data frame, every other rank allocates the same
 * shape, the packed block travels over NVLink (ncclBroadcast on the pointers
 * bnl_data_device_view returns, factor block and numeric block), and
 * bnl_data_commit finds the NA flags, checks the codes and forms the Gaussian
 * moments. Layout: factor column s (registration order among the factors) at
 * u8 + s * stride, stride = u8_bytes / #factors, codes 0-based, NA and padding
 * 0xFF; numeric column likewise in the f64 block. No counterpart in the
 * reference (its cluster workers each receive the data frame from R). Requests
 * on a data set that was allocated but not committed treat every factor as
 * possibly holding NA. */
int bnl_data_alloc(bnl_ctx *ctx, int ncols, const int *is_numeric, const int *nlevels, int64_t nobs,
                   bnl_data *handle);
int bnl_data_device_view(bnl_ctx *ctx, bnl_data data, int dev_index, void **u8, size_t *u8_bytes,
                         void **f64, size_t *f64_bytes);
int bnl_data_commit(bnl_ctx *ctx, bnl_data data);
int bnl_data_release(bnl_ctx *ctx, bnl_data handle);
/* Drop every registered data set; meant to be called from reset_test_counter()
 * (src/test.counter.c:26), which every learner calls first. */
int bnl_data_flush(bnl_ctx *ctx);

/* ---- discrete: tables, scores, tests ----------------------------------- */

/* Raw contingency tables for n units. Unit u counts x[u] (rows), y[u] (columns;
 * -1 for none) within the configurations of z_idx[z_off[u] .. z_off[u+1]).
 * tables + table_off[u] receives n[z][x][y] as ((z * llx) + x) * lly + y;
 * table_off has n + 1 entries. Rows with an NA in any involved column are
 * skipped. Replaces fill_1d/2d/3d_table (src/core/contingency.tables.c:20-34,
 * 93-130, 198-223) after c_fast_config (src/core/sets.c:162-218). */
int bnl_counts(bnl_ctx *ctx, bnl_data data, int n, const int *x, const int *y,
               const int *z_off, const int *z_idx, int32_t *tables,
               const int64_t *table_off);

/* Node scores for n (target, parent set) units; parents of unit u are
 * pa_idx[pa_off[u] .. pa_off[u+1]). k_or_iss is the penalty coefficient k for
 * AIC/BIC and the imaginary sample size for BDe/BDs (uniform graph prior); it is
 * ignored for BDJ and K2, whose per-cell hyperparameter is fixed (1/2 and 1,
 * src/scores/per.node.score.c:233-246). NAL / PNAL (the scores bnlearn switches
 * to on incomplete data, src/scores/discrete.nal.c:8-68) average the
 * log-likelihood over the locally complete rows; k_or_iss is PNAL's k.
 * nparams may be NULL. Replaces c_per_node_score() for these score_e values
 * (src/scores/per.node.score.c:51-534 -> loglik_dnode,
 * src/scores/discrete.loglikelihood.c:76-118; dirichlet_node,
 * src/scores/dirichlet.posterior.c:152-202), i.e. the body of the candidate
 * loop in score_cache_fill() (src/learning/score/hc.cache.lookup.c:54-117). */
int bnl_score_discrete(bnl_ctx *ctx, bnl_data data, int score_e, double k_or_iss,
                       int n, const int *target, const int *pa_off,
                       const int *pa_idx, double *score, double *nparams);

/* Interventional data for BNL_MBDE (src/scores/dirichlet.posterior.c:39-60,
 * 102-135: rows on which the node itself was intervened upon are left out of its
 * counts). mask_col is a registered factor column with ONE level whose NA rows are
 * the interventions on `node` (R: factor(ifelse(seq_len(n) %in% exp[[node]], NA, 1)));
 * -1 clears it. bnl_score_discrete(BNL_MBDE, ...) then counts every unit of that
 * node with the mask as an extra conditioning column -- a level-1 column adds
 * nothing to the cell index and its NA rows are skipped like any other NA -- and
 * scores the tables as BDe. Nodes without a mask score as plain BDe. */
int bnl_data_set_interventions(bnl_ctx *ctx, bnl_data data, int node, int mask_col);

/* Discrete CI tests x[u] _||_ y[u] | z (mi, mi-adf, x2, x2-adf): statistic (MI
 * scaled to G^2), degrees of freedom and p-value = pchisq(stat, df, upper).
 * Replaces c_chisqtest / c_cchisqtest (src/tests/discrete/discrete.tests.c:11-94),
 * discrete_df / discrete_cdf (src/tests/discrete/df.adjust.c:4-90) and the
 * pchisq call in ut_discrete, ct_discrete, ast_discrete, rrd_discrete
 * (src/tests/patterns/utest.c:23, ctest.c:11, allsubs.test.c:52,
 * roundrobin.test.c:32). */
int bnl_ci_discrete(bnl_ctx *ctx, bnl_data data, int test_e, int n, const int *x,
                    const int *y, const int *z_off, const int *z_idx,
                    double *stat, double *df, double *pvalue);

/* The statistics of bnl_ci_discrete (kind 0, code = test_e: stat / df / p-value
 * in o0 / o1 / o2) or bnl_score_discrete (kind 1, code = score_e, arg = k or
 * iss: score / nparams in o0 / o1) from tables the caller already holds, laid
 * out as bnl_counts returns them: unit u has llx[u] x lly[u] x llz[u] cells at
 * tables + table_off[u] (lly = 1 for scores), nz[u] conditioning columns or
 * parents behind the llz[u] strata, and nrows is the row count of the data (the
 * penalty of nal / pnal uses it). This is the second half of a row-sharded run:
 * every GPU counts its block of rows with bnl_counts, the int32 tables are summed
 * (ncclAllReduce over NVLink, or MPI, or on the host) and one rank turns them
 * into statistics -- the same mi_kernel / cdlik / cdpost arithmetic on the same
 * n[z][x][y], so the result is that of an unsharded run. Also the hook for
 * callers that fill tables themselves (src/core/contingency.tables.c). */
int bnl_stat_from_tables(bnl_ctx *ctx, int kind, int code, double arg, int64_t nrows, int n,
                         const int *llx, const int *lly, const int *llz, const int *nz,
                         const int32_t *tables, const int64_t *table_off, double *o0, double *o1,
                         double *o2);

/* ---- permutation tests -------------------------------------------------- */

/* The caller's uniform generator: returns the next U(0, 1) draw of its stream.
 * In the reference this is R's unif_rand() between GetRNGstate() / PutRNGstate()
 * (src/tests/permutation/discrete.monte.carlo.c:49, 144); the binding passes a
 * one-line wrapper, so the library advances R's own stream draw by draw. */
typedef double (*bnl_unif_fn)(void *state);

/* A generator for callers without one: splitmix64 mapped to (0, 1); state is a
 * uint64_t the caller seeds. */
double bnl_unif_splitmix64(void *state);

/* Discrete Monte Carlo permutation (mc-mi, mc-x2, mc-jt), sequential (smc-*) and
 * semiparametric (sp-mi, sp-x2) tests. Replaces c_mcarlo / c_cmcarlo
 * (src/tests/permutation/discrete.monte.carlo.c:16-159, 162-308) behind
 * ut_dperm / ct_dperm (src/tests/patterns/utest.c:347, ctest.c). The observed
 * tables of all n tests are counted in one device pass; the B random tables per
 * test (Patefield's algorithm given the margins, c_rcont2 in
 * src/core/contingency.tables.c:288) are drawn on the host from `unif`, test after
 * test in the order given, which is the order the reference consumes its stream.
 * alpha is the early-stopping threshold of the sequential tests (ignored otherwise,
 * where the reference's callers pass 1). stat as in the reference's htest; df is
 * the estimated one for sp-*, NaN otherwise (may be NULL). */
int bnl_ci_discrete_mc(bnl_ctx *ctx, bnl_data data, int test_e, int B, double alpha, int n, const int *x,
                       const int *y, const int *z_off, const int *z_idx, bnl_unif_fn unif,
                       void *unif_state, double *stat, double *df, double *pvalue);

/* The permutation stage alone, over count tables the caller already holds
 * (n[z][x][y] per test, as bnl_counts returns them; nz[u] != 0 marks a conditional
 * test): for rows sharded over GPUs -- bnl_counts per shard, all-reduce, then this
 * on one rank. Runs on the host; ctx may be NULL. */
int bnl_mc_from_tables(bnl_ctx *ctx, int test_e, int B, double alpha, int n, const int *llx, const int *lly,
                       const int *nz, const int32_t *tables, const int64_t *table_off, bnl_unif_fn unif,
                       void *unif_state, double *stat, double *df, double *pvalue);

/* Gaussian permutation tests (mc-cor, mc-zf, mc-mi-g and their smc-* forms); a test
 * with a missing value in one of its columns runs on its complete rows, as
 * ut_gperm / ct_gperm build them. Replaces c_gauss_mcarlo / c_gauss_cmcarlo
 * (src/tests/permutation/gaussian.monte.carlo.c:24-129, 132-273). Permutations of
 * y are drawn on the host from `unif` exactly as RandomPermutation() draws them
 * (src/core/sampling.c:12-28), in batches that never run past the reference's
 * early exit, so the caller's stream ends where the reference leaves it; the
 * permuted cross-products (the O(B n k) part) are formed on the device. The
 * sums are reduced in parallel, so a permuted value equal to the observed one to
 * the last bit may compare differently from the reference's sequential long
 * double sum. stat as bnl_ci_gaussian returns it. */
int bnl_ci_gaussian_mc(bnl_ctx *ctx, bnl_data data, int test_e, int B, double alpha, int n, const int *x,
                       const int *y, const int *z_off, const int *z_idx, bnl_unif_fn unif,
                       void *unif_state, double *stat, double *pvalue);

/* ---- Gaussian ---------------------------------------------------------- */

/* Gaussian CI tests (cor, zf, mi-g), served from the centred cross-product
 * matrix that bnl_data_register() forms once per data set. stat is the
 * (partial) correlation for cor and the transformed statistic for zf / mi-g, as
 * in the reference's htest. Replaces ut/ct/ast/rrd_gaustests_complete
 * (src/tests/patterns/utest.c:65, ctest.c:58, allsubs.test.c:150,
 * roundrobin.test.c:106) over c_covmat (src/core/covariance.matrix.c:65-97) and
 * c_fast_pcor (src/core/correlation.c:43-92). */
int bnl_ci_gaussian(bnl_ctx *ctx, bnl_data data, int test_e, int n, const int *x,
                    const int *y, const int *z_off, const int *z_idx,
                    double *stat, double *df, double *pvalue);

/* Every pair i < j of the listed numeric columns tested marginally (cor, zf, mi-g):
 * what level 0 of pc.stable asks for, one .Call per pair in the reference
 * (pc.stable.backend, R/pcalgo.R:8-28 -> allsubs_test -> ut_gaustests_complete,
 * src/tests/patterns/utest.c:65-129). No index arrays go up and no per-pair
 * triples come down: pvalue (may be NULL) receives the ncols x ncols symmetric
 * matrix of p-values (row-major, diagonal 0), and the pairs with p <= alpha -- the
 * arcs that stay in the skeleton -- are written to keep_i / keep_j in row-major
 * order of the upper triangle (positions in cols[]); *nkeep is their number, of
 * which at most cap are written (call again with larger arrays if it exceeds cap).
 * Same p-values as bnl_ci_gaussian on the same pairs. Complete data only.
 * test_counter += ncols (ncols - 1) / 2. */
int bnl_ci_gaussian_allpairs(bnl_ctx *ctx, bnl_data data, int test_e, double alpha, int ncols,
                             const int *cols, double *pvalue, int *keep_i, int *keep_j, long long cap,
                             long long *nkeep, double *test_counter);

/* Gaussian node scores (loglik-g, aic-g, bic-g) from the same matrix. Replaces
 * loglik_gnode (src/scores/gaussian.loglikelihood.c:85-114) over c_ols
 * (src/math/least.squares.c:866-932). */
int bnl_score_gaussian(bnl_ctx *ctx, bnl_data data, int score_e, double k, int n,
                       const int *target, const int *pa_off, const int *pa_idx,
                       double *score, double *nparams);

/* Conditional Gaussian node scores (loglik-cg, aic-cg, bic-cg) on a data set
 * that mixes factors and numeric columns: loglik_cgnode (src/scores/
 * cg.loglikelihood.c:119-199). Discrete nodes with discrete parents take the
 * discrete score, numeric nodes with numeric parents the Gaussian one, a
 * discrete node with a numeric parent scores -Inf, and a numeric node with at
 * least one discrete parent is regressed on its numeric parents within every
 * configuration of the discrete ones (ccgloglik :10-117 over c_cls,
 * src/math/conditional.least.squares.c): nparams = configurations x (numeric
 * parents + 2). Complete data only; at most 7 numeric and 8 discrete parents. */
int bnl_score_cg(bnl_ctx *ctx, bnl_data data, int score_e, double k, int n, const int *target,
                 const int *pa_off, const int *pa_idx, double *score, double *nparams);

/* The conditional Gaussian mutual information test mi-cg on mixed data
 * (ut_micg_complete, src/tests/patterns/utest.c:195-273; c_micg_chisq /
 * ct_micg_complete, src/tests/patterns/ctest.c:232-334): x[u], y[u] and the
 * conditioning columns may be factors or numeric. Two factors with discrete
 * conditioning reduce to mi, two numeric columns without conditioning to mi-g;
 * every other case is 2 (L1 - L0) for two conditional Gaussian log-likelihoods
 * (or the chain rule of c_cmicg_unroll when two factors are tested given numeric
 * columns), with the reference's degrees of freedom. Complete data only. */
int bnl_ci_cg(bnl_ctx *ctx, bnl_data data, int n, const int *x, const int *y, const int *z_off,
              const int *z_idx, double *stat, double *df, double *pvalue);

/* Copy the p x p covariance matrix (divisor nobs - 1, column-major, p = number
 * of numeric columns in registration order) and the column means to the host.
 * Replaces c_meanvec + c_covmat over all numeric columns. */
int bnl_gaussian_moments(bnl_ctx *ctx, bnl_data data, double *mean, double *cov);

/* Rows sharded over GPUs, Gaussian half (the discrete half is bnl_counts +
 * all-reduce + bnl_stat_from_tables): each rank registers its block of rows and
 * reads its moments with bnl_gaussian_moments; the ranks pool them -- n = sum n_r,
 * mean = sum n_r mean_r / n, (n - 1) cov = sum [(n_r - 1) cov_r + n_r (mean_r - mean)
 * (mean_r - mean)'], two sum all-reduces over p and p^2 doubles -- and hand the result
 * back here. Every Gaussian test and score of the data set is then served from the
 * pooled moments with nobs_total rows (complete data only). What needs a pass over
 * the rows themselves is refused on such a data set with BNL_EUNSUPPORTED instead
 * of answering from this rank's rows alone: mi-g-sh (shrinkage intensity), the
 * permutation tests, bnl_score_cg and bnl_ci_cg. Replaces c_covmat over the whole
 * data (src/core/covariance.matrix.c:65-97). */
int bnl_data_set_moments(bnl_ctx *ctx, bnl_data data, int64_t nobs_total, const double *mean,
                         const double *cov);

/* ---- call sites: drop-in bodies for the reference's per-call loops ----------
 * These keep the reference's semantics that consume the numbers: evaluation
 * order, early exit, min/max p-value tracking, and the exact test-counter
 * arithmetic ("metrics" row). `test_counter` points to the
 * caller's counter (the reference's global `double test_counter`,
 * src/test.counter.c:4) and is incremented exactly as the reference would. */

/* indep_test(..., learning = TRUE) (src/tests/rinterface/indep.test.c:5-19 ->
 * utest.c:9-102 / ctest.c:10-141): nx variables x[] each tested against y given
 * sx[0..nsx); pvalue[nx]. stat/df (may be NULL) receive the LAST test's values,
 * as the htest path does. test_counter += nx. */
int bnl_indep_test(bnl_ctx *ctx, bnl_data data, int test_e, int nx, const int *x, int y,
                   const int *sx, int nsx, double *pvalue, double *stat, double *df,
                   double *test_counter);

/* allsubs_test (src/tests/rinterface/allsubs.test.c:10-210 over ast_discrete /
 * ast_gaustests_complete, src/tests/patterns/allsubs.test.c:52-287) for n
 * independent requests at once (n = 1 is the reference's call). Request r tests
 * x[r] vs y[r] given every subset, of sizes minsize..maxsize, of the non-fixed
 * part of sx[sx_off[r] .. sx_off[r+1]) (the first nfixed[r] entries are fixed
 * and always conditioned on), in lexicographic order, stopping at the first
 * p-value > alpha. out[3r..3r+2] = p.value, min.p.value, max.p.value;
 * dsep[dsep_off[r]..] receives the separating set (column indices) and ndsep[r]
 * its size, or -1 when none was found (dsep_off[r+1] - dsep_off[r] must be >=
 * sx_off[r+1] - sx_off[r]). test_counter is advanced by exactly the number of
 * tests the serial code would have run (the marginal test counts twice,
 * rinterface/allsubs.test.c:23-30). */
int bnl_allsubs_test(bnl_ctx *ctx, bnl_data data, int test_e, int n, const int *x, const int *y,
                     const int *sx_off, const int *sx, const int *nfixed, int minsize,
                     int maxsize, double alpha, double *out, const int *dsep_off, int *dsep,
                     int *ndsep, double *test_counter);

/* The whole of hill.climbing(optimized = TRUE, restart = 0) in one call
 * (R/hill-climbing.R:2-318: the repeat loop around score_cache_fill and
 * hc_opt_step, src/learning/score/hc.cache.lookup.c:10-363, with
 * arcs.to.be.added -> hc_to_be_added :373-489): the score cache lives in the
 * library between iterations and only the deltas of the nodes touched by the
 * last move are recomputed. The first nnodes registered columns are the nodes.
 * start / blmat / wlmat: nnodes x nnodes column-major 0/1 matrices (NULL = empty
 * graph / no blacklist / no whitelist); maxp: maximum number of parents
 * (INFINITY for none); max_iter <= 0: no limit. Outputs: amat (column-major),
 * node_scores[nnodes] (the reference scores of the last network), iterations,
 * and test_counter advanced as score_delta() would (score.delta.c:278). Moves,
 * ties (first candidate better by more than sqrt(eps) wins, additions before
 * removals before reversals) and the rescoring of singular nodes
 * (R/hill-climbing.R:229-256) follow the reference. */
int bnl_hc(bnl_ctx *ctx, bnl_data data, int score_e, double k_or_iss, int nnodes,
           const int *start, const int *blmat, const int *wlmat, double maxp,
           long long max_iter, int *amat, double *node_scores, long long *iterations,
           double *test_counter);

/* tabu.search(optimized = TRUE) in one call (R/tabu.R:2-300 around
 * score_cache_fill and tabu_step, src/learning/score/tabu.c:93-435): arguments
 * as bnl_hc plus the length of the tabu list and the number of consecutive
 * non-improving iterations allowed (max.tabu). Returns the best network seen when
 * the search ended on a non-improving stretch, like the reference; node_scores
 * are those of the LAST network. Not restated: the ranking of networks that hold
 * singular (-Inf) nodes, R/tabu.R:74-117. */
int bnl_tabu(bnl_ctx *ctx, bnl_data data, int score_e, double k_or_iss, int nnodes,
             const int *start, const int *blmat, const int *wlmat, double maxp,
             long long max_iter, int tabu, int max_tabu, int *amat, double *node_scores,
             long long *iterations, double *test_counter);

/* The skeleton phase of pc.stable in one call (pc.stable.backend,
 * R/pcalgo.R:2-148, one allsubs_test .Call per node pair and level there): at
 * every level all pairs see the skeleton of the previous level, so a level is
 * two batched bnl_allsubs_test calls. The first nnodes registered columns are
 * the nodes; max_sx < 0: no limit on the conditioning-set size. Outputs: adj
 * (nnodes x nnodes, symmetric 0/1), pvalue (same shape, may be NULL), and for
 * the pairs i < j in row-major order ndsep[pair] (-1: not separated) with the
 * separating sets in dsep_idx[dsep_off[pair] .. dsep_off[pair + 1]) (dsep_off
 * has npairs + 1 entries; dsep_cap entries are available in dsep_idx).
 * test_counter advances as in the pair-by-pair loop. */
int bnl_pc_stable(bnl_ctx *ctx, bnl_data data, int test_e, double alpha, int nnodes, int max_sx,
                  int *adj, double *pvalue, int *ndsep, long long *dsep_off, int *dsep_idx,
                  long long dsep_cap, double *test_counter);

/* estimate_mi_matrix (src/learning/local/mi.matrix.c:184-252 over
 * mi_matrix_discrete :40-137 / mi_matrix_gaussian :141-181), the coefficient
 * matrix behind aracne() and chow.liu(): for the ncols listed columns (all
 * factors with estimator BNL_MI, all numeric with BNL_MI_G) the pairwise mutual
 * information mi_kernel()/nobs, cmi_kernel()/nobs given the factor column `cond`
 * (TAN's class variable; -1 for none), or -log(1 - r^2)/2. out has
 * ncols (ncols - 1) / 2 entries, pair (i, j > i) at the position UTREL() gives it
 * (row-major over the strict upper triangle, src/core/uppertriangular.h:32-43). */
int bnl_mi_matrix(bnl_ctx *ctx, bnl_data data, int estimator_e, int ncols, const int *cols,
                  int cond, double *out);

/* roundrobin_test (src/tests/rinterface/roundrobin.test.c:10-154 over rrd_discrete
 * / rrd_gaustests_complete, src/tests/patterns/roundrobin.test.c:32-219): x vs
 * each non-fixed z[i] given the other not-yet-dropped z; a variable found
 * independent (p > alpha) is dropped from the later conditioning sets.
 * is_fixed[nz] flags fixed variables; pvalue has nz - #fixed entries, zero for
 * tests never run. test_counter += nz - #fixed. */
int bnl_roundrobin_test(bnl_ctx *ctx, bnl_data data, int test_e, int x, const int *z, int nz,
                        const int *is_fixed, double alpha, double *pvalue,
                        double *test_counter);

/* The three call sites above with a permutation test (mc-*, smc-*, sp-*; ut_dperm /
 * ct_dperm / ut_gperm / ct_gperm, ast_dperm / ast_gperm, rrd_dperm / rrd_gperm in
 * src/tests/patterns/): B permutations per test and the caller's uniform stream
 * (see bnl_ci_discrete_mc). alpha is also the early-stopping threshold of the smc-*
 * tests, as in the reference's callers. The tests are evaluated strictly in the
 * reference's order and nothing is drawn past an early exit, so the stream ends
 * where the reference leaves it; allsubs therefore takes one request per call. */
int bnl_indep_test_mc(bnl_ctx *ctx, bnl_data data, int test_e, int B, double alpha, bnl_unif_fn unif,
                      void *unif_state, int nx, const int *x, int y, const int *sx, int nsx,
                      double *pvalue, double *stat, double *df, double *test_counter);
int bnl_allsubs_test_mc(bnl_ctx *ctx, bnl_data data, int test_e, int B, bnl_unif_fn unif, void *unif_state,
                        int x, int y, const int *sx, int nsx, int nfixed, int minsize, int maxsize,
                        double alpha, double *out, int *dsep, int *ndsep, double *test_counter);
int bnl_roundrobin_test_mc(bnl_ctx *ctx, bnl_data data, int test_e, int B, bnl_unif_fn unif,
                           void *unif_state, int x, const int *z, int nz, const int *is_fixed, double alpha,
                           double *pvalue, double *test_counter);

/* score_cache_fill (src/learning/score/hc.cache.lookup.c:10-128): the score delta
 * of adding (amat[i,j] == 0) or dropping (amat[i,j] == 1) arc i -> j for every
 * j in updated[], i != j, not blacklisted, with the score-equivalence copy rule
 * (:84-89) and robust_score_difference (src/learning/score/score.delta.c:234-258).
 * Node v is data column v; amat, blmat and cache are nnodes x nnodes,
 * column-major ([i + j * nnodes] is arc i -> j); cache is updated in place;
 * reference[nnodes] holds the current per-node scores. Parents are taken in
 * increasing node order with the candidate appended last (score.delta.c:44-51).
 * test_counter += number of deltas actually scored (score.delta.c:278). */
int bnl_score_cache_fill(bnl_ctx *ctx, bnl_data data, int score_e, double k_or_iss,
                         int nnodes, const int *amat, const int *blmat, const int *updated,
                         int nupdated, int equivalence, const double *reference,
                         double *cache, double *test_counter);

/* ---- allocation-free entry point --------------------------------------------
 * This one has no counterpart in the reference. It is for callers that must not
 * let a library allocate device memory: code running inside a framework that
 * owns the card through its own pool (PyTorch's caching allocator, RMM, a
 * serving process that reserved all memory up front), where a hidden cudaMalloc
 * can fail or, at best, synchronises the whole device; and tight loops that
 * cannot afford that synchronisation. Such callers keep their data on the device
 * already, so everything here is a DEVICE pointer they own.
 *
 * One discrete CI test (mi, mi-adf, x2, x2-adf) of column 0 (x) against column 1
 * (y) given columns 2..k-1. d_cols is a HOST array of k device pointers to
 * packed uint8 columns (level code - 1, NA = 0xFF; 16-byte aligned, nobs bytes
 * each); nlevels is a host array. d_out receives 4 doubles on the device:
 * statistic, df, p-value, number of complete rows. The work is queued on
 * `stream` (a cudaStream_t, or NULL) and is NOT synchronised. On error the
 * message goes to errbuf.
 *
 * The count table lives in a fixed array inside the library instead of a
 * cudaMalloc'ed buffer, hence:
 *   - at most 65536 table cells (llx * lly * prod(levels of z)); larger tests
 *     return BNL_EUNSUPPORTED -- use bnl_ci_discrete();
 *   - each GPU that loads the library holds that array (256 KiB) until exit;
 *   - all calls share it: it is cleared at the start of every call, and calls
 *     whose GPU work may overlap (other streams, other threads) race on it, so
 *     keep them ordered;
 *   - the first call on each device looks up the array's device address; issue
 *     that first call outside any CUDA stream capture.
 * No cudaMalloc, cudaFree or allocating library routine is called. */
int bnl_ci_discrete_noalloc(int test_e, int k, const uint8_t *const *d_cols, const int *nlevels,
                            int64_t nobs, double *d_out, void *stream, char *errbuf, int errlen);

#ifdef __cplusplus
}
#endif

#endif
