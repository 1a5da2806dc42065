// bnl_internal.h -- host-side state behind the C ABI (include/bnl_b200.h).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <map>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/bnl_b200.h"

#define BNL_MAXK 16            // columns per unit: x, y and up to 14 conditioning / 15 parents
#define BNL_NA_U8 0xFFu        // NA and row padding in the packed uint8 layout
#define BNL_MAX_LEVELS 127     // bit 7 of a packed code marks NA
// Factors with 128..32767 levels are kept as uint16 (bit 15 marks NA) in a block
// of their own; a unit that touches one carries BNL_WIDE_BIT in that col[] entry
// and is counted by the generic kernel.
#define BNL_MAX_LEVELS_WIDE 32767
#define BNL_WIDE_BIT (1 << 30)
#define BNL_NA_U16 0xFFFFu
#define BNL_COL_ALIGN 256      // bytes; every packed column starts 256-byte aligned

// one unit of discrete work as the kernels see it.
struct BnlUnit {
  long long table_off;    // offset (int32 elements) of this unit's table in the scratch
  int cells;              // llx * lly * llz
  int ncols;              // columns involved
  int llx, lly, llz;      // table is n[z][x][y]
  int has_na;             // some involved column holds an NA
  int mode;               // 1: count ROW PAIRS (tables of <= 16 cells, no NA), see count_small_kernel
  int pad;
  int col[BNL_MAXK];      // packed-column slot, order: y (if any), x, z_1, z_2, ...
  int mul[BNL_MAXK];      // contribution of each column to the cell index
};

// per-device resources.
struct BnlDevice {
  int id = 0;
  cudaStream_t stream = nullptr;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;       // whole call
  cudaEvent_t evc0 = nullptr, evc1 = nullptr, eve1 = nullptr;   // count kernel, epilogue
  // grow-only scratch
  BnlUnit *d_units = nullptr; size_t units_cap = 0;
  int *d_tables = nullptr; size_t tables_cap = 0;        // int32 elements
  double *d_out = nullptr; size_t out_cap = 0;           // doubles
  int *d_gidx = nullptr; size_t gidx_cap = 0;            // Gaussian unit descriptors
  int *d_order = nullptr; size_t order_cap = 0;          // launch order -> unit slot
  int *h_order = nullptr; size_t h_order_cap = 0;        // pinned
  int2 *d_tiles = nullptr; size_t tiles_cap = 0;         // candidate-group tiles
  int2 *h_tiles = nullptr; size_t h_tiles_cap = 0;       // pinned
  uint16_t *d_basecols = nullptr; size_t basecols_cap = 0;  // base-index columns of candidate runs (count_based_kernel)
  int *d_runfirst = nullptr; size_t runfirst_cap = 0;    // first unit of each such run
  int *h_runfirst = nullptr; size_t h_runfirst_cap = 0;  // pinned
  BnlUnit *h_units = nullptr; size_t h_units_cap = 0;    // pinned
  double *h_out = nullptr; size_t h_out_cap = 0;         // pinned
  int *h_tables = nullptr; size_t h_tables_cap = 0;      // pinned: tables on their way back (bnl_counts)
  // all-pairs contraction (bnl_pairs.cu): the indicator cross-product matrix and its plan
  int *d_pairG = nullptr; size_t pairG_cap = 0;
  int *d_pairmeta = nullptr; size_t pairmeta_cap = 0;
  int *h_pairmeta = nullptr; size_t h_pairmeta_cap = 0;  // pinned
  bool pair_attr_done = false;
  bool p32_attr_done = false;                            // count_pair32_kernel opted in to 128 KB of shared memory
  int *h_gidx = nullptr; size_t h_gidx_cap = 0;          // pinned
  // registration scratch (grow-only) and a small cache of released data-set
  // buffers: cudaMalloc / cudaFree synchronise the device and can cost tens of
  // milliseconds for hundreds of MB, which would dominate re-registration
  // (boot.strength registers a fresh data frame per replicate).
  int *stage = nullptr; size_t stage_cap = 0;            // two int32 columns (double buffer)
  cudaStream_t stream2 = nullptr;                        // host-packed columns travel here
  cudaEvent_t ev_stage[2] = {nullptr, nullptr};
  uint8_t *h_pack = nullptr; size_t h_pack_cap = 0;      // pinned ring of host-packed columns
  uint8_t *h_nstage = nullptr; size_t h_nstage_cap = 0;  // pinned chunks for numeric columns from pageable memory
  std::vector<cudaEvent_t> ev_nstage;                    // one per chunk
  int *d_flags = nullptr; size_t flags_cap = 0;          // bad-code flag + per-column NA flags
  double *d_regret = nullptr;                            // NML regret table, (K + 1) x (N + 1), see bnl_regret_table()
  std::vector<std::pair<void *, size_t>> pool;           // released blocks (pointer, bytes)
  int sm_count = 148;
  int max_smem = 48 * 1024;
  bool attr_done = false;   // per-device kernel attributes set
  bool gattr_done = false;
  bool epi_attr_done = false;                            // epilogue opted in to > 48 KB of shared memory
};

// one device's copy of a data set.
struct BnlDeviceData {
  uint8_t *u8 = nullptr;       // nd columns, stride_u8 bytes apart
  double *f64 = nullptr;       // nc columns, stride_f64 doubles apart
  double *mean = nullptr;      // nc
  double *cov = nullptr;       // nc x nc, column-major, divisor nobs - 1
  uint16_t *u16 = nullptr;     // nw wide factor columns, stride_u16 elements apart
  size_t u16_bytes = 0;
  uint8_t *u4 = nullptr;       // nibble copy of the u8 block (two codes per byte, stride_u8 / 2 apart), built on first use
  size_t u4_bytes = 0;
  bool u4_refused = false;     // no room for the nibble copy on this device: not asked again
  int small_calls = 0;         // counting calls seen so far (the nibble copy waits for the second)
  double *bge_nu = nullptr;    // nc prior means of the BGe score (null: the column means)
  size_t u8_bytes = 0, f64_bytes = 0, mean_bytes = 0, cov_bytes = 0, nu_bytes = 0;
};

struct BnlDataSet {
  int ncols = 0;
  long long nobs = 0;
  int nd = 0, nc = 0, nw = 0;           // uint8 factor / numeric / uint16 factor column counts
  size_t stride_u8 = 0, stride_f64 = 0, stride_u16 = 0;
  std::vector<int> is_numeric, nlevels, slot;
  std::vector<int> is_wide;             // per column: a factor kept as uint16 (slot indexes that block)
  std::vector<int> slot_has_na;         // per packed factor column: any NA?
  std::vector<int> wide_has_na;         // the same for the uint16 block
  std::vector<BnlDeviceData> dev;       // one per context device
  std::vector<double> h_mean, h_cov;    // host copies of the Gaussian moments
  bool any_nan = false;                 // some numeric column holds a missing value
  bool committed = true;                // false between bnl_data_alloc and bnl_data_commit
  double bge_mu = 1.0, bge_w = -1.0;    // BGe imaginary sample sizes (bge_w < 0: ncols + 2)
  double ebic_gamma = 0.5;              // extended BIC (the front end's default)
  long long gauss_nobs = 0;             // > 0: rows behind moments set by bnl_data_set_moments (row shards pooled)
  std::vector<int> mask_col;            // per column: its intervention mask column (mBDe), -1 none; empty: none set
};

// bnl_score_cache_fill's candidate lists, kept between calls (cleared, not freed): the first
// iteration of hc on 1000 nodes lists 499 500 candidates, 14 MB of index arrays whose fresh
// pages cost more than filling them.
struct BnlCacheFillScratch {
  std::vector<int> target, pa_off, pa_idx, ci, cj;
  std::vector<std::pair<int, int>> copies;
  std::vector<double> score;
};

struct bnl_ctx {
  BnlCacheFillScratch cf;
  std::vector<BnlDevice> devs;
  std::map<bnl_data, BnlDataSet> data;
  bnl_data next_handle = 1;
  std::string err;
  std::mutex err_mutex;
  double last_ms = 0.0, last_count_ms = 0.0, last_epi_ms = 0.0;
  int last_launches = 0;
  std::vector<double> mc_fact;            // log(k!) for the permutation tests, grown on demand
};

int bnl_fail(bnl_ctx *ctx, int code, const char *fmt, ...);

#define BNL_CUDA(ctx, call)                                                      \
  do {                                                                           \
    cudaError_t e__ = (call);                                                    \
    if (e__ != cudaSuccess)                                                      \
      return bnl_fail((ctx), BNL_ECUDA, "%s failed: %s (%s:%d)", #call,          \
                      cudaGetErrorString(e__), __FILE__, __LINE__);              \
  } while (0)

// grow-only device / pinned buffers.
template <typename T>
int bnl_reserve_device(bnl_ctx *ctx, T **p, size_t *cap, size_t need) {
  if (need <= *cap) return BNL_OK;
  if (*p) cudaFree(*p);
  *p = nullptr; *cap = 0;
  size_t want = need + need / 4 + 64;
  cudaError_t e = cudaMalloc((void **)p, want * sizeof(T));
  if (e != cudaSuccess)
    return bnl_fail(ctx, BNL_ENOMEM, "cudaMalloc of %zu bytes failed: %s", want * sizeof(T),
                    cudaGetErrorString(e));
  *cap = want;
  return BNL_OK;
}

template <typename T>
int bnl_reserve_pinned(bnl_ctx *ctx, T **p, size_t *cap, size_t need) {
  if (need <= *cap) return BNL_OK;
  if (*p) cudaFreeHost(*p);
  *p = nullptr; *cap = 0;
  size_t want = need + need / 4 + 64;
  cudaError_t e = cudaMallocHost((void **)p, want * sizeof(T));
  if (e != cudaSuccess)
    return bnl_fail(ctx, BNL_ENOMEM, "cudaMallocHost of %zu bytes failed: %s", want * sizeof(T),
                    cudaGetErrorString(e));
  *cap = want;
  return BNL_OK;
}

// pooled device blocks (bnl_context.cu).
int bnl_pool_alloc(bnl_ctx *ctx, BnlDevice &dev, void **p, size_t bytes);
void bnl_pool_free(BnlDevice &dev, void *p, size_t bytes);

// defined in bnl_gaussian.cu: forms means and the covariance matrix on a device.
int bnl_gaussian_prepare(bnl_ctx *ctx, BnlDataSet &ds, int dev_index);

// discrete tests computed from the count tables by epilogue_kernel.
static inline bool bnl_is_discrete_test(int t) {
  return t == BNL_MI || t == BNL_MI_ADF || t == BNL_X2 || t == BNL_X2_ADF || t == BNL_JT ||
         t == BNL_MI_SH;
}

// statistic codes used inside the library only (bnl_mi_matrix): the mutual
// information COEFFICIENT instead of the G^2 statistic.
#define BNL_MI_COEF 90
#define BNL_MI_G_COEF 91
int bnl_ci_discrete_any(bnl_ctx *ctx, bnl_data data, int test_e, int n, const int *x, const int *y,
                        const int *z_off, const int *z_idx, double *stat, double *df, double *pvalue);
int bnl_ci_gaussian_any(bnl_ctx *ctx, bnl_data data, int test_e, int n, const int *x, const int *y,
                        const int *z_off, const int *z_idx, double *stat, double *df, double *pvalue);

int bnl_ci_gaussian_mc_ex(bnl_ctx *ctx, bnl_data data, int test_e, int B, double alpha, int n, const int *x,
                          const int *y, const int *z_off, const int *z_idx, bnl_unif_fn unif, void *unif_state,
                          double *stat, double *pvalue, const int *mask_cols, int nmask);

int bnl_score_cache_fill_np(bnl_ctx *ctx, bnl_data data, int score_e, double k_or_iss, int nnodes,
                            const int *amat, const int *blmat, const int *updated, int nupdated,
                            int equivalence, const double *reference, double *cache,
                            double *test_counter, const int *nparents);

// fNML / qNML: the multinomial regret table of src/scores/nml.regret.table.c
// (n <= 1000, k <= 100), built on the host at first use and copied to `dev`.
#define BNL_REGRET_N 1000
#define BNL_REGRET_K 100
int bnl_regret_table(bnl_ctx *ctx, BnlDevice &dev);

// Host-side time accounting for development (BNL_TRACE=1: totals per label go to stderr at
// bnl_shutdown). Zero cost when off apart from one branch per scope.
struct BnlTraceSlot { const char *label; double seconds; long long calls; };
extern BnlTraceSlot g_bnl_trace[32];
extern int g_bnl_trace_on;             // -1 unknown, 0 off, 1 on
int bnl_trace_slot(const char *label);
struct BnlTraceScope {
  int slot; double t0;
  explicit BnlTraceScope(const char *label);
  ~BnlTraceScope();
};
#define BNL_TRACE(label) BnlTraceScope bnl_trace_scope_##__LINE__(label)

// per-call timing of one device block.
struct BnlTiming { float total = 0.f, count = 0.f, epi = 0.f; int launches = 0; };

// All-pairs contraction on the tensor cores (bnl_pairs.cu, kernels in bnl_pairs.cuh): decide whether the pair units of a batch are worth one
// contraction, run it and fill their tables. picked[] lists positions in h_units.
bool bnl_pairs_eligible(const BnlUnit &u, const std::vector<int> &lev_by_slot);
// used[slot]: the packed columns the units touch; picked: positions of the pair units in
// the batch's unit array (null: all nunits units, in order); cc: those units' column pairs
// (host copy), to leave out output tiles that none of them needs.
int bnl_pairs_run(bnl_ctx *ctx, BnlDevice &dev, const BnlDataSet &ds, const BnlDeviceData &dd,
                  const std::vector<char> &used, const std::vector<int> *picked, int nunits,
                  const std::vector<int> &lev_by_slot, BnlTiming *tm, const int2 *cc = nullptr);
// estimated seconds of the contraction for these units, and of counting them one by one
void bnl_pairs_cost(const BnlDevice &dev, const BnlDataSet &ds, const std::vector<char> &used, long long nunits,
                    const std::vector<int> &lev_by_slot, double *tensor_s, double *atomic_s,
                    const int2 *cc = nullptr);

// split [0, n) into ndev contiguous blocks.
static inline void bnl_block(int n, int ndev, int d, int *begin, int *end) {
  long long b = (long long)n * d / ndev, e = (long long)n * (d + 1) / ndev;
  *begin = (int)b; *end = (int)e;
}
