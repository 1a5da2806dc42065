// bnl_discrete.cu -- host side of the discrete entry points: unit validation,
// batching, multi-device dispatch, launches.
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <memory>
#include <thread>

#include "bnl_discrete_kernels.cuh"

namespace {

// tuning knobs for experiments (read once): BNL_PAIR=0 disables row-pair
// counting, BNL_CHUNK_ROWS overrides the rows per CTA of the small-table kernel,
// BNL_L2_MB the L2 budget of the chunking rule.
long long env_ll(const char *name, long long dflt) {
  const char *v = getenv(name);
  return (v && *v) ? atoll(v) : dflt;
}
const long long kKnobPair = env_ll("BNL_PAIR", 1);
const long long kKnobChunkRows = env_ll("BNL_CHUNK_ROWS", 0);
const long long kKnobL2MB = env_ll("BNL_L2_MB", 192);
const int kKnobMidThreads = (int)env_ll("BNL_MID_THREADS", 1024);
const long long kKnobGroup = env_ll("BNL_GROUP", 1);
const long long kKnobGroupMinCols = env_ll("BNL_GROUP_MINCOLS", 3);
// conditional tests that share (y, x[, z1]) and differ in the last conditioning column: 0 off,
// 1-3: runs whose tables fit 16 / 32 / 64 KiB of lane counters per unit
const long long kKnobTestGroup = env_ll("BNL_TEST_GROUP", 0);
const long long kKnobGroupWaves = env_ll("BNL_GROUP_WAVES", 0);   // 0: pick row chunks by the cost model
const long long kKnobGroupUnits = env_ll("BNL_GROUP_UNITS", BNL_GROUP_MAX);
// DRAM -> L2 prefetch of the candidates' own columns, in thread-passes ahead (0: off). Two passes
// ahead is worth ~9 % on the base-index kernel with one unit per CTA and nothing on the group
// kernel (profiles/r1b_l2_prefetch_sweep.log).
const int kKnobPrefetch = (int)env_ll("BNL_L2_PREFETCH", 2);
const int kKnobGroupPrefetch = (int)env_ll("BNL_GROUP_L2_PREFETCH", 0);
const long long kKnobGroupMaxCells = env_ll("BNL_GROUP_MAXCELLS", 0);   // 0: whatever four tables per SM allow
// units per CTA of the base-index kernel. One unit and 1024 threads beat tiles of 2-8 units of
// 256-512 threads that share the base column (profiles/r1b_l2_prefetch_sweep.log).
const long long kKnobBasedNuCap = env_ll("BNL_BASED_NU_CAP", 1);
const long long kKnobBasedMerge = env_ll("BNL_BASED_MERGE", 2);
const long long kKnobBasedU8 = env_ll("BNL_BASED_U8", 1);
const long long kKnobBasedNT = env_ll("BNL_BASED_NT", 512);
const long long kKnobBased = env_ll("BNL_BASED", 1);
const long long kKnobBasedMinCols = env_ll("BNL_BASED_MINCOLS", 4);
const long long kKnobBasedWaves = env_ll("BNL_BASED_WAVES", 0);
const size_t kMaxBaseScratch = (size_t)1 << 30;   // bytes of base-index columns per launch
// all-pairs contraction on the tensor cores (bnl_pairs.cu): 0 off, 1 when the cost model prefers it,
// 2 whenever eligible. Read at every call, so that a test can switch it inside one process.
long long knob_pairs() { return env_ll("BNL_PAIRS", 1); }
// nibble copy of the factor columns for the small-table kernel (count_small_kernel<false>): 0 off;
// otherwise the least number of rows from which it is made (first use on a data set and device).
long long knob_nibbles() { return env_ll("BNL_NIBBLES", 262144); }
// ... and the number of earlier counting calls on the data set before it is made: a data set that is
// registered, counted once and released (bench.py's end-to-end step) would pay for the copy (0.1 ms
// per 370 MB) about what it saves.
long long knob_nibbles_after() { return env_ll("BNL_NIBBLES_AFTER", 1); }
// tables of at most 32 cells from the nibble copy as row pairs (count_pair32_kernel): 0 off, 1 the units
// that no group takes, 2 also the candidates of a score batch that the group kernel would take
long long knob_pair32() { return env_ll("BNL_PAIR32", 1); }

const long long kMaxBatchCells = 48ll << 20;   // int32 cells of table scratch per launch (192 MiB)
const int kMaxBatchUnits = 1 << 19;           // the first hc iteration on 1000 nodes (499 500 units) is one batch

BnlDataSet *find_data(bnl_ctx *ctx, bnl_data h) {
  auto it = ctx->data.find(h);
  return it == ctx->data.end() ? nullptr : &it->second;
}

// Build the kernel-side description of one unit. Columns: y (optional), x, then z.
int make_unit(bnl_ctx *ctx, const BnlDataSet &ds, int x, int y, const int *z, int nz, BnlUnit *u) {
  if (nz + 2 > BNL_MAXK)
    return bnl_fail(ctx, BNL_EUNSUPPORTED, "a unit may involve at most %d columns (got %d)",
                    BNL_MAXK, nz + 2);
  auto check = [&](int c) {
    return c >= 0 && c < ds.ncols && !ds.is_numeric[c];
  };
  if (!check(x) || (y >= 0 && !check(y)))
    return bnl_fail(ctx, BNL_EINVAL, "unit refers to column %d / %d: out of range or not a factor", x, y);
  memset(u, 0, sizeof(*u));
  int k = 0;
  long long mul = 1;
  // packed slot of a column, with BNL_WIDE_BIT for the uint16 block; NA flag alongside
  auto place = [&](int c, long long m) {
    const bool wide = ds.is_wide[c] != 0;
    u->col[k] = ds.slot[c] | (wide ? BNL_WIDE_BIT : 0);
    u->mul[k] = (int)m;
    k++;
    if (wide) { u->pad = 1; if (!ds.wide_has_na.empty() && ds.wide_has_na[ds.slot[c]]) u->has_na = 1; }
    else if (!ds.slot_has_na.empty() && ds.slot_has_na[ds.slot[c]]) u->has_na = 1;
  };
  u->llx = ds.nlevels[x];
  u->lly = y >= 0 ? ds.nlevels[y] : 1;
  if (y >= 0) { place(y, 1); mul = u->lly; }
  place(x, mul);
  mul *= u->llx;
  if (mul >= INT32_MAX)
    return bnl_fail(ctx, BNL_EOVERFLOW, "attempting to create a factor with more than INT_MAX levels");
  long long llz = 1;
  for (int j = 0; j < nz; j++) {
    if (!check(z[j]))
      return bnl_fail(ctx, BNL_EINVAL, "unit refers to column %d: out of range or not a factor", z[j]);
    llz *= ds.nlevels[z[j]];
    // the reference refuses configuration counts that do not fit an int
    // (src/core/sets.c:182-183).
    if (llz >= INT32_MAX || mul * ds.nlevels[z[j]] >= INT32_MAX)
      return bnl_fail(ctx, BNL_EOVERFLOW, "attempting to create a factor with more than INT_MAX levels");
    place(z[j], mul);
    mul *= ds.nlevels[z[j]];
  }
  u->ncols = k;
  u->llz = (int)llz;
  u->cells = (int)mul;
  if (mul > kMaxBatchCells)
    return bnl_fail(ctx, BNL_EUNSUPPORTED, "a single table of %lld cells exceeds the engine's scratch limit", mul);
  return BNL_OK;
}

// shared memory of the epilogue: llx + lly margins per warp (tests only). Up to
// kEpiMaxMargins ints per warp; the opt-in above 48 KB is set once per device.
const int kEpiMaxMargins = 10240;
template <typename Units>
int epilogue_stride(bnl_ctx *ctx, BnlDevice &dev, const Units &units, int n, int kind, int *stride) {
  int need = 2 * BNL_MAX_LEVELS + 2;
  if (kind == 0)
    for (int i = 0; i < n; i++) need = std::max(need, units[i].llx + units[i].lly);
  if (need > kEpiMaxMargins)
    return bnl_fail(ctx, BNL_EUNSUPPORTED, "a test on factors with %d levels in total exceeds the epilogue's limit of %d",
                    need, kEpiMaxMargins);
  *stride = (need + 31) & ~31;
  if ((size_t)*stride * BNL_EPI_WARPS * sizeof(int) > 48 * 1024 && !dev.epi_attr_done) {
    BNL_CUDA(ctx, cudaFuncSetAttribute(epilogue_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       kEpiMaxMargins * BNL_EPI_WARPS * (int)sizeof(int)));
    BNL_CUDA(ctx, cudaFuncSetAttribute(epilogue_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       kEpiMaxMargins * BNL_EPI_WARPS * (int)sizeof(int)));
    dev.epi_attr_done = true;
  }
  return BNL_OK;
}

// One warp can do a unit's whole epilogue, in the order the four-warp CTA would
// (epilogue_kernel<EXT, true>): tests in a single stratum, and node scores that take
// the one-lane-per-stratum route with at most 32 strata or have a single stratum.
bool epilogue_warp_ok(int kind, int code, int llx, int lly, int llz) {
  if (kind == 0) return llz == 1 && code != BNL_JT && code != BNL_MI_SH;
  if (code == BNL_BDS) return false;
  return llz == 1 || (llx * lly <= 16 && llz <= 32);
}

struct Request {
  int kind;        // 0 tests, 1 scores, 2 raw tables
  int code;
  double arg;
  int n;
  std::vector<BnlUnit> units;
  double *o0, *o1, *o2;          // per-unit outputs (may be null)
  int32_t *tables;               // kind 2
  const int64_t *table_off;
};

// Run units [b, e) of the request on device slot d, in sub-batches bounded by
// the scratch size. Everything is queued on the device's stream; the caller
// synchronises. Sub-batches reuse the pinned staging buffers, so each one is
// synchronised before the next is staged.
int run_block(bnl_ctx *ctx, BnlDataSet &ds, int d, Request &rq, int b, int e, BnlTiming *tm) {
  BnlDevice &dev = ctx->devs[d];
  BnlDeviceData &dd = ds.dev[d];
  BNL_CUDA(ctx, cudaSetDevice(dev.id));
  if (!dev.attr_done) {
    BNL_CUDA(ctx, cudaFuncSetAttribute(count_units_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       dev.max_smem - 4096));
    BNL_CUDA(ctx, cudaFuncSetAttribute(count_small_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       dev.max_smem - 4096));
    BNL_CUDA(ctx, cudaFuncSetAttribute(count_small_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       dev.max_smem - 4096));
    BNL_CUDA(ctx, cudaFuncSetAttribute(count_group_kernel<512, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       dev.max_smem - 1024));
    BNL_CUDA(ctx, cudaFuncSetAttribute(count_group_kernel<512, 16384>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       dev.max_smem - 1024));
    BNL_CUDA(ctx, cudaFuncSetAttribute(count_group_kernel<512, 32768>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       dev.max_smem - 1024));
    BNL_CUDA(ctx, cudaFuncSetAttribute(count_lane_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       dev.max_smem - 4096));
    BNL_CUDA(ctx, cudaFuncSetAttribute(count_based_kernel<5, 8, 256>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       dev.max_smem - 4096));
    BNL_CUDA(ctx, cudaFuncSetAttribute(count_based_kernel<5, 4, 512>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       dev.max_smem - 4096));
    BNL_CUDA(ctx, cudaFuncSetAttribute(count_based_kernel<5, 1, 1024>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       dev.max_smem - 4096));
    BNL_CUDA(ctx, cudaFuncSetAttribute(count_based_kernel<5, 1, 512, true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       dev.max_smem - 4096));
    BNL_CUDA(ctx, cudaFuncSetAttribute(count_based_kernel<5, 1, 1024, true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       dev.max_smem - 4096));
    BNL_CUDA(ctx, cudaFuncSetAttribute(count_based_kernel<4, 1, 1024>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       dev.max_smem - 4096));
    BNL_CUDA(ctx, cudaFuncSetAttribute(count_based_kernel<0, 1, 1024>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       dev.max_smem - 4096));
    BNL_CUDA(ctx, cudaFuncSetAttribute(count_based_kernel<3, 1, 1024>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       dev.max_smem - 4096));
    dev.attr_done = true;
  }
  const int smem_cap_cells = (dev.max_smem - 8192) / 4;
  const long long groups = (ds.nobs + 15) / 16;
  std::vector<int> pair_lev;
  std::vector<char> pair_used;
  int pos = b;
  while (pos < e) {
    BNL_TRACE("run_block: whole sub-batch");
    long long cells = 0;
    int cnt = 0;
    while (pos + cnt < e && cnt < kMaxBatchUnits && cells + rq.units[pos + cnt].cells <= kMaxBatchCells) {
      cells += rq.units[pos + cnt].cells;
      cnt++;
    }
    if (cnt == 0) return bnl_fail(ctx, BNL_EUNSUPPORTED, "table too large for the scratch buffer");
    int rc;
    std::unique_ptr<BnlTraceScope> tr_cls(new BnlTraceScope("run_block: classify"));
    if ((rc = bnl_reserve_pinned(ctx, &dev.h_units, &dev.h_units_cap, (size_t)cnt))) return rc;
    if ((rc = bnl_reserve_device(ctx, &dev.d_units, &dev.units_cap, (size_t)cnt))) return rc;
    if ((rc = bnl_reserve_device(ctx, &dev.d_tables, &dev.tables_cap, (size_t)cells))) return rc;
    if ((rc = bnl_reserve_device(ctx, &dev.d_out, &dev.out_cap, (size_t)cnt * 4))) return rc;
    if ((rc = bnl_reserve_pinned(ctx, &dev.h_out, &dev.h_out_cap, (size_t)cnt * 4))) return rc;

    long long off = 0;
    for (int i = 0; i < cnt; i++) {
      dev.h_units[i] = rq.units[pos + i];
      dev.h_units[i].table_off = off;
      off += dev.h_units[i].cells;
    }
    // four size classes, one launch each (results keep the caller's order: the
    // kernels reach their units through the `order` map).
    //   S : <= 256 cells    64 lane counters per cell, byte-packed indices   (count_small_kernel<false>)
    //   M1: <= mid_cap      64 lane counters per cell, 16-bit packed indices (count_small_kernel<true>)
    //   M2: <= lane_cap     32 lane counters per cell, 16-bit packed indices (count_lane_kernel<true>)
    //   L : the rest        one shared (or global) table per CTA             (count_units_kernel)
    const int mid_cap = (dev.max_smem - 8192) / 256;
    const int lane_cap = (dev.max_smem - 8192) / 128;
    // candidate groups: runs of consecutive units that share every column but the
    // last (score_cache_fill's candidates of one node), small tables, no NA. They
    // are cut into tiles of up to BNL_GROUP_MAX units whose tables fit one SM.
    std::vector<char> grouped(cnt, 0);
    // Units that are one- or two-way tables of small factors: when the batch asks for
    // many pairs of few columns (first hc iteration, pc.stable level 0, mi.matrix) they
    // all come out of ONE indicator cross-product on the tensor cores (bnl_pairs.cuh)
    // instead of one row pass each.
    std::vector<int> pair_units;
    const long long kKnobPairs = knob_pairs();
    if (kKnobPairs && ds.nobs >= 4096) {
      std::vector<int> lev_by_slot(ds.nd, 0);
      for (int c = 0; c < ds.ncols; c++)
        if (!ds.is_numeric[c] && !ds.is_wide[c]) lev_by_slot[ds.slot[c]] = ds.nlevels[c];
      for (int i = 0; i < cnt; i++)
        if (bnl_pairs_eligible(dev.h_units[i], lev_by_slot)) pair_units.push_back(i);
      bool take = pair_units.size() >= 32;
      pair_used.assign(ds.nd, 0);
      for (int i : pair_units)
        for (int k = 0; k < dev.h_units[i].ncols; k++) pair_used[dev.h_units[i].col[k]] = 1;
      if (take) {
        double tensor_s, atomic_s;
        bnl_pairs_cost(dev, ds, pair_used, (long long)pair_units.size(), lev_by_slot, &tensor_s, &atomic_s);
        take = kKnobPairs >= 2 ? tensor_s < 1e29 : tensor_s < 0.8 * atomic_s;
      }
      if (!take) pair_units.clear();
      for (int i : pair_units) grouped[i] = 3;
      pair_lev.swap(lev_by_slot);
    }
    std::vector<std::vector<int>> runs, bruns;
    // base-column classes: counters per cell / units per tile (count_based_kernel)
    const int kBasedClasses = 6;
    const int based_cap[kBasedClasses] = {(dev.max_smem - 8192) / 128 / 8,
                                          (dev.max_smem - 8192) / 128 / 4, (dev.max_smem - 8192) / 128 / 2,
                                          (dev.max_smem - 8192) / 128, (dev.max_smem - 8192) / 64,
                                          (dev.max_smem - 8192) / 32};
    int based_nu[kBasedClasses] = {8, 4, 2, 1, 1, 1};
    for (int k = 0; k < kBasedClasses; k++) based_nu[k] = std::min(based_nu[k], (int)kKnobBasedNuCap);
    std::vector<int> bcls[kBasedClasses], based_run(cnt, -1), based_lc(cnt, 5);
    int merged_bytes = 0;                     // largest table of the merged launch (BNL_BASED_MERGE=2)
    std::vector<int2> btiles[kBasedClasses];
    int bmax[kBasedClasses] = {0, 0, 0, 0, 0, 0};
    const size_t bstride = (size_t)((ds.nobs + 15) / 16) * 16 + 128;   // uint16 elements per base column
    const size_t brun_bytes = bstride * sizeof(uint16_t);
    std::vector<int2> tiles;
    std::vector<int> gorder;
    int gstride = 0;
    // seven 32 KiB tables must fit: the group kernel has ~0.5 KiB of static shared memory
    const int group_smem = dev.max_smem - 1024;
    // Score requests only: on pc.stable's mixed test batches (many tiny tables that
    // count row pairs, 29 units per distinct column kept hot in L2) the extra launch
    // and the lost pair mode cost more than the shared loads save (config B: 2.58 ms
    // ungrouped, 2.64-2.79 ms grouped).
    // Row chunks of the one-CTA-per-SM kernels: nt tiles of `upt` units each are cut
    // into ch chunks; the launch takes waves x (tile time / ch + fixed cost of a CTA:
    // clearing and collapsing ~200 KB of tables, ~4 us). 166 tiles in one chunk would
    // run as two waves, the second with 18 CTAs.
    auto model_chunks = [&](long long nt, double upt, long long maxch, double *cost) {
      const double unit_us = (double)ds.nobs / 25e3, fixed_us = 4.0;    // ~25 G row-units/s per SM
      long long best = 1;
      double bc = 1e300;
      for (long long ch = 1; ch <= std::max(1ll, maxch); ch++) {
        const long long waves = (nt * ch + dev.sm_count - 1) / dev.sm_count;
        const double c = (double)waves * (upt * unit_us / (double)ch + fixed_us);
        if (c < bc) { bc = c; best = ch; }
      }
      if (cost) *cost = bc;
      return best;
    };
    // (candidates of at most 32 cells go to the row-pair kernel instead once the data set has its nibble copy)
    std::vector<unsigned char> nib_on(ds.nd, 0);
    for (int c = 0; c < ds.ncols; c++)
      if (!ds.is_numeric[c] && !ds.is_wide[c]) nib_on[ds.slot[c]] = ds.nlevels[c] <= 15;
    const bool nib_seen_before = dd.u4 || dd.small_calls >= knob_nibbles_after();   // (the copy waits for a data set's second call)
    dd.small_calls++;
    const bool p32_scores = knob_pair32() >= 2 && nib_seen_before && !dd.u4_refused && knob_nibbles() > 0 &&
                            ds.nobs >= knob_nibbles() && dev.max_smem >= 132 * 1024;
    if (kKnobGroup && rq.kind == 1 && ds.nobs >= 65536) {
      auto same_base = [&](const BnlUnit &a, const BnlUnit &b) {
        // at least two shared columns (a node with one or more parents): for
        // two-column units the per-unit kernel with its row-pair mode is faster
        // (scripts/group_sweep.sh: 0.244 vs 0.297 ms on 999 candidates x 1 M rows).
        if (a.ncols != b.ncols || a.ncols < kKnobGroupMinCols || a.has_na || b.has_na || a.pad || b.pad) return false;
        for (int k = 0; k + 1 < a.ncols; k++)
          if (a.col[k] != b.col[k] || a.mul[k] != b.mul[k]) return false;
        return a.mul[a.ncols - 1] == b.mul[b.ncols - 1];
      };
      // maximal runs of consecutive units with the same base columns; inside a run
      // the small tables (at least four of them per SM) go to the group kernel, the
      // mid-size ones to the base-column kernel (count_based_kernel).
      int gcap = group_smem / (4 * 256);
      if (kKnobGroupMaxCells > 0) gcap = std::min(gcap, (int)kKnobGroupMaxCells);
      const int bcap = based_cap[kBasedClasses - 1];       // 8 lane counters per cell
      int i = 0;
      while (i < cnt) {
        int j = i + 1;
        while (j < cnt && same_base(dev.h_units[i], dev.h_units[j])) j++;
        if (j - i >= 4) {
          std::vector<int> small, big;
          for (int u = i; u < j; u++) {
            const BnlUnit &hu = dev.h_units[u];
            if (p32_scores && hu.cells <= 32 && !hu.has_na && !hu.pad) {
              bool ok = true;
              for (int k = 0; k < hu.ncols; k++) ok = ok && nib_on[hu.col[k]];
              if (ok) continue;                      // left to the row-pair kernel (count_pair32_kernel)
            }
            const bool based_ok = kKnobBased && hu.ncols >= kKnobBasedMinCols && hu.cells <= bcap;
            // tables past 32 KiB at 64 counters per cell: four of them per SM in the group
            // kernel, eight at 32 counters per cell in the base-index kernel (3 % faster)
            if (hu.cells <= gcap && !(based_ok && hu.cells > 128)) small.push_back(u);
            else if (based_ok) big.push_back(u);
          }
          if (small.size() >= 4) runs.push_back(std::move(small));
          // the base-index column costs 2 bytes per row of scratch and one extra pass
          if (big.size() >= 4 && (bruns.size() + 1) * brun_bytes <= kMaxBaseScratch && bruns.size() < 65000)
            bruns.push_back(std::move(big));          // (grid.y of combine_base_kernel)
        }
        i = j;
      }
      // units per tile: more units share more loads, but the CTAs (tiles x row
      // chunks, one per SM at a time) should also fill whole waves -- 999
      // candidates make 125 tiles of 8 (84 % of 148 SMs) or 143 tiles of 7 (97 %).
      // one table stride for the whole launch (count_group_kernel): the largest table.
      for (auto &r : runs)
        for (int u : r) gstride = std::max(gstride, dev.h_units[u].cells * 256);
      // rounded up to 16 or 32 KiB where possible: those strides are compiled in, which
      // turns the table offset into an immediate of the atomics' address.
      if (gstride > 0 && gstride <= 16384) gstride = 16384;
      else if (gstride > 16384 && gstride <= 32768) gstride = 32768;
      const bool small_block = (unsigned long long)ds.stride_u8 * (unsigned long long)ds.nd < (1ull << 35);
      const int fit = gstride > 0 && small_block ? std::min(BNL_GROUP_MAX, group_smem / gstride) : 0;
      if (fit < 4) runs.clear();               // tables too large to share an SM usefully
      int best_u = std::min((int)kKnobGroupUnits, fit);
      if (!getenv("BNL_GROUP_UNITS") && !runs.empty()) {
        double best_cost = 1e300;
        for (int u = fit; u >= std::min(fit, 5); u--) {
          long long nt = 0;
          for (auto &r : runs) nt += ((long long)r.size() + u - 1) / u;
          double cost;
          if (kKnobGroupWaves) {
            const long long ch = std::max(1ll, (kKnobGroupWaves * dev.sm_count + nt / 2) / nt);
            const long long ctas = nt * ch;
            const double eff = (double)ctas / (double)(((ctas + dev.sm_count - 1) / dev.sm_count) * dev.sm_count);
            cost = (1.0 + 0.08 * (double)(BNL_GROUP_MAX - u)) / eff;
          } else {
            // fewer units per tile: the shared columns are loaded more often per unit
            model_chunks(nt, (double)u * (1.0 + 0.08 * (double)(BNL_GROUP_MAX - u)), std::max(1ll, groups / (32 * 512)), &cost);
          }
          if (cost < best_cost) { best_cost = cost; best_u = u; }
        }
      }
      for (auto &r : runs) {
        size_t t0 = 0;
        while (t0 < r.size()) {
          const size_t t1 = std::min(r.size(), t0 + (size_t)best_u);
          tiles.push_back(make_int2((int)gorder.size(), (int)(t1 - t0)));
          for (size_t q = t0; q < t1; q++) { gorder.push_back(r[q]); grouped[r[q]] = 1; }
          t0 = t1;
        }
      }
      // (the register-operand shift costs ~3 %: only when 16- or 8-counter tables are present)
      bool mixed_lc = false;
      for (auto &r : bruns)
        for (int u : r) mixed_lc = mixed_lc || dev.h_units[u].cells > based_cap[3];
      // base-column runs: per class, tiles of units of one run
      for (size_t r = 0; r < bruns.size(); r++) {
        std::vector<int> members[kBasedClasses];
        for (int u : bruns[r]) {
          const int c = dev.h_units[u].cells;
          int k = 0;
          while (c > based_cap[k]) k++;
          // one unit per CTA: every table with 32 counters per cell runs the same kernel, one
          // CTA per SM whatever its size, so they share one launch (no partial wave in between)
          if (kKnobBasedMerge && based_nu[k] == 1 && k < 3) k = 3;
          based_lc[u] = k <= 3 ? 5 : (k == 4 ? 4 : 3);
          // ... and with the counters per cell read from the unit, the 16- and 8-counter tables too
          if (kKnobBasedMerge >= 2 && based_nu[k] == 1 && mixed_lc) {
            merged_bytes = std::max(merged_bytes, c << (based_lc[u] + 2));
            k = 3;
          }
          members[k].push_back(u);
          bmax[k] = std::max(bmax[k], c);
          grouped[u] = 2;
          based_run[u] = (int)r;
        }
        for (int k = 0; k < kBasedClasses; k++)
          for (size_t t0 = 0; t0 < members[k].size(); t0 += based_nu[k]) {
            const size_t t1 = std::min(members[k].size(), t0 + (size_t)based_nu[k]);
            btiles[k].push_back(make_int2((int)bcls[k].size(), (int)(t1 - t0)));
            for (size_t q = t0; q < t1; q++) bcls[k].push_back(members[k][q]);
          }
      }
    }
    // Conditional TESTS of a pc.stable level arrive in the same shape: the tests of one edge
    // (x, y) against each neighbour z -- or against (z1, z) for each z -- share every column but
    // the last (ct_discrete forms the configuration of the shared part once for all of them,
    // src/tests/patterns/ctest.c:11-55). Runs are short (2-5 units), so one run = one tile of
    // count_group_kernel; three launches by table size (16, 32, 64 KiB of lane counters per unit).
    const int kTestClasses = 3;
    const int tg_stride[kTestClasses] = {16384, 32768, 65536};
    std::vector<int2> ttiles[kTestClasses];
    std::vector<int> torder[kTestClasses];
    if (kKnobTestGroup && rq.kind == 0 && ds.nobs >= 65536 &&
        (unsigned long long)ds.stride_u8 * (unsigned long long)ds.nd < (1ull << 35)) {
      auto same_tbase = [&](const BnlUnit &a, const BnlUnit &b) {
        if (a.ncols != b.ncols || a.ncols < 3 || a.has_na || b.has_na || a.pad || b.pad) return false;
        for (int k = 0; k + 1 < a.ncols; k++)
          if (a.col[k] != b.col[k] || a.mul[k] != b.mul[k]) return false;
        return a.mul[a.ncols - 1] == b.mul[b.ncols - 1];
      };
      int i = 0;
      while (i < cnt) {
        int j = i + 1;
        if (!grouped[i])
          while (j < cnt && j - i < BNL_GROUP_MAX && !grouped[j] && same_tbase(dev.h_units[i], dev.h_units[j])) j++;
        if (j - i >= 2) {
          int cmaxr = 0;
          for (int u = i; u < j; u++) cmaxr = std::max(cmaxr, dev.h_units[u].cells);
          int k = 0;
          while (k < kTestClasses && cmaxr * 256 > tg_stride[k]) k++;
          // the run's tables must fit one SM next to the kernel's own ~1 KiB
          if (k < kTestClasses && k < kKnobTestGroup && (j - i) * tg_stride[k] <= group_smem) {
            ttiles[k].push_back(make_int2((int)torder[k].size(), j - i));
            for (int u = i; u < j; u++) { torder[k].push_back(u); grouped[u] = 4; }
          }
        }
        i = j;
      }
    }
    std::vector<int> cls[4], p32;             // p32: units of <= 32 cells counted as row pairs from the nibble copy
    int cmax[4] = {0, 0, 0, 0};
    // byte base indices (count_based_kernel<.., B8>): when EVERY base-column run of the sub-batch has at
    // most 256 base cells (the candidates of one node, or of the two nodes a reversal touches, usually
    // agree); one unit per CTA only; BNL_BASED_U8=0 keeps them 16-bit
    bool all_u8 = kKnobBasedU8 && kKnobBasedNuCap == 1 && !bruns.empty();
    for (int i = 0; i < cnt && all_u8; i++)
      if (grouped[i] == 2 && dev.h_units[i].mul[dev.h_units[i].ncols - 1] > 256) all_u8 = false;
    for (int i = 0; i < cnt; i++) {
      BnlUnit &hu = dev.h_units[i];
      hu.mode = grouped[i] == 2 ? (16 + based_run[i]) | (based_lc[i] << 24) : 0;   // base-column units: run, counters per cell
      // ... and bit 30: the run's base table (every column but the candidate's) has at most 256 cells,
      // its indices are bytes (one unit per CTA only; BNL_BASED_U8=0 keeps them 16-bit)
      if (grouped[i] == 2 && all_u8) hu.mode |= 1 << 30;
      if (grouped[i]) continue;
      const int c = hu.cells;
      // a unit with a uint16 column is counted by the generic kernel, whatever its size
      const int k = hu.pad ? 3 : (c <= 256 ? 0 : (c <= mid_cap ? 1 : (c <= lane_cap ? 2 : 3)));
      cls[k].push_back(i);
      cmax[k] = std::max(cmax[k], c);
    }
    // tiny tables without NA count row pairs (16 x 16 pair table = 256 slots per
    // CTA). Worth it when they make up at least half of the small-table launch; a few of
    // them among bigger tables would only cost an extra launch or extra shared
    // memory (scripts/microbench_c.py).
    if (kKnobPair && ds.nobs >= 4096 && !cls[0].empty()) {
      size_t eligible = 0;
      for (int i : cls[0]) eligible += dev.h_units[i].cells <= 16 && !dev.h_units[i].has_na;
      if (eligible * 2 >= cls[0].size()) {
        for (int i : cls[0])
          if (dev.h_units[i].cells <= 16 && !dev.h_units[i].has_na) dev.h_units[i].mode = 1;
        cmax[0] = 256;
      }
    }
    // The rest of the small-table class reads the nibble copy of its columns when every one of
    // them has one (at most 15 levels) and none holds NA: half the bytes through the LSU pipe.
    // The copy is made the first time a data set of enough rows is counted this way.
    {
      const long long nib_min = knob_nibbles();
      if (nib_min > 0 && ds.nobs >= nib_min && !cls[0].empty() && nib_seen_before) {
        std::vector<unsigned char> on(ds.nd, 0);
        for (int c = 0; c < ds.ncols; c++)
          if (!ds.is_numeric[c] && !ds.is_wide[c]) on[ds.slot[c]] = ds.nlevels[c] <= 15;
        int wanted = 0;
        for (int i : cls[0]) {
          BnlUnit &hu = dev.h_units[i];
          if (hu.mode != 0 || hu.has_na) continue;
          bool ok = true;
          for (int k = 0; k < hu.ncols; k++) ok = ok && on[hu.col[k]];
          if (ok) { hu.mode = 2; wanted++; }
        }
        if (wanted && !dd.u4 && !dd.u4_refused) {
          // not at the price of the data set's other needs: the copy must fit twice over in what is free
          size_t free_b = 0, total_b = 0;
          cudaMemGetInfo(&free_b, &total_b);
          for (auto &blk : dev.pool) free_b += blk.second;
          if ((ds.stride_u8 / 2) * (size_t)ds.nd * 2 > free_b) dd.u4_refused = true;
        }
        if (wanted && !dd.u4 && dd.u4_refused) {
          for (int i : cls[0]) if (dev.h_units[i].mode == 2) dev.h_units[i].mode = 0;
          wanted = 0;
        }
        if (wanted && !dd.u4) {
          BNL_TRACE("run_block: nibble copy made");
          const size_t stride4 = ds.stride_u8 / 2;
          dd.u4_bytes = stride4 * ds.nd;
          if ((rc = bnl_pool_alloc(ctx, dev, (void **)&dd.u4, dd.u4_bytes))) return rc;
          if ((rc = bnl_reserve_device(ctx, &dev.d_flags, &dev.flags_cap, (size_t)ds.nd / 4 + 2))) return rc;
          const unsigned char *d_on = reinterpret_cast<const unsigned char *>(dev.d_flags);
          BNL_CUDA(ctx, cudaMemcpyAsync(dev.d_flags, on.data(), (size_t)ds.nd, cudaMemcpyHostToDevice, dev.stream));
          const unsigned gx = (unsigned)std::max<long long>(1, std::min<long long>(((long long)(stride4 / 16) + 255) / 256,
                                                                                   (long long)dev.sm_count * 4));
          bnl_nibbles_kernel<<<dim3(gx, (unsigned)ds.nd), 256, 0, dev.stream>>>(dd.u8, ds.stride_u8, d_on, dd.u4, stride4);
          BNL_CUDA(ctx, cudaGetLastError());
          BNL_CUDA(ctx, cudaStreamSynchronize(dev.stream));     // `on` is a local
        }
        // ... and those of at most 32 cells count row pairs there (count_pair32_kernel), in a launch of their own
        if (wanted && dd.u4 && knob_pair32() && dev.max_smem >= 132 * 1024) {
          std::vector<int> keep;
          for (int i : cls[0]) {
            BnlUnit &hu = dev.h_units[i];
            if (hu.mode == 2 && hu.cells <= 32) { hu.mode = 3; p32.push_back(i); }
            else keep.push_back(i);
          }
          if (p32.size() < 8) {                       // not worth a launch
            for (int i : p32) dev.h_units[i].mode = 2;
            p32.clear();
          } else {
            cls[0].swap(keep);
            cmax[0] = 0;
            for (int i : cls[0]) cmax[0] = std::max(cmax[0], dev.h_units[i].mode == 1 ? 256 : dev.h_units[i].cells);
          }
        }
      }
    }
    if ((rc = bnl_reserve_pinned(ctx, &dev.h_order, &dev.h_order_cap, (size_t)cnt))) return rc;
    if ((rc = bnl_reserve_device(ctx, &dev.d_order, &dev.order_cap, (size_t)cnt))) return rc;
    int n_ungrouped = 0, n_before_tests = 0, n_before_p32 = 0;
    {
      int o = 0;
      for (int k = 0; k < 4; k++) for (int i : cls[k]) dev.h_order[o++] = i;
      n_ungrouped = o;
      for (int i : gorder) dev.h_order[o++] = i;       // group members follow, in tile order
      for (int k = 0; k < kBasedClasses; k++) for (int i : bcls[k]) dev.h_order[o++] = i;   // then the base-column classes
      n_before_tests = o;
      for (int k = 0; k < kTestClasses; k++) for (int i : torder[k]) dev.h_order[o++] = i;  // then the runs of tests
      n_before_p32 = o;
      for (int i : p32) dev.h_order[o++] = i;                                                // then the 32 x 32 row-pair units
    }
    if (!bruns.empty()) {
      if ((rc = bnl_reserve_pinned(ctx, &dev.h_runfirst, &dev.h_runfirst_cap, bruns.size()))) return rc;
      if ((rc = bnl_reserve_device(ctx, &dev.d_runfirst, &dev.runfirst_cap, bruns.size()))) return rc;
      if ((rc = bnl_reserve_device(ctx, &dev.d_basecols, &dev.basecols_cap, bruns.size() * bstride))) return rc;
      for (size_t r = 0; r < bruns.size(); r++) dev.h_runfirst[r] = bruns[r][0];
    }
    const size_t ngtiles = tiles.size();
    for (int k = 0; k < kBasedClasses; k++) tiles.insert(tiles.end(), btiles[k].begin(), btiles[k].end());
    const size_t n_tiles_before_tests = tiles.size();
    for (int k = 0; k < kTestClasses; k++) tiles.insert(tiles.end(), ttiles[k].begin(), ttiles[k].end());
    if (!tiles.empty()) {
      if ((rc = bnl_reserve_pinned(ctx, &dev.h_tiles, &dev.h_tiles_cap, tiles.size()))) return rc;
      if ((rc = bnl_reserve_device(ctx, &dev.d_tiles, &dev.tiles_cap, tiles.size()))) return rc;
      memcpy(dev.h_tiles, tiles.data(), tiles.size() * sizeof(int2));
    }
    // Rows per CTA. Every CTA pays a fixed price (clearing and collapsing its
    // replicated table), so chunks should be as long as load balance allows: aim
    // for ~8 waves of CTAs. CTAs run chunk-major, so the rows [chunk] of the
    // columns that several units share should survive in L2 between their uses;
    // columns used by a single unit just stream through. The kernels are bound by
    // shared-memory atomics, not DRAM, so the L2 budget is deliberately loose
    // (scripts/microbench_c.py).
    std::vector<unsigned char> uses(ds.nd);
    auto shared_cols = [&](const std::vector<int> &members) {
      std::fill(uses.begin(), uses.end(), 0);
      for (int i : members)
        for (int k = 0; k < dev.h_units[i].ncols; k++) {
          unsigned char &u = uses[dev.h_units[i].col[k]];
          if (u < 2) u++;
        }
      int n = 0;
      for (unsigned char u : uses) n += u >= 2;
      return std::max(n, 1);
    };
    auto pick_chunks = [&](const std::vector<int> &members, int ctas_per_sm) {
      const long long nunits = (long long)members.size();
      const long long want = (8ll * dev.sm_count * ctas_per_sm + nunits - 1) / nunits;
      long long rows = std::max(16384ll, ds.nobs / std::max(1ll, want));
      if (kKnobChunkRows) rows = kKnobChunkRows;
      const long long l2_rows = std::max(16384ll, (kKnobL2MB << 20) / shared_cols(members));
      rows = std::min(rows, l2_rows);
      long long ch = (ds.nobs + rows - 1) / rows;
      ch = std::min(ch, std::max(1ll, groups / 256));
      return (int)std::max(1ll, ch);
    };

    tr_cls.reset();
    std::unique_ptr<BnlTraceScope> tr_enq(new BnlTraceScope("run_block: enqueue"));
    BNL_CUDA(ctx, cudaEventRecord(dev.ev0, dev.stream));
    BNL_CUDA(ctx, cudaMemcpyAsync(dev.d_units, dev.h_units, (size_t)cnt * sizeof(BnlUnit),
                                  cudaMemcpyHostToDevice, dev.stream));
    BNL_CUDA(ctx, cudaMemcpyAsync(dev.d_order, dev.h_order, (size_t)cnt * sizeof(int),
                                  cudaMemcpyHostToDevice, dev.stream));
    if (!tiles.empty())
      BNL_CUDA(ctx, cudaMemcpyAsync(dev.d_tiles, dev.h_tiles, tiles.size() * sizeof(int2),
                                    cudaMemcpyHostToDevice, dev.stream));
    if (!bruns.empty())
      BNL_CUDA(ctx, cudaMemcpyAsync(dev.d_runfirst, dev.h_runfirst, bruns.size() * sizeof(int),
                                    cudaMemcpyHostToDevice, dev.stream));
    BNL_CUDA(ctx, cudaMemsetAsync(dev.d_tables, 0, (size_t)cells * sizeof(int), dev.stream));
    BNL_CUDA(ctx, cudaEventRecord(dev.evc0, dev.stream));
    if (!pair_units.empty() &&
        (rc = bnl_pairs_run(ctx, dev, ds, dd, pair_used, &pair_units, (int)pair_units.size(), pair_lev, tm))) return rc;
    if (groups > 0 && !bruns.empty()) {
      // base indices of every run (complete 16-row groups), then one launch per class
      const long long full = ds.nobs / 16;
      const unsigned bx = (unsigned)std::max(1ll, std::min<long long>((full + 255) / 256,
                                                                    (4ll * dev.sm_count * 8 + (long long)bruns.size() - 1) /
                                                                        (long long)bruns.size()));
      combine_base_kernel<<<dim3(bx, (unsigned)bruns.size()), 256, 0, dev.stream>>>(
          dd.u8, ds.stride_u8, full, dev.d_units, dev.d_runfirst, dev.d_basecols, bstride);
      BNL_CUDA(ctx, cudaGetLastError());
      tm->launches++;
      int bbase = n_ungrouped + (int)gorder.size();
      size_t tbase = ngtiles;
#define BNL_BASED_LAUNCH(K, LC, NU, NT) BNL_BASED_LAUNCH8(K, LC, NU, NT, false)
#define BNL_BASED_LAUNCH8(K, LC, NU, NT, B8)                                                                  \
      if (!bcls[K].empty()) {                                                                                  \
        const long long nt = (long long)btiles[K].size();                                                      \
        long long ch = kKnobBasedWaves ? std::max(1ll, (kKnobBasedWaves * dev.sm_count + nt / 2) / nt)         \
                                       : model_chunks(nt, (double)bcls[K].size() / (double)nt * 1.5,           \
                                                      ds.nobs / 131072, nullptr);                              \
        ch = std::max(1ll, std::min(ch, std::max(1ll, ds.nobs / 131072)));                                     \
        const int ts = ((bmax[K] << ((LC ? LC : 5) + 2)) + 255) & ~255;                                                   \
        count_based_kernel<LC, NU, NT, B8><<<(unsigned)(nt * ch), NT, (size_t)ts * based_nu[K], dev.stream>>>(  \
            dd.u8, ds.stride_u8, ds.nobs, dev.d_basecols, bstride, dev.d_units, dev.d_order + bbase,           \
            dev.d_tiles + tbase, dev.d_tables, (int)nt, (int)ch, ts, kKnobPrefetch);                                          \
        BNL_CUDA(ctx, cudaGetLastError());                                                                     \
        tm->launches++;                                                                                        \
        bbase += (int)bcls[K].size();                                                                          \
        tbase += btiles[K].size();                                                                             \
      }
      // one unit per tile: the 1024-thread kernel
      if (merged_bytes > 0) bmax[3] = (merged_bytes + 127) >> 7;      // stride of the merged launch, in 128-byte cells
      if (based_nu[0] == 1) { BNL_BASED_LAUNCH(0, 5, 1, 1024) } else { BNL_BASED_LAUNCH(0, 5, 8, 256) }
      if (based_nu[1] == 1) { BNL_BASED_LAUNCH(1, 5, 1, 1024) } else { BNL_BASED_LAUNCH(1, 5, 4, 512) }
      if (based_nu[2] == 1) { BNL_BASED_LAUNCH(2, 5, 1, 1024) } else { BNL_BASED_LAUNCH(2, 5, 4, 512) }
      if (merged_bytes > 0) { BNL_BASED_LAUNCH(3, 0, 1, 1024) }
      // byte base indices: two CTAs of 512 threads per SM when two of the launch's largest table fit -- one
      // counts while the other clears or collapses its table (0.299 -> 0.289 ms on a 3-parent batch; with
      // 16-bit indices the two CTAs get in each other's way, 0.340 -> 0.354)
      else if (all_u8 && kKnobBasedNT <= 512 && (size_t)bmax[3] * 128 * 2 + 4096 <= (size_t)dev.max_smem) { BNL_BASED_LAUNCH8(3, 5, 1, 512, true) }
      else if (all_u8) { BNL_BASED_LAUNCH8(3, 5, 1, 1024, true) }
      else { BNL_BASED_LAUNCH(3, 5, 1, 1024) }
      BNL_BASED_LAUNCH(4, 4, 1, 1024)
      BNL_BASED_LAUNCH(5, 3, 1, 1024)
#undef BNL_BASED_LAUNCH
#undef BNL_BASED_LAUNCH8
    }
    if (groups > 0 && ngtiles > 0) {
      // one CTA per SM (its tables fill the shared memory): ~8 waves of CTAs.
      int smem = 0;
      for (size_t t = 0; t < ngtiles; t++) smem = std::max(smem, tiles[t].y * gstride);
      const long long nt = (long long)ngtiles;
      long long want = kKnobGroupWaves ? std::max(1ll, (kKnobGroupWaves * dev.sm_count + nt / 2) / nt)
                                       : model_chunks(nt, (double)gorder.size() / (double)nt, groups / (32 * 512), nullptr);
      if (kKnobChunkRows) want = (ds.nobs + kKnobChunkRows - 1) / kKnobChunkRows;
      // at least 32 groups of 16 rows per thread-pass and chunk, so that clearing and
      // collapsing the tables stays a small part of a CTA's life.
      int ch = (int)std::max(1ll, std::min(want, std::max(1ll, groups / (32 * 512))));
      const unsigned ggrid = (unsigned)ngtiles * ch;
#define BNL_GROUP_LAUNCH(TS)                                                                          \
      count_group_kernel<512, TS><<<ggrid, 512, (size_t)smem, dev.stream>>>(                            \
          dd.u8, ds.stride_u8, ds.nobs, dev.d_units, dev.d_order + n_ungrouped, dev.d_tiles, dev.d_tables, \
          (int)ngtiles, ch, gstride, kKnobGroupPrefetch)
      if (gstride == 16384) BNL_GROUP_LAUNCH(16384);
      else if (gstride == 32768) BNL_GROUP_LAUNCH(32768);
      else BNL_GROUP_LAUNCH(0);
#undef BNL_GROUP_LAUNCH
      BNL_CUDA(ctx, cudaGetLastError());
      tm->launches++;
    }
    if (groups > 0) {
      int obase = n_before_tests;
      size_t tbase = n_tiles_before_tests;
      for (int k = 0; k < kTestClasses; k++) {
        if (ttiles[k].empty()) continue;
        const long long nt = (long long)ttiles[k].size();
        int smem = 0;
        for (auto &t : ttiles[k]) smem = std::max(smem, t.y * tg_stride[k]);
        const long long want = model_chunks(nt, (double)torder[k].size() / (double)nt, groups / (32 * 512), nullptr);
        const int ch = (int)std::max(1ll, std::min(want, std::max(1ll, groups / (32 * 512))));
        const unsigned ggrid = (unsigned)nt * ch;
#define BNL_TGROUP_LAUNCH(TS)                                                                           \
        count_group_kernel<512, TS><<<ggrid, 512, (size_t)smem, dev.stream>>>(                            \
            dd.u8, ds.stride_u8, ds.nobs, dev.d_units, dev.d_order + obase, dev.d_tiles + tbase, dev.d_tables, \
            (int)nt, ch, tg_stride[k], kKnobGroupPrefetch)
        if (k == 0) BNL_TGROUP_LAUNCH(16384);
        else if (k == 1) BNL_TGROUP_LAUNCH(32768);
        else BNL_TGROUP_LAUNCH(0);
#undef BNL_TGROUP_LAUNCH
        BNL_CUDA(ctx, cudaGetLastError());
        tm->launches++;
        obase += (int)torder[k].size();
        tbase += ttiles[k].size();
      }
    }
    if (groups > 0 && !p32.empty()) {
      if (!dev.p32_attr_done) {
        BNL_CUDA(ctx, cudaFuncSetAttribute(count_pair32_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 128 * 1024));
        dev.p32_attr_done = true;
      }
      // one CTA of 1024 threads per SM: ~8 waves of CTAs, chunks of at least 2^17 rows
      long long ch = std::max(1ll, (8ll * dev.sm_count + (long long)p32.size() - 1) / (long long)p32.size());
      ch = std::max(1ll, std::min(ch, std::max(1ll, ds.nobs / 131072)));
      count_pair32_kernel<<<(unsigned)(p32.size() * ch), 1024, 128 * 1024, dev.stream>>>(
          dd.u8, ds.stride_u8, dd.u4, ds.stride_u8 / 2, ds.nobs, dev.d_units, dev.d_order + n_before_p32, dev.d_tables,
          (int)p32.size(), (int)ch);
      BNL_CUDA(ctx, cudaGetLastError());
      tm->launches++;
    }
    if (groups > 0) {
      int base = 0;
      if (!cls[0].empty()) {
        const int ch = pick_chunks(cls[0], 3);
        count_small_kernel<false><<<(unsigned)cls[0].size() * ch, BNL_CNT_THREADS, (size_t)cmax[0] * 256, dev.stream>>>(
            dd.u8, ds.stride_u8, ds.nobs, dev.d_units, dev.d_order + base, dev.d_tables, (int)cls[0].size(), ch,
            dd.u4, ds.stride_u8 / 2);
        BNL_CUDA(ctx, cudaGetLastError());
        tm->launches++;
      }
      base += (int)cls[0].size();
      if (!cls[1].empty()) {
        const int ch = pick_chunks(cls[1], 1);
        count_small_kernel<true><<<(unsigned)cls[1].size() * ch, kKnobMidThreads, (size_t)cmax[1] * 256, dev.stream>>>(
            dd.u8, ds.stride_u8, ds.nobs, dev.d_units, dev.d_order + base, dev.d_tables, (int)cls[1].size(), ch);
        BNL_CUDA(ctx, cudaGetLastError());
        tm->launches++;
      }
      base += (int)cls[1].size();
      if (!cls[2].empty()) {
        const int ch = pick_chunks(cls[2], 1);
        count_lane_kernel<true><<<(unsigned)cls[2].size() * ch, kKnobMidThreads, (size_t)cmax[2] * 128, dev.stream>>>(
            dd.u8, ds.stride_u8, groups, dev.d_units, dev.d_order + base, dev.d_tables, (int)cls[2].size(), ch);
        BNL_CUDA(ctx, cudaGetLastError());
        tm->launches++;
      }
      base += (int)cls[2].size();
      if (!cls[3].empty()) {
        int smem_cells = 0;
        for (int i : cls[3])
          if (dev.h_units[i].cells <= smem_cap_cells) smem_cells = std::max(smem_cells, dev.h_units[i].cells);
        int ch = (4 * dev.sm_count + (int)cls[3].size() - 1) / (int)cls[3].size();
        ch = (int)std::max(1ll, std::min<long long>(ch, std::max(1ll, groups / 256)));
        count_units_kernel<<<(unsigned)cls[3].size() * ch, BNL_CNT_THREADS, (size_t)smem_cells * sizeof(int), dev.stream>>>(
            dd.u8, ds.stride_u8, groups, dev.d_units, dev.d_order + base, dev.d_tables, (int)cls[3].size(), ch,
            smem_cells, dd.u16, ds.stride_u16);
        BNL_CUDA(ctx, cudaGetLastError());
        tm->launches++;
      }
    }
    BNL_CUDA(ctx, cudaEventRecord(dev.evc1, dev.stream));
    if (rq.kind == 2) {
      BNL_CUDA(ctx, cudaEventRecord(dev.ev1, dev.stream));
      // tables back to the caller in ONE copy: the scratch tables of a sub-batch are
      // contiguous. When the caller's offsets are contiguous too the copy lands in
      // place; otherwise it goes to a pinned buffer and is scattered on the host.
      bool contiguous = true;
      for (int i = 0; i < cnt && contiguous; i++)
        contiguous = rq.table_off[pos + i + 1] - rq.table_off[pos + i] == (int64_t)dev.h_units[i].cells;
      if (contiguous) {
        BNL_CUDA(ctx, cudaMemcpyAsync(rq.tables + rq.table_off[pos], dev.d_tables, (size_t)cells * sizeof(int),
                                      cudaMemcpyDeviceToHost, dev.stream));
        BNL_CUDA(ctx, cudaStreamSynchronize(dev.stream));
      } else {
        if ((rc = bnl_reserve_pinned(ctx, &dev.h_tables, &dev.h_tables_cap, (size_t)cells))) return rc;
        BNL_CUDA(ctx, cudaMemcpyAsync(dev.h_tables, dev.d_tables, (size_t)cells * sizeof(int), cudaMemcpyDeviceToHost,
                                      dev.stream));
        BNL_CUDA(ctx, cudaStreamSynchronize(dev.stream));
        for (int i = 0; i < cnt; i++)
          memcpy(rq.tables + rq.table_off[pos + i], dev.h_tables + dev.h_units[i].table_off,
                 (size_t)dev.h_units[i].cells * sizeof(int));
      }
    } else {
      if (rq.kind == 1 && (rq.code == BNL_FNML || rq.code == BNL_QNML) && (rc = bnl_regret_table(ctx, dev)))
        return rc;
      int mstride = 0;
      if ((rc = epilogue_stride(ctx, dev, dev.h_units, cnt, rq.kind, &mstride))) return rc;
      const size_t esmem = (size_t)mstride * BNL_EPI_WARPS * sizeof(int);
      if (rq.kind == 0 && (rq.code == BNL_JT || rq.code == BNL_MI_SH))
        epilogue_kernel<true><<<cnt, BNL_EPI_THREADS, esmem, dev.stream>>>(dev.d_units, dev.d_tables, dev.d_out,
                                                                         rq.kind, rq.code, rq.arg, ds.nobs,
                                                                         dev.d_regret, mstride);
      else
        epilogue_kernel<false><<<cnt, BNL_EPI_THREADS, esmem, dev.stream>>>(dev.d_units, dev.d_tables, dev.d_out,
                                                                          rq.kind, rq.code, rq.arg, ds.nobs,
                                                                          dev.d_regret, mstride);
      BNL_CUDA(ctx, cudaGetLastError());
      tm->launches++;
      BNL_CUDA(ctx, cudaEventRecord(dev.eve1, dev.stream));
      BNL_CUDA(ctx, cudaMemcpyAsync(dev.h_out, dev.d_out, (size_t)cnt * 4 * sizeof(double),
                                    cudaMemcpyDeviceToHost, dev.stream));
      BNL_CUDA(ctx, cudaEventRecord(dev.ev1, dev.stream));
      tr_enq.reset();
      {
        BNL_TRACE("run_block: sync wait");
        BNL_CUDA(ctx, cudaStreamSynchronize(dev.stream));
      }
      for (int i = 0; i < cnt; i++) {
        if (rq.o0) rq.o0[pos + i] = dev.h_out[4 * i + 0];
        if (rq.o1) rq.o1[pos + i] = dev.h_out[4 * i + 1];
        if (rq.o2) rq.o2[pos + i] = dev.h_out[4 * i + 2];
      }
    }
    tr_enq.reset();
    float t = 0.f;
    BNL_CUDA(ctx, cudaEventElapsedTime(&t, dev.ev0, dev.ev1));
    tm->total += t;
    BNL_CUDA(ctx, cudaEventElapsedTime(&t, dev.evc0, dev.evc1));
    tm->count += t;
    if (rq.kind != 2) {
      BNL_CUDA(ctx, cudaEventElapsedTime(&t, dev.evc1, dev.eve1));
      tm->epi += t;
    }
    pos += cnt;
  }
  return BNL_OK;
}


// ---------------------------------------------------------------------------
// Batches made ONLY of one- and two-way tables of small factors, large enough for
// the tensor-core contraction (bnl_pairs.cu): the first score_cache_fill of hc(),
// pc.stable's level 0, mi.matrix. No BnlUnit is built on the host; the caller's
// index arrays go up as they are and build_pair_units_kernel writes the
// descriptors. *handled = false leaves the request to the general path.
// kind 0: tests, 1: scores, 2: tables (as Request).
int compact_pairs(bnl_ctx *ctx, BnlDataSet &ds, int kind, int code, double arg, int n, const int *x, const int *y,
                  const int *z_off, const int *z_idx, double *o0, double *o1, double *o2, int32_t *tables,
                  const int64_t *table_off, bool *handled) {
  *handled = false;
  BNL_TRACE("compact_pairs: whole call");
  const long long knob = knob_pairs();
  if (!knob || n < 4096 || n > kMaxBatchUnits || ctx->devs.size() != 1 || ds.nobs < 4096 || ds.nd == 0) return BNL_OK;
  if (kind == 0 && (code == BNL_JT || code == BNL_MI_SH)) return BNL_OK;
  std::vector<int> lev_by_slot(ds.nd, 0);
  for (int c = 0; c < ds.ncols; c++)
    if (!ds.is_numeric[c] && !ds.is_wide[c]) lev_by_slot[ds.slot[c]] = ds.nlevels[c];
  auto slot_of = [&](int c) {            // packed slot of an eligible column, or -1
    if (c < 0 || c >= ds.ncols || ds.is_numeric[c] || ds.is_wide[c]) return -1;
    const int l = ds.nlevels[c];
    return (l >= 2 && l <= 8) ? ds.slot[c] : -1;
  };
  BnlDevice &dev = ctx->devs[0];
  int rc;
  // staging: int2 cc[n] | long long off[n] | int lev_by_slot[nd]
  const size_t stage_ints = (size_t)n * 4;
  if ((rc = bnl_reserve_pinned(ctx, &dev.h_gidx, &dev.h_gidx_cap, stage_ints + (size_t)ds.nd + 8))) return rc;
  int2 *h_cc = reinterpret_cast<int2 *>(dev.h_gidx);
  long long *h_off = reinterpret_cast<long long *>(dev.h_gidx + (size_t)n * 2);
  std::vector<char> used(ds.nd, 0);
  long long cells = 0;
  bool warp_ok = true;
  std::unique_ptr<BnlTraceScope> tr_prep(new BnlTraceScope("compact_pairs: index arrays"));
  for (int u = 0; u < n; u++) {
    const int nz = z_off[u + 1] - z_off[u];
    int c0, c1 = -1;
    if (kind == 1) {                     // node x, at most one parent
      if (nz > 1) return BNL_OK;
      c0 = slot_of(x[u]);
      if (nz == 1 && (c1 = slot_of(z_idx[z_off[u]])) < 0) return BNL_OK;
    } else {                             // x against y (make_unit puts y first), no conditioning
      if (nz != 0) return BNL_OK;
      const int yy = y ? y[u] : -1;
      if (yy >= 0) { c0 = slot_of(yy); c1 = slot_of(x[u]); if (c1 < 0) return BNL_OK; }
      else c0 = slot_of(x[u]);
    }
    if (c0 < 0 || c0 == c1) return BNL_OK;
    used[c0] = 1;
    if (c1 >= 0) used[c1] = 1;
    h_cc[u] = make_int2(c0, c1);
    h_off[u] = cells;
    const int l0 = lev_by_slot[c0], l1 = c1 >= 0 ? lev_by_slot[c1] : 1;
    if (kind == 2 && table_off[u + 1] - table_off[u] < (int64_t)l0 * l1)
      return bnl_fail(ctx, BNL_EINVAL, "bnl_counts: table_off leaves %lld cells for unit %d, which needs %d",
                      (long long)(table_off[u + 1] - table_off[u]), u, l0 * l1);
    if (kind != 2) {
      const bool sc = kind == 1 || c1 < 0;
      warp_ok = warp_ok && epilogue_warp_ok(kind, code, sc ? l0 : l1, sc ? 1 : l0, sc ? l1 : 1);
    }
    cells += (long long)l0 * l1;
  }
  if (cells > kMaxBatchCells) return BNL_OK;
  tr_prep.reset();
  {
    BNL_TRACE("compact_pairs: cost model");
    double tensor_s, atomic_s;
    bnl_pairs_cost(dev, ds, used, n, lev_by_slot, &tensor_s, &atomic_s, h_cc);
    if (!(knob >= 2 ? tensor_s < 1e29 : tensor_s < 0.8 * atomic_s)) return BNL_OK;
  }
  *handled = true;
  BnlDeviceData &dd = ds.dev[0];
  BNL_CUDA(ctx, cudaSetDevice(dev.id));
  if ((rc = bnl_reserve_device(ctx, &dev.d_gidx, &dev.gidx_cap, stage_ints + (size_t)ds.nd + 8))) return rc;
  if ((rc = bnl_reserve_device(ctx, &dev.d_units, &dev.units_cap, (size_t)n))) return rc;
  if ((rc = bnl_reserve_device(ctx, &dev.d_tables, &dev.tables_cap, (size_t)cells))) return rc;
  if ((rc = bnl_reserve_device(ctx, &dev.d_out, &dev.out_cap, (size_t)n * 4))) return rc;
  if ((rc = bnl_reserve_pinned(ctx, &dev.h_out, &dev.h_out_cap, (size_t)n * 4))) return rc;
  // the per-slot level counts ride behind the staged arrays
  memcpy(dev.h_gidx + stage_ints, lev_by_slot.data(), (size_t)ds.nd * sizeof(int));
  int *d_lev = dev.d_gidx + stage_ints;
  BnlTiming tm;
  BNL_CUDA(ctx, cudaEventRecord(dev.ev0, dev.stream));
  BNL_CUDA(ctx, cudaMemcpyAsync(dev.d_gidx, dev.h_gidx, (stage_ints + (size_t)ds.nd) * sizeof(int), cudaMemcpyHostToDevice,
                                dev.stream));
  build_pair_units_kernel<<<(n + 255) / 256, 256, 0, dev.stream>>>(
      reinterpret_cast<const int2 *>(dev.d_gidx), reinterpret_cast<const long long *>(dev.d_gidx + (size_t)n * 2), d_lev,
      kind == 1, n, dev.d_units);
  BNL_CUDA(ctx, cudaGetLastError());
  tm.launches++;
  BNL_CUDA(ctx, cudaEventRecord(dev.evc0, dev.stream));
  { BNL_TRACE("compact_pairs: bnl_pairs_run (host)"); if ((rc = bnl_pairs_run(ctx, dev, ds, dd, used, nullptr, n, lev_by_slot, &tm, h_cc))) return rc; }
  BNL_CUDA(ctx, cudaEventRecord(dev.evc1, dev.stream));
  if (kind == 2) {
    BNL_CUDA(ctx, cudaEventRecord(dev.ev1, dev.stream));
    if ((rc = bnl_reserve_pinned(ctx, &dev.h_tables, &dev.h_tables_cap, (size_t)cells))) return rc;
    BNL_CUDA(ctx, cudaMemcpyAsync(dev.h_tables, dev.d_tables, (size_t)cells * sizeof(int), cudaMemcpyDeviceToHost, dev.stream));
    BNL_CUDA(ctx, cudaStreamSynchronize(dev.stream));
    for (int u = 0; u < n; u++) {
      const long long c = (u + 1 < n ? h_off[u + 1] : cells) - h_off[u];
      memcpy(tables + table_off[u], dev.h_tables + h_off[u], (size_t)c * sizeof(int));
    }
  } else {
    if (kind == 1 && (code == BNL_FNML || code == BNL_QNML) && (rc = bnl_regret_table(ctx, dev))) return rc;
    const int mstride = (2 * BNL_MAX_LEVELS + 2 + 31) & ~31;
    const size_t esmem = (size_t)mstride * BNL_EPI_WARPS * sizeof(int);
    if (warp_ok)
      epilogue_kernel<false, true><<<(n + BNL_EPI_WARPS - 1) / BNL_EPI_WARPS, BNL_EPI_THREADS, esmem, dev.stream>>>(
          dev.d_units, dev.d_tables, dev.d_out, kind, code, arg, ds.nobs, dev.d_regret, mstride, n);
    else
      epilogue_kernel<false><<<n, BNL_EPI_THREADS, esmem, dev.stream>>>(dev.d_units, dev.d_tables, dev.d_out, kind, code,
                                                                       arg, ds.nobs, dev.d_regret, mstride);
    BNL_CUDA(ctx, cudaGetLastError());
    tm.launches++;
    BNL_CUDA(ctx, cudaEventRecord(dev.eve1, dev.stream));
    BNL_CUDA(ctx, cudaMemcpyAsync(dev.h_out, dev.d_out, (size_t)n * 4 * sizeof(double), cudaMemcpyDeviceToHost, dev.stream));
    BNL_CUDA(ctx, cudaEventRecord(dev.ev1, dev.stream));
    { BNL_TRACE("compact_pairs: sync wait"); BNL_CUDA(ctx, cudaStreamSynchronize(dev.stream)); }
    BNL_TRACE("compact_pairs: outputs");
    for (int u = 0; u < n; u++) {
      if (o0) o0[u] = dev.h_out[4 * (size_t)u];
      if (o1) o1[u] = dev.h_out[4 * (size_t)u + 1];
      if (o2) o2[u] = dev.h_out[4 * (size_t)u + 2];
    }
  }
  float t = 0.f;
  BNL_CUDA(ctx, cudaEventElapsedTime(&t, dev.ev0, dev.ev1));
  ctx->last_ms = t;
  BNL_CUDA(ctx, cudaEventElapsedTime(&t, dev.evc0, dev.evc1));
  ctx->last_count_ms = t;
  ctx->last_epi_ms = 0.0;
  if (kind != 2) {
    BNL_CUDA(ctx, cudaEventElapsedTime(&t, dev.evc1, dev.eve1));
    ctx->last_epi_ms = t;
  }
  ctx->last_launches = tm.launches;
  return BNL_OK;
}

int run_request(bnl_ctx *ctx, BnlDataSet &ds, Request &rq) {
  const int ndev = (int)ctx->devs.size();
  ctx->last_ms = ctx->last_count_ms = ctx->last_epi_ms = 0.0;
  ctx->last_launches = 0;
  if (rq.n == 0) return BNL_OK;
  std::vector<int> rcs(ndev, BNL_OK);
  std::vector<BnlTiming> tm(ndev);
  if (ndev == 1) {
    rcs[0] = run_block(ctx, ds, 0, rq, 0, rq.n, &tm[0]);
  } else {
    // several devices: contiguous blocks of units, one host thread per device so
    // that the blocks overlap (each thread only touches its own device's state;
    // bnl_fail() serialises writes to the error string).
    std::vector<std::thread> workers;
    for (int d = 0; d < ndev; d++) {
      int b, e;
      bnl_block(rq.n, ndev, d, &b, &e);
      if (b >= e) continue;
      workers.emplace_back([&, d, b, e]() { rcs[d] = run_block(ctx, ds, d, rq, b, e, &tm[d]); });
    }
    for (auto &w : workers) w.join();
    cudaSetDevice(ctx->devs[0].id);
  }
  for (int d = 0; d < ndev; d++) {
    ctx->last_ms = std::max(ctx->last_ms, (double)tm[d].total);
    ctx->last_count_ms = std::max(ctx->last_count_ms, (double)tm[d].count);
    ctx->last_epi_ms = std::max(ctx->last_epi_ms, (double)tm[d].epi);
    ctx->last_launches += tm[d].launches;
  }
  for (int d = 0; d < ndev; d++)
    if (rcs[d] != BNL_OK) return rcs[d];
  return BNL_OK;
}

}  // namespace

extern "C" int bnl_counts(bnl_ctx *ctx, bnl_data data, int n, const int *x, const int *y,
                          const int *z_off, const int *z_idx, int32_t *tables,
                          const int64_t *table_off) {
  if (!ctx) return BNL_EINVAL;
  BnlDataSet *ds = find_data(ctx, data);
  if (!ds) return bnl_fail(ctx, BNL_EINVAL, "unknown data handle");
  if (n < 0 || (n > 0 && (!x || !z_off || !tables || !table_off)))
    return bnl_fail(ctx, BNL_EINVAL, "bnl_counts: bad arguments");
  {
    bool handled = false;
    int rc = compact_pairs(ctx, *ds, 2, 0, 0.0, n, x, y, z_off, z_idx, nullptr, nullptr, nullptr, tables, table_off, &handled);
    if (rc || handled) return rc;
  }
  Request rq{2, 0, 0.0, n, {}, nullptr, nullptr, nullptr, tables, table_off};
  rq.units.resize(n);
  for (int u = 0; u < n; u++) {
    int nz = z_off[u + 1] - z_off[u];
    int rc = make_unit(ctx, *ds, x[u], y ? y[u] : -1, z_idx ? z_idx + z_off[u] : nullptr, nz, &rq.units[u]);
    if (rc) return rc;
    if (table_off[u + 1] - table_off[u] < rq.units[u].cells)
      return bnl_fail(ctx, BNL_EINVAL, "bnl_counts: table_off leaves %lld cells for unit %d, which needs %d",
                      (long long)(table_off[u + 1] - table_off[u]), u, rq.units[u].cells);
  }
  return run_request(ctx, *ds, rq);
}

// Statistics from tables the caller already holds (device 0 only: the tables are
// a few MB at most, the row pass that produced them is what shards).
extern "C" int bnl_stat_from_tables(bnl_ctx *ctx, int kind, int code, double arg, int64_t nrows, int n,
                                    const int *llx, const int *lly, const int *llz, const int *nz,
                                    const int32_t *tables, const int64_t *table_off, double *o0,
                                    double *o1, double *o2) {
  if (!ctx || ctx->devs.empty()) return BNL_EINVAL;
  if (n < 0 || (n > 0 && (!llx || !lly || !llz || !nz || !tables || !table_off || !o0)))
    return bnl_fail(ctx, BNL_EINVAL, "bnl_stat_from_tables: bad arguments");
  if (kind == 0 ? !bnl_is_discrete_test(code)
                : !(code == BNL_LOGLIK || code == BNL_AIC || code == BNL_BIC || code == BNL_BDE || code == BNL_BDS ||
                    code == BNL_BDJ || code == BNL_K2 || code == BNL_FNML || code == BNL_QNML || code == BNL_NAL ||
                    code == BNL_PNAL))
    return bnl_fail(ctx, BNL_EUNSUPPORTED, "statistic %d (kind %d) is not computed from count tables here", code, kind);
  if (n == 0) return BNL_OK;
  BnlDevice &dev = ctx->devs[0];
  BNL_CUDA(ctx, cudaSetDevice(dev.id));
  int rc;
  long long cells = 0;
  std::vector<BnlUnit> units(n);
  for (int u = 0; u < n; u++) {
    if (llx[u] < 1 || lly[u] < 1 || llz[u] < 1 || llx[u] > BNL_MAX_LEVELS_WIDE || lly[u] > BNL_MAX_LEVELS_WIDE ||
        nz[u] < 0)
      return bnl_fail(ctx, BNL_EINVAL, "bnl_stat_from_tables: bad table shape for unit %d", u);
    const long long c = (long long)llx[u] * lly[u] * llz[u];
    if (table_off[u + 1] - table_off[u] < c || c > kMaxBatchCells)
      return bnl_fail(ctx, BNL_EINVAL, "bnl_stat_from_tables: table_off leaves too few cells for unit %d", u);
    memset(&units[u], 0, sizeof(BnlUnit));
    units[u].llx = llx[u]; units[u].lly = lly[u]; units[u].llz = llz[u];
    units[u].cells = (int)c;
    // the epilogue only asks whether the unit had conditioning columns / parents.
    units[u].ncols = (kind == 0 ? 2 : 1) + nz[u];
    units[u].table_off = cells;
    cells += c;
  }
  if (cells > kMaxBatchCells)
    return bnl_fail(ctx, BNL_EUNSUPPORTED, "bnl_stat_from_tables: %lld cells exceed one batch; split the call", cells);
  if ((rc = bnl_reserve_device(ctx, &dev.d_units, &dev.units_cap, (size_t)n))) return rc;
  if ((rc = bnl_reserve_device(ctx, &dev.d_tables, &dev.tables_cap, (size_t)cells))) return rc;
  if ((rc = bnl_reserve_device(ctx, &dev.d_out, &dev.out_cap, (size_t)n * 4))) return rc;
  if ((code == BNL_FNML || code == BNL_QNML) && kind == 1 && (rc = bnl_regret_table(ctx, dev))) return rc;
  BNL_CUDA(ctx, cudaMemcpyAsync(dev.d_units, units.data(), (size_t)n * sizeof(BnlUnit), cudaMemcpyHostToDevice,
                                dev.stream));
  for (int u = 0; u < n; u++)
    BNL_CUDA(ctx, cudaMemcpyAsync(dev.d_tables + units[u].table_off, tables + table_off[u],
                                  (size_t)units[u].cells * sizeof(int), cudaMemcpyHostToDevice, dev.stream));
  int mstride = 0;
  if ((rc = epilogue_stride(ctx, dev, units, n, kind, &mstride))) return rc;
  const size_t esmem = (size_t)mstride * BNL_EPI_WARPS * sizeof(int);
  if (kind == 0 && (code == BNL_JT || code == BNL_MI_SH))
    epilogue_kernel<true><<<n, BNL_EPI_THREADS, esmem, dev.stream>>>(dev.d_units, dev.d_tables, dev.d_out, kind,
                                                                   code, arg, nrows, dev.d_regret, mstride);
  else
    epilogue_kernel<false><<<n, BNL_EPI_THREADS, esmem, dev.stream>>>(dev.d_units, dev.d_tables, dev.d_out, kind,
                                                                    code, arg, nrows, dev.d_regret, mstride);
  BNL_CUDA(ctx, cudaGetLastError());
  std::vector<double> out((size_t)n * 4);
  BNL_CUDA(ctx, cudaMemcpyAsync(out.data(), dev.d_out, out.size() * sizeof(double), cudaMemcpyDeviceToHost,
                                dev.stream));
  BNL_CUDA(ctx, cudaStreamSynchronize(dev.stream));
  for (int u = 0; u < n; u++) {
    o0[u] = out[4 * (size_t)u];
    if (o1) o1[u] = out[4 * (size_t)u + 1];
    if (o2) o2[u] = out[4 * (size_t)u + 2];
  }
  return BNL_OK;
}

extern "C" int bnl_ci_discrete(bnl_ctx *ctx, bnl_data data, int test_e, int n, const int *x,
                               const int *y, const int *z_off, const int *z_idx,
                               double *stat, double *df, double *pvalue) {
  if (!ctx) return BNL_EINVAL;
  BnlDataSet *ds = find_data(ctx, data);
  if (!ds) return bnl_fail(ctx, BNL_EINVAL, "unknown data handle");
  if (!bnl_is_discrete_test(test_e))
    return bnl_fail(ctx, BNL_EUNSUPPORTED, "test %d is not a discrete asymptotic test handled here", test_e);
  return bnl_ci_discrete_any(ctx, data, test_e, n, x, y, z_off, z_idx, stat, df, pvalue);
}

// Same request with the library-internal statistic codes allowed (BNL_MI_COEF).
int bnl_ci_discrete_any(bnl_ctx *ctx, bnl_data data, int test_e, int n, const int *x, const int *y,
                        const int *z_off, const int *z_idx, double *stat, double *df,
                        double *pvalue) {
  BnlDataSet *ds = find_data(ctx, data);
  if (!ds) return bnl_fail(ctx, BNL_EINVAL, "unknown data handle");
  if (n < 0 || (n > 0 && (!x || !y || !z_off)))
    return bnl_fail(ctx, BNL_EINVAL, "bnl_ci_discrete: bad arguments");
  {
    bool handled = false;
    int rc = compact_pairs(ctx, *ds, 0, test_e, 0.0, n, x, y, z_off, z_idx, stat, df, pvalue, nullptr, nullptr, &handled);
    if (rc || handled) return rc;
  }
  Request rq{0, test_e, 0.0, n, {}, stat, df, pvalue, nullptr, nullptr};
  rq.units.resize(n);
  for (int u = 0; u < n; u++) {
    int nz = z_off[u + 1] - z_off[u];
    int rc = make_unit(ctx, *ds, x[u], y[u], z_idx ? z_idx + z_off[u] : nullptr, nz, &rq.units[u]);
    if (rc) return rc;
  }
  return run_request(ctx, *ds, rq);
}

extern "C" int bnl_score_discrete(bnl_ctx *ctx, bnl_data data, int score_e, double k_or_iss,
                                  int n, const int *target, const int *pa_off,
                                  const int *pa_idx, double *score, double *nparams) {
  if (!ctx) return BNL_EINVAL;
  BnlDataSet *ds = find_data(ctx, data);
  if (!ds) return bnl_fail(ctx, BNL_EINVAL, "unknown data handle");
  if (score_e == BNL_EBIC) {
    // extended BIC (src/scores/per.node.score.c:116-134): BIC minus
    // 4 gamma #parents log(ncol(data)); gamma comes from bnl_set_ebic_gamma.
    int rc = bnl_score_discrete(ctx, data, BNL_BIC, k_or_iss, n, target, pa_off, pa_idx, score, nparams);
    if (rc) return rc;
    const double pen = 4.0 * ds->ebic_gamma * log((double)ds->ncols);
    for (int u = 0; u < n; u++) score[u] -= pen * (double)(pa_off[u + 1] - pa_off[u]);
    return BNL_OK;
  }
  if (score_e == BNL_MBDE) {
    // mBDe: BDe over the rows on which the node was not intervened upon; the node's
    // mask column (one level, NA on those rows) goes first among the parents, so that
    // score_cache_fill's candidates still share every column but the last.
    if (n < 0 || (n > 0 && (!target || !pa_off || !score)))
      return bnl_fail(ctx, BNL_EINVAL, "bnl_score_discrete: bad arguments");
    std::vector<int> off(n + 1, 0), idx;
    for (int u = 0; u < n; u++) {
      const int t = target[u];
      if (t >= 0 && t < (int)ds->mask_col.size() && ds->mask_col[t] >= 0) idx.push_back(ds->mask_col[t]);
      for (int k = pa_off[u]; k < pa_off[u + 1]; k++) idx.push_back(pa_idx[k]);
      off[u + 1] = (int)idx.size();
    }
    return bnl_score_discrete(ctx, data, BNL_BDE, k_or_iss, n, target, off.data(), idx.empty() ? nullptr : idx.data(),
                              score, nparams);
  }
  if (score_e != BNL_LOGLIK && score_e != BNL_AIC && score_e != BNL_BIC && score_e != BNL_BDE &&
      score_e != BNL_BDS && score_e != BNL_BDJ && score_e != BNL_K2 && score_e != BNL_NAL &&
      score_e != BNL_PNAL && score_e != BNL_FNML && score_e != BNL_QNML)
    return bnl_fail(ctx, BNL_EUNSUPPORTED, "score %d is not a discrete score handled here", score_e);
  if (n < 0 || (n > 0 && (!target || !pa_off || !score)))
    return bnl_fail(ctx, BNL_EINVAL, "bnl_score_discrete: bad arguments");
  {
    bool handled = false;
    int rc = compact_pairs(ctx, *ds, 1, score_e, k_or_iss, n, target, nullptr, pa_off, pa_idx, score, nparams, nullptr,
                           nullptr, nullptr, &handled);
    if (rc || handled) return rc;
  }
  Request rq{1, score_e, k_or_iss, n, {}, score, nparams, nullptr, nullptr, nullptr};
  {
    BNL_TRACE("score_discrete: make_unit");
    rq.units.resize(n);
    // score_cache_fill's candidates of one node share the node and every parent but the last:
    // the shared part is described once and each candidate adds its own column to a copy
    BnlUnit base;
    int base_u = -1;                        // unit whose node and leading parents `base` describes
    for (int u = 0; u < n; u++) {
      const int np = pa_off[u + 1] - pa_off[u];
      const int *z = pa_idx ? pa_idx + pa_off[u] : nullptr;
      bool shared = false;
      if (np >= 2 && base_u >= 0 && target[base_u] == target[u] && pa_off[base_u + 1] - pa_off[base_u] == np)
        shared = memcmp(pa_idx + pa_off[base_u], z, sizeof(int) * (size_t)(np - 1)) == 0;
      if (np >= 2 && !shared) {
        int rc = make_unit(ctx, *ds, target[u], -1, z, np - 1, &base);
        if (rc) return rc;
        base_u = u;
        shared = true;
      }
      int rc = BNL_OK;
      if (shared) {
        const int c = z[np - 1];
        BnlUnit &un = rq.units[u];
        const long long lev = (c >= 0 && c < ds->ncols && !ds->is_numeric[c]) ? ds->nlevels[c] : -1;
        // anything out of the ordinary (a bad column, a wide one, sizes near the limits) takes the full route
        if (lev < 0 || ds->is_wide[c] || np + 2 > BNL_MAXK || (long long)base.llz * lev >= INT32_MAX ||
            (long long)base.cells * lev >= INT32_MAX || (long long)base.cells * lev > kMaxBatchCells) {
          rc = make_unit(ctx, *ds, target[u], -1, z, np, &un);
        } else {
          un = base;
          un.col[un.ncols] = ds->slot[c];
          un.mul[un.ncols] = base.cells;
          un.ncols++;
          if (!ds->slot_has_na.empty() && ds->slot_has_na[ds->slot[c]]) un.has_na = 1;
          un.llz = (int)(base.llz * lev);
          un.cells = (int)(base.cells * lev);
        }
      } else {
        rc = make_unit(ctx, *ds, target[u], -1, z, np, &rq.units[u]);
      }
      if (rc) return rc;
    }
  }
  BNL_TRACE("score_discrete: run_request");
  return run_request(ctx, *ds, rq);
}

extern "C" int bnl_data_set_interventions(bnl_ctx *ctx, bnl_data data, int node, int mask_col) {
  if (!ctx) return BNL_EINVAL;
  BnlDataSet *ds = find_data(ctx, data);
  if (!ds) return bnl_fail(ctx, BNL_EINVAL, "unknown data handle");
  if (node < 0 || node >= ds->ncols || ds->is_numeric[node])
    return bnl_fail(ctx, BNL_EINVAL, "bnl_data_set_interventions: node %d is not a factor column", node);
  if (mask_col >= 0 && (mask_col >= ds->ncols || ds->is_numeric[mask_col] || ds->nlevels[mask_col] != 1))
    return bnl_fail(ctx, BNL_EINVAL, "bnl_data_set_interventions: the mask must be a factor column with one level");
  if (ds->mask_col.empty()) ds->mask_col.assign(ds->ncols, -1);
  ds->mask_col[node] = mask_col < 0 ? -1 : mask_col;
  return BNL_OK;
}
