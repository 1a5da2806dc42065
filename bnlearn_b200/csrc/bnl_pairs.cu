// bnl_pairs.cu -- host side of the all-pairs contraction (bnl_pairs.cuh): which
// units of a batch it serves, the tile / split plan, launches.
#include <math.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>

#include "bnl_pairs.cuh"

namespace {

long long env_ll(const char *name, long long dflt) {
  const char *v = getenv(name);
  return (v && *v) ? atoll(v) : dflt;
}
// effective multiply-accumulate rate of pair_onehot_kernel on one B200 (measured,
// profiles/): the producers' code expansion, not the tensor pipe, sets it.
const double kPairMacRate = (double)env_ll("BNL_PAIRS_GMACS", 750000) * 1e9;
const double kAtomicRate = 4.2e12;        // row updates per second of the per-unit counting kernels

struct Plan {
  std::vector<int> slots;                 // selected packed columns, ascending
  std::vector<int> first;                 // first indicator of each
  std::vector<int> eff;                   // indicators of each: its levels, minus the last one when it holds no NA
  int L = 0, ldg = 0, bn = 256;
  std::vector<int2> tiles;
  int total_stages = 0, splits = 1, per_split = 0;
};

// cc (optional): the batch's units as {first column, second column or -1} packed slots.
// When they are far from all pairs of the selected columns -- a rank's share of a
// sharded first iteration: the targets of its block against every other node -- only
// the output tiles that hold one of their tables are computed.
void make_plan(const BnlDevice &dev, const BnlDataSet &ds, const std::vector<char> &used,
               const std::vector<int> &lev_by_slot, Plan *p, const int2 *cc = nullptr, int ncc = 0) {
  // A column without NA needs no indicator for its LAST level: the rows that are in none of the
  // others are in it, so pair_extract_kernel gets those cells from the margins (exact: integers).
  // A three-level factor costs two rows of H instead of three -- 1000 factors of 2-4 levels: L from
  // 3000 to 2000, 156 output tiles to 72.
  // A producer thread owns a fixed number of column slots, 228 per tile in the widest variant (3 x
  // 76 column lanes, A and B rows together): no 128-row block of H may hold more than 72 columns
  // (+ one that straddles in from each side). One-row columns -- NA-free binary factors -- would
  // exceed that, so a block that is full is closed and the next column starts the next block
  // (the rows left over are never written and stay zero).
  const int kMaxColsPerBlock = 72;
  int blk = 0, blk_cols = 0, prev_end = -1;
  for (int s = 0; s < ds.nd; s++)
    if (used[s]) {
      const bool na = ds.slot_has_na.empty() || ds.slot_has_na[s];
      const int e = lev_by_slot[s] - (na ? 0 : 1);
      if (p->L / PAIR_BM != blk) { blk = p->L / PAIR_BM; blk_cols = prev_end >= blk * PAIR_BM ? 1 : 0; }
      if (blk_cols >= kMaxColsPerBlock) { p->L = (blk + 1) * PAIR_BM; blk++; blk_cols = 0; }
      p->slots.push_back(s);
      p->first.push_back(p->L);
      p->eff.push_back(e);
      p->L += e;
      prev_end = p->L - 1;
      blk_cols++;
    }
  p->ldg = (p->L + 255) & ~255;
  p->bn = p->L <= 128 ? 128 : 256;
  const int nbm = (p->L + PAIR_BM - 1) / PAIR_BM, nbn = (p->L + p->bn - 1) / p->bn;
  std::vector<char> need;
  const double all_pairs = 0.5 * (double)p->slots.size() * ((double)p->slots.size() - 1.0);
  if (cc && (double)ncc < 0.9 * all_pairs && nbm * nbn > 1) {
    need.assign((size_t)nbm * nbn, 0);
    std::vector<int> first_by_slot(ds.nd, 0), eff_by_slot(ds.nd, 1);
    for (size_t c = 0; c < p->slots.size(); c++) { first_by_slot[p->slots[c]] = p->first[c]; eff_by_slot[p->slots[c]] = p->eff[c]; }
    for (int u = 0; u < ncc; u++) {
      int s0 = cc[u].x, s1 = cc[u].y < 0 ? cc[u].x : cc[u].y;
      if (first_by_slot[s0] > first_by_slot[s1]) std::swap(s0, s1);     // smaller indicator first: G's upper triangle
      const int a0 = first_by_slot[s0], a1 = a0 + eff_by_slot[s0] - 1, b0 = first_by_slot[s1], b1 = b0 + eff_by_slot[s1] - 1;
      for (int bm = a0 / PAIR_BM; bm <= a1 / PAIR_BM; bm++)
        for (int bn = b0 / p->bn; bn <= b1 / p->bn; bn++) need[(size_t)bm * nbn + bn] = 1;
      // the one-way counts on the diagonal, which rebuild the cells of a dropped last level
      for (int bm = a0 / PAIR_BM; bm <= a1 / PAIR_BM; bm++)
        for (int bn = a0 / p->bn; bn <= a1 / p->bn; bn++) need[(size_t)bm * nbn + bn] = 1;
      for (int bm = b0 / PAIR_BM; bm <= b1 / PAIR_BM; bm++)
        for (int bn = b0 / p->bn; bn <= b1 / p->bn; bn++) need[(size_t)bm * nbn + bn] = 1;
    }
  }
  for (int bm = 0; bm < nbm; bm++)
    for (int bn = 0; bn < nbn; bn++)
      if (bn * p->bn + p->bn > bm * PAIR_BM && (need.empty() || need[(size_t)bm * nbn + bn]))
        p->tiles.push_back(make_int2(bm, bn));                                       // reaches the upper triangle
  p->total_stages = (int)((ds.nobs + PAIR_KROWS - 1) / PAIR_KROWS);
  // row splits: whole waves of CTAs (one per SM), each long enough to pay for its
  // prologue and for draining the accumulator (~20 stages' worth) -- and short enough
  // that the slabs of rows the resident CTAs walk through stay in L2: every tile of a
  // slab re-reads the slab's columns (with 72 tiles and two splits of 500 MB each the
  // kernel pulled 14.8 GB from DRAM for 1.0 GB of data and ran at the HBM rate).
  const long long nt = (long long)p->tiles.size();
  const double slab_budget = (double)env_ll("BNL_PAIRS_SLAB_MB", 40) * 1e6;   // bytes of packed columns per slab; 2-3 slabs are live at a time
  const int s_min = (int)std::min<double>(std::max(1, p->total_stages / 32),
                                          std::ceil((double)ds.nobs * (double)p->slots.size() / slab_budget));
  double best = 1e300;
  for (int s = std::max(1, s_min); s <= 4096 && s <= std::max(1, p->total_stages / 32); s++) {
    const long long waves = (nt * s + dev.sm_count - 1) / dev.sm_count;
    const double cost = (double)waves * ((double)((p->total_stages + s - 1) / s) + 20.0);
    if (cost < best - 1e-9) { best = cost; p->splits = s; }
  }
  p->per_split = (p->total_stages + p->splits - 1) / p->splits;
  p->splits = (p->total_stages + p->per_split - 1) / p->per_split;
}

}  // namespace

bool bnl_pairs_eligible(const BnlUnit &u, const std::vector<int> &lev_by_slot) {
  if (u.ncols < 1 || u.ncols > 2 || u.pad) return false;          // pad: a uint16 column
  for (int k = 0; k < u.ncols; k++) {
    const int l = lev_by_slot[u.col[k]];
    if (l < 2 || l > PAIR_MAX_LEVELS) return false;
  }
  return u.ncols == 1 || u.col[0] != u.col[1];
}

void bnl_pairs_cost(const BnlDevice &dev, const BnlDataSet &ds, const std::vector<char> &used, long long nunits,
                    const std::vector<int> &lev_by_slot, double *tensor_s, double *atomic_s, const int2 *cc) {
  Plan p;
  make_plan(dev, ds, used, lev_by_slot, &p, cc, cc ? (int)nunits : 0);
  const double macs = (double)p.tiles.size() * PAIR_BM * p.bn * (double)ds.nobs;
  const double hbm = (double)p.slots.size() * (double)ds.nobs / 5e12;
  *tensor_s = std::max(macs / kPairMacRate, hbm) + 40e-6;
  // G is ldg x ldg ints: beyond 16384 indicator rows (1 GiB) the contraction is not offered
  if (p.ldg > 16384) *tensor_s = 1e30;
  // the kernel addresses the packed block with 32-bit offsets in 16-byte units
  if ((unsigned long long)ds.stride_u8 * (unsigned long long)ds.nd >= (1ull << 35)) *tensor_s = 1e30;
  *atomic_s = (double)nunits * (double)ds.nobs / kAtomicRate + 10e-6;
}

int bnl_pairs_run(bnl_ctx *ctx, BnlDevice &dev, const BnlDataSet &ds, const BnlDeviceData &dd,
                  const std::vector<char> &used, const std::vector<int> *picked, int nunits,
                  const std::vector<int> &lev_by_slot, BnlTiming *tm, const int2 *cc) {
  Plan p;
  make_plan(dev, ds, used, lev_by_slot, &p, cc, cc ? nunits : 0);
  const int nc = (int)p.slots.size(), nt = (int)p.tiles.size(), np = picked ? (int)picked->size() : 0;
  // one int block: cols (int4 per column) | ind2col [ldg] | slot_first [nd] + slot_eff [nd] | tiles (int2) | ids [np]
  const size_t n_cols = (size_t)nc * 4, n_ind = (size_t)p.ldg, n_sf = 2 * (((size_t)ds.nd + 1) & ~(size_t)1), n_tiles = (size_t)nt * 2;
  const size_t total = n_cols + n_ind + n_sf + n_tiles + (size_t)np + 16;
  int rc;
  if ((rc = bnl_reserve_pinned(ctx, &dev.h_pairmeta, &dev.h_pairmeta_cap, total))) return rc;
  if ((rc = bnl_reserve_device(ctx, &dev.d_pairmeta, &dev.pairmeta_cap, total))) return rc;
  if ((rc = bnl_reserve_device(ctx, &dev.d_pairG, &dev.pairG_cap, (size_t)p.ldg * p.ldg))) return rc;
  int *h = dev.h_pairmeta;
  int *h_cols = h, *h_ind = h_cols + n_cols, *h_sf = h_ind + n_ind, *h_tiles = h_sf + n_sf, *h_ids = h_tiles + n_tiles;
  int *h_se = h_sf + n_sf / 2;
  for (int s = 0; s < ds.nd; s++) { h_sf[s] = -1; h_se[s] = 0; }
  for (int c = 0; c < nc; c++) {
    const int s = p.slots[c], l = p.eff[c];
    h_cols[4 * c] = s; h_cols[4 * c + 1] = p.first[c]; h_cols[4 * c + 2] = l; h_cols[4 * c + 3] = lev_by_slot[s];
    // rows of this column, and the unused rows up to the next column's first (the tail of a closed
    // block: a tile that ends there must still find its LAST column through ind2col)
    const int until = c + 1 < nc ? p.first[c + 1] : p.ldg;
    for (int r = p.first[c]; r < until; r++) h_ind[r] = c;
    h_sf[s] = p.first[c];
    h_se[s] = l;
  }
  memcpy(h_tiles, p.tiles.data(), n_tiles * sizeof(int));
  for (int i = 0; i < np; i++) h_ids[i] = (*picked)[i];
  int *d = dev.d_pairmeta;
  int *d_cols = d, *d_ind = d_cols + n_cols, *d_sf = d_ind + n_ind, *d_tiles = d_sf + n_sf, *d_ids = d_tiles + n_tiles;
  BNL_CUDA(ctx, cudaMemcpyAsync(d, h, total * sizeof(int), cudaMemcpyHostToDevice, dev.stream));
  const int atomic_out = p.splits > 1;
  if (atomic_out)
    BNL_CUDA(ctx, cudaMemsetAsync(dev.d_pairG, 0, (size_t)p.ldg * p.ldg * sizeof(int), dev.stream));
  // variants: <indicator columns per tile, columns per producer thread, stages of loads in flight>.
  // One output tile of <= 44 columns (a pc.stable level 0 on a few dozen factors) keeps 8 stages
  // in flight per thread; the widest tile needs 5 columns per thread and affords 3.
  const size_t smem256 = (size_t)PAIR_STAGES * (PAIR_BM + 256) * 128 + 1024;
  // NA bytes between valid codes need one more instruction per word (selector4); factors
  // with more than four levels need the second half of the level stores
  bool any_na = false, wide = false;
  for (int s : p.slots) {
    any_na = any_na || ds.slot_has_na.empty() || ds.slot_has_na[s];
  }
  for (int l : p.eff) wide = wide || l > 4;
  const unsigned grid = (unsigned)nt * (unsigned)p.splits;
#define BNL_PAIR_GO(BN, ITEMS, DEPTH, NA, WIDE, NPW, SUP, NSTG, SMEM)                                                        \
  do {                                                                                                            \
    auto kern = pair_onehot_kernel<BN, ITEMS, DEPTH, NA, WIDE, NPW, SUP, NSTG>;                                              \
    BNL_CUDA(ctx, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(SMEM)));          \
    kern<<<grid, (NPW + 1) * 32, SMEM, dev.stream>>>(dd.u8, ds.stride_u8, reinterpret_cast<const int4 *>(d_cols), \
                                                     d_ind, p.L, reinterpret_cast<const int2 *>(d_tiles), nt,     \
                                                     p.per_split, p.total_stages, dev.d_pairG, p.ldg, atomic_out); \
  } while (0)
#define BNL_PAIR_LAUNCH(BN, ITEMS, DEPTH, NPW, SUP, NSTG, SMEM)                                                   \
  do {                                                                                                            \
    if (any_na && wide) BNL_PAIR_GO(BN, ITEMS, DEPTH, true, true, NPW, SUP, NSTG, SMEM);                          \
    else if (any_na) BNL_PAIR_GO(BN, ITEMS, DEPTH, true, false, NPW, SUP, NSTG, SMEM);                            \
    else if (wide) BNL_PAIR_GO(BN, ITEMS, DEPTH, false, true, NPW, SUP, NSTG, SMEM);                              \
    else BNL_PAIR_GO(BN, ITEMS, DEPTH, false, false, NPW, SUP, NSTG, SMEM);                                       \
  } while (0)
  // columns per producer thread: the tile's columns over 4 * NPW column lanes (19 + 1
  // warps measured 4 % faster than 11 + 1 on the wide tile, 15 + 1 and 23 + 1 slower)
  // two K-blocks per full / empty handshake (SUP = 2): a handshake has a floor of several hundred
  // cycles whatever the work behind it (config B level 0: 0.225 -> 0.173 ms, config C 3.58 -> 3.48 ms)
  // The single-tile launches (L <= 128) keep 8 buffers of B rows only and shake hands every 4 K-blocks.
  const size_t smem_diag = (size_t)8 * 128 * 128 + 1024;
  if (p.bn == 128 && nc <= 44) BNL_PAIR_LAUNCH(128, 1, 2, 11, 4, 8, smem_diag);
  else if (p.bn == 128) BNL_PAIR_LAUNCH(128, 2, 2, 19, 2, 8, smem_diag);              // <= 74 columns over 76 lanes
  else BNL_PAIR_LAUNCH(256, 3, 2, 19, 2, PAIR_STAGES, smem256);                       // <= 218 columns over 76 lanes
#undef BNL_PAIR_GO
#undef BNL_PAIR_LAUNCH
  BNL_CUDA(ctx, cudaGetLastError());
  // picked == null: every unit of the batch, in order
  pair_extract_kernel<<<(nunits + 255) / 256, 256, 0, dev.stream>>>(dev.d_units, picked ? d_ids : nullptr, nunits, d_sf,
                                                                  d_sf + n_sf / 2, (int)ds.nobs, dev.d_pairG, p.ldg,
                                                                  dev.d_tables);
  BNL_CUDA(ctx, cudaGetLastError());
  tm->launches += 2;
  return BNL_OK;
}
