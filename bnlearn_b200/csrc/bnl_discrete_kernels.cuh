// bnl_discrete_kernels.cuh -- counting kernels and fused statistic epilogues for
// discrete data (uint8 packed columns).
//
// count_units_kernel   joint count table of one unit per CTA (x row-chunk):
//                      replaces c_fast_config + fill_{1,2,3}d_table
//                      (src/core/sets.c:162-218, src/core/contingency.tables.c).
// epilogue_kernel      margins, statistic / score, df and p-value per unit:
//                      replaces mi/x2/cmi/cx2_kernel + discrete_(c)df + pchisq
//                      (src/tests/discrete/*.c) and dlik/cdlik/dpost/cdpost
//                      (src/scores/discrete.loglikelihood.c, dirichlet.posterior.c).
#pragma once
#include "bnl_internal.h"
#include "bnl_special.cuh"

#define BNL_CNT_THREADS 256
#define BNL_EPI_THREADS 128
#define BNL_EPI_WARPS (BNL_EPI_THREADS / 32)

__device__ __forceinline__ uint4 ldg_stream(const uint4 *p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
  return r;
}

// DRAM -> L2 ahead of the register pipeline: a column that only one CTA reads
// (a candidate of a score_cache_fill batch) otherwise pays the full DRAM latency
// inside the one-group-ahead register prefetch.
__device__ __forceinline__ void prefetch_l2(const void *p) {
  asm volatile("prefetch.global.L2 [%0];" :: "l"(p));
}

// ---------------------------------------------------------------------------
// Generic counting: one CTA = (unit, row chunk). Each thread takes 16 rows at a
// time (one 128-bit load per column), forms the 16 cell indices in registers and
// updates the CTA's shared-memory table (or the global one when the table does
// not fit). NA / padding bytes have bit 7 set and veto their row.
__global__ void __launch_bounds__(BNL_CNT_THREADS)
count_units_kernel(const uint8_t *__restrict__ u8, size_t stride, long long groups,
                   const BnlUnit *__restrict__ units, const int *__restrict__ order,
                   int *__restrict__ tables, int nunits, int chunks, int smem_cells,
                   const uint16_t *__restrict__ u16, size_t stride16) {
  extern __shared__ int hist[];
  __shared__ BnlUnit su;
  const int chunk = blockIdx.x / nunits, u = order[blockIdx.x % nunits];
  if (threadIdx.x < sizeof(BnlUnit) / sizeof(int))
    reinterpret_cast<int *>(&su)[threadIdx.x] = reinterpret_cast<const int *>(&units[u])[threadIdx.x];
  __syncthreads();
  const int cells = su.cells, ncols = su.ncols;
  const bool in_smem = cells <= smem_cells;
  int *gtab = tables + su.table_off;
  int *tab = in_smem ? hist : gtab;
  if (in_smem) {
    for (int c = threadIdx.x; c < cells; c += blockDim.x) hist[c] = 0;
    __syncthreads();
  }

  const long long per = (groups + chunks - 1) / chunks;
  const long long g0 = per * chunk;
  long long g1 = g0 + per;
  if (g1 > groups) g1 = groups;

  for (long long g = g0 + threadIdx.x; g < g1; g += blockDim.x) {
    int idx[16];
    unsigned na[4] = {0u, 0u, 0u, 0u};
#pragma unroll
    for (int j = 0; j < 16; j++) idx[j] = 0;
    for (int k = 0; k < ncols; k++) {
      const int m = su.mul[k];
      if (su.col[k] & BNL_WIDE_BIT) {
        // uint16 column (128..32767 levels): 16 rows = two 128-bit loads; bit 15 marks NA
        const uint4 *p16 = reinterpret_cast<const uint4 *>(u16 + (size_t)(su.col[k] & ~BNL_WIDE_BIT) * stride16) + 2 * g;
        const uint4 a = ldg_stream(p16), b = ldg_stream(p16 + 1);
        const unsigned hw[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
#pragma unroll
        for (int h = 0; h < 8; h++) {
          const unsigned lo = hw[h] & 0xFFFFu, hi = hw[h] >> 16;
          const int j0 = 2 * h, j1 = 2 * h + 1;
          if (lo & 0x8000u) na[j0 >> 2] |= 0x80u << (8 * (j0 & 3));
          if (hi & 0x8000u) na[j1 >> 2] |= 0x80u << (8 * (j1 & 3));
          idx[j0] += (int)lo * m;
          idx[j1] += (int)hi * m;
        }
        continue;
      }
      const uint4 w = ldg_stream(reinterpret_cast<const uint4 *>(u8 + (size_t)su.col[k] * stride) + g);
      const unsigned ww[4] = {w.x, w.y, w.z, w.w};
#pragma unroll
      for (int q = 0; q < 4; q++) {
        na[q] |= ww[q] & 0x80808080u;
#pragma unroll
        for (int j = 0; j < 4; j++)
          idx[q * 4 + j] += (int)__byte_perm(ww[q], 0, 0x4440 + j) * m;
      }
    }
#pragma unroll
    for (int q = 0; q < 4; q++)
#pragma unroll
      for (int j = 0; j < 4; j++)
        if (!(na[q] & (0x80u << (8 * j)))) atomicAdd(&tab[idx[q * 4 + j]], 1);
  }

  if (in_smem) {
    __syncthreads();
    for (int c = threadIdx.x; c < cells; c += blockDim.x) {
      const int v = hist[c];
      if (v) atomicAdd(&gtab[c], v);
    }
  }
}

// ---------------------------------------------------------------------------
// Small tables (the common case: a handful of 2-5 level factors): one CTA =
// (unit, row chunk) again, but the CTA's table is replicated 32 times in shared
// memory, one copy per LANE, laid out cell-major: word (cell * 32 + lane). A warp's
// 32 updates then always fall in 32 distinct banks, whatever the data, so every
// warp-wide atomic is a single conflict-free wavefront (the generic kernel above
// spends ~9.5 wavefronts per warp on same-address / same-bank collisions; see
// profiles/). Cell indices for 4 rows are formed in one packed 32-bit register
// (bytes when the table has <= 256 cells, 16-bit halves otherwise), so the
// arithmetic is one IMAD per column per 4 (or 2) rows. Rows with an NA fall to a
// scalar path. `order` maps the launch's CTA groups to unit slots.
template <bool IDX16>
__global__ void __launch_bounds__(1024)
count_lane_kernel(const uint8_t *__restrict__ u8, size_t stride, long long groups,
                  const BnlUnit *__restrict__ units, const int *__restrict__ order,
                  int *__restrict__ tables, int nunits, int chunks) {
  extern __shared__ int hist[];
  __shared__ BnlUnit su;
  const int chunk = blockIdx.x / nunits, u = order[blockIdx.x % nunits];
  if (threadIdx.x < sizeof(BnlUnit) / sizeof(int))
    reinterpret_cast<int *>(&su)[threadIdx.x] = reinterpret_cast<const int *>(&units[u])[threadIdx.x];
  __syncthreads();
  const int cells = su.cells, ncols = su.ncols;
  for (int c = threadIdx.x; c < cells * 32; c += blockDim.x) hist[c] = 0;
  __syncthreads();
  int *hl = hist + (threadIdx.x & 31);

  const long long per = (groups + chunks - 1) / chunks;
  const long long g0 = per * chunk;
  long long g1 = g0 + per;
  if (g1 > groups) g1 = groups;

  for (long long g = g0 + threadIdx.x; g < g1; g += blockDim.x) {
    constexpr int NI = IDX16 ? 8 : 4;
    unsigned idx[NI];
#pragma unroll
    for (int j = 0; j < NI; j++) idx[j] = 0u;
    unsigned na = 0u;
    for (int kb = 0; kb < ncols; kb += 4) {
      uint4 w[4];
#pragma unroll
      for (int kk = 0; kk < 4; kk++)
        if (kb + kk < ncols)
          w[kk] = ldg_stream(reinterpret_cast<const uint4 *>(u8 + (size_t)su.col[kb + kk] * stride) + g);
#pragma unroll
      for (int kk = 0; kk < 4; kk++)
        if (kb + kk < ncols) {
          const unsigned m = (unsigned)su.mul[kb + kk];
          na |= (w[kk].x | w[kk].y | w[kk].z | w[kk].w) & 0x80808080u;
          const unsigned ww[4] = {w[kk].x, w[kk].y, w[kk].z, w[kk].w};
#pragma unroll
          for (int q = 0; q < 4; q++) {
            if (IDX16) {
              idx[2 * q] += __byte_perm(ww[q], 0, 0x4140) * m;
              idx[2 * q + 1] += __byte_perm(ww[q], 0, 0x4342) * m;
            } else {
              idx[q] += ww[q] * m;
            }
          }
        }
    }
    if (na == 0u) {
#pragma unroll
      for (int j = 0; j < NI; j++) {
        if (IDX16) {
          atomicAdd(hl + ((idx[j] & 0xFFFFu) << 5), 1);
          atomicAdd(hl + ((idx[j] >> 16) << 5), 1);
        } else {
          atomicAdd(hl + ((idx[j] & 0xFFu) << 5), 1);
          atomicAdd(hl + (((idx[j] >> 8) & 0xFFu) << 5), 1);
          atomicAdd(hl + (((idx[j] >> 16) & 0xFFu) << 5), 1);
          atomicAdd(hl + ((idx[j] >> 24) << 5), 1);
        }
      }
    } else {
      // some row of this group has an NA (or is padding): redo the 16 rows one by
      // one; the packed sums above may have carried across lanes.
      for (int r = 0; r < 16; r++) {
        int cell = 0;
        bool bad = false;
        for (int k = 0; k < ncols; k++) {
          const unsigned b = u8[(size_t)su.col[k] * stride + (size_t)g * 16 + r];
          bad |= b >= 0x80u;
          cell += (int)b * su.mul[k];
        }
        if (!bad) atomicAdd(hl + (cell << 5), 1);
      }
    }
  }

  __syncthreads();
  int *gtab = tables + su.table_off;
  for (int c = threadIdx.x; c < cells; c += blockDim.x) {
    int sum = 0;
#pragma unroll
    for (int l = 0; l < 32; l++) sum += hist[c * 32 + ((l + threadIdx.x) & 31)];
    if (sum) atomicAdd(&gtab[c], sum);
  }
}

// ---------------------------------------------------------------------------
// Tables of at most 256 cells (byte-packed indices) or 880 cells (16-bit packed)
// -- almost every test and score delta on 2-5 level factors -- get the tightest loop. Same lane-replicated idea, with three more
// steps taken after profiling the kernel above (profiles/r1_v2_*):
//   * a cell owns 256 bytes of shared memory (64 counters: lane 0..31 for even
//     warps, 32..63 for odd warps), so the byte offset of a row's counter is
//     (cell << 8) | (slot << 2): ONE byte-permute of the packed index register
//     with the thread's constant slot byte, then the atomic. 2 instructions/row.
//   * units whose columns hold no NA (flags computed when the data set is
//     registered) skip the NA bookkeeping; only the last, partial 16-row group
//     is done row by row.
//   * CTAs are ordered chunk-major: neighbouring CTAs work on the same rows of
//     different units, so a row chunk of every column is fetched from HBM once
//     and then served from L2 to all units that touch it.
__device__ __forceinline__ void small_bump(int *hist, unsigned v, unsigned slot4) {
  // v holds 4 packed cell indices; selector 0x55j4: byte0 = slot byte, byte1 = cell.
  atomicAdd(reinterpret_cast<int *>(reinterpret_cast<char *>(hist) + __byte_perm(v, slot4, 0x5504)), 1);
  atomicAdd(reinterpret_cast<int *>(reinterpret_cast<char *>(hist) + __byte_perm(v, slot4, 0x5514)), 1);
  atomicAdd(reinterpret_cast<int *>(reinterpret_cast<char *>(hist) + __byte_perm(v, slot4, 0x5524)), 1);
  atomicAdd(reinterpret_cast<int *>(reinterpret_cast<char *>(hist) + __byte_perm(v, slot4, 0x5534)), 1);
}

__device__ __forceinline__ void small_rows_scalar(const uint8_t *__restrict__ u8, size_t stride,
                                                  const BnlUnit &su, int *hist, unsigned slot4,
                                                  long long row0, int nrows) {
  for (int r = 0; r < nrows; r++) {
    unsigned cell = 0;
    bool bad = false;
    for (int k = 0; k < su.ncols; k++) {
      const unsigned b = u8[(size_t)su.col[k] * stride + (size_t)(row0 + r)];
      bad |= b >= 0x80u;
      cell += b * (unsigned)su.mul[k];
    }
    if (!bad) atomicAdd(reinterpret_cast<int *>(reinterpret_cast<char *>(hist) + ((cell << 8) | slot4)), 1);
  }
}

// Row-pair variant for tables of at most 16 cells: two consecutive rows share one
// update of a 16 x 16 "pair table" (index = cell_a + 16 * cell_b), which halves
// the shared-memory atomics -- the resource that bounds this kernel. v | (v >> 4)
// leaves the pair indices of rows (0,1) and (2,3) in bytes 0 and 2.
__device__ __forceinline__ void pair_bump(int *hist, unsigned v, unsigned slot4) {
  const unsigned t = v | (v >> 4);
  atomicAdd(reinterpret_cast<int *>(reinterpret_cast<char *>(hist) + __byte_perm(t, slot4, 0x5504)), 1);
  atomicAdd(reinterpret_cast<int *>(reinterpret_cast<char *>(hist) + __byte_perm(t, slot4, 0x5524)), 1);
}

// 16-bit variant of small_bump for tables of 257..880 cells: v holds 2 packed
// indices; the offset is (index << 8) | slot byte, still one PRMT.
__device__ __forceinline__ void mid_bump(int *hist, unsigned v, unsigned slot4) {
  atomicAdd(reinterpret_cast<int *>(reinterpret_cast<char *>(hist) + __byte_perm(v, slot4, 0x5104)), 1);
  atomicAdd(reinterpret_cast<int *>(reinterpret_cast<char *>(hist) + __byte_perm(v, slot4, 0x5324)), 1);
}

template <int NC, bool NA, bool W16, bool PAIR>
__device__ __forceinline__ void small_rows(const uint8_t *__restrict__ u8, size_t stride,
                                           const BnlUnit &su, int *hist, unsigned slot4,
                                           long long g0, long long g1) {
  const uint4 *p[NC];
  unsigned m[NC];
#pragma unroll
  for (int k = 0; k < NC; k++) {
    p[k] = reinterpret_cast<const uint4 *>(u8 + (size_t)su.col[k] * stride);
    m[k] = (unsigned)su.mul[k];
  }
  for (long long g = g0 + threadIdx.x; g < g1; g += blockDim.x) {
    uint4 w[NC];
#pragma unroll
    for (int k = 0; k < NC; k++) w[k] = ldg_stream(p[k] + g);
    unsigned na = 0u;
    if (NA) {
#pragma unroll
      for (int k = 0; k < NC; k++) na |= w[k].x | w[k].y | w[k].z | w[k].w;
    }
    if (NA && (na & 0x80808080u) != 0u) {
      small_rows_scalar(u8, stride, su, hist, slot4, g * 16, 16);
      continue;
    }
    if (!W16) {
      uint4 v = w[0];                    // the first column always has multiplier 1
#pragma unroll
      for (int k = 1; k < NC; k++) {
        v.x += w[k].x * m[k]; v.y += w[k].y * m[k]; v.z += w[k].z * m[k]; v.w += w[k].w * m[k];
      }
      if (PAIR) {
        pair_bump(hist, v.x, slot4);
        pair_bump(hist, v.y, slot4);
        pair_bump(hist, v.z, slot4);
        pair_bump(hist, v.w, slot4);
      } else {
        small_bump(hist, v.x, slot4);
        small_bump(hist, v.y, slot4);
        small_bump(hist, v.z, slot4);
        small_bump(hist, v.w, slot4);
      }
    } else {
      unsigned v[8];
      const unsigned w0[4] = {w[0].x, w[0].y, w[0].z, w[0].w};
#pragma unroll
      for (int q = 0; q < 4; q++) {
        v[2 * q] = __byte_perm(w0[q], 0, 0x4140);
        v[2 * q + 1] = __byte_perm(w0[q], 0, 0x4342);
      }
#pragma unroll
      for (int k = 1; k < NC; k++) {
        const unsigned wk[4] = {w[k].x, w[k].y, w[k].z, w[k].w};
#pragma unroll
        for (int q = 0; q < 4; q++) {
          v[2 * q] += __byte_perm(wk[q], 0, 0x4140) * m[k];
          v[2 * q + 1] += __byte_perm(wk[q], 0, 0x4342) * m[k];
        }
      }
#pragma unroll
      for (int j = 0; j < 8; j++) mid_bump(hist, v[j], slot4);
    }
  }
}

// generic column count (> 4 columns).
template <bool NA, bool W16, bool PAIR>
__device__ __forceinline__ void small_rows_any(const uint8_t *__restrict__ u8, size_t stride,
                                               const BnlUnit &su, int *hist, unsigned slot4,
                                               long long g0, long long g1) {
  const int ncols = su.ncols;
  for (long long g = g0 + threadIdx.x; g < g1; g += blockDim.x) {
    unsigned v[W16 ? 8 : 4];
#pragma unroll
    for (int j = 0; j < (W16 ? 8 : 4); j++) v[j] = 0u;
    unsigned na = 0u;
    for (int k = 0; k < ncols; k++) {
      const uint4 w = ldg_stream(reinterpret_cast<const uint4 *>(u8 + (size_t)su.col[k] * stride) + g);
      const unsigned m = (unsigned)su.mul[k];
      const unsigned ww[4] = {w.x, w.y, w.z, w.w};
      if (NA) na |= w.x | w.y | w.z | w.w;
#pragma unroll
      for (int q = 0; q < 4; q++) {
        if (W16) {
          v[2 * q] += __byte_perm(ww[q], 0, 0x4140) * m;
          v[2 * q + 1] += __byte_perm(ww[q], 0, 0x4342) * m;
        } else {
          v[q] += ww[q] * m;
        }
      }
    }
    if (NA && (na & 0x80808080u) != 0u) {
      small_rows_scalar(u8, stride, su, hist, slot4, g * 16, 16);
      continue;
    }
#pragma unroll
    for (int j = 0; j < (W16 ? 8 : 4); j++) {
      if (W16) mid_bump(hist, v[j], slot4);
      else if (PAIR) pair_bump(hist, v[j], slot4);
      else small_bump(hist, v[j], slot4);
    }
  }
}

// The same from the NIBBLE copy of the columns (bnl_nibbles_kernel: two codes per byte, factors
// of at most 15 levels, units without NA): 32 rows per 128-bit load, so half the bytes cross
// the LSU pipe that the atomics already keep busy. Byte j of a column holds row 2j in its low
// nibble and row 2j + 1 in the high one; a table does not care in which order rows arrive.
// NC = 0: the column count is the unit's (more than 4 columns).
template <int NC>
__device__ __forceinline__ void small_rows_nib(const uint8_t *__restrict__ u4, size_t stride4,
                                               const BnlUnit &su, int *hist, unsigned slot4,
                                               long long g0, long long g1) {
  constexpr int NCA = NC ? NC : 1;
  const uint4 *p[NCA];
  unsigned m[NCA];
#pragma unroll
  for (int k = 0; k < NCA; k++) {
    p[k] = reinterpret_cast<const uint4 *>(u4 + (size_t)su.col[k] * stride4);
    m[k] = (unsigned)su.mul[k];
  }
  const unsigned M = 0x0F0F0F0Fu;
  for (long long g = g0 + threadIdx.x; g < g1; g += blockDim.x) {
    uint4 lo, hi;
    if (NC) {
      uint4 w[NCA];
#pragma unroll
      for (int k = 0; k < NCA; k++) w[k] = ldg_stream(p[k] + g);
      lo = make_uint4(w[0].x & M, w[0].y & M, w[0].z & M, w[0].w & M);       // the first column has multiplier 1
      hi = make_uint4((w[0].x >> 4) & M, (w[0].y >> 4) & M, (w[0].z >> 4) & M, (w[0].w >> 4) & M);
#pragma unroll
      for (int k = 1; k < NCA; k++) {
        lo.x += (w[k].x & M) * m[k]; lo.y += (w[k].y & M) * m[k]; lo.z += (w[k].z & M) * m[k]; lo.w += (w[k].w & M) * m[k];
        hi.x += ((w[k].x >> 4) & M) * m[k]; hi.y += ((w[k].y >> 4) & M) * m[k];
        hi.z += ((w[k].z >> 4) & M) * m[k]; hi.w += ((w[k].w >> 4) & M) * m[k];
      }
    } else {
      lo = hi = make_uint4(0u, 0u, 0u, 0u);
      for (int k = 0; k < su.ncols; k++) {
        const uint4 w = ldg_stream(reinterpret_cast<const uint4 *>(u4 + (size_t)su.col[k] * stride4) + g);
        const unsigned mk = (unsigned)su.mul[k];
        lo.x += (w.x & M) * mk; lo.y += (w.y & M) * mk; lo.z += (w.z & M) * mk; lo.w += (w.w & M) * mk;
        hi.x += ((w.x >> 4) & M) * mk; hi.y += ((w.y >> 4) & M) * mk; hi.z += ((w.z >> 4) & M) * mk; hi.w += ((w.w >> 4) & M) * mk;
      }
    }
    small_bump(hist, lo.x, slot4); small_bump(hist, hi.x, slot4);
    small_bump(hist, lo.y, slot4); small_bump(hist, hi.y, slot4);
    small_bump(hist, lo.z, slot4); small_bump(hist, hi.z, slot4);
    small_bump(hist, lo.w, slot4); small_bump(hist, hi.w, slot4);
  }
}

// two codes per byte: 0xFF (NA, padding) becomes 0xF. grid (32-row groups, columns); only the
// columns flagged in `on` (at most 15 levels) are written.
__global__ void __launch_bounds__(256)
bnl_nibbles_kernel(const uint8_t *__restrict__ u8, size_t stride, const unsigned char *__restrict__ on,
                   uint8_t *__restrict__ u4, size_t stride4) {
  const int c = blockIdx.y;
  if (!on[c]) return;
  const uint4 *src = reinterpret_cast<const uint4 *>(u8 + (size_t)c * stride);
  uint4 *dst = reinterpret_cast<uint4 *>(u4 + (size_t)c * stride4);
  const long long n = (long long)(stride4 / 16);
  for (long long g = blockIdx.x * (long long)blockDim.x + threadIdx.x; g < n; g += (long long)gridDim.x * blockDim.x) {
    const uint4 a = ldg_stream(src + 2 * g), b = ldg_stream(src + 2 * g + 1);
    const unsigned w[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
    unsigned t[8];
#pragma unroll
    for (int i = 0; i < 8; i++) { const unsigned q = w[i] & 0x0F0F0F0Fu; t[i] = q | (q >> 4); }
    dst[g] = make_uint4(__byte_perm(t[0], t[1], 0x6420), __byte_perm(t[2], t[3], 0x6420),
                        __byte_perm(t[4], t[5], 0x6420), __byte_perm(t[6], t[7], 0x6420));
  }
}

template <bool NA, bool W16, bool PAIR>
__device__ __forceinline__ void small_dispatch(const uint8_t *__restrict__ u8, size_t stride,
                                               const BnlUnit &su, int *hist, unsigned slot4,
                                               long long g0, long long g1) {
  switch (su.ncols) {
    case 1: small_rows<1, NA, W16, PAIR>(u8, stride, su, hist, slot4, g0, g1); break;
    case 2: small_rows<2, NA, W16, PAIR>(u8, stride, su, hist, slot4, g0, g1); break;
    case 3: small_rows<3, NA, W16, PAIR>(u8, stride, su, hist, slot4, g0, g1); break;
    case 4: small_rows<4, NA, W16, PAIR>(u8, stride, su, hist, slot4, g0, g1); break;
    default:
      // nodes with 3-6 parents (mid-size tables only; in the small-table kernel the
      // extra instantiations cost config B 1 ms, scripts/group_sweep.sh): all loads of
      // a 16-row group issued together -- the generic loop uses each column as it
      // arrives, one round trip after the other.
      if (W16 && su.ncols == 5) small_rows<5, NA, W16, false>(u8, stride, su, hist, slot4, g0, g1);
      else if (W16 && su.ncols == 6) small_rows<6, NA, W16, false>(u8, stride, su, hist, slot4, g0, g1);
      else if (W16 && su.ncols == 7) small_rows<7, NA, W16, false>(u8, stride, su, hist, slot4, g0, g1);
      else if (W16 && su.ncols == 8) small_rows<8, NA, W16, false>(u8, stride, su, hist, slot4, g0, g1);
      else small_rows_any<NA, W16, PAIR>(u8, stride, su, hist, slot4, g0, g1);
  }
}

// W16 = false: tables <= 256 cells, 256 threads. W16 = true: 257..880 cells (the
// table fills most of the SM's shared memory, so one CTA of 1024 threads per SM).
template <bool W16>
__global__ void __launch_bounds__(W16 ? 1024 : BNL_CNT_THREADS)
count_small_kernel(const uint8_t *__restrict__ u8, size_t stride, long long nobs,
                   const BnlUnit *__restrict__ units, const int *__restrict__ order,
                   int *__restrict__ tables, int nunits, int chunks,
                   const uint8_t *__restrict__ u4 = nullptr, size_t stride4 = 0) {
  extern __shared__ __align__(256) int hist[];      // cells x 64 counters
  __shared__ BnlUnit su;
  const int chunk = blockIdx.x / nunits, u = order[blockIdx.x % nunits];
  if (threadIdx.x < sizeof(BnlUnit) / sizeof(int))
    reinterpret_cast<int *>(&su)[threadIdx.x] = reinterpret_cast<const int *>(&units[u])[threadIdx.x];
  __syncthreads();
  const int cells = su.cells;
  const bool pair = !W16 && su.mode == 1;            // pair table: slot a + 16 b, a, b < cells
  if (pair) {
    // only the cells x cells corner of the 16 x 16 pair table is ever touched.
    for (int e = threadIdx.x; e < cells * cells * 64; e += blockDim.x) {
      const int pc = e >> 6, a = pc % cells, b = pc / cells;
      hist[(a + 16 * b) * 64 + (e & 63)] = 0;
    }
  } else {
    for (int c = threadIdx.x; c < cells * 64; c += blockDim.x) hist[c] = 0;
  }
  __syncthreads();
  const unsigned slot4 = ((threadIdx.x & 31u) + 32u * ((threadIdx.x >> 5) & 1u)) << 2;

  int *gtab = tables + su.table_off;
  if (!W16 && su.mode == 2) {
    // nibble copy: complete 32-row groups; the rows left over, once, from the byte columns
    const long long full4 = nobs / 32;
    const long long per4 = (full4 + chunks - 1) / chunks;
    const long long h0 = per4 * chunk;
    long long h1 = h0 + per4;
    if (h1 > full4) h1 = full4;
    switch (su.ncols) {
      case 1: small_rows_nib<1>(u4, stride4, su, hist, slot4, h0, h1); break;
      case 2: small_rows_nib<2>(u4, stride4, su, hist, slot4, h0, h1); break;
      case 3: small_rows_nib<3>(u4, stride4, su, hist, slot4, h0, h1); break;
      case 4: small_rows_nib<4>(u4, stride4, su, hist, slot4, h0, h1); break;
      default: small_rows_nib<0>(u4, stride4, su, hist, slot4, h0, h1);
    }
    if (chunk == chunks - 1 && threadIdx.x == 0 && full4 * 32 < nobs)
      small_rows_scalar(u8, stride, su, hist, slot4, full4 * 32, (int)(nobs - full4 * 32));
    __syncthreads();
    for (int c = threadIdx.x; c < cells; c += blockDim.x) {
      int sum = 0;
#pragma unroll
      for (int l = 0; l < 64; l++) sum += hist[c * 64 + ((l + threadIdx.x) & 63)];
      if (sum) atomicAdd(&gtab[c], sum);
    }
    return;
  }
  const long long full = nobs / 16;                  // complete 16-row groups
  const long long per = (full + chunks - 1) / chunks;
  const long long g0 = per * chunk;
  long long g1 = g0 + per;
  if (g1 > full) g1 = full;

  if (pair) small_dispatch<false, false, true>(u8, stride, su, hist, slot4, g0, g1);
  else if (su.has_na) small_dispatch<true, W16, false>(u8, stride, su, hist, slot4, g0, g1);
  else small_dispatch<false, W16, false>(u8, stride, su, hist, slot4, g0, g1);
  // the partial group at the end of the data, once (pair mode: straight into the
  // global table, its shared table does not hold single rows).
  if (chunk == chunks - 1 && threadIdx.x == 0 && full * 16 < nobs) {
    if (pair) {
      for (long long r = full * 16; r < nobs; r++) {
        int cell = 0;
        for (int k = 0; k < su.ncols; k++) cell += (int)u8[(size_t)su.col[k] * stride + (size_t)r] * su.mul[k];
        atomicAdd(&gtab[cell], 1);
      }
    } else {
      small_rows_scalar(u8, stride, su, hist, slot4, full * 16, (int)(nobs - full * 16));
    }
  }

  __syncthreads();
  if (!pair) {
    for (int c = threadIdx.x; c < cells; c += blockDim.x) {
      int sum = 0;
#pragma unroll
      for (int l = 0; l < 64; l++) sum += hist[c * 64 + ((l + threadIdx.x) & 63)];
      if (sum) atomicAdd(&gtab[c], sum);
    }
    return;
  }
  // pair mode: collapse the lane copies, then each row of a pair adds its margin.
  __shared__ int psum[256];
  for (int e = threadIdx.x; e < cells * cells; e += blockDim.x) {
    const int c = (e % cells) + 16 * (e / cells);
    int sum = 0;
#pragma unroll
    for (int l = 0; l < 64; l++) sum += hist[c * 64 + ((l + threadIdx.x) & 63)];
    psum[c] = sum;
  }
  __syncthreads();
  for (int c = threadIdx.x; c < cells; c += blockDim.x) {
    int sum = 0;
    for (int o = 0; o < cells; o++) sum += psum[c + 16 * o] + psum[o + 16 * c];
    if (sum) atomicAdd(&gtab[c], sum);
  }
}

// ---------------------------------------------------------------------------
// Candidate groups. A score_cache_fill batch evaluates parents(t) + i for every
// candidate i: the units share every column but the last (the candidate, whose
// multiplier is the number of base cells). One CTA takes up to BNL_GROUP_MAX such
// units over a row chunk: the shared columns are loaded and combined into a packed
// base index ONCE per 16 rows, then each candidate costs one 128-bit load, four
// IMADs and the 16 (PRMT + ATOMS) updates of its own lane-replicated table. This
// removes the repeated loads of the target / parent columns (every unit re-reading
// its columns through L2 is what caps the per-unit kernel, at 86 % of the L2
// fabric) and brings the candidate batch to 0.57 of the HBM roofline, from 0.41
// (DESIGN.md section 3.1).
// tiles[] = {first position in `order`, number of units}.
#define BNL_GROUP_MAX 8

// The loads of the next 16-row group are issued before the current group is
// counted (software pipelining in registers): with the tables filling the shared
// memory only one CTA fits an SM, and without the overlap every warp sat out a
// full DRAM round trip per group (ncu: long-scoreboard stalls on top, LSU pipe
// under 50 %). NT threads per CTA; fewer threads leave the registers for the
// second set of loads. Up to BNL_GROUP_PB shared columns are prefetched raw.
#define BNL_GROUP_PB 4

struct GroupRegs {
  uint4 base[BNL_GROUP_PB];
  uint4 cand[BNL_GROUP_MAX];
};

template <int NT, int TSTRIDE>
__global__ void __launch_bounds__(NT)
count_group_kernel(const uint8_t *__restrict__ u8, size_t stride, long long nobs,
                   const BnlUnit *__restrict__ units, const int *__restrict__ order,
                   const int2 *__restrict__ tiles, int *__restrict__ tables, int ntiles, int chunks,
                   int tstride, int pf) {
  // unit j's table starts at j * tstride bytes: the same stride for every tile of
  // the launch (a kernel parameter, so the offset lives in a uniform register and
  // the update is PRMT + ATOMS [R + UR] with no address add in between).
  extern __shared__ __align__(256) int hist[];
  __shared__ BnlUnit su;                               // first unit of the tile: the shared columns
  __shared__ int s_cells[BNL_GROUP_MAX];
  __shared__ unsigned s_q[BNL_GROUP_MAX], s_qb[BNL_GROUP_PB];   // column starts in 16-byte units
  __shared__ long long s_tab[BNL_GROUP_MAX];
  if (TSTRIDE > 0) tstride = TSTRIDE;                    // compile-time stride: ATOMS [R + UR + imm]
  const unsigned stride16 = (unsigned)(stride >> 4);
  const int chunk = blockIdx.x / ntiles;
  const int2 tile = tiles[blockIdx.x % ntiles];
  const int nu = tile.y;
  if (threadIdx.x < sizeof(BnlUnit) / sizeof(int))
    reinterpret_cast<int *>(&su)[threadIdx.x] =
        reinterpret_cast<const int *>(&units[order[tile.x]])[threadIdx.x];
  if (threadIdx.x < nu) {
    const BnlUnit &uj = units[order[tile.x + threadIdx.x]];
    s_q[threadIdx.x] = (unsigned)uj.col[uj.ncols - 1] * stride16;
    s_cells[threadIdx.x] = uj.cells;
    s_tab[threadIdx.x] = uj.table_off;
  }
  __syncthreads();
  if (threadIdx.x < BNL_GROUP_PB) s_qb[threadIdx.x] = (unsigned)su.col[threadIdx.x] * stride16;
  const int total_quads = (nu * tstride) >> 4;
  for (int c = threadIdx.x; c < total_quads; c += NT) reinterpret_cast<uint4 *>(hist)[c] = make_uint4(0, 0, 0, 0);
  __syncthreads();

  const unsigned slot4 = ((threadIdx.x & 31u) + 32u * ((threadIdx.x >> 5) & 1u)) << 2;
  const int nbase = su.ncols - 1;                      // shared columns
  const int npre = nbase < BNL_GROUP_PB ? nbase : BNL_GROUP_PB;
  const unsigned cmul = (unsigned)su.mul[nbase];       // = number of base cells
  const long long full = nobs / 16;
  const long long per = (full + chunks - 1) / chunks;
  const long long g0 = per * chunk;
  long long g1 = g0 + per;
  if (g1 > full) g1 = full;

  // 32-bit offsets (16-byte units) from the start of the packed block: one add and
  // one widening multiply-add per load instead of 64-bit pointer arithmetic per
  // column and iteration (the host only groups when the block is below 64 GB).
  const uint4 *base4 = reinterpret_cast<const uint4 *>(u8);
  auto load = [&](GroupRegs &R, long long g) {
    const unsigned g32 = (unsigned)g;
#pragma unroll
    for (int k = 0; k < BNL_GROUP_PB; k++)
      if (k < npre) R.base[k] = ldg_stream(base4 + (s_qb[k] + g32));
#pragma unroll
    for (int j = 0; j < BNL_GROUP_MAX; j++)
      if (j < nu) R.cand[j] = ldg_stream(base4 + (s_q[j] + g32));
    // the candidates' own columns, `pf` thread-passes further on (one 128-byte line
    // covers 8 threads' loads: every 8th thread asks)
    if (pf > 0 && (threadIdx.x & 7) == 0 && g + (long long)pf * NT < g1) {
      const unsigned gp = g32 + (unsigned)pf * NT;
#pragma unroll
      for (int j = 0; j < BNL_GROUP_MAX; j++)
        if (j < nu) prefetch_l2(base4 + (s_q[j] + gp));
    }
  };

  GroupRegs cur, nxt;
  long long g = g0 + threadIdx.x;
  if (g < g1) load(cur, g);
  while (g < g1) {
    const long long gn = g + NT;
    if (gn < g1) load(nxt, gn);
    uint4 vb = make_uint4(0, 0, 0, 0);
#pragma unroll
    for (int k = 0; k < BNL_GROUP_PB; k++)
      if (k < npre) {
        const unsigned m = (unsigned)su.mul[k];
        vb.x += cur.base[k].x * m; vb.y += cur.base[k].y * m; vb.z += cur.base[k].z * m; vb.w += cur.base[k].w * m;
      }
    for (int k = BNL_GROUP_PB; k < nbase; k++) {       // wide parent sets: the rest is loaded here
      const uint4 w = ldg_stream(reinterpret_cast<const uint4 *>(u8 + (size_t)su.col[k] * stride) + g);
      const unsigned m = (unsigned)su.mul[k];
      vb.x += w.x * m; vb.y += w.y * m; vb.z += w.z * m; vb.w += w.w * m;
    }
#pragma unroll
    for (int j = 0; j < BNL_GROUP_MAX; j++)
      if (j < nu) {
        int *hj = reinterpret_cast<int *>(reinterpret_cast<char *>(hist) + j * tstride);
        small_bump(hj, vb.x + cur.cand[j].x * cmul, slot4);
        small_bump(hj, vb.y + cur.cand[j].y * cmul, slot4);
        small_bump(hj, vb.z + cur.cand[j].z * cmul, slot4);
        small_bump(hj, vb.w + cur.cand[j].w * cmul, slot4);
      }
    cur = nxt;
    g = gn;
  }
  // the partial group at the end of the data, once, straight into the global tables.
  if (chunk == chunks - 1 && threadIdx.x < nu && full * 16 < nobs) {
    const int j = threadIdx.x;
    for (long long r = full * 16; r < nobs; r++) {
      int cell = 0;
      for (int k = 0; k < nbase; k++) cell += (int)u8[(size_t)su.col[k] * stride + (size_t)r] * su.mul[k];
      cell += (int)u8[(size_t)s_q[j] * 16 + (size_t)r] * (int)cmul;
      atomicAdd(&tables[s_tab[j] + cell], 1);
    }
  }
  __syncthreads();
  for (int j = 0; j < nu; j++) {
    const int *hj = reinterpret_cast<const int *>(reinterpret_cast<const char *>(hist) + j * tstride);
    int *gtab = tables + s_tab[j];
    for (int c = threadIdx.x; c < s_cells[j]; c += NT) {
      int sum = 0;
#pragma unroll
      for (int l = 0; l < 64; l++) sum += hj[c * 64 + ((l + threadIdx.x) & 63)];
      if (sum) atomicAdd(&gtab[c], sum);
    }
  }
}

// ---------------------------------------------------------------------------
// Candidate runs with mid-size tables (a node that already has 2-5 parents: 257
// to ~7000 cells). The tables are too large to share an SM, so the group kernel
// above does not apply and every unit would re-read its 4-8 columns through L2.
// Instead the shared columns are combined ONCE per run into a uint16 column of
// base indices (the reference does the same for a test's conditioning set:
// src/tests/patterns/ctest.c:18-19 forms the configuration of Z once for all x),
// and each candidate becomes a two-column unit: 3 bytes per row instead of
// ncols, one PRMT + IMAD per two rows of index arithmetic.
__global__ void __launch_bounds__(256)
combine_base_kernel(const uint8_t *__restrict__ u8, size_t stride, long long full,
                    const BnlUnit *__restrict__ units, const int *__restrict__ run_first,
                    uint16_t *__restrict__ basecols, size_t bstride) {
  __shared__ int s_col[BNL_MAXK], s_mul[BNL_MAXK];
  __shared__ int s_nbase;
  const BnlUnit &un = units[run_first[blockIdx.y]];
  if (threadIdx.x == 0) s_nbase = un.ncols - 1;
  if (threadIdx.x < BNL_MAXK) { s_col[threadIdx.x] = un.col[threadIdx.x]; s_mul[threadIdx.x] = un.mul[threadIdx.x]; }
  __syncthreads();
  const int nbase = s_nbase;
  uint4 *out = reinterpret_cast<uint4 *>(basecols + (size_t)blockIdx.y * bstride);
  if ((un.mode >> 30) & 1) {
    // at most 256 base cells: the index fits a byte, 16 rows per 128-bit word (the candidate
    // kernel then loads 2 bytes per row instead of 3, which takes it to the update rate)
    for (long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x; g < full;
         g += (long long)gridDim.x * blockDim.x) {
      uint4 v = make_uint4(0u, 0u, 0u, 0u);
      for (int k = 0; k < nbase; k++) {
        const uint4 w = ldg_stream(reinterpret_cast<const uint4 *>(u8 + (size_t)s_col[k] * stride) + g);
        const unsigned m = (unsigned)s_mul[k];
        v.x += w.x * m; v.y += w.y * m; v.z += w.z * m; v.w += w.w * m;       // bytewise: every sum stays below 256
      }
      out[g] = v;
    }
    return;
  }
  for (long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x; g < full;
       g += (long long)gridDim.x * blockDim.x) {
    unsigned v[8];
#pragma unroll
    for (int j = 0; j < 8; j++) v[j] = 0u;
    for (int k = 0; k < nbase; k++) {
      const uint4 w = ldg_stream(reinterpret_cast<const uint4 *>(u8 + (size_t)s_col[k] * stride) + g);
      const unsigned m = (unsigned)s_mul[k];
      const unsigned ww[4] = {w.x, w.y, w.z, w.w};
#pragma unroll
      for (int q = 0; q < 4; q++) {
        v[2 * q] += __byte_perm(ww[q], 0, 0x4140) * m;
        v[2 * q + 1] += __byte_perm(ww[q], 0, 0x4342) * m;
      }
    }
    out[2 * g] = make_uint4(v[0], v[1], v[2], v[3]);
    out[2 * g + 1] = make_uint4(v[4], v[5], v[6], v[7]);
  }
}

// LC = log2 of the lane counters per cell: 5 (one counter per lane, no bank
// collisions; <= ~1780 cells), 4 or 3 (<= ~7000 cells; lanes 16 or 8 apart then
// share a counter, still far fewer collisions than one shared table). A CTA takes
// a tile of up to NUMAX units of one run (tables of <= ~445 cells: four per SM,
// <= ~890: two), so the base column is loaded once per tile: (2 + nu) / nu bytes
// per row-unit through L2. tiles[] = {first position in `order`, units}.
#define BNL_BASED_MAX 8

// LC = 0: the counters per cell are the unit's own (bits 24.. of BnlUnit::mode), so that
// tables of every size class can share one launch; the bytes per cell are then a register operand.
// `pre` is the 32-bit shared address of this lane's counter of cell 0 and `v` holds two 16-bit cell
// indices. A byte-permute takes each half (the compiler cannot fold it into the shift-and-mask it
// otherwise builds), cell * bytes + pre is one IMAD and red.shared.add 1 an ATOMS.POPC.INC: three
// instructions per update instead of four -- and the loop is issue-bound, not only update-bound
// (scripts/based_bytes_probe.cu: 0.235 -> 0.224 ms on 999 units x 1 M rows).
template <int LC>
__device__ __forceinline__ void based_bump(unsigned pre, unsigned v, unsigned bytes) {
  if (LC) bytes = 1u << (LC + 2);
  asm volatile("red.shared.add.u32 [%0], 1;" ::"r"(__byte_perm(v, 0, 0x4410) * bytes + pre));
  asm volatile("red.shared.add.u32 [%0], 1;" ::"r"(__byte_perm(v, 0, 0x4432) * bytes + pre));
}

template <int NUMAX>
struct BasedRegs { uint4 b0, b1, c[NUMAX]; };

// NT threads; the loads of the next TWO 16-row groups are in flight while the
// current one is counted (one CTA per SM: nothing else hides the round trip, and
// one group ahead left the kernel latency-bound at 40 GB/s per SM).
// B8: the base indices of every run of the launch are bytes (combine_base_kernel, at most 256 base
// cells): 2 bytes per row instead of 3, which takes the loop to the update rate.
template <int LC, int NUMAX, int NT, bool B8 = false>
__global__ void __launch_bounds__(NT, (NUMAX == 1 && NT == 512) ? 2 : 1)
count_based_kernel(const uint8_t *__restrict__ u8, size_t stride, long long nobs,
                   const uint16_t *__restrict__ basecols, size_t bstride,
                   const BnlUnit *__restrict__ units, const int *__restrict__ order,
                   const int2 *__restrict__ tiles, int *__restrict__ tables, int ntiles, int chunks,
                   int tstride, int pf) {
  extern __shared__ __align__(256) int hist[];        // unit j: tstride bytes from j * tstride
  __shared__ BnlUnit su;                               // first unit of the tile: run, shared columns
  __shared__ int s_cells[BNL_BASED_MAX], s_col[BNL_BASED_MAX];
  __shared__ long long s_tab[BNL_BASED_MAX];
  const int chunk = blockIdx.x / ntiles;
  const int2 tile = tiles[blockIdx.x % ntiles];
  const int nu = NUMAX == 1 ? 1 : tile.y;
  if (threadIdx.x < sizeof(BnlUnit) / sizeof(int))
    reinterpret_cast<int *>(&su)[threadIdx.x] =
        reinterpret_cast<const int *>(&units[order[tile.x]])[threadIdx.x];
  if (threadIdx.x < nu) {
    const BnlUnit &uj = units[order[tile.x + threadIdx.x]];
    s_col[threadIdx.x] = uj.col[uj.ncols - 1];
    s_cells[threadIdx.x] = uj.cells;
    s_tab[threadIdx.x] = uj.table_off;
  }
  __syncthreads();
  const int lc = LC ? LC : ((su.mode >> 24) & 0x3F);
  constexpr bool base8 = B8;
  static_assert(!B8 || NUMAX == 1, "byte base indices: one unit per CTA");
  // one unit per CTA: only its own table is cleared (the launch's stride is the largest one)
  const int quads = NUMAX == 1 ? ((s_cells[0] << (lc + 2)) + 15) >> 4 : (nu * tstride) >> 4;
  for (int c = threadIdx.x; c < quads; c += NT) reinterpret_cast<uint4 *>(hist)[c] = make_uint4(0, 0, 0, 0);
  __syncthreads();
  const unsigned slot4 = (threadIdx.x & ((1u << lc) - 1u)) << 2;
  const unsigned pre0 = (unsigned)__cvta_generic_to_shared(hist) + slot4;      // this lane's counter of cell 0, unit 0
  const unsigned cbytes = 1u << (lc + 2);
  const int last = su.ncols - 1;
  const unsigned cmul = (unsigned)su.mul[last];
  const uint4 *bp = reinterpret_cast<const uint4 *>(basecols + (size_t)((su.mode & 0xFFFFFF) - 16) * bstride);
  const uint4 *cp[NUMAX];
#pragma unroll
  for (int j = 0; j < NUMAX; j++)
    cp[j] = reinterpret_cast<const uint4 *>(u8 + (size_t)s_col[j < nu ? j : 0] * stride);

  const long long full = nobs / 16;
  const long long per = (full + chunks - 1) / chunks;
  const long long g0 = per * chunk;
  long long g1 = g0 + per;
  if (g1 > full) g1 = full;

  auto load = [&](BasedRegs<NUMAX> &R, long long g) {
    if (base8) R.b0 = ldg_stream(bp + g);
    else {
      R.b0 = ldg_stream(bp + 2 * g);
      R.b1 = ldg_stream(bp + 2 * g + 1);
    }
#pragma unroll
    for (int j = 0; j < NUMAX; j++)
      if (j < nu) R.c[j] = ldg_stream(cp[j] + g);
    if (pf > 0 && (threadIdx.x & 7) == 0 && g + (long long)pf * NT < g1) {
#pragma unroll
      for (int j = 0; j < NUMAX; j++)
        if (j < nu) prefetch_l2(cp[j] + g + (long long)pf * NT);
    }
  };
  auto count = [&](const BasedRegs<NUMAX> &R) {
    if (base8) {                       // (one unit per CTA)
      const unsigned b4[4] = {R.b0.x, R.b0.y, R.b0.z, R.b0.w};
      const unsigned cw[4] = {R.c[0].x, R.c[0].y, R.c[0].z, R.c[0].w};
#pragma unroll
      for (int q = 0; q < 4; q++) {
        based_bump<LC>(pre0, __byte_perm(b4[q], 0, 0x4140) + __byte_perm(cw[q], 0, 0x4140) * cmul, cbytes);
        based_bump<LC>(pre0, __byte_perm(b4[q], 0, 0x4342) + __byte_perm(cw[q], 0, 0x4342) * cmul, cbytes);
      }
      return;
    }
    const unsigned bw[8] = {R.b0.x, R.b0.y, R.b0.z, R.b0.w, R.b1.x, R.b1.y, R.b1.z, R.b1.w};
#pragma unroll
    for (int j = 0; j < NUMAX; j++)
      if (j < nu) {
        const unsigned pj = pre0 + (unsigned)(j * tstride);
        const unsigned cw[4] = {R.c[j].x, R.c[j].y, R.c[j].z, R.c[j].w};
#pragma unroll
        for (int q = 0; q < 4; q++) {
          based_bump<LC>(pj, bw[2 * q] + __byte_perm(cw[q], 0, 0x4140) * cmul, cbytes);
          based_bump<LC>(pj, bw[2 * q + 1] + __byte_perm(cw[q], 0, 0x4342) * cmul, cbytes);
        }
      }
  };
  BasedRegs<NUMAX> r0, r1, r2;
  long long g = g0 + threadIdx.x;
  if (g < g1) load(r0, g);
  if (g + NT < g1) load(r1, g + NT);
#define BNL_BASED_STEP(A, C)                  \
  if (g < g1) {                               \
    if (g + 2 * NT < g1) load(C, g + 2 * NT); \
    count(A);                                 \
    g += NT;                                  \
  }
  while (g < g1) {
    BNL_BASED_STEP(r0, r2)
    BNL_BASED_STEP(r1, r0)
    BNL_BASED_STEP(r2, r1)
  }
#undef BNL_BASED_STEP
  // the partial group at the end of the data, once, from the original columns.
  if (chunk == chunks - 1 && threadIdx.x < nu && full * 16 < nobs) {
    const int j = threadIdx.x;
    for (long long r = full * 16; r < nobs; r++) {
      int cell = 0;
      for (int k = 0; k < last; k++) cell += (int)u8[(size_t)su.col[k] * stride + (size_t)r] * su.mul[k];
      cell += (int)u8[(size_t)s_col[j] * stride + (size_t)r] * (int)cmul;
      atomicAdd(&tables[s_tab[j] + cell], 1);
    }
  }
  __syncthreads();
  const int ncnt = 1 << lc;
  for (int j = 0; j < nu; j++) {
    const int *hj = reinterpret_cast<const int *>(reinterpret_cast<const char *>(hist) + j * tstride);
    int *gtab = tables + s_tab[j];
    for (int c = threadIdx.x; c < s_cells[j]; c += NT) {
      int sum = 0;
#pragma unroll 8
      for (int l = 0; l < ncnt; l++) sum += hj[(c << lc) + ((l + threadIdx.x) & (ncnt - 1))];
      if (sum) atomicAdd(&gtab[c], sum);
    }
  }
}

// ---------------------------------------------------------------------------
// Row pairs for tables of at most 32 cells, from the nibble copy of the columns. The update
// rate of shared memory bounds every counting kernel here (16 lanes/clk/SM), so the cheapest
// row is the one that shares its update: byte j of a nibble column holds rows 2j and 2j + 1, and
// the pair (cell a of the even row, cell b of the odd row) bumps slot a + 32 b of a 32 x 32
// pair table -- one update per TWO rows. The table keeps one counter per lane and slot
// (1024 slots x 128 bytes = 128 KB: one CTA of 1024 threads per SM); at the end cell c gets
// the sum of row c and of column c of the pair table. Units without NA whose columns all
// have a nibble copy; the rows past the last 32-row group are counted singly by one thread.
// A third of config B's conditional tests (tables of 17 - 32 cells) come here, and the
// few of <= 16 cells that the 16 x 16 pair mode of count_small_kernel does not take.
template <int NC>
__device__ __forceinline__ void pair32_rows(const uint8_t *__restrict__ u4, size_t stride4, const BnlUnit &su,
                                            unsigned pre, long long g0, long long g1) {
  constexpr int NCA = NC ? NC : 1;
  const uint4 *p[NCA];
  unsigned m[NCA];
#pragma unroll
  for (int k = 0; k < NCA; k++) {
    p[k] = reinterpret_cast<const uint4 *>(u4 + (size_t)su.col[k] * stride4);
    m[k] = (unsigned)su.mul[k];
  }
  const unsigned M = 0x0F0F0F0Fu;
  auto pairs = [&](unsigned lo, unsigned hi) {
    // bytes of lo: cells of rows 0, 2, 4, 6 of these 8; of hi: rows 1, 3, 5, 7. Two 16-bit slots per register.
    based_bump<5>(pre, __byte_perm(lo, 0, 0x4140) + __byte_perm(hi, 0, 0x4140) * 32u, 128u);
    based_bump<5>(pre, __byte_perm(lo, 0, 0x4342) + __byte_perm(hi, 0, 0x4342) * 32u, 128u);
  };
  uint4 w[NCA], wn[NCA];
  long long g = g0 + threadIdx.x;
  if (NC && g < g1) {
#pragma unroll
    for (int k = 0; k < NCA; k++) w[k] = ldg_stream(p[k] + g);
  }
  for (; g < g1; g += blockDim.x) {
    uint4 lo, hi;
    if (NC) {
      const long long gn = g + blockDim.x;             // the next group's loads in flight while this one is counted
      if (gn < g1) {
#pragma unroll
        for (int k = 0; k < NCA; k++) wn[k] = ldg_stream(p[k] + gn);
      }
      lo = make_uint4(w[0].x & M, w[0].y & M, w[0].z & M, w[0].w & M);       // the first column has multiplier 1
      hi = make_uint4((w[0].x >> 4) & M, (w[0].y >> 4) & M, (w[0].z >> 4) & M, (w[0].w >> 4) & M);
#pragma unroll
      for (int k = 1; k < NCA; k++) {
        lo.x += (w[k].x & M) * m[k]; lo.y += (w[k].y & M) * m[k]; lo.z += (w[k].z & M) * m[k]; lo.w += (w[k].w & M) * m[k];
        hi.x += ((w[k].x >> 4) & M) * m[k]; hi.y += ((w[k].y >> 4) & M) * m[k];
        hi.z += ((w[k].z >> 4) & M) * m[k]; hi.w += ((w[k].w >> 4) & M) * m[k];
      }
#pragma unroll
      for (int k = 0; k < NCA; k++) w[k] = wn[k];
    } else {
      lo = hi = make_uint4(0u, 0u, 0u, 0u);
      for (int k = 0; k < su.ncols; k++) {
        const uint4 x = ldg_stream(reinterpret_cast<const uint4 *>(u4 + (size_t)su.col[k] * stride4) + g);
        const unsigned mk = (unsigned)su.mul[k];
        lo.x += (x.x & M) * mk; lo.y += (x.y & M) * mk; lo.z += (x.z & M) * mk; lo.w += (x.w & M) * mk;
        hi.x += ((x.x >> 4) & M) * mk; hi.y += ((x.y >> 4) & M) * mk; hi.z += ((x.z >> 4) & M) * mk; hi.w += ((x.w >> 4) & M) * mk;
      }
    }
    pairs(lo.x, hi.x); pairs(lo.y, hi.y); pairs(lo.z, hi.z); pairs(lo.w, hi.w);
  }
}

__global__ void __launch_bounds__(1024)
count_pair32_kernel(const uint8_t *__restrict__ u8, size_t stride, const uint8_t *__restrict__ u4, size_t stride4,
                    long long nobs, const BnlUnit *__restrict__ units, const int *__restrict__ order,
                    int *__restrict__ tables, int nunits, int chunks) {
  extern __shared__ __align__(256) int hist[];        // 1024 slots x 32 lane counters
  __shared__ BnlUnit su;
  __shared__ int psum[1024];
  const int chunk = blockIdx.x / nunits, u = order[blockIdx.x % nunits];
  if (threadIdx.x < sizeof(BnlUnit) / sizeof(int))
    reinterpret_cast<int *>(&su)[threadIdx.x] = reinterpret_cast<const int *>(&units[u])[threadIdx.x];
  __syncthreads();
  const int cells = su.cells;
  // slots a + 32 b with b < cells: the first `cells` rows of 32 slots
  for (int q = threadIdx.x; q < cells * 32 * 8; q += blockDim.x) reinterpret_cast<uint4 *>(hist)[q] = make_uint4(0, 0, 0, 0);
  __syncthreads();
  const unsigned pre = (unsigned)__cvta_generic_to_shared(hist) + ((threadIdx.x & 31u) << 2);
  const long long full4 = nobs / 32;
  const long long per4 = (full4 + chunks - 1) / chunks;
  const long long h0 = per4 * chunk;
  long long h1 = h0 + per4;
  if (h1 > full4) h1 = full4;
  switch (su.ncols) {
    case 1: pair32_rows<1>(u4, stride4, su, pre, h0, h1); break;
    case 2: pair32_rows<2>(u4, stride4, su, pre, h0, h1); break;
    case 3: pair32_rows<3>(u4, stride4, su, pre, h0, h1); break;
    case 4: pair32_rows<4>(u4, stride4, su, pre, h0, h1); break;
    default: pair32_rows<0>(u4, stride4, su, pre, h0, h1);
  }
  int *gtab = tables + su.table_off;
  // the rows left over, once, singly, straight into the global table (the pair table holds no single rows)
  if (chunk == chunks - 1 && threadIdx.x == 0)
    for (long long r = full4 * 32; r < nobs; r++) {
      int cell = 0;
      for (int k = 0; k < su.ncols; k++) cell += (int)u8[(size_t)su.col[k] * stride + (size_t)r] * su.mul[k];
      atomicAdd(&gtab[cell], 1);
    }
  __syncthreads();
  // lane copies of every slot of the cells x cells corner, then each row of a pair adds its margin
  for (int e = threadIdx.x; e < cells * cells; e += blockDim.x) {
    const int slot = (e % cells) + 32 * (e / cells);
    int sum = 0;
#pragma unroll
    for (int l = 0; l < 32; l++) sum += hist[slot * 32 + ((l + threadIdx.x) & 31)];
    psum[slot] = sum;
  }
  __syncthreads();
  for (int c = threadIdx.x; c < cells; c += blockDim.x) {
    int sum = 0;
    for (int o = 0; o < cells; o++) sum += psum[c + 32 * o] + psum[o + 32 * c];
    if (sum) atomicAdd(&gtab[c], sum);
  }
}

// ---------------------------------------------------------------------------
// Unit descriptors of a batch made only of one- and two-way tables, written on the
// device: the host sends 8 bytes of column slots and the table offset per unit
// instead of a 160-byte BnlUnit (499 500 units in the first hc iteration: 80 MB of
// descriptors built, copied twice and uploaded were most of the call's wall time).
// cc = {first column, second column or -1} in make_unit's order; score_like: the
// first column is the node and the second its parent (a stratum per parent level),
// otherwise the first is y and the second x of a two-way table in one stratum.
__global__ void __launch_bounds__(256)
build_pair_units_kernel(const int2 *__restrict__ cc, const long long *__restrict__ off,
                        const int *__restrict__ lev_by_slot, int score_like, int n, BnlUnit *__restrict__ units) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n) return;
  const int2 c = cc[t];
  const int l0 = lev_by_slot[c.x], l1 = c.y >= 0 ? lev_by_slot[c.y] : 1;
  BnlUnit u;
  int *w = reinterpret_cast<int *>(&u);
#pragma unroll
  for (int i = 0; i < (int)(sizeof(BnlUnit) / sizeof(int)); i++) w[i] = 0;
  u.table_off = off[t];
  u.cells = l0 * l1;
  u.ncols = c.y >= 0 ? 2 : 1;
  if (score_like || c.y < 0) { u.llx = l0; u.lly = 1; u.llz = l1; }
  else { u.lly = l0; u.llx = l1; u.llz = 1; }
  u.col[0] = c.x; u.mul[0] = 1;
  if (c.y >= 0) { u.col[1] = c.y; u.mul[1] = l0; }
  static_assert(sizeof(BnlUnit) % 8 == 0, "BnlUnit is copied in 8-byte words");
  long long *dst = reinterpret_cast<long long *>(units + t);      // 168 bytes: 8-byte aligned, not 16
  const long long *src = reinterpret_cast<const long long *>(&u);
#pragma unroll
  for (int i = 0; i < (int)(sizeof(BnlUnit) / 8); i++) dst[i] = src[i];
}

// ---------------------------------------------------------------------------
// Epilogue. kind 0: CI test (code = test_e); kind 1: node score (code = score_e).
// One CTA per unit, one warp per stratum (configuration of Z / of the parents).

__device__ __forceinline__ double warp_sum_d(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

__device__ __forceinline__ long long warp_sum_ll(long long v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

__device__ __forceinline__ void warp_merge_kahan(bnl::KahanSum &k) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const double os = __shfl_xor_sync(0xffffffffu, k.s, o);
    const double oc = __shfl_xor_sync(0xffffffffu, k.c, o);
    k.merge(os, oc);
  }
}

// Multinomial NML regret log C(n, k) (nml_regret, src/scores/nml.regret.c:36-62):
// exact table for n <= 1000 and k <= 100, else the smaller of the two
// Szpankowski-style approximations (:6-33).
__device__ __forceinline__ double nml_regret(const double *__restrict__ tab, double n, double k) {
  if (n == 1.0) return log(k);
  if (k == 1.0 || n == 0.0) return 0.0;
  if (n <= (double)BNL_REGRET_N && k <= (double)BNL_REGRET_K)
    return tab[(int)k + (size_t)(int)n * (BNL_REGRET_K + 1)];
  const double hK = 0.5 * k, hKh = hK - 0.5;
  const double lg = bnl::lgamma_pos(hK), ratio = exp(lg - bnl::lgamma_pos(hKh));
  const double gterm = 0.70710678118654752440 - ratio * k / 3.0 / sqrt(n);
  const double a1 = hKh * (log(n) - 0.69314718055994530942) + 0.572364942924700087071713675677 - lg +
                    0.5 - gterm * gterm + (3.0 + k * (k - 2.0) * (2.0 * k + 1.0)) / (36.0 * n);
  const double a = k / n, Ca = 0.5 * (1.0 + sqrt(1.0 + 4.0 / a));
  const double a2 = n * (log(a) + (a + 2.0) * log(Ca) - 1.0 / Ca) - 0.5 * log(Ca + 2.0 / a);
  return a1 < a2 ? a1 : a2;
}

// EXT compiles in the Jonckheere-Terpstra and shrinkage-MI branches; the common
// statistics keep their own, leaner instantiation (the extra code costs registers).
// WPU ("warp per unit"): batches of many tiny tables -- the first hc iteration asks for
// 499 500 of them -- give every warp of the CTA its own unit instead of a CTA each
// (CTA scheduling, not arithmetic, was the cost: 2.9 ms). The host only selects it when
// one warp does all the work of the unit anyway (bnl_epilogue_warp_ok), so the lanes
// see the same cells and the sums are formed in the same order: identical results.
template <bool EXT, bool WPU = false>
__global__ void __launch_bounds__(BNL_EPI_THREADS)
epilogue_kernel(const BnlUnit *__restrict__ units, const int *__restrict__ tables,
                double *__restrict__ out, int kind, int code, double arg, long long nrows,
                const double *__restrict__ regret, int marg_stride, int nunits = 0) {
  // margins of one stratum per warp: llx + lly ints, in dynamic shared memory so
  // that tests on factors with thousands of levels (uint16 columns) fit as well.
  extern __shared__ int s_marg_dyn[];
  __shared__ double s_s[BNL_EPI_WARPS], s_c[BNL_EPI_WARPS], s_df[BNL_EPI_WARPS];
  __shared__ long long s_n[BNL_EPI_WARPS];
  __shared__ int s_q;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int unit = WPU ? blockIdx.x * BNL_EPI_WARPS + warp : blockIdx.x;
  if (WPU && unit >= nunits) return;
  const BnlUnit U = units[unit];
  const int *T = tables + U.table_off;
  // position of this warp / thread among those that share the unit
  const int sw = WPU ? 0 : warp, nw = WPU ? 1 : BNL_EPI_WARPS;
  const int tid = WPU ? lane : (int)threadIdx.x, nthr = WPU ? 32 : BNL_EPI_THREADS;
  const int llx = U.llx, lly = U.lly, llz = U.llz, sl = llx * lly;
  int *marg = s_marg_dyn + (size_t)warp * marg_stride;

  // BDs: the prior is spread over the OBSERVED parent configurations only
  // (src/scores/dirichlet.posterior.c:180 -> all_levels = FALSE).
  int q = llz;
  if (!WPU && kind == 1 && code == BNL_BDS) {
    if (threadIdx.x == 0) s_q = 0;
    __syncthreads();
    int mine = 0;
    for (int k = warp; k < llz; k += BNL_EPI_WARPS) {
      long long nk = 0;
      for (int c = lane; c < sl; c += 32) nk += T[(size_t)k * sl + c];
      nk = warp_sum_ll(nk);
      mine += nk > 0;
    }
    if (lane == 0 && mine) atomicAdd(&s_q, mine);
    __syncthreads();
    q = s_q;
  }

  // mi-sh: the James-Stein shrinkage intensity needs the whole table first
  // (mi_lambda, src/tests/discrete/shrinkage.c:138-184): two block-wide passes,
  // the complete-row count, then sum (n/N)^2 and sum (target - n/N)^2.
  double sh_lambda = 0.0, sh_target = 0.0, sh_n = 0.0;
  if (EXT && !WPU && kind == 0 && code == BNL_MI_SH) {
    const long long cells = (long long)sl * llz;
    long long loc = 0;
    for (long long c = threadIdx.x; c < cells; c += BNL_EPI_THREADS) loc += T[c];
    loc = warp_sum_ll(loc);
    if (lane == 0) s_n[warp] = loc;
    __syncthreads();
    long long ntot = 0;
    for (int w = 0; w < BNL_EPI_WARPS; w++) ntot += s_n[w];
    __syncthreads();
    sh_n = (double)ntot;
    sh_target = 1.0 / ((double)llx * (double)lly * (double)llz);
    bnl::KahanSum lnum, lden;
    if (ntot > 1) {
      for (long long c = threadIdx.x; c < cells; c += BNL_EPI_THREADS) {
        const double f = (double)T[c] / sh_n;
        lnum.add(f * f);
        lden.add((sh_target - f) * (sh_target - f));
      }
    }
    warp_merge_kahan(lnum);
    warp_merge_kahan(lden);
    if (lane == 0) { s_s[warp] = lnum.s; s_c[warp] = lnum.c; s_df[warp] = lden.value(); }
    __syncthreads();
    bnl::KahanSum tn;
    double td = 0.0;
    for (int w = 0; w < BNL_EPI_WARPS; w++) { tn.merge(s_s[w], s_c[w]); td += s_df[w]; }
    __syncthreads();
    sh_lambda = td == 0.0 ? 1.0 : (1.0 - tn.value()) / ((sh_n - 1.0) * td);
    sh_lambda = fmin(fmax(sh_lambda, 0.0), 1.0);
  }

  bnl::KahanSum acc;
  double dfacc = 0.0;
  long long nobs = 0;
  const bool per_cell = kind == 1 && (code == BNL_BDJ || code == BNL_K2);
  const bool is_bd = kind == 1 && (code == BNL_BDE || code == BNL_BDS || per_cell);
  // BDe / BDs: alpha = iss / (llx q), prior mass iss / q per parent configuration.
  // BDJ / K2 (per-cell hyperparameter a = 1/2 or 1): alpha = a and a * llx per
  // configuration -- except that for a node WITHOUT parents the reference uses
  // llx itself, whatever a is (src/scores/dirichlet.posterior.c:19-25 vs :81-88).
  const double pc = code == BNL_K2 ? 1.0 : 0.5;
  double alpha = 1.0, aq = 1.0;
  if (per_cell) { alpha = pc; aq = U.ncols == 1 ? (double)llx : pc * (double)llx; }
  else if (is_bd && q > 0) { alpha = arg / ((double)llx * (double)q); aq = arg / (double)q; }
  const double lg_alpha = is_bd ? bnl::lgamma_pos(alpha) : 0.0;
  const double lg_aq = is_bd ? bnl::lgamma_pos(aq) : 0.0;

  // Scores of narrow nodes (a stratum = one parent configuration holds only llx
  // cells): one LANE per stratum, otherwise a warp would run with 2-4 active lanes
  // through the lgamma / log terms.
  const bool lane_strata = kind == 1 && sl <= 16;
  if (lane_strata) {
    for (int k = tid; k < llz; k += nthr) {
      const int *S = T + (size_t)k * sl;
      long long nk = 0;
      for (int c = 0; c < sl; c++) nk += S[c];
      nobs += nk;
      if (is_bd) {
        if (q == 0 || (code == BNL_BDS && nk == 0)) continue;
        for (int c = 0; c < sl; c++) acc.add(bnl::lgamma_pos((double)S[c] + alpha) - lg_alpha);
        acc.add(lg_aq - bnl::lgamma_pos((double)nk + aq));
      } else if (nk > 0) {
        if (code == BNL_QNML && U.ncols > 1) {          // cdlik_diff(): sum n log n - n_j log n_j
          for (int c = 0; c < sl; c++) { const int n = S[c]; if (n) acc.add((double)n * log((double)n)); }
          acc.add(-(double)nk * log((double)nk));
        } else {
          for (int c = 0; c < sl; c++) {
            const int n = S[c];
            if (n) acc.add((double)n * log((double)n / (double)nk));
          }
          if (code == BNL_FNML || code == BNL_QNML) acc.add(-nml_regret(regret, (double)nk, (double)llx));
        }
      }
    }
  }
  for (int k = lane_strata ? llz : sw; k < llz; k += nw) {
    const int *S = T + (size_t)k * sl;
    if (kind == 0) {
      for (int i = lane; i < llx + lly; i += 32) marg[i] = 0;
      __syncwarp();
      long long nk = 0;
      for (int c = lane; c < sl; c += 32) {
        const int n = S[c];
        if (n) {
          atomicAdd(&marg[c / lly], n);
          atomicAdd(&marg[llx + c % lly], n);
          nk += n;
        }
      }
      nk = warp_sum_ll(nk);
      __syncwarp();
      if (code == BNL_MI_ADF || code == BNL_X2_ADF) {
        int alx = 0, aly = 0;
        for (int i = lane; i < llx; i += 32) alx += marg[i] > 0;
        for (int j = lane; j < lly; j += 32) aly += marg[llx + j] > 0;
        alx = (int)warp_sum_ll(alx);
        aly = (int)warp_sum_ll(aly);
        if (alx < 1) alx = 1;
        if (aly < 1) aly = 1;
        if (lane == 0) dfacc += (double)(alx - 1) * (double)(aly - 1);
      }
      if (lane == 0) nobs += nk;
      if (EXT && code == BNL_JT) {
        // Jonckheere-Terpstra (src/tests/discrete/jonckheere.c:98-172): the pair
        // (i > j) of x levels is one lane's work, with a running prefix over the
        // y levels instead of the reference's inner re-summation. Every partial
        // value is a multiple of 1/4, so the sums are exact below 2^51.
        if (nk > 0) {
          double st = 0.0;
          const int npairs = llx * (llx - 1) / 2;
          for (int pr = lane; pr < npairs; pr += 32) {
            int i = (int)((1.0 + sqrt(1.0 + 8.0 * (double)pr)) * 0.5);
            while (i * (i - 1) / 2 > pr) i--;
            while ((i + 1) * i / 2 <= pr) i++;
            const int j = pr - i * (i - 1) / 2;
            double w = 0.0, wi = 0.0;
            for (int t = 0; t < lly; t++) {
              const double a = (double)S[i * lly + t], b = (double)S[j * lly + t];
              w += (wi + (a + b + 1.0) / 2.0) * a;
              wi += a + b;
            }
            const double ni = (double)marg[i];
            st += w - ni * (ni + 1.0) / 2.0;
          }
          st = warp_sum_d(st);
          if (lane == 0) {
            const double dn = (double)nk;
            double mean = dn * dn;       // in double: the reference's int product overflows past 46340 rows
            for (int i = 0; i < llx; i++) mean -= (double)marg[i] * (double)marg[i];
            mean /= 4.0;
            double U1 = dn * (dn - 1.0) * (2.0 * dn + 5.0), U2a = 0.0, U2b = 0.0, U3a = 0.0, U3b = 0.0;
            for (int i = 0; i < llx; i++) { const double v = (double)marg[i]; U1 -= v * (v - 1.0) * (2.0 * v + 5.0); }
            for (int j = 0; j < lly; j++) { const double v = (double)marg[llx + j]; U1 -= v * (v - 1.0) * (2.0 * v + 5.0); }
            for (int i = 0; i < llx; i++) { const double v = (double)marg[i]; U2a += v * (v - 1.0) * (v - 2.0); U3a += v * (v - 1.0); }
            for (int j = 0; j < lly; j++) { const double v = (double)marg[llx + j]; U2b += v * (v - 1.0) * (v - 2.0); U3b += v * (v - 1.0); }
            const double var = U1 / 72.0 + U2a * U2b / (36.0 * dn * (dn - 1.0) * (dn - 2.0)) +
                               U3a * U3b / (8.0 * dn * (dn - 1.0));
            acc.add(st - mean);
            // c_jt() divides by the variance as it is; c_cjt() leaves out NaN and
            // tiny per-stratum variances (:64-65).
            if (U.ncols == 2 || (!isnan(var) && var > bnl::kSqrtEps)) dfacc += var;
          }
        }
        __syncwarp();
        continue;
      }
      if (EXT && code == BNL_MI_SH) {
        // shrunk cell probabilities lambda * target + (1 - lambda) n / N and their
        // margins (src/tests/discrete/shrinkage.c:45-66, 112-150).
        if (sh_n > 1.0) {
          const double lt = sh_lambda * sh_target, om = (1.0 - sh_lambda) / sh_n;
          const double pk = (double)sl * lt + om * (double)nk;
          const bool cond = U.ncols > 2;
          if (!(cond && pk == 0.0)) {
            for (int c = lane; c < sl; c += 32) {
              const double pc = lt + om * (double)S[c];
              const double pi = (double)lly * lt + om * (double)marg[c / lly];
              const double pj = (double)llx * lt + om * (double)marg[llx + c % lly];
              if (cond) { if (pc > 0.0) acc.add(pc * log(pc * pk / (pi * pj))); }
              else if (pc != 0.0) acc.add(pc * log(pc / (pi * pj)));
            }
          }
        }
        __syncwarp();
        continue;
      }
      if (nk > 0) {
        const double dnk = (double)nk;
        if (code == BNL_MI || code == BNL_MI_ADF || code == BNL_MI_COEF) {
          for (int c = lane; c < sl; c += 32) {
            const int n = S[c];
            if (n) {
              const double ni = (double)marg[c / lly], nj = (double)marg[llx + c % lly];
              acc.add((double)n * log((double)n * dnk / (ni * nj)));
            }
          }
        } else {
          for (int c = lane; c < sl; c += 32) {
            const double e = (double)marg[c / lly] * (double)marg[llx + c % lly] / dnk;
            if (e != 0.0) {
              const double d = (double)S[c] - e;
              acc.add(d * d / e);
            }
          }
        }
      }
      __syncwarp();
    } else {
      long long nk = 0;
      for (int c = lane; c < sl; c += 32) nk += S[c];
      nk = warp_sum_ll(nk);
      if (lane == 0) nobs += nk;
      if (is_bd) {
        if (q == 0 || (code == BNL_BDS && nk == 0)) continue;
        for (int c = lane; c < sl; c += 32)
          acc.add(bnl::lgamma_pos((double)S[c] + alpha) - lg_alpha);
        if (lane == 0) acc.add(lg_aq - bnl::lgamma_pos((double)nk + aq));
      } else if (nk > 0) {
        if (code == BNL_QNML && U.ncols > 1) {
          for (int c = lane; c < sl; c += 32) { const int n = S[c]; if (n) acc.add((double)n * log((double)n)); }
          if (lane == 0) acc.add(-(double)nk * log((double)nk));
        } else {
          for (int c = lane; c < sl; c += 32) {
            const int n = S[c];
            if (n) acc.add((double)n * log((double)n / (double)nk));
          }
          if (lane == 0 && (code == BNL_FNML || code == BNL_QNML))
            acc.add(-nml_regret(regret, (double)nk, (double)llx));
        }
      }
    }
  }

  warp_merge_kahan(acc);
  dfacc = warp_sum_d(dfacc);
  nobs = warp_sum_ll(nobs);
  bnl::KahanSum tot;
  double df = 0.0;
  long long n = 0;
  if (WPU) {
    if (lane != 0) return;
    // what thread 0 does below when the other three warps held nothing
    tot.merge(acc.s, acc.c);
    for (int w = 1; w < BNL_EPI_WARPS; w++) tot.merge(0.0, 0.0);
    df = 0.0 + dfacc;
    n = nobs;
  } else {
    if (lane == 0) { s_s[warp] = acc.s; s_c[warp] = acc.c; s_df[warp] = dfacc; s_n[warp] = nobs; }
    __syncthreads();
    if (threadIdx.x != 0) return;
    for (int w = 0; w < BNL_EPI_WARPS; w++) { tot.merge(s_s[w], s_c[w]); df += s_df[w]; n += s_n[w]; }
  }
  double *o = out + (size_t)unit * 4;
  if (kind == 0) {
    const bool adf = code == BNL_MI_ADF || code == BNL_X2_ADF;
    const double chan = df;
    if (!adf) df = (double)(llx - 1) * (double)(lly - 1) * (double)llz;
    double stat = 0.0;
    if (n > 0 && !(adf && (double)n < 5.0 * (double)llx * (double)lly * (double)llz)) {
      if (code == BNL_MI || code == BNL_MI_ADF) stat = tot.value() / (double)n * (2.0 * (double)n);
      else if (code == BNL_MI_COEF) stat = tot.value() / (double)n;   // mi_kernel() / nobs
      else stat = tot.value();
    }
    if (EXT && code == BNL_JT) {
      // standardised statistic, two-sided normal p-value (patterns/utest.c:50-55).
      const double tv = chan;          // the variance travelled in the df channel
      if (U.ncols == 2) stat = n <= 1 ? 0.0 : (tv < bnl::kSqrtEps ? 0.0 : tot.value() / sqrt(tv));
      else stat = tv < bnl::kSqrtEps ? 0.0 : tot.value() / sqrt(tv);
      o[0] = stat; o[1] = 0.0; o[2] = bnl::pnorm_two_sided(stat); o[3] = (double)n;
      return;
    }
    if (EXT && code == BNL_MI_SH) stat = n <= 1 ? 0.0 : tot.value() * (2.0 * (double)n);
    o[0] = stat; o[1] = df; o[2] = bnl::pchisq_upper(stat, df); o[3] = (double)n;
  } else {
    const double nparams = (double)(llx - 1) * (double)llz;
    double s = tot.value();
    if (code == BNL_AIC || code == BNL_BIC) s -= arg * nparams;
    // qNML with parents: - regret(N, llx q) + regret(N, q) over ALL q parent
    // configurations (cqnml, src/scores/normalized.maximum.likelihood.c:113-130).
    if (code == BNL_QNML && U.ncols > 1)
      s += nml_regret(regret, (double)n, (double)llz) - nml_regret(regret, (double)n, (double)llx * (double)llz);
    // node-average likelihood: per locally complete row, penalty per data row.
    if (code == BNL_NAL || code == BNL_PNAL)
      s = n == 0 ? -INFINITY : s / (double)n - (code == BNL_PNAL ? arg : 0.0) / (double)nrows * nparams;
    o[0] = s; o[1] = nparams; o[2] = 0.0; o[3] = (double)n;
  }
}
