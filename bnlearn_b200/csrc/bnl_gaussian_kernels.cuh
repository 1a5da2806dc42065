// bnl_gaussian_kernels.cuh -- Gaussian side: column means, the centred
// cross-product (Gram) matrix on the fp64 tensor cores, and batched small
// solves (partial correlations, OLS residual variances) served from it.
//
//   col_sum_kernel / gram_dmma_kernel / gram_finish_kernel
//        replace c_meanvec + c_covmat recomputed per test from the raw columns
//        (src/core/moments.c:34-45, src/core/covariance.matrix.c:65-97).
//   gauss_ci_kernel    replaces c_fast_cor / c_fast_pcor + cor_*_trans + pt /
//        pnorm / pchisq (src/core/correlation.c:22-92,
//        src/tests/gaussian/gaussian.tests.c:4-22, df.adjust.c:4-28).
//   gauss_score_kernel replaces glik / cglik over c_ols
//        (src/scores/gaussian.loglikelihood.c:8-83, src/math/least.squares.c).
#pragma once
#include "bnl_internal.h"
#include "bnl_special.cuh"

#define GRAM_TILE 128         // C tile edge per CTA
#define GRAM_KB 32            // rows staged per pipeline stage
#define GRAM_LD (GRAM_KB + 4) // smem row stride in doubles: conflict-free 64-bit fragment loads
#define GRAM_STAGES 3
#define GRAM_THREADS 512
#define GRAM_SMEM (GRAM_STAGES * 2 * GRAM_TILE * GRAM_LD * 8 + 64)   // + one mbarrier per stage
#define BNL_GW (BNL_MAXK + 1) // ints per Gaussian unit descriptor: count, then column slots

// ---------------------------------------------------------------------------
// per-column partial sums: grid (nc, splits); partial[split * nc + col].
__global__ void __launch_bounds__(256)
col_sum_kernel(const double *__restrict__ X, size_t stride, long long nobs, int nc,
               double *__restrict__ partial) {
  __shared__ double ss[8], sc[8];
  const int col = blockIdx.x, split = blockIdx.y, nsplit = gridDim.y;
  const long long per = (nobs + nsplit - 1) / nsplit;
  const long long r0 = per * split;
  long long r1 = r0 + per;
  if (r1 > nobs) r1 = nobs;
  const double *x = X + (size_t)col * stride;
  bnl::KahanSum acc;
  for (long long r = r0 + threadIdx.x; r < r1; r += blockDim.x) acc.add(x[r]);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const double os = __shfl_xor_sync(0xffffffffu, acc.s, o);
    const double oc = __shfl_xor_sync(0xffffffffu, acc.c, o);
    acc.merge(os, oc);
  }
  if ((threadIdx.x & 31) == 0) { ss[threadIdx.x >> 5] = acc.s; sc[threadIdx.x >> 5] = acc.c; }
  __syncthreads();
  if (threadIdx.x == 0) {
    bnl::KahanSum t;
    for (int w = 0; w < 8; w++) t.merge(ss[w], sc[w]);
    partial[(size_t)split * nc + col] = t.value();
  }
}

__global__ void col_mean_kernel(const double *__restrict__ partial, int nsplit, int nc,
                                long long nobs, double *__restrict__ mean) {
  const int col = blockIdx.x * blockDim.x + threadIdx.x;
  if (col >= nc) return;
  bnl::KahanSum t;
  for (int s = 0; s < nsplit; s++) t.add(partial[(size_t)s * nc + col]);
  mean[col] = nobs > 0 ? t.value() / (double)nobs : 0.0;
}

// ---------------------------------------------------------------------------
// Gram matrix of the centred columns on the fp64 tensor cores (DMMA,
// mma.sync.m8n8k4.f64; tcgen05 has no fp64 kind). The reference centres too
// (src/core/covariance.matrix.c:87); rows past nobs and columns past p
// contribute nothing.
__device__ __forceinline__ void dmma_m8n8k4(double &c0, double &c1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
               : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}

namespace bnlgram {
__device__ __forceinline__ unsigned su32(const void *p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(unsigned long long *b, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(su32(b)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(unsigned long long *b, unsigned bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(su32(b)), "r"(bytes) : "memory");
}
// one bulk copy (the TMA engine, UBLKCP in the SASS) global -> shared, completion counted in bytes on `b`
__device__ __forceinline__ void bulk_g2s(void *dst, const void *src, unsigned bytes, unsigned long long *b) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(su32(dst)), "l"(src), "r"(bytes), "r"(su32(b)) : "memory");
}
// bounded: a protocol slip must end in a trap, not in a hung device
__device__ __forceinline__ void mbar_wait(unsigned long long *b, unsigned parity) {
  for (int tries = 0; tries < (1 << 26); tries++) {
    unsigned done;
    asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
                 : "=r"(done) : "r"(su32(b)), "r"(parity) : "memory");
    if (done) return;
  }
  __trap();
}
}  // namespace bnlgram

// X - mean for rows [0, nrows_pad), zero past nobs: the operand the Gram kernel reads when
// there is room for it (one pass at the HBM rate; the subtraction then costs the DMMA pipe
// nothing).
__global__ void __launch_bounds__(256)
gram_centre_kernel(const double *__restrict__ X, size_t stride, long long nobs,
                   long long nrows_pad, const double *__restrict__ mean, double *__restrict__ out) {
  const int col = blockIdx.y;
  const double m = mean[col];
  const double2 *src = reinterpret_cast<const double2 *>(X + (size_t)col * stride);
  double2 *dst = reinterpret_cast<double2 *>(out + (size_t)col * (size_t)nrows_pad);
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < nrows_pad / 2;
       i += (long long)gridDim.x * blockDim.x) {
    const long long r = 2 * i;
    double2 v = make_double2(0.0, 0.0);
    if (r + 1 < nobs) { v = src[i]; v.x -= m; v.y -= m; }
    else if (r < nobs) { v.x = X[(size_t)col * stride + r] - m; }
    dst[i] = v;
  }
}

// One CTA = one upper-triangular 128 x 128 tile of C over one slice of the rows of the view
// it is given; 16 warps in a 4 x 4 grid, each a 32 x 32 sub-tile = 4 x 4 DMMA tiles. The rows
// stream through a 3-stage ring of 32-row stages filled by the bulk-copy engine: one
// cp.async.bulk of 256 bytes per column and stage, issued by 256 threads (one column each),
// their bytes counted on the stage's mbarrier. (With cp.async -- 4 LDGSTS of 16 bytes per
// thread and operand -- the copies cost the DMMA pipe 8 %; the bulk engine's cost is within
// 1.5 % of no copies at all, profiles/r2_gram_probe.txt.)
//   PRE = true   X is gram_centre_kernel's output for the slab: centred, zero past nobs, `rows`
//                a multiple of GRAM_KB -- the inner loop is LDS.64 + DMMA only.
//   PRE = false  X is the raw data (no scratch memory to be had): after a stage has landed
//                every thread centres and masks 1/512 of it in shared memory, then the barrier.
// The warp grid is laid out so that the four warps of an all-padding column block (p = 2000:
// columns 2016-2047 of the last tile column) sit on four different schedulers and just skip.
template <bool PRE>
__global__ void __launch_bounds__(GRAM_THREADS, 1)
gram_dmma_kernel(const double *__restrict__ X, size_t stride, long long rows, int p,
                 const double *__restrict__ mean, double *__restrict__ partial, int ntile) {
  extern __shared__ __align__(16) double gsm_tiles[];
  typedef double (*tile_t)[GRAM_TILE][GRAM_LD];
  tile_t As = reinterpret_cast<tile_t>(gsm_tiles);
  tile_t Bs = reinterpret_cast<tile_t>(gsm_tiles + (size_t)GRAM_STAGES * GRAM_TILE * GRAM_LD);
  unsigned long long *full = reinterpret_cast<unsigned long long *>(gsm_tiles + (size_t)2 * GRAM_STAGES * GRAM_TILE * GRAM_LD);

  int t = blockIdx.x, ti = 0;
  while (t >= ntile - ti) { t -= ntile - ti; ti++; }
  const int tj = ti + t;
  const bool diag = ti == tj;
  const int nsplit = gridDim.y, split = blockIdx.y;
  const long long steps = (rows + GRAM_KB - 1) / GRAM_KB;
  const long long per = (steps + nsplit - 1) / nsplit;
  const long long s0 = per * split;
  long long s1 = s0 + per;
  if (s1 > steps) s1 = steps;
  const int nst = (int)(s1 > s0 ? s1 - s0 : 0);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // a diagonal tile needs only the 10 blocks on and above its diagonal: warps 0-9 take them
  // (3, 3, 2, 2 per scheduler instead of 4 each), gram_finish_kernel mirrors the rest
  int bi = warp & 3, bj = warp >> 2;
  if (diag && warp < 10) { bi = (int)((0x3221110000ull >> (4 * warp)) & 15); bj = (int)((0x3323213210ull >> (4 * warp)) & 15); }
  const int wm = bi * 32, wn = bj * 32;
  const int g = lane >> 2, q = lane & 3;
  // nothing this warp could compute is wanted: it only copies and synchronises
  const bool idle = (diag && warp >= 10) || ti * GRAM_TILE + wm >= p || tj * GRAM_TILE + wn >= p;

  if (threadIdx.x == 0) {
    for (int i = 0; i < GRAM_STAGES; i++) bnlgram::mbar_init(&full[i], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();

  // copies: threads 0-127 one column of the A operand each, 128-255 of the B operand
  const int myc = threadIdx.x & 127;
  const bool copier = threadIdx.x < (diag ? 128 : 256);
  int mycol = (threadIdx.x < 128 ? ti : tj) * GRAM_TILE + myc;
  if (mycol >= p) mycol = p - 1;
  const double *mysrc = X + (size_t)mycol * stride;
  const unsigned stage_bytes = (diag ? 1u : 2u) * GRAM_TILE * GRAM_KB * 8;
  auto expect = [&](int sl) { if (threadIdx.x == 0) bnlgram::mbar_expect_tx(&full[sl % GRAM_STAGES], stage_bytes); };
  auto issue = [&](int sl) {
    if (!copier) return;
    const int buf = sl % GRAM_STAGES;
    double *dst = threadIdx.x < 128 ? &As[buf][myc][0] : &Bs[buf][myc][0];
    if (!PRE) asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // this thread's centring stores, then the engine's
    bnlgram::bulk_g2s(dst, mysrc + (s0 + sl) * GRAM_KB, GRAM_KB * 8, &full[buf]);
  };

  // PRE = false: this thread's share of a landed stage, GRAM_CREPS 16-byte chunks per operand
  constexpr int CPC = GRAM_KB / 2, CREPS = GRAM_TILE * CPC / GRAM_THREADS;
  double cma[CREPS], cmb[CREPS];
#pragma unroll
  for (int rep = 0; rep < CREPS; rep++) {
    const int c = (threadIdx.x + rep * GRAM_THREADS) / CPC;
    int col = ti * GRAM_TILE + c; if (col >= p) col = p - 1;
    cma[rep] = PRE ? 0.0 : mean[col];
    col = tj * GRAM_TILE + c; if (col >= p) col = p - 1;
    cmb[rep] = PRE ? 0.0 : mean[col];
  }
  auto centre_own = [&](int sl) {
    const long long r0 = (s0 + sl) * GRAM_KB;
    const int buf = sl % GRAM_STAGES;
#pragma unroll
    for (int rep = 0; rep < CREPS; rep++) {
      const int idx = threadIdx.x + rep * GRAM_THREADS;
      const int c = idx / CPC, r2 = (idx % CPC) * 2;
      const bool l0 = r0 + r2 < rows, l1 = r0 + r2 + 1 < rows;
      double2 v = *reinterpret_cast<double2 *>(&As[buf][c][r2]);
      v.x = l0 ? v.x - cma[rep] : 0.0; v.y = l1 ? v.y - cma[rep] : 0.0;
      *reinterpret_cast<double2 *>(&As[buf][c][r2]) = v;
      if (!diag) {
        double2 w = *reinterpret_cast<double2 *>(&Bs[buf][c][r2]);
        w.x = l0 ? w.x - cmb[rep] : 0.0; w.y = l1 ? w.y - cmb[rep] : 0.0;
        *reinterpret_cast<double2 *>(&Bs[buf][c][r2]) = w;
      }
    }
  };

  double acc[4][4][2];
#pragma unroll
  for (int i = 0; i < 4; i++)
#pragma unroll
    for (int j = 0; j < 4; j++) acc[i][j][0] = acc[i][j][1] = 0.0;

  for (int s = 0; s < GRAM_STAGES - 1; s++) if (s < nst) expect(s);
  __syncthreads();
  for (int s = 0; s < GRAM_STAGES - 1; s++) if (s < nst) issue(s);
  for (int s = 0; s < nst; s++) {
    const int buf = s % GRAM_STAGES;
    bnlgram::mbar_wait(&full[buf], (unsigned)((s / GRAM_STAGES) & 1));
    if (!PRE) centre_own(s);
    if (s + GRAM_STAGES - 1 < nst) expect(s + GRAM_STAGES - 1);
    __syncthreads();                    // stage s centred and visible; everyone is done reading stage s - 1
    if (s + GRAM_STAGES - 1 < nst) issue(s + GRAM_STAGES - 1);
    if (idle) continue;
    const double (*Ap)[GRAM_LD] = As[buf];
    const double (*Bp)[GRAM_LD] = diag ? As[buf] : Bs[buf];
#pragma unroll
    for (int kk = 0; kk < GRAM_KB; kk += 4) {
      double a[4], b[4];
#pragma unroll
      for (int i = 0; i < 4; i++) a[i] = Ap[wm + i * 8 + g][kk + q];
#pragma unroll
      for (int j = 0; j < 4; j++) b[j] = Bp[wn + j * 8 + g][kk + q];
#pragma unroll
      for (int i = 0; i < 4; i++)
#pragma unroll
        for (int j = 0; j < 4; j++) dmma_m8n8k4(acc[i][j][0], acc[i][j][1], a[i], b[j]);
    }
  }

  double *out = partial + (size_t)split * p * p;
#pragma unroll
  for (int i = 0; i < 4; i++)
#pragma unroll
    for (int j = 0; j < 4; j++) {
      const int row = ti * GRAM_TILE + wm + i * 8 + g;
      const int col = tj * GRAM_TILE + wn + j * 8 + q * 2;
      if (row < p && col < p) out[(size_t)row + (size_t)col * p] = acc[i][j][0];
      if (row < p && col + 1 < p) out[(size_t)row + (size_t)(col + 1) * p] = acc[i][j][1];
    }
}

// sum the row-slice partials, divide by nobs - 1, mirror the upper triangle.
__global__ void gram_finish_kernel(const double *__restrict__ partial, int nsplit, int p,
                                   long long nobs, double *__restrict__ cov) {
  const long long e = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (e >= (long long)p * p) return;
  const int i = (int)(e % p), j = (int)(e / p);
  if (i > j) return;                             // only blocks on or above the diagonal are written
  bnl::KahanSum t;
  for (int s = 0; s < nsplit; s++) t.add(partial[(size_t)s * p * p + e]);
  const double v = nobs > 1 ? t.value() / (double)(nobs - 1) : 0.0;
  cov[e] = v;
  if (i < j) cov[(size_t)j + (size_t)i * p] = v;
}

// ---------------------------------------------------------------------------
// Batched small solves. One thread per unit; its k x k work matrices live in
// shared memory, interleaved across the CTA's threads (element e of thread t at
// [e * blockDim.x + t]) so that every access is bank-conflict free.

struct SmemMat {
  double *base; int ld;   // ld = blockDim.x
  __device__ __forceinline__ double &operator()(int e) const { return base[(size_t)e * ld]; }
};

// cyclic Jacobi eigen-decomposition of the symmetric k x k matrix A (destroyed:
// its diagonal ends up holding the eigenvalues), eigenvectors in the columns of V.
__device__ void jacobi_eigen(SmemMat A, SmemMat V, int k) {
  for (int i = 0; i < k; i++)
    for (int j = 0; j < k; j++) V(i + j * k) = (i == j) ? 1.0 : 0.0;
  for (int sweep = 0; sweep < 60; sweep++) {
    double off = 0.0, diag = 0.0;
    for (int i = 0; i < k; i++) {
      const double d = A(i + i * k);
      diag += d * d;
      for (int j = i + 1; j < k; j++) { const double o = A(i + j * k); off += o * o; }
    }
    if (off == 0.0 || off <= 1e-40 * diag) break;
    for (int r = 0; r < k - 1; r++)
      for (int c = r + 1; c < k; c++) {
        const double apq = A(r + c * k);
        if (apq == 0.0) continue;
        const double theta = (A(c + c * k) - A(r + r * k)) / (2.0 * apq);
        const double t = (theta >= 0.0 ? 1.0 : -1.0) / (fabs(theta) + sqrt(theta * theta + 1.0));
        const double cs = 1.0 / sqrt(t * t + 1.0), sn = t * cs;
        for (int m = 0; m < k; m++) {
          const double amr = A(m + r * k), amc = A(m + c * k);
          A(m + r * k) = cs * amr - sn * amc;
          A(m + c * k) = sn * amr + cs * amc;
        }
        for (int m = 0; m < k; m++) {
          const double arm = A(r + m * k), acm = A(c + m * k);
          A(r + m * k) = cs * arm - sn * acm;
          A(c + m * k) = sn * arm + cs * acm;
        }
        for (int m = 0; m < k; m++) {
          const double vmr = V(m + r * k), vmc = V(m + c * k);
          V(m + r * k) = cs * vmr - sn * vmc;
          V(m + c * k) = sn * vmr + cs * vmc;
        }
      }
  }
}

// Units arrive exactly as the caller passed them (x[], y[], CSR z): `slot` maps a
// data-set column to its numeric slot (-1 for factors). A bad column raises *err
// (1 + unit index) and the unit is skipped. out: 4 doubles per unit.
__global__ void gauss_ci_kernel(const int *__restrict__ ux, const int *__restrict__ uy,
                                const int *__restrict__ zoff, const int *__restrict__ zidx,
                                const int *__restrict__ slot, int ncols, int n,
                                const double *__restrict__ cov, int p, long long nobs, int test,
                                int kmax, double *__restrict__ out, int *__restrict__ err,
                                const double *__restrict__ ucov, const long long *__restrict__ unobs,
                                const double *__restrict__ ushr, int shr_chunks) {
  extern __shared__ double gsm[];
  const int u = blockIdx.x * blockDim.x + threadIdx.x;
  if (u >= n) return;
  const int z0 = zoff[u], nz = zoff[u + 1] - z0, k = nz + 2;
  double *o = out + (size_t)u * 4;
  const double tol = bnl::kSqrtEps;
  int d[BNL_MAXK + 1];
  bool bad = k > BNL_MAXK || nz < 0;
  if (!bad) {
    for (int i = 0; i < k; i++) {
      const int c = i == 0 ? ux[u] : (i == 1 ? uy[u] : zidx[z0 + i - 2]);
      const int sl = (c >= 0 && c < ncols) ? slot[c] : -1;
      bad |= sl < 0;
      d[1 + i] = sl;
    }
  }
  if (bad) { atomicCAS(err, 0, u + 1); o[0] = o[1] = o[2] = o[3] = 0.0; return; }
  // incomplete data: this unit's own covariance over its complete rows
  // (ct_gaustests_with_missing, src/tests/patterns/ctest.c:144-229).
  if (ucov) {
    cov = ucov + (size_t)u * kmax * kmax; p = kmax; nobs = unobs[u];
    for (int i = 0; i < k; i++) d[1 + i] = i;
  }

  double df = (test == BNL_COR) ? (double)(nobs - nz - 2)
            : (test == BNL_ZF) ? (double)(nobs - nz - 3) : 1.0;
  if (df < 1.0 || (ucov && nobs == 0)) { o[0] = 0.0; o[1] = 0.0; o[2] = 1.0; o[3] = 1.0; return; }   // o[3]: R warning

  // mi-g-sh: shrinkage intensity from the third pass (chunk order), on the log
  // scale as the reference does (src/tests/gaussian/shrinkage.c:40-50, 88-99).
  double lambda = 0.0;
  if (test == BNL_MI_G_SH) {
    double num = 0.0;
    for (int c = 0; c < shr_chunks; c++) num += ushr[(size_t)u * shr_chunks + c];
    double den = 0.0;
    for (int i = 0; i < k; i++)
      for (int j = i + 1; j < k; j++) { const double c = cov[(size_t)d[1 + i] + (size_t)d[1 + j] * p]; den += c * c; }
    if (num > tol) {
      const double nc = (double)nobs;
      lambda = exp(log(num) - log(den) + log(nc) - 3.0 * log(nc - 1.0));
      lambda = fmin(fmax(lambda, 0.0), 1.0);
      if (isnan(lambda)) lambda = 0.0;
    }
  }

  double r;
  if (nz == 0) {
    // c_fast_cor(): cov / sqrt(var_x var_y) with the guard on the SUMS of squares.
    const double sxx = cov[(size_t)d[1] + (size_t)d[1] * p] * (double)(nobs - 1);
    const double syy = cov[(size_t)d[2] + (size_t)d[2] * p] * (double)(nobs - 1);
    const double sxy = cov[(size_t)d[1] + (size_t)d[2] * p] * (double)(nobs - 1);
    r = (sxx < tol || syy < tol) ? 0.0 : sxy / sqrt(sxx * syy);
    if (test == BNL_MI_G_SH) {          // ut_gaustests: the coefficient itself is shrunk (patterns/utest.c:110-116)
      if (r > 1.0) r = 1.0; else if (r < -1.0) r = -1.0;
      r *= 1.0 - lambda;
    }
  } else {
    SmemMat A{gsm + threadIdx.x, (int)blockDim.x};
    SmemMat V{gsm + (size_t)kmax * kmax * blockDim.x + threadIdx.x, (int)blockDim.x};
    for (int i = 0; i < k; i++)
      for (int j = 0; j < k; j++)       // covmat_shrink(): off-diagonal entries scaled by 1 - lambda
        A(i + j * k) = cov[(size_t)d[1 + i] + (size_t)d[1 + j] * p] * (i == j ? 1.0 : 1.0 - lambda);
    jacobi_eigen(A, V, k);
    // pseudo-inverse entries with singular values <= dim * d_max * eps dropped
    // (src/core/correlation.c:69-83).
    double dmax = 0.0;
    for (int i = 0; i < k; i++) dmax = fmax(dmax, fabs(A(i + i * k)));
    const double sv_tol = (double)k * dmax * tol * tol;
    double k11 = 0.0, k12 = 0.0, k22 = 0.0;
    for (int i = 0; i < k; i++) {
      const double lam = A(i + i * k);
      if (fabs(lam) > sv_tol) {
        const double v0 = V(0 + i * k), v1 = V(1 + i * k);
        k11 += v0 * v0 / lam; k12 += v0 * v1 / lam; k22 += v1 * v1 / lam;
      }
    }
    r = (k11 < tol || k22 < tol) ? 0.0 : -k12 / sqrt(k11 * k22);
  }
  double clamped = 0.0;
  if (r > 1.0) { r = 1.0; clamped = 2.0; } else if (r < -1.0) { r = -1.0; clamped = 2.0; }

  double stat, pv;
  if (test == BNL_COR) {
    const double t = r * sqrt(df) / sqrt(1.0 - r * r);
    stat = r;
    pv = bnl::pt_two_sided(t, df);
  } else if (test == BNL_ZF) {
    stat = log((1.0 + r) / (1.0 - r)) / 2.0 * sqrt(df);
    pv = bnl::pnorm_two_sided(stat);
  } else if (test == BNL_MI_G_COEF) {     // cor_mi_trans() alone, for the coefficient matrix
    stat = -0.5 * log(1.0 - r * r);
    pv = 0.0;
  } else {
    stat = 2.0 * (double)nobs * (-0.5 * log(1.0 - r * r));
    pv = bnl::pchisq_upper(stat, df);
  }
  o[0] = stat; o[1] = df; o[2] = pv; o[3] = clamped;
}

// ---------------------------------------------------------------------------
// Every pair of a column list tested marginally: level 0 of pc.stable on numeric
// data (R/pcalgo.R:8-28 -> ut_gaustests_complete, src/tests/patterns/utest.c:65-129).
// Nothing per pair crosses PCIe in either direction: the pairs are enumerated here,
// the arithmetic is gauss_ci_kernel's marginal branch to the letter (same p-values),
// and what goes back is the p-value matrix (optional) and the pairs that stay
// adjacent (p <= alpha), compacted in row-major order of the upper triangle.
// One CTA per first column i; keep[i * n + j] flags the surviving j > i.
__global__ void __launch_bounds__(256)
gauss_allpairs_kernel(const double *__restrict__ cov, int p, const int *__restrict__ slots, int n,
                      long long nobs, int test, double alpha, double *__restrict__ pmat,
                      unsigned char *__restrict__ keep, int *__restrict__ rowcount) {
  __shared__ int s_cnt;
  const int i = blockIdx.x;
  if (threadIdx.x == 0) s_cnt = 0;
  __syncthreads();
  const double tol = bnl::kSqrtEps;
  const double df = (test == BNL_COR) ? (double)(nobs - 2) : (test == BNL_ZF) ? (double)(nobs - 3) : 1.0;
  const int si = slots[i];
  const double sxx = cov[(size_t)si + (size_t)si * p] * (double)(nobs - 1);
  int mine = 0;
  for (int j = i + 1 + threadIdx.x; j < n; j += blockDim.x) {
    const int sj = slots[j];
    double pv = 1.0;
    if (df >= 1.0) {
      const double syy = cov[(size_t)sj + (size_t)sj * p] * (double)(nobs - 1);
      const double sxy = cov[(size_t)si + (size_t)sj * p] * (double)(nobs - 1);
      double r = (sxx < tol || syy < tol) ? 0.0 : sxy / sqrt(sxx * syy);
      if (r > 1.0) r = 1.0; else if (r < -1.0) r = -1.0;
      if (test == BNL_COR) pv = bnl::pt_two_sided(r * sqrt(df) / sqrt(1.0 - r * r), df);
      else if (test == BNL_ZF) pv = bnl::pnorm_two_sided(log((1.0 + r) / (1.0 - r)) / 2.0 * sqrt(df));
      else pv = bnl::pchisq_upper(2.0 * (double)nobs * (-0.5 * log(1.0 - r * r)), df);
    }
    if (pmat) { pmat[(size_t)i * n + j] = pv; pmat[(size_t)j * n + i] = pv; }
    const bool k = pv <= alpha;
    keep[(size_t)i * n + j] = k;
    mine += k;
  }
  if (pmat && threadIdx.x == 0) pmat[(size_t)i * n + i] = 0.0;
  if (mine) atomicAdd(&s_cnt, mine);
  __syncthreads();
  if (threadIdx.x == 0) rowcount[i] = s_cnt;
}

// one warp per row: position of its first survivor = sum of the earlier rows' counts
__global__ void __launch_bounds__(128)
gauss_allpairs_compact_kernel(const unsigned char *__restrict__ keep, const int *__restrict__ rowcount, int n,
                              long long cap, int *__restrict__ out_i, int *__restrict__ out_j,
                              long long *__restrict__ total) {
  const int i = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (i >= n) return;
  long long off = 0;
  for (int r = lane; r < i; r += 32) off += rowcount[r];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) off += __shfl_xor_sync(0xffffffffu, off, o);
  if (i == n - 1 && lane == 0) *total = off + rowcount[i];
  for (int j0 = i + 1; j0 < n; j0 += 32) {
    const int j = j0 + lane;
    const bool k = j < n && keep[(size_t)i * n + j];
    const unsigned m = __ballot_sync(0xffffffffu, k);
    const long long pos = off + __popc(m & ((1u << lane) - 1u));
    if (k && pos < cap) { out_i[pos] = i; out_j[pos] = j; }
    off += __popc(m);
  }
}

// Gaussian node score from the moments. Unit u: target[u] and the CSR parent list.
// S(i, j) below is the centred cross-product (n - 1) * cov of the unit's columns.
__global__ void gauss_score_kernel(const int *__restrict__ target, const int *__restrict__ poff,
                                   const int *__restrict__ pidx, const int *__restrict__ slot,
                                   int ncols, int n, const double *__restrict__ cov,
                                   const double *__restrict__ mean, int p, long long nobs,
                                   int score, double kpen, int kmax, double *__restrict__ out,
                                   int *__restrict__ err, const double *__restrict__ ucov,
                                   const double *__restrict__ umean,
                                   const long long *__restrict__ unobs, double bge_w,
                                   const double *__restrict__ bge_nu) {
  extern __shared__ double gsm[];
  const int u = blockIdx.x * blockDim.x + threadIdx.x;
  if (u >= n) return;
  const long long nrows = nobs;        // the penalty of nal-g / pnal-g uses the full row count
  const int p0 = poff[u], np = poff[u + 1] - p0, k = np + 1;
  double *o = out + (size_t)u * 4;
  const double tol = bnl::kSqrtEps;
  int d[BNL_MAXK + 1];
  bool bad = k > BNL_MAXK || np < 0;
  if (!bad) {
    for (int i = 0; i < k; i++) {
      const int c = i == 0 ? target[u] : pidx[p0 + i - 1];
      const int sl = (c >= 0 && c < ncols) ? slot[c] : -1;
      bad |= sl < 0;
      d[1 + i] = sl;
    }
  }
  if (bad) { atomicCAS(err, 0, u + 1); o[0] = o[1] = o[2] = o[3] = 0.0; return; }
  if (ucov) {                          // moments over this unit's complete rows (c_ols(..., missing = TRUE))
    cov = ucov + (size_t)u * kmax * kmax; mean = umean + (size_t)u * kmax; p = kmax; nobs = unobs[u];
    for (int i = 0; i < k; i++) d[1 + i] = i;
  }
  const double nm1 = (double)(nobs - 1);
  auto C = [&](int i, int j) { return cov[(size_t)d[1 + i] + (size_t)d[1 + j] * p]; };

  if (score == BNL_BGE) {
    // Bayesian Gaussian equivalent score (wpost / cwpost, src/scores/
    // wishart.posterior.c:10-106): kpen = iss.mu, bge_w = iss.w, bge_nu = prior
    // means by numeric slot; ncols columns in the data frame. The reference takes
    // log|R| and log|R~| (R without the node's row and column) from two LU
    // factorisations; both are SPD here, so one Cholesky with the node LAST gives
    // log|R~| from the leading block and log|R| = log|R~| + log(Schur complement).
    const double N = (double)nobs, amu = kpen, aw = bge_w, nc = (double)ncols, pp = (double)np;
    const double t = amu * (aw - nc - 1.0) / (amu + 1.0);
    const double shrink = N * amu / (N + amu);
    double lp = 0.5 * (log(amu) - log(N + amu));
    lp += bnl::lgamma_pos(0.5 * (N + aw - nc + pp + 1.0)) - bnl::lgamma_pos(0.5 * (aw - nc + pp + 1.0));
    lp -= 0.5 * N * 1.1447298858494001741434273513531;       // log(pi)
    if (np == 0) {
      const double dm = mean[d[1]] - bge_nu[d[1]];
      const double r = t + C(0, 0) * nm1 + shrink * dm * dm;
      lp += 0.5 * (aw - nc + 1.0) * log(t) - 0.5 * (N + aw - nc + 1.0) * log(r);
    } else {
      lp += 0.5 * (aw - nc + pp + 1.0) * (pp + 1.0) * log(t) - 0.5 * (aw - nc + pp) * pp * log(t);
      SmemMat R{gsm + threadIdx.x, (int)blockDim.x};
      auto at = [&](int i) { return i == k - 1 ? 0 : i + 1; };     // position i of R -> unit column
      for (int i = 0; i < k; i++)
        for (int j = 0; j < k; j++) {
          const int a = at(i), b = at(j);
          R(i + j * k) = C(a, b) * nm1 + (i == j ? t : 0.0) +
                         shrink * (mean[d[1 + a]] - bge_nu[d[1 + a]]) * (mean[d[1 + b]] - bge_nu[d[1 + b]]);
        }
      double ld_tilde = 0.0, ld_last = 0.0;
      bool singular = false;
      for (int j = 0; j < k; j++) {                                 // right-looking Cholesky, lower part
        double piv = R(j + j * k);
        if (!(piv > 0.0)) { singular = true; break; }
        if (j < k - 1) ld_tilde += log(piv); else ld_last = log(piv);
        const double inv = 1.0 / piv;
        for (int i = j + 1; i < k; i++) {
          const double f = R(i + j * k) * inv;
          for (int l = j + 1; l <= i; l++) R(i + l * k) -= f * R(l + j * k);
        }
      }
      if (singular) lp = -INFINITY;
      else lp += 0.5 * (N + aw - nc + pp) * ld_tilde - 0.5 * (N + aw - nc + pp + 1.0) * (ld_tilde + ld_last);
    }
    o[0] = lp; o[1] = (double)(np + 2); o[2] = 0.0; o[3] = 0.0;
    return;
  }

  double sse = 0.0;      // residual sum of squares of the fit
  double sd;
  if (nobs == 0) sd = NAN;
  else if (nobs <= np + 1) sd = 0.0;
  else {
    if (np == 0) {
      sse = C(0, 0) * nm1;
    } else if (np == 1) {
      // c_ols1(): singular iff the regressor's variance is below sqrt(eps).
      const bool singular = fabs(C(1, 1)) < tol;
      sse = singular ? C(0, 0) * nm1 : (C(0, 0) - C(0, 1) / C(1, 1) * C(0, 1)) * nm1;
    } else if (np == 2) {
      // c_ols2(): the three collinearity cases of src/math/least.squares.c:171-173.
      const double c11 = C(1, 1), c22 = C(2, 2), c12 = C(1, 2), c01 = C(0, 1), c02 = C(0, 2);
      const bool s1 = fabs(c11) < tol;
      const bool s2 = (fabs(c22) < tol) || (fabs(c12) / sqrt(c11 * c22) > 1.0 - tol);
      double fit;
      if (s1 && !s2) fit = c02 / c22 * c02;
      else if (!s1 && s2) fit = c01 / c11 * c01;
      else if (s1 && s2) fit = 0.0;
      else {
        const double den = c11 * c22 - c12 * c12;
        const double b1 = (c22 * c01 - c12 * c02) / den, b2 = (c11 * c02 - c12 * c01) / den;
        fit = b1 * c01 + b2 * c02;
      }
      sse = (C(0, 0) - fit) * nm1;
    } else {
      // general case: the reference runs R's dqrdc2 on [1, parents] with
      // tolerance sqrt(eps) (src/math/linear.algebra.c:140, 197): a column whose
      // norm after projecting out the accepted ones falls below tol * its
      // original (uncentred) norm is set aside. Same rule here, as an outer-
      // product Cholesky on the centred cross-products (the intercept is the
      // centring), parents in order, target last.
      SmemMat S{gsm + threadIdx.x, (int)blockDim.x};
      for (int i = 0; i < k; i++)
        for (int j = 0; j < k; j++) S(i + j * k) = C(i, j) * nm1;
      for (int l = 1; l < k; l++) {
        const double m = mean[d[1 + l]];
        const double orig = sqrt(C(l, l) * nm1 + (double)nobs * m * m);
        const double piv = S(l + l * k);
        if (!(piv > 0.0) || sqrt(piv) < orig * tol) continue;     // dropped column
        for (int i = 0; i < k; i++) {
          if (i == l) continue;
          const double f = S(i + l * k) / piv;
          if (f == 0.0) continue;
          for (int j = 0; j < k; j++)
            if (j != l) S(i + j * k) -= f * S(l + j * k);
        }
        for (int i = 0; i < k; i++)
          if (i != l) { S(i + l * k) = 0.0; S(l + i * k) = 0.0; }
      }
      sse = S(0);
    }
    if (sse < 0.0) sse = 0.0;
    sd = sqrt(sse / (double)(nobs - np - 1));
  }

  double res;
  if (!(sd >= tol)) res = (sd < tol) ? -INFINITY : NAN;
  else res = -(double)nobs * (bnl::kLnSqrt2Pi + log(sd)) - 0.5 * sse / (sd * sd);
  const double nparams = (double)(np + 2);
  if (score == BNL_AIC_G || score == BNL_BIC_G) res -= kpen * nparams;
  // node-average likelihood (src/scores/gaussian.nal.c:8-78): singular or empty -> -Inf.
  if (score == BNL_NAL_G || score == BNL_PNAL_G)
    res = (nobs == 0 || !(sd >= tol)) ? -INFINITY
          : res / (double)nobs - (score == BNL_PNAL_G ? kpen : 0.0) / (double)nrows * nparams;
  o[0] = res; o[1] = nparams; o[2] = 0.0; o[3] = 0.0;
}

// ---------------------------------------------------------------------------
// Incomplete numeric data: moments over the rows that are complete for the
// columns of ONE unit (c_covmat_with_missing, src/core/covariance.matrix.c:
// 100-184; c_cor_with_missing, src/core/correlation.c:96-155). Two passes per
// unit, as in the reference: count + sums -> means, then centred cross-products.
// A CTA handles (row chunk, unit); partial results are reduced in chunk order by
// a second kernel, so the result does not depend on scheduling.

#define BNL_MASK_THREADS 256

// columns of unit u in the order the consumers use: x, y, z... (tests) or
// target, parents... (scores). Returns k, or -1 for a bad unit.
__device__ __forceinline__ int masked_unit_cols(int u, const int *ua, const int *ub, const int *off,
                                                const int *idx, const int *slot, int ncols,
                                                int *cols) {
  const int lead = ub ? 2 : 1;
  const int o0 = off[u], nz = off[u + 1] - o0, k = nz + lead;
  if (nz < 0 || k > BNL_MAXK) return -1;
  for (int i = 0; i < k; i++) {
    const int c = i == 0 ? ua[u] : ((ub && i == 1) ? ub[u] : idx[o0 + i - lead]);
    const int sl = (c >= 0 && c < ncols) ? slot[c] : -1;
    if (sl < 0) return -1;
    cols[i] = sl;
  }
  return k;
}

__device__ __forceinline__ double block_sum_256(double v, double *sm) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  __syncthreads();
  if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = v;
  __syncthreads();
  double t = 0.0;
  if (threadIdx.x < 32) {
    t = threadIdx.x < BNL_MASK_THREADS / 32 ? sm[threadIdx.x] : 0.0;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
  }
  return t;      // valid in thread 0
}

// pass 1: partial[(u * nchunks + chunk) * (BNL_MAXK + 1) + {0: count, 1 + i: sum of column i}]
__global__ void __launch_bounds__(BNL_MASK_THREADS)
masked_sums_kernel(const double *__restrict__ data, size_t stride, long long nobs, long long chunk_rows,
                   const int *__restrict__ ua, const int *__restrict__ ub, const int *__restrict__ off,
                   const int *__restrict__ idx, const int *__restrict__ slot, int ncols,
                   double *__restrict__ partial) {
  __shared__ double sm[BNL_MASK_THREADS / 32];
  const int u = blockIdx.y, chunk = blockIdx.x, nchunks = gridDim.x;
  int cols[BNL_MAXK];
  const int k = masked_unit_cols(u, ua, ub, off, idx, slot, ncols, cols);
  double *out = partial + ((size_t)u * nchunks + chunk) * (BNL_MAXK + 1);
  if (k < 0) { if (threadIdx.x <= BNL_MAXK) out[threadIdx.x] = 0.0; return; }
  const long long r0 = chunk * chunk_rows, r1 = min(nobs, r0 + chunk_rows);
  double cnt = 0.0, sum[BNL_MAXK];
#pragma unroll
  for (int i = 0; i < BNL_MAXK; i++) sum[i] = 0.0;
  for (long long r = r0 + threadIdx.x; r < r1; r += BNL_MASK_THREADS) {
    double v[BNL_MAXK];
    bool ok = true;
#pragma unroll
    for (int i = 0; i < BNL_MAXK; i++)
      if (i < k) { v[i] = data[stride * cols[i] + r]; ok &= !isnan(v[i]); }
    if (!ok) continue;
    cnt += 1.0;
#pragma unroll
    for (int i = 0; i < BNL_MAXK; i++) if (i < k) sum[i] += v[i];
  }
  double t = block_sum_256(cnt, sm);
  if (threadIdx.x == 0) out[0] = t;
#pragma unroll
  for (int i = 0; i < BNL_MAXK; i++) {
    if (i < k) {                       // k is uniform over the CTA
      t = block_sum_256(sum[i], sm);
      if (threadIdx.x == 0) out[1 + i] = t;
    }
  }
}

// means and complete-row counts from the pass-1 partials (chunk order).
__global__ void masked_means_kernel(const double *__restrict__ partial, int nchunks, int n, int kmax,
                                    double *__restrict__ umean, long long *__restrict__ unobs) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  const int u = t / (kmax + 1), i = t % (kmax + 1);
  if (u >= n) return;
  double s = 0.0;
  for (int c = 0; c < nchunks; c++) s += partial[((size_t)u * nchunks + c) * (BNL_MAXK + 1) + i];
  if (i == 0) { unobs[u] = (long long)s; return; }
  double nc = 0.0;
  for (int c = 0; c < nchunks; c++) nc += partial[((size_t)u * nchunks + c) * (BNL_MAXK + 1)];
  umean[(size_t)u * kmax + i - 1] = nc > 0.0 ? s / nc : 0.0;
}

// pass 2: centred cross-products of the complete rows; KB bounds the unit width
// so that the KB (KB + 1) / 2 accumulators stay in registers for the common
// narrow units. partial2[(u * nchunks + chunk) * npairs + pair(i <= j)].
template <int KB>
__global__ void __launch_bounds__(BNL_MASK_THREADS)
masked_cross_kernel(const double *__restrict__ data, size_t stride, long long nobs, long long chunk_rows,
                    const int *__restrict__ ua, const int *__restrict__ ub, const int *__restrict__ off,
                    const int *__restrict__ idx, const int *__restrict__ slot, int ncols, int kmax,
                    const double *__restrict__ umean, double *__restrict__ partial2) {
  __shared__ double sm[BNL_MASK_THREADS / 32];
  constexpr int NP = KB * (KB + 1) / 2;
  const int u = blockIdx.y, chunk = blockIdx.x, nchunks = gridDim.x;
  int cols[BNL_MAXK];
  const int k = masked_unit_cols(u, ua, ub, off, idx, slot, ncols, cols);
  double *out = partial2 + ((size_t)u * nchunks + chunk) * NP;
  if (k < 0 || k > KB) { for (int p = threadIdx.x; p < NP; p += BNL_MASK_THREADS) out[p] = 0.0; return; }
  double m[KB], acc[NP];
#pragma unroll
  for (int i = 0; i < KB; i++) m[i] = i < k ? umean[(size_t)u * kmax + i] : 0.0;
#pragma unroll
  for (int p = 0; p < NP; p++) acc[p] = 0.0;
  const long long r0 = chunk * chunk_rows, r1 = min(nobs, r0 + chunk_rows);
  for (long long r = r0 + threadIdx.x; r < r1; r += BNL_MASK_THREADS) {
    double d[KB];
    bool ok = true;
#pragma unroll
    for (int i = 0; i < KB; i++) {
      d[i] = 0.0;
      if (i < k) { const double v = data[stride * cols[i] + r]; ok &= !isnan(v); d[i] = v - m[i]; }
    }
    if (!ok) continue;
    int p = 0;
#pragma unroll
    for (int i = 0; i < KB; i++)
#pragma unroll
      for (int j = i; j < KB; j++, p++) acc[p] += d[i] * d[j];
  }
#pragma unroll
  for (int p = 0; p < NP; p++) {
    const double t = block_sum_256(acc[p], sm);
    if (threadIdx.x == 0) out[p] = t;
  }
}

// covariance (divisor nc - 1, as the reference: 0 / 0 when a single row is
// complete) of every unit, column-major kmax x kmax.
__global__ void masked_cov_kernel(const double *__restrict__ partial2, int nchunks, int n, int kmax, int kb,
                                  const long long *__restrict__ unobs, double *__restrict__ ucov) {
  const int np = kb * (kb + 1) / 2;
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  const int u = t / np, p = t % np;
  if (u >= n) return;
  int i = 0, rem = p;
  while (rem >= kb - i) { rem -= kb - i; i++; }
  const int j = i + rem;
  if (j >= kmax) return;
  double s = 0.0;
  for (int c = 0; c < nchunks; c++) s += partial2[((size_t)u * nchunks + c) * np + p];
  const long long nc = unobs[u];
  const double v = nc == 0 ? 0.0 : s / (double)(nc - 1);
  ucov[(size_t)u * kmax * kmax + i + (size_t)j * kmax] = v;
  ucov[(size_t)u * kmax * kmax + j + (size_t)i * kmax] = v;
}

// Third pass for the shrinkage test mi-g-sh: the sum over column pairs i < j and
// complete rows of ((x_i - m_i)(x_j - m_j) - s_ij)^2, the numerator of the
// James-Stein intensity. s_ij is the covariance with divisor nc for conditional
// tests (covmat_lambda, src/tests/gaussian/shrinkage.c:60-106) and with divisor
// nc - 1 for marginal ones (cor_lambda, :15-57). partial3[u * nchunks + chunk].
template <int KB>
__global__ void __launch_bounds__(BNL_MASK_THREADS)
masked_shrink_kernel(const double *__restrict__ data, size_t stride, long long nobs, long long chunk_rows,
                     const int *__restrict__ ua, const int *__restrict__ ub, const int *__restrict__ off,
                     const int *__restrict__ idx, const int *__restrict__ slot, int ncols, int kmax,
                     const double *__restrict__ umean, const double *__restrict__ ucov,
                     const long long *__restrict__ unobs, double *__restrict__ partial3) {
  __shared__ double sm[BNL_MASK_THREADS / 32];
  const int u = blockIdx.y, chunk = blockIdx.x, nchunks = gridDim.x;
  int cols[BNL_MAXK];
  const int k = masked_unit_cols(u, ua, ub, off, idx, slot, ncols, cols);
  double *out = partial3 + (size_t)u * nchunks + chunk;
  if (k < 0 || k > KB) { if (threadIdx.x == 0) *out = 0.0; return; }
  const double nc = (double)unobs[u];
  const double scale = k == 2 ? 1.0 : (nc - 1.0) / nc;
  double m[KB];
#pragma unroll
  for (int i = 0; i < KB; i++) m[i] = i < k ? umean[(size_t)u * kmax + i] : 0.0;
  const double *cv = ucov + (size_t)u * kmax * kmax;
  double acc = 0.0;
  const long long r0 = chunk * chunk_rows, r1 = min(nobs, r0 + chunk_rows);
  for (long long r = r0 + threadIdx.x; r < r1; r += BNL_MASK_THREADS) {
    double d[KB];
    bool ok = true;
#pragma unroll
    for (int i = 0; i < KB; i++) {
      d[i] = 0.0;
      if (i < k) { const double v = data[stride * cols[i] + r]; ok &= !isnan(v); d[i] = v - m[i]; }
    }
    if (!ok) continue;
#pragma unroll
    for (int i = 0; i < KB; i++)
#pragma unroll
      for (int j = i + 1; j < KB; j++)
        if (j < k) { const double t = d[i] * d[j] - cv[i + (size_t)j * kmax] * scale; acc += t * t; }
  }
  const double t = block_sum_256(acc, sm);
  if (threadIdx.x == 0) *out = t;
}

// ---------------------------------------------------------------------------
// Conditional Gaussian nodes (src/scores/cg.loglikelihood.c:10-117 over c_cls,
// src/math/conditional.least.squares.c): a numeric node regressed on its numeric
// parents separately within every configuration of its discrete parents. The
// moments of every (unit, configuration) are gathered in two row passes -- counts
// and sums, then centred cross-products -- into
//   mom[off + s * M + 0]             rows of configuration s
//   mom[off + s * M + 1 + i]         mean of numeric column i (0 = the node), sums until pass 1 is finalised
//   mom[off + s * M + 1 + K + pair]  centred cross-products, pairs i <= j in row-major order
// with K numeric columns and M = 1 + K + K (K + 1) / 2. A CTA works on one row
// chunk of one unit and accumulates in shared memory when the unit's block fits
// (BNL_CG_SMEM_DOUBLES), then adds its block to the global one; double atomics, so
// the last bits depend on the order of arrival.

#define BNL_CG_MAXG 8            // numeric columns per unit (node + 7 numeric parents)
#define BNL_CG_MAXD 8            // discrete parents per unit
#define BNL_CG_SMEM_DOUBLES 5120
#define BNL_CG_THREADS 256

struct BnlCgUnit {
  long long off;                 // first double of the unit's block in `mom`
  int nconfig, K, nd, dwide;     // dwide: bit j set when discrete parent j lives in the uint16 block
  int gslot[BNL_CG_MAXG];        // numeric slots, node first
  int dslot[BNL_CG_MAXD];        // packed factor slots
  int dmul[BNL_CG_MAXD];         // configuration multipliers (first parent fastest)
};

template <int PASS>
__global__ void __launch_bounds__(BNL_CG_THREADS)
cg_moments_kernel(const double *__restrict__ f64, size_t fstride, const uint8_t *__restrict__ u8,
                  size_t ustride, long long nobs, long long chunk_rows,
                  const BnlCgUnit *__restrict__ units, double *__restrict__ mom, int *__restrict__ err,
                  const uint16_t *__restrict__ u16, size_t wstride) {
  extern __shared__ double cgsm[];
  const BnlCgUnit U = units[blockIdx.y];
  const int K = U.K, M = 1 + K + K * (K + 1) / 2;
  const long long total = (long long)U.nconfig * M;
  const bool staged = total <= BNL_CG_SMEM_DOUBLES;
  double *acc = staged ? cgsm : mom + U.off;
  if (staged) {
    for (int i = threadIdx.x; i < (int)total; i += BNL_CG_THREADS) cgsm[i] = 0.0;
    __syncthreads();
  }
  const double *gm = mom + U.off;           // finalised means (pass 2 only)
  const long long r0 = blockIdx.x * chunk_rows, r1 = min(nobs, r0 + chunk_rows);
  for (long long r = r0 + threadIdx.x; r < r1; r += BNL_CG_THREADS) {
    int s = 0;
    bool bad = false;
    for (int j = 0; j < U.nd; j++) {
      if ((U.dwide >> j) & 1) {             // a factor with 128..32767 levels: bit 15 marks NA
        const unsigned b = u16[wstride * U.dslot[j] + r];
        bad |= b >= 0x8000u;
        s += (int)b * U.dmul[j];
      } else {
        const unsigned b = u8[ustride * U.dslot[j] + r];
        bad |= b >= 0x80u;
        s += (int)b * U.dmul[j];
      }
    }
    double v[BNL_CG_MAXG];
#pragma unroll
    for (int i = 0; i < BNL_CG_MAXG; i++)
      if (i < K) { v[i] = f64[fstride * U.gslot[i] + r]; bad |= isnan(v[i]); }
    if (bad) { *err = 1; continue; }        // the complete-data code path: missing values are refused
    double *a = acc + (size_t)s * M;
    if (PASS == 1) {
      atomicAdd(a, 1.0);
#pragma unroll
      for (int i = 0; i < BNL_CG_MAXG; i++) if (i < K) atomicAdd(a + 1 + i, v[i]);
    } else {
      const double *m = gm + (size_t)s * M + 1;
#pragma unroll
      for (int i = 0; i < BNL_CG_MAXG; i++) if (i < K) v[i] -= m[i];
      int p = 1 + K;
#pragma unroll
      for (int i = 0; i < BNL_CG_MAXG; i++)
#pragma unroll
        for (int j = i; j < BNL_CG_MAXG; j++)
          if (j < K) atomicAdd(a + p++, v[i] * v[j]);
    }
  }
  if (staged) {
    __syncthreads();
    double *g = mom + U.off;
    for (int i = threadIdx.x; i < (int)total; i += BNL_CG_THREADS) {
      const int w = i % M;
      const bool mine = PASS == 1 ? w <= K : w > K;
      if (mine && cgsm[i] != 0.0) atomicAdd(g + i, cgsm[i]);
    }
  }
}

// sums -> means, once per (unit, configuration).
__global__ void cg_means_kernel(const BnlCgUnit *__restrict__ units, int n, double *__restrict__ mom) {
  const int u = blockIdx.y;
  if (u >= n) return;
  const BnlCgUnit U = units[u];
  const int K = U.K, M = 1 + K + K * (K + 1) / 2;
  for (int s = blockIdx.x * blockDim.x + threadIdx.x; s < U.nconfig; s += gridDim.x * blockDim.x) {
    double *a = mom + U.off + (size_t)s * M;
    if (a[0] > 0.0) for (int i = 0; i < K; i++) a[1 + i] /= a[0];
  }
}

// Conditional least squares of one configuration from its moments: residual sum
// of squares of node ~ 1 + numeric parents. c_cls1 / c_cls2 test collinearity on
// the SUMS of squares (not on variances, unlike c_ols1 / c_ols2); three or more
// regressors go through c_qr, i.e. dqrdc2's column-dropping rule as in
// gauss_score_kernel. S is the K x K centred cross-product matrix (row-major,
// K <= BNL_CG_MAXG), mean[] the column means, ns the rows of the configuration.
__device__ double cg_stratum_sse(double *S, const double *mean, double ns, int K) {
  const double tol = bnl::kSqrtEps;
  const int np = K - 1;
  auto A = [&](int i, int j) -> double & { return S[i * K + j]; };
  if (np == 0) return A(0, 0);
  if (np == 1) return fabs(A(1, 1)) < tol ? A(0, 0) : A(0, 0) - A(0, 1) / A(1, 1) * A(0, 1);
  if (np == 2) {
    const double c11 = A(1, 1), c22 = A(2, 2), c12 = A(1, 2), c01 = A(0, 1), c02 = A(0, 2);
    const bool s1 = fabs(c11) < tol;
    const bool s2 = (fabs(c22) < tol) || (fabs(c12) / sqrt(c11 * c22) > 1.0 - tol);
    double fit;
    if (s1 && !s2) fit = c02 / c22 * c02;
    else if (!s1 && s2) fit = c01 / c11 * c01;
    else if (s1 && s2) fit = 0.0;
    else {
      const double den = c11 * c22 - c12 * c12;
      const double b1 = (c22 * c01 - c12 * c02) / den, b2 = (c11 * c02 - c12 * c01) / den;
      fit = b1 * c01 + b2 * c02;
    }
    return A(0, 0) - fit;
  }
  for (int l = 1; l < K; l++) {
    const double orig = sqrt(A(l, l) + ns * mean[l] * mean[l]);
    const double piv = A(l, l);
    if (!(piv > 0.0) || sqrt(piv) < orig * tol) continue;          // dropped column
    for (int i = 0; i < K; i++) {
      if (i == l) continue;
      const double f = A(i, l) / piv;
      if (f == 0.0) continue;
      for (int j = 0; j < K; j++) if (j != l) A(i, j) -= f * A(l, j);
    }
    for (int i = 0; i < K; i++) if (i != l) { A(i, l) = 0.0; A(l, i) = 0.0; }
  }
  return A(0, 0);
}

// One thread per unit: log-likelihood = sum over the observed configurations;
// -Inf as soon as one of them is singular (sd < sqrt(eps), which includes
// configurations with no more rows than coefficients).
__global__ void cg_score_kernel(const BnlCgUnit *__restrict__ units, int n, const double *__restrict__ mom,
                                double *__restrict__ out, int nal, double kpen, long long nobs) {
  const int u = blockIdx.x * blockDim.x + threadIdx.x;
  if (u >= n) return;
  const BnlCgUnit U = units[u];
  const int K = U.K, np = K - 1, M = 1 + K + K * (K + 1) / 2;
  const double tol = bnl::kSqrtEps;
  bnl::KahanSum tot;
  bool singular = false;
  double nc = 0.0;
  double S[BNL_CG_MAXG * BNL_CG_MAXG], mean[BNL_CG_MAXG];
  for (int s = 0; s < U.nconfig && !singular; s++) {
    const double *a = mom + U.off + (size_t)s * M;
    const double ns = a[0];
    if (ns == 0.0) {
      // sd = NA: no row refers to it. The node-average likelihood refuses the node
      // (c_fast_ccgnal, src/scores/cg.nal.c:52-58), the log-likelihood skips it.
      if (nal) singular = true;
      continue;
    }
    if (ns <= (double)(np + 1)) { singular = true; break; }       // sd = 0
    nc += ns;
    int p = 1 + K;
    for (int i = 0; i < K; i++) {
      mean[i] = a[1 + i];
      for (int j = i; j < K; j++) { S[i * K + j] = a[p]; S[j * K + i] = a[p]; p++; }
    }
    double sse = cg_stratum_sse(S, mean, ns, K);
    if (sse < 0.0) sse = 0.0;
    const double sd = sqrt(sse / (ns - (double)np - 1.0));
    if (!(sd >= tol)) { singular = true; break; }
    tot.add(-ns * (bnl::kLnSqrt2Pi + log(sd)) - 0.5 * sse / (sd * sd));
  }
  const double npar = (double)U.nconfig * (double)(np + 2);
  double res = singular ? -INFINITY : tot.value();
  // nal-cg / pnal-cg: averaged over the locally complete rows, penalty scaled by the
  // full row count (src/scores/cg.nal.c:68-71).
  if (nal && !singular) res = res / nc - kpen / (double)nobs * npar;
  out[(size_t)u * 4 + 0] = res;
  out[(size_t)u * 4 + 1] = npar;
  out[(size_t)u * 4 + 2] = 0.0;
  out[(size_t)u * 4 + 3] = 0.0;
}

// Units with at most BNL_CG_SMALL_CFG configurations and BNL_CG_SMALL_K numeric
// columns (a numeric node with one small discrete parent -- most candidates of an
// hc iteration on mixed data): the rows of a warp would collide on 2-4 addresses
// in the atomic kernel, so here every thread keeps the accumulators of ALL the
// configurations in registers and adds each row to its own configuration through
// a predicate. One block reduction per CTA at the end, no atomics in the row loop.
#define BNL_CG_SMALL_CFG 4
#define BNL_CG_SMALL_K 4

template <int PASS>
__global__ void __launch_bounds__(BNL_CG_THREADS)
cg_moments_small_kernel(const double *__restrict__ f64, size_t fstride, const uint8_t *__restrict__ u8,
                        size_t ustride, long long nobs, long long chunk_rows,
                        const BnlCgUnit *__restrict__ units, double *__restrict__ mom,
                        int *__restrict__ err) {
  constexpr int KS = BNL_CG_SMALL_K, NC = BNL_CG_SMALL_CFG;
  constexpr int NA = PASS == 1 ? 1 + KS : KS * (KS + 1) / 2;
  __shared__ double s_red[BNL_CG_THREADS / 32];
  __shared__ double s_mean[NC * KS];
  const BnlCgUnit U = units[blockIdx.y];
  const int K = U.K, M = 1 + K + K * (K + 1) / 2, nconfig = U.nconfig;
  if (PASS == 2) {
    if (threadIdx.x < NC * KS) {
      const int c = threadIdx.x / KS, i = threadIdx.x % KS;
      s_mean[threadIdx.x] = (c < nconfig && i < K) ? mom[U.off + (size_t)c * M + 1 + i] : 0.0;
    }
    __syncthreads();
  }
  double acc[NC][NA];
#pragma unroll
  for (int c = 0; c < NC; c++)
#pragma unroll
    for (int a = 0; a < NA; a++) acc[c][a] = 0.0;
  bool bad_any = false;
  const long long r0 = blockIdx.x * chunk_rows, r1 = min(nobs, r0 + chunk_rows);
  for (long long r = r0 + threadIdx.x; r < r1; r += BNL_CG_THREADS) {
    int s = 0;
    bool bad = false;
    for (int j = 0; j < U.nd; j++) {
      const unsigned b = u8[ustride * U.dslot[j] + r];
      bad |= b >= 0x80u;
      s += (int)b * U.dmul[j];
    }
    double v[KS];
#pragma unroll
    for (int i = 0; i < KS; i++) {
      v[i] = 0.0;
      if (i < K) { v[i] = f64[fstride * U.gslot[i] + r]; bad |= isnan(v[i]); }
    }
    if (bad) { bad_any = true; continue; }
    if (PASS == 2) {
#pragma unroll
      for (int i = 0; i < KS; i++) v[i] -= s_mean[s * KS + i];     // columns beyond K stay 0
    }
#pragma unroll
    for (int c = 0; c < NC; c++) {
      const double w = s == c ? 1.0 : 0.0;
      if (PASS == 1) {
        acc[c][0] += w;
#pragma unroll
        for (int i = 0; i < KS; i++) acc[c][1 + i] += w * v[i];
      } else {
        int p = 0;
#pragma unroll
        for (int i = 0; i < KS; i++)
#pragma unroll
          for (int j = i; j < KS; j++, p++) acc[c][p] += w * (v[i] * v[j]);
      }
    }
  }
  if (bad_any) *err = 1;
  // block reduction of every accumulator; thread 0 adds the CTA's share to the unit's block.
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
#pragma unroll
  for (int c = 0; c < NC; c++) {
#pragma unroll
    for (int a = 0; a < NA; a++) {
      double t = acc[c][a];
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
      __syncthreads();
      if (lane == 0) s_red[warp] = t;
      __syncthreads();
      if (threadIdx.x == 0 && c < nconfig) {
        double tot = 0.0;
        for (int w = 0; w < BNL_CG_THREADS / 32; w++) tot += s_red[w];
        // slot of accumulator `a` in the unit's own layout (K columns, not KS)
        int dst = -1;
        if (PASS == 1) { if (a <= K) dst = a; }
        else {
          int i = 0, rem = a;
          while (rem >= KS - i) { rem -= KS - i; i++; }
          const int j = i + rem;
          if (j < K) dst = 1 + K + (i * K - i * (i - 1) / 2 + (j - i));
        }
        if (dst >= 0 && tot != 0.0) atomicAdd(mom + U.off + (size_t)c * M + dst, tot);
      }
    }
  }
}
