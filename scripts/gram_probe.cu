// gram_probe.cu -- variants of the Gram kernel's inner structure, timed on random
// columns (p x n doubles, column-major) against each other. Development probe for
// bnlearn_b200/csrc/bnl_gaussian_kernels.cuh (gram_dmma_kernel); nothing here ships.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -o scripts/_bin/gram_probe scripts/gram_probe.cu
//   scripts/_bin/gram_probe [p=2000] [n=262144]
#include <cstdio>
#include <cstdlib>
#include <cmath>
#include <vector>
#include <cuda_runtime.h>

#define TILE 128
#define THREADS 512

__device__ __forceinline__ void dmma(double &c0, double &c1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
               : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}
__device__ __forceinline__ void cp16(void *smem, const void *gmem) {
  const unsigned d = (unsigned)__cvta_generic_to_shared(smem);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" :: "r"(d), "l"(gmem));
}
__device__ __forceinline__ void cp_commit() { asm volatile("cp.async.commit_group;"); }
template <int N> __device__ __forceinline__ void cp_wait() { asm volatile("cp.async.wait_group %0;" :: "n"(N)); }

// CENTRE 0: means subtracted when fragments are read (the shipped kernel), 1: each thread
// centres the chunks it copied itself before the stage's barrier, 2: not at all (bound only).
template <int KB, int STAGES, int CENTRE>
__global__ void __launch_bounds__(THREADS, 1)
gram_variant(const double *__restrict__ X, size_t stride, long long nobs, int p,
             const double *__restrict__ mean, double *__restrict__ partial, int ntile) {
  constexpr int LD = KB + 4;
  constexpr int CPC = KB / 2;                       // 16-byte chunks per column per stage
  constexpr int REPS = TILE * CPC / THREADS;
  extern __shared__ __align__(16) double sm[];
  typedef double (*tile_t)[TILE][LD];
  tile_t As = reinterpret_cast<tile_t>(sm);
  tile_t Bs = reinterpret_cast<tile_t>(sm + (size_t)STAGES * TILE * LD);

  int t = blockIdx.x, ti = 0;
  while (t >= ntile - ti) { t -= ntile - ti; ti++; }
  const int tj = ti + t;
  const bool diag = ti == tj;
  const int nsplit = gridDim.y, split = blockIdx.y;
  const long long steps = (nobs + KB - 1) / KB;
  const long long per = (steps + nsplit - 1) / nsplit;
  const long long s0 = per * split;
  long long s1 = s0 + per;
  if (s1 > steps) s1 = steps;
  const int nst = (int)(s1 > s0 ? s1 - s0 : 0);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int wm = (warp >> 2) * 32, wn = (warp & 3) * 32;
  const int g = lane >> 2, q = lane & 3;

  double cma[REPS], cmb[REPS];                      // CENTRE 1: means of the columns this thread copies
#pragma unroll
  for (int rep = 0; rep < REPS; rep++) {
    const int c = (threadIdx.x + rep * THREADS) / CPC;
    int col = ti * TILE + c; if (col >= p) col = p - 1;
    cma[rep] = (CENTRE == 1 || CENTRE == 4) ? mean[col] : 0.0;
    col = tj * TILE + c; if (col >= p) col = p - 1;
    cmb[rep] = (CENTRE == 1 || CENTRE == 4) ? mean[col] : 0.0;
  }

  auto issue = [&](int s_local) {
    const long long r0 = (s0 + s_local) * KB;
    const int buf = s_local % STAGES;
#pragma unroll
    for (int rep = 0; rep < REPS; rep++) {
      const int idx = threadIdx.x + rep * THREADS;
      const int c = idx / CPC, r2 = (idx % CPC) * 2;
      int col = ti * TILE + c; if (col >= p) col = p - 1;
      cp16(&As[buf][c][r2], X + (size_t)col * stride + r0 + r2);
      if (!diag) {
        col = tj * TILE + c; if (col >= p) col = p - 1;
        cp16(&Bs[buf][c][r2], X + (size_t)col * stride + r0 + r2);
      }
    }
  };
  auto centre_own = [&](int s_local) {
    const long long r0 = (s0 + s_local) * KB;
    const int buf = s_local % STAGES;
#pragma unroll
    for (int rep = 0; rep < REPS; rep++) {
      const int idx = threadIdx.x + rep * THREADS;
      const int c = idx / CPC, r2 = (idx % CPC) * 2;
      const bool l0 = r0 + r2 < nobs, l1 = r0 + r2 + 1 < nobs;
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

  double ma[4], mb[4];
#pragma unroll
  for (int i = 0; i < 4; i++) {
    int col = ti * TILE + wm + i * 8 + g;
    ma[i] = mean[col < p ? col : p - 1];
    col = tj * TILE + wn + i * 8 + g;
    mb[i] = mean[col < p ? col : p - 1];
  }
  double acc[4][4][2];
#pragma unroll
  for (int i = 0; i < 4; i++)
#pragma unroll
    for (int j = 0; j < 4; j++) acc[i][j][0] = acc[i][j][1] = 0.0;

#pragma unroll
  for (int s = 0; s < STAGES - 1; s++) {
    if (s < nst && CENTRE != 3) issue(s);
    cp_commit();
  }
  if (CENTRE == 4) { cp_wait<STAGES - 2>(); if (nst > 0) centre_own(0); }
  for (int s = 0; s < nst; s++) {
    if (CENTRE != 4) cp_wait<STAGES - 2>();
    if (CENTRE == 1) centre_own(s);
    __syncthreads();
    if (s + STAGES - 1 < nst && CENTRE != 3) issue(s + STAGES - 1);
    cp_commit();
    const int buf = s % STAGES;
    const double (*Ap)[LD] = As[buf];
    const double (*Bp)[LD] = diag ? As[buf] : Bs[buf];
    const long long r0 = (s0 + s) * KB;
    const bool ragged = r0 + KB > nobs;
#pragma unroll
    for (int kk = 0; kk < KB; kk += 4) {
      double a[4], b[4];
      if (CENTRE == 4 && kk == KB / 2) {          // own chunks of the next stage: landed by now, centred between the DMMAs
        cp_wait<STAGES - 2>();
        if (s + 1 < nst) centre_own(s + 1);
      }
      if (CENTRE == 0) {
        const bool live = !ragged || (r0 + kk + q < nobs);
#pragma unroll
        for (int i = 0; i < 4; i++) a[i] = live ? Ap[wm + i * 8 + g][kk + q] - ma[i] : 0.0;
#pragma unroll
        for (int j = 0; j < 4; j++) b[j] = live ? Bp[wn + j * 8 + g][kk + q] - mb[j] : 0.0;
      } else {
#pragma unroll
        for (int i = 0; i < 4; i++) a[i] = Ap[wm + i * 8 + g][kk + q];
#pragma unroll
        for (int j = 0; j < 4; j++) b[j] = Bp[wn + j * 8 + g][kk + q];
      }
#pragma unroll
      for (int i = 0; i < 4; i++)
#pragma unroll
        for (int j = 0; j < 4; j++) dmma(acc[i][j][0], acc[i][j][1], a[i], b[j]);
    }
  }
  cp_wait<0>();

  double *out = partial + (size_t)split * p * p;
#pragma unroll
  for (int i = 0; i < 4; i++)
#pragma unroll
    for (int j = 0; j < 4; j++) {
      const int row = ti * TILE + wm + i * 8 + g;
      const int col = tj * TILE + wn + j * 8 + q * 2;
      if (row < p && col < p) out[(size_t)row + (size_t)col * p] = acc[i][j][0];
      if (row < p && col + 1 < p) out[(size_t)row + (size_t)(col + 1) * p] = acc[i][j][1];
    }
}

// 8 warps of 64 x 32 (8 x 4 DMMA tiles each), 256 threads: twice the independent DMMAs per warp and
// k-step, fragments of the next k-step loaded while the current one runs (DB = 1).
template <int KB, int STAGES, int CENTRE, int DB>
__global__ void __launch_bounds__(256, 1)
gram_fat(const double *__restrict__ X, size_t stride, long long nobs, int p,
             const double *__restrict__ mean, double *__restrict__ partial, int ntile) {
  constexpr int LD = KB + 4;
  constexpr int CPC = KB / 2;                       // 16-byte chunks per column per stage
  constexpr int REPS = TILE * CPC / 256;
  extern __shared__ __align__(16) double sm[];
  typedef double (*tile_t)[TILE][LD];
  tile_t As = reinterpret_cast<tile_t>(sm);
  tile_t Bs = reinterpret_cast<tile_t>(sm + (size_t)STAGES * TILE * LD);

  int t = blockIdx.x, ti = 0;
  while (t >= ntile - ti) { t -= ntile - ti; ti++; }
  const int tj = ti + t;
  const bool diag = ti == tj;
  const int nsplit = gridDim.y, split = blockIdx.y;
  const long long steps = (nobs + KB - 1) / KB;
  const long long per = (steps + nsplit - 1) / nsplit;
  const long long s0 = per * split;
  long long s1 = s0 + per;
  if (s1 > steps) s1 = steps;
  const int nst = (int)(s1 > s0 ? s1 - s0 : 0);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int wm = (warp >> 2) * 64, wn = (warp & 3) * 32;
  const int g = lane >> 2, q = lane & 3;

  double cma[REPS], cmb[REPS];                      // CENTRE 1: means of the columns this thread copies
#pragma unroll
  for (int rep = 0; rep < REPS; rep++) {
    const int c = (threadIdx.x + rep * 256) / CPC;
    int col = ti * TILE + c; if (col >= p) col = p - 1;
    cma[rep] = CENTRE == 1 ? mean[col] : 0.0;
    col = tj * TILE + c; if (col >= p) col = p - 1;
    cmb[rep] = CENTRE == 1 ? mean[col] : 0.0;
  }

  auto issue = [&](int s_local) {
    const long long r0 = (s0 + s_local) * KB;
    const int buf = s_local % STAGES;
#pragma unroll
    for (int rep = 0; rep < REPS; rep++) {
      const int idx = threadIdx.x + rep * 256;
      const int c = idx / CPC, r2 = (idx % CPC) * 2;
      int col = ti * TILE + c; if (col >= p) col = p - 1;
      cp16(&As[buf][c][r2], X + (size_t)col * stride + r0 + r2);
      if (!diag) {
        col = tj * TILE + c; if (col >= p) col = p - 1;
        cp16(&Bs[buf][c][r2], X + (size_t)col * stride + r0 + r2);
      }
    }
  };
  auto centre_own = [&](int s_local) {
    const long long r0 = (s0 + s_local) * KB;
    const int buf = s_local % STAGES;
#pragma unroll
    for (int rep = 0; rep < REPS; rep++) {
      const int idx = threadIdx.x + rep * 256;
      const int c = idx / CPC, r2 = (idx % CPC) * 2;
      const bool l0 = r0 + r2 < nobs, l1 = r0 + r2 + 1 < nobs;
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

  double acc[8][4][2];
#pragma unroll
  for (int i = 0; i < 8; i++)
#pragma unroll
    for (int j = 0; j < 4; j++) acc[i][j][0] = acc[i][j][1] = 0.0;

#pragma unroll
  for (int s = 0; s < STAGES - 1; s++) {
    if (s < nst) issue(s);
    cp_commit();
  }
  for (int s = 0; s < nst; s++) {
    cp_wait<STAGES - 2>();
    if (CENTRE == 1) centre_own(s);
    __syncthreads();
    if (s + STAGES - 1 < nst) issue(s + STAGES - 1);
    cp_commit();
    const int buf = s % STAGES;
    const double (*Ap)[LD] = As[buf];
    const double (*Bp)[LD] = diag ? As[buf] : Bs[buf];
    double a[2][8], b[2][4];
    auto frag = [&](int kk, int w) {
#pragma unroll
      for (int i = 0; i < 8; i++) a[w][i] = Ap[wm + i * 8 + g][kk + q];
#pragma unroll
      for (int j = 0; j < 4; j++) b[w][j] = Bp[wn + j * 8 + g][kk + q];
    };
    if (DB) frag(0, 0);
#pragma unroll
    for (int kk = 0; kk < KB; kk += 4) {
      const int w = DB ? (kk >> 2) & 1 : 0;
      if (DB) { if (kk + 4 < KB) frag(kk + 4, w ^ 1); } else frag(kk, 0);
#pragma unroll
      for (int i = 0; i < 8; i++)
#pragma unroll
        for (int j = 0; j < 4; j++) dmma(acc[i][j][0], acc[i][j][1], a[w][i], b[w][j]);
    }
  }
  cp_wait<0>();

  double *out = partial + (size_t)split * p * p;
#pragma unroll
  for (int i = 0; i < 8; i++)
#pragma unroll
    for (int j = 0; j < 4; j++) {
      const int row = ti * TILE + wm + i * 8 + g;
      const int col = tj * TILE + wn + j * 8 + q * 2;
      if (row < p && col < p) out[(size_t)row + (size_t)col * p] = acc[i][j][0];
      if (row < p && col + 1 < p) out[(size_t)row + (size_t)(col + 1) * p] = acc[i][j][1];
    }
}

// Stage copies by the bulk-copy engine: one cp.async.bulk of KB * 8 bytes per column and stage
// (256 threads issue one each), completion counted in bytes on one mbarrier per stage.
__device__ __forceinline__ unsigned su32(const void *p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mb_init(unsigned long long *b, int c) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(su32(b)), "r"(c)); }
__device__ __forceinline__ void mb_expect(unsigned long long *b, unsigned bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(su32(b)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void *dst, const void *src, unsigned bytes, unsigned long long *b) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               :: "r"(su32(dst)), "l"(src), "r"(bytes), "r"(su32(b)) : "memory");
}
__device__ __forceinline__ void mb_wait(unsigned long long *b, unsigned parity) {
  for (int tries = 0; tries < (1 << 26); tries++) {
    unsigned done;
    asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
                 : "=r"(done) : "r"(su32(b)), "r"(parity) : "memory");
    if (done) return;
  }
  __trap();
}

template <int KB, int STAGES, int CENTRE>
__global__ void __launch_bounds__(THREADS, 1)
gram_bulk(const double *__restrict__ X, size_t stride, long long nobs, int p,
          const double *__restrict__ mean, double *__restrict__ partial, int ntile) {
  constexpr int LD = KB + 4;
  constexpr int CPC = KB / 2;
  constexpr int REPS = TILE * CPC / THREADS;
  extern __shared__ __align__(16) double sm[];
  typedef double (*tile_t)[TILE][LD];
  tile_t As = reinterpret_cast<tile_t>(sm);
  tile_t Bs = reinterpret_cast<tile_t>(sm + (size_t)STAGES * TILE * LD);
  unsigned long long *full = reinterpret_cast<unsigned long long *>(sm + (size_t)2 * STAGES * TILE * LD);

  int t = blockIdx.x, ti = 0;
  while (t >= ntile - ti) { t -= ntile - ti; ti++; }
  const int tj = ti + t;
  const bool diag = ti == tj;
  const int nsplit = gridDim.y, split = blockIdx.y;
  const long long steps = (nobs + KB - 1) / KB;
  const long long per = (steps + nsplit - 1) / nsplit;
  const long long s0 = per * split;
  long long s1 = s0 + per;
  if (s1 > steps) s1 = steps;
  const int nst = (int)(s1 > s0 ? s1 - s0 : 0);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int wm = (warp >> 2) * 32, wn = (warp & 3) * 32;
  const int g = lane >> 2, q = lane & 3;
  if (threadIdx.x == 0) {
    for (int i = 0; i < STAGES; i++) mb_init(&full[i], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();

  // my column of the copies: threads 0-127 the A operand, 128-255 the B operand
  const int myc = threadIdx.x & 127;
  const bool copier = threadIdx.x < (diag ? 128 : 256);
  int mycol = (threadIdx.x < 128 ? ti : tj) * TILE + myc;
  if (mycol >= p) mycol = p - 1;
  const double *mysrc = X + (size_t)mycol * stride;
  const unsigned stage_bytes = (diag ? 1u : 2u) * TILE * KB * 8;

  double cma[REPS], cmb[REPS];
#pragma unroll
  for (int rep = 0; rep < REPS; rep++) {
    const int c = (threadIdx.x + rep * THREADS) / CPC;
    int col = ti * TILE + c; if (col >= p) col = p - 1;
    cma[rep] = CENTRE == 1 ? mean[col] : 0.0;
    col = tj * TILE + c; if (col >= p) col = p - 1;
    cmb[rep] = CENTRE == 1 ? mean[col] : 0.0;
  }
  auto expect = [&](int s_local) { if (threadIdx.x == 0) mb_expect(&full[s_local % STAGES], stage_bytes); };
  auto issue = [&](int s_local) {
    if (!copier) return;
    const long long r0 = (s0 + s_local) * KB;
    const int buf = s_local % STAGES;
    double *dst = threadIdx.x < 128 ? &As[buf][myc][0] : &Bs[buf][myc][0];
    if (CENTRE == 1) asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // generic writes (centring) before async writes
    bulk_g2s(dst, mysrc + r0, KB * 8, &full[buf]);
  };
  auto centre_own = [&](int s_local) {
    const long long r0 = (s0 + s_local) * KB;
    const int buf = s_local % STAGES;
#pragma unroll
    for (int rep = 0; rep < REPS; rep++) {
      const int idx = threadIdx.x + rep * THREADS;
      const int c = idx / CPC, r2 = (idx % CPC) * 2;
      const bool l0 = r0 + r2 < nobs, l1 = r0 + r2 + 1 < nobs;
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

  for (int s = 0; s < STAGES - 1; s++) if (s < nst) expect(s);
  __syncthreads();
  for (int s = 0; s < STAGES - 1; s++) if (s < nst) issue(s);
  for (int s = 0; s < nst; s++) {
    const int buf = s % STAGES;
    mb_wait(&full[buf], (unsigned)((s / STAGES) & 1));
    if (CENTRE == 1) centre_own(s);
    if (s + STAGES - 1 < nst) expect(s + STAGES - 1);
    __syncthreads();
    if (s + STAGES - 1 < nst) issue(s + STAGES - 1);
    const double (*Ap)[LD] = As[buf];
    const double (*Bp)[LD] = diag ? As[buf] : Bs[buf];
#pragma unroll
    for (int kk = 0; kk < KB; kk += 4) {
      double a[4], b[4];
#pragma unroll
      for (int i = 0; i < 4; i++) a[i] = Ap[wm + i * 8 + g][kk + q];
#pragma unroll
      for (int j = 0; j < 4; j++) b[j] = Bp[wn + j * 8 + g][kk + q];
#pragma unroll
      for (int i = 0; i < 4; i++)
#pragma unroll
        for (int j = 0; j < 4; j++) dmma(acc[i][j][0], acc[i][j][1], a[i], b[j]);
    }
  }
  double *out = partial + (size_t)split * p * p;
#pragma unroll
  for (int i = 0; i < 4; i++)
#pragma unroll
    for (int j = 0; j < 4; j++) {
      const int row = ti * TILE + wm + i * 8 + g;
      const int col = tj * TILE + wn + j * 8 + q * 2;
      if (row < p && col < p) out[(size_t)row + (size_t)col * p] = acc[i][j][0];
      if (row < p && col + 1 < p) out[(size_t)row + (size_t)(col + 1) * p] = acc[i][j][1];
    }
}

__global__ void fill(double *X, size_t stride, long long nobs, int p) {
  const long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  const int c = blockIdx.y;
  if (i >= nobs) return;
  unsigned long long z = (unsigned long long)c * 0x9e3779b97f4a7c15ull + (unsigned long long)i * 0xbf58476d1ce4e5b9ull + 12345;
  z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ull; z = (z ^ (z >> 27)) * 0x94d049bb133111ebull; z ^= z >> 31;
  X[(size_t)c * stride + i] = (double)(z >> 11) * (1.0 / 9007199254740992.0) * 4.0 + (c % 7) - 3.0;
}
__global__ void means(const double *X, size_t stride, long long nobs, double *mean) {
  __shared__ double s[256];
  const int c = blockIdx.x;
  double a = 0;
  for (long long i = threadIdx.x; i < nobs; i += 256) a += X[(size_t)c * stride + i];
  s[threadIdx.x] = a; __syncthreads();
  for (int o = 128; o > 0; o >>= 1) { if (threadIdx.x < o) s[threadIdx.x] += s[threadIdx.x + o]; __syncthreads(); }
  if (threadIdx.x == 0) mean[c] = s[0] / (double)nobs;
}
__global__ void sum_partials(const double *partial, int nsplit, int p, double *out) {
  const long long e = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (e >= (long long)p * p) return;
  double t = 0;
  for (int s = 0; s < nsplit; s++) t += partial[(size_t)s * p * p + e];
  out[e] = t;
}

struct Env { double *X, *mean, *partial, *ref, *got; size_t stride; long long n; int p, ntile, ntri, nsplit; };

template <int KB, int STAGES, int CENTRE>
void run(const Env &E, const char *name, bool is_ref) {
  const size_t smem = (size_t)STAGES * 2 * TILE * (KB + 4) * 8;
  cudaFuncSetAttribute(gram_variant<KB, STAGES, CENTRE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  float best = 1e30f;
  for (int rep = 0; rep < 3; rep++) {
    cudaEventRecord(e0);
    gram_variant<KB, STAGES, CENTRE><<<dim3(E.ntri, E.nsplit), THREADS, smem>>>(E.X, E.stride, E.n, E.p, E.mean, E.partial, E.ntile);
    cudaEventRecord(e1);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("%s: %s\n", name, cudaGetErrorString(e)); return; }
    float ms; cudaEventElapsedTime(&ms, e0, e1); best = ms < best ? ms : best;
  }
  const long long el = (long long)E.p * E.p;
  sum_partials<<<(unsigned)((el + 255) / 256), 256>>>(E.partial, E.nsplit, E.p, is_ref ? E.ref : E.got);
  cudaDeviceSynchronize();
  double worst = 0;
  if (!is_ref) {
    std::vector<double> r(el), g(el);
    cudaMemcpy(r.data(), E.ref, el * 8, cudaMemcpyDeviceToHost);
    cudaMemcpy(g.data(), E.got, el * 8, cudaMemcpyDeviceToHost);
    for (int j = 0; j < E.p; j++) for (int i = 0; i <= j; i++) {
      const double a = r[(size_t)i + (size_t)j * E.p], b = g[(size_t)i + (size_t)j * E.p];
      const double d = fabs(a - b) / fmax(1.0, fmax(fabs(a), fabs(b)));
      if (d > worst) worst = d;
    }
  }
  const double alg = (double)E.n * E.p * (E.p + 1), act = (double)E.ntri * TILE * TILE * 2.0 * E.n;
  printf("%-44s smem %6zu  %8.3f ms  %6.2f TFLOP/s algorithmic  %6.2f issued  max rel diff %.2e\n", name, smem, best,
         alg / best / 1e9, act / best / 1e9, worst);
}

template <int KB, int STAGES, int CENTRE>
void run_bulk(const Env &E, const char *name) {
  const size_t smem = (size_t)STAGES * 2 * TILE * (KB + 4) * 8 + 64;
  cudaFuncSetAttribute(gram_bulk<KB, STAGES, CENTRE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  float best = 1e30f;
  for (int rep = 0; rep < 3; rep++) {
    cudaEventRecord(e0);
    gram_bulk<KB, STAGES, CENTRE><<<dim3(E.ntri, E.nsplit), THREADS, smem>>>(E.X, E.stride, E.n, E.p, E.mean, E.partial, E.ntile);
    cudaEventRecord(e1);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("%s: %s\n", name, cudaGetErrorString(e)); return; }
    float ms; cudaEventElapsedTime(&ms, e0, e1); best = ms < best ? ms : best;
  }
  const long long el = (long long)E.p * E.p;
  sum_partials<<<(unsigned)((el + 255) / 256), 256>>>(E.partial, E.nsplit, E.p, E.got);
  cudaDeviceSynchronize();
  double worst = 0;
  std::vector<double> r(el), g(el);
  cudaMemcpy(r.data(), E.ref, el * 8, cudaMemcpyDeviceToHost);
  cudaMemcpy(g.data(), E.got, el * 8, cudaMemcpyDeviceToHost);
  for (int j = 0; j < E.p; j++) for (int i = 0; i <= j; i++) {
    const double a = r[(size_t)i + (size_t)j * E.p], b = g[(size_t)i + (size_t)j * E.p];
    const double d = fabs(a - b) / fmax(1.0, fmax(fabs(a), fabs(b)));
    if (d > worst) worst = d;
  }
  const double alg = (double)E.n * E.p * (E.p + 1), act = (double)E.ntri * TILE * TILE * 2.0 * E.n;
  printf("%-44s smem %6zu  %8.3f ms  %6.2f TFLOP/s algorithmic  %6.2f issued  max rel diff %.2e\n", name, smem, best,
         alg / best / 1e9, act / best / 1e9, worst);
}

template <int KB, int STAGES, int CENTRE, int DB>
void run_fat(const Env &E, const char *name) {
  const size_t smem = (size_t)STAGES * 2 * TILE * (KB + 4) * 8;
  cudaFuncSetAttribute(gram_fat<KB, STAGES, CENTRE, DB>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  float best = 1e30f;
  for (int rep = 0; rep < 3; rep++) {
    cudaEventRecord(e0);
    gram_fat<KB, STAGES, CENTRE, DB><<<dim3(E.ntri, E.nsplit), 256, smem>>>(E.X, E.stride, E.n, E.p, E.mean, E.partial, E.ntile);
    cudaEventRecord(e1);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("%s: %s\n", name, cudaGetErrorString(e)); return; }
    float ms; cudaEventElapsedTime(&ms, e0, e1); best = ms < best ? ms : best;
  }
  const long long el = (long long)E.p * E.p;
  sum_partials<<<(unsigned)((el + 255) / 256), 256>>>(E.partial, E.nsplit, E.p, E.got);
  cudaDeviceSynchronize();
  double worst = 0;
  std::vector<double> r(el), g(el);
  cudaMemcpy(r.data(), E.ref, el * 8, cudaMemcpyDeviceToHost);
  cudaMemcpy(g.data(), E.got, el * 8, cudaMemcpyDeviceToHost);
  for (int j = 0; j < E.p; j++) for (int i = 0; i <= j; i++) {
    const double a = r[(size_t)i + (size_t)j * E.p], b = g[(size_t)i + (size_t)j * E.p];
    const double d = fabs(a - b) / fmax(1.0, fmax(fabs(a), fabs(b)));
    if (d > worst) worst = d;
  }
  const double alg = (double)E.n * E.p * (E.p + 1), act = (double)E.ntri * TILE * TILE * 2.0 * E.n;
  printf("%-44s smem %6zu  %8.3f ms  %6.2f TFLOP/s algorithmic  %6.2f issued  max rel diff %.2e\n", name, smem, best,
         alg / best / 1e9, act / best / 1e9, worst);
}

int main(int argc, char **argv) {
  Env E;
  E.p = argc > 1 ? atoi(argv[1]) : 2000;
  E.n = argc > 2 ? atoll(argv[2]) : 262144;
  E.stride = (size_t)((E.n + 63) / 64 * 64 + 64);
  E.ntile = (E.p + TILE - 1) / TILE; E.ntri = E.ntile * (E.ntile + 1) / 2;
  long long ns = (12ll * 148 + E.ntri - 1) / E.ntri;
  if (ns > E.n / 8192) ns = E.n / 8192 > 0 ? E.n / 8192 : 1;
  E.nsplit = (int)ns;
  cudaMalloc(&E.X, E.stride * E.p * 8); cudaMalloc(&E.mean, E.p * 8);
  cudaMalloc(&E.partial, (size_t)E.nsplit * E.p * E.p * 8);
  cudaMalloc(&E.ref, (size_t)E.p * E.p * 8); cudaMalloc(&E.got, (size_t)E.p * E.p * 8);
  cudaMemset(E.X, 0, E.stride * E.p * 8);
  fill<<<dim3((unsigned)((E.n + 255) / 256), E.p), 256>>>(E.X, E.stride, E.n, E.p);
  means<<<E.p, 256>>>(E.X, E.stride, E.n, E.mean);
  cudaMemset(E.partial, 0, (size_t)E.nsplit * E.p * E.p * 8);
  printf("p %d  n %lld  tiles %d  splits %d\n", E.p, E.n, E.ntri, E.nsplit);
  run<16, 4, 0>(E, "KB16 x4, centred at fragment read (shipped)", true);
  run<16, 4, 2>(E, "KB16 x4, not centred (bound only)", false);
  run<16, 4, 1>(E, "KB16 x4, centred in smem by the copier", false);
  run<32, 3, 1>(E, "KB32 x3, centred in smem by the copier", false);
  run<32, 3, 0>(E, "KB32 x3, centred at fragment read", false);
  run<32, 3, 2>(E, "KB32 x3, not centred (bound only)", false);
  run<16, 4, 3>(E, "KB16 x4, no copies at all (bound only)", false);
  run<32, 3, 3>(E, "KB32 x3, no copies at all (bound only)", false);
  run<16, 4, 4>(E, "KB16 x4, next stage centred mid-loop", false);
  run<32, 3, 4>(E, "KB32 x3, next stage centred mid-loop", false);
  run<16, 5, 4>(E, "KB16 x5, next stage centred mid-loop", false);
  run<24, 4, 4>(E, "KB24 x4, next stage centred mid-loop", false);
  run_bulk<16, 4, 2>(E, "bulk copies KB16 x4, not centred");
  run_bulk<32, 3, 2>(E, "bulk copies KB32 x3, not centred");
  run_bulk<16, 4, 1>(E, "bulk copies KB16 x4, smem-centred");
  run_bulk<32, 3, 1>(E, "bulk copies KB32 x3, smem-centred");
  run_fat<16, 4, 1, 0>(E, "fat 8 warps KB16 x4, smem-centred");
  run_fat<16, 4, 1, 1>(E, "fat 8 warps KB16 x4, smem-centred, frag DB");
  run_fat<32, 3, 1, 1>(E, "fat 8 warps KB32 x3, smem-centred, frag DB");
  run_fat<32, 3, 2, 1>(E, "fat 8 warps KB32 x3, not centred, frag DB");
  return 0;
}
