// bnl_noalloc.cu -- one discrete CI test without any allocation.
//
// Who needs this: code that lives inside somebody else's memory regime. A host
// framework that hands out device memory from its own pool (PyTorch's caching
// allocator, RMM, a serving runtime that reserved the whole card at start-up)
// cannot let a library call cudaMalloc behind its back: the call may fail outright
// because the pool already owns the memory, and even when it succeeds it
// synchronises the device and stalls every other stream of the process. The same
// holds for a latency-sensitive loop that must not hit the allocator at all.
// Such a caller already has its packed columns and its result buffer on the
// device; all the engine is missing is the count table it normally keeps in a
// cudaMalloc'ed scratch buffer. Here that table lives in a file-scope
// __device__ array of fixed size instead.
//
// Consequences of the fixed array, all documented in include/bnl_b200.h too:
//   * tables larger than BNL_NOALLOC_MAX_CELLS are refused (BNL_EUNSUPPORTED);
//   * every GPU that loads the library carries the array (256 KiB) until exit;
//   * every call uses the SAME array: it is cleared at the start of each call, and
//     two calls whose GPU work overlaps (different streams, different threads)
//     race on it -- serialise them;
//   * the first call on a device looks the array's device address up
//     (cudaGetSymbolAddress); make that first call outside any stream capture.
//
// The kernels are variants of count_units_kernel / epilogue_kernel
// (bnl_discrete_kernels.cuh): same row loop, same statistic, one unit, the unit
// descriptor passed by value, columns given as separate device pointers.
#include <stdarg.h>
#include <stdio.h>
#include <string.h>

#include <mutex>

#include "bnl_internal.h"
#include "bnl_special.cuh"

#define BNL_NOALLOC_MAX_CELLS 65536
#define NOALLOC_SMEM_CELLS 8192
#define NOALLOC_MAX_DEVICES 64

// the scratch: one count table. In HOST code this name is only a stand-in for the
// device object -- its device address comes from cudaGetSymbolAddress below.
static __device__ __align__(256) int g_noalloc_table[BNL_NOALLOC_MAX_CELLS];

namespace {

struct NoallocUnit {
  const uint8_t *col[BNL_MAXK];   // y, x, z_1, ... (device pointers, 16-byte aligned)
  int mul[BNL_MAXK];
  int ncols, cells, llx, lly, llz;
};

__global__ void __launch_bounds__(256)
noalloc_count_kernel(NoallocUnit U, long long nobs, int *__restrict__ table) {
  __shared__ int hist[NOALLOC_SMEM_CELLS];
  const bool in_smem = U.cells <= NOALLOC_SMEM_CELLS;
  int *tab = in_smem ? hist : table;
  if (in_smem) {
    for (int c = threadIdx.x; c < U.cells; c += blockDim.x) hist[c] = 0;
    __syncthreads();
  }
  const long long full = nobs / 16;
  for (long long g = blockIdx.x * (long long)blockDim.x + threadIdx.x; g < full;
       g += (long long)gridDim.x * blockDim.x) {
    int idx[16];
    unsigned na[4] = {0u, 0u, 0u, 0u};
#pragma unroll
    for (int j = 0; j < 16; j++) idx[j] = 0;
    for (int k = 0; k < U.ncols; k++) {
      const uint4 w = __ldg(reinterpret_cast<const uint4 *>(U.col[k]) + g);
      const int m = U.mul[k];
      const unsigned ww[4] = {w.x, w.y, w.z, w.w};
#pragma unroll
      for (int q = 0; q < 4; q++) {
        na[q] |= ww[q] & 0x80808080u;
#pragma unroll
        for (int j = 0; j < 4; j++) idx[q * 4 + j] += (int)__byte_perm(ww[q], 0, 0x4440 + j) * m;
      }
    }
#pragma unroll
    for (int q = 0; q < 4; q++)
#pragma unroll
      for (int j = 0; j < 4; j++)
        if (!(na[q] & (0x80u << (8 * j)))) atomicAdd(&tab[idx[q * 4 + j]], 1);
  }
  // rows past the last complete group of 16, one by one (no read beyond nobs).
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    for (long long r = full * 16; r < nobs; r++) {
      int cell = 0;
      bool bad = false;
      for (int k = 0; k < U.ncols; k++) {
        const unsigned b = U.col[k][r];
        bad |= b >= 0x80u;
        cell += (int)b * U.mul[k];
      }
      if (!bad) atomicAdd(&tab[cell], 1);
    }
  }
  if (in_smem) {
    __syncthreads();
    for (int c = threadIdx.x; c < U.cells; c += blockDim.x) {
      const int v = hist[c];
      if (v) atomicAdd(&table[c], v);
    }
  }
}

// one CTA; names the scratch array directly (device code may).
__global__ void __launch_bounds__(128)
noalloc_epilogue_kernel(NoallocUnit U, int test, double *__restrict__ out) {
  __shared__ int s_marg[4][2 * BNL_MAX_LEVELS + 2];
  __shared__ double s_s[4], s_c[4], s_df[4];
  __shared__ long long s_n[4];
  const int *T = g_noalloc_table;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int llx = U.llx, lly = U.lly, llz = U.llz, sl = llx * lly;
  int *marg = s_marg[warp];
  bnl::KahanSum acc;
  double dfacc = 0.0;
  long long nobs = 0;
  for (int k = warp; k < llz; k += 4) {
    const int *S = T + (size_t)k * sl;
    for (int i = lane; i < llx + lly; i += 32) marg[i] = 0;
    __syncwarp();
    long long nk = 0;
    for (int c = lane; c < sl; c += 32) {
      const int n = S[c];
      if (n) { atomicAdd(&marg[c / lly], n); atomicAdd(&marg[llx + c % lly], n); nk += n; }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) nk += __shfl_xor_sync(0xffffffffu, nk, o);
    __syncwarp();
    if (test == BNL_MI_ADF || test == BNL_X2_ADF) {
      int alx = 0, aly = 0;
      for (int i = lane; i < llx; i += 32) alx += marg[i] > 0;
      for (int j = lane; j < lly; j += 32) aly += marg[llx + j] > 0;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        alx += __shfl_xor_sync(0xffffffffu, alx, o);
        aly += __shfl_xor_sync(0xffffffffu, aly, o);
      }
      if (alx < 1) alx = 1;
      if (aly < 1) aly = 1;
      if (lane == 0) dfacc += (double)(alx - 1) * (double)(aly - 1);
    }
    if (lane == 0) nobs += nk;
    if (nk > 0) {
      const double dnk = (double)nk;
      for (int c = lane; c < sl; c += 32) {
        const double ni = (double)marg[c / lly], nj = (double)marg[llx + c % lly];
        if (test == BNL_MI || test == BNL_MI_ADF) {
          const int n = S[c];
          if (n) acc.add((double)n * log((double)n * dnk / (ni * nj)));
        } else {
          const double e = ni * nj / dnk;
          if (e != 0.0) { const double d = (double)S[c] - e; acc.add(d * d / e); }
        }
      }
    }
    __syncwarp();
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const double os = __shfl_xor_sync(0xffffffffu, acc.s, o);
    const double oc = __shfl_xor_sync(0xffffffffu, acc.c, o);
    acc.merge(os, oc);
    dfacc += __shfl_xor_sync(0xffffffffu, dfacc, o);
    nobs += __shfl_xor_sync(0xffffffffu, nobs, o);
  }
  if (lane == 0) { s_s[warp] = acc.s; s_c[warp] = acc.c; s_df[warp] = dfacc; s_n[warp] = nobs; }
  __syncthreads();
  if (threadIdx.x != 0) return;
  bnl::KahanSum tot;
  double df = 0.0;
  long long n = 0;
  for (int w = 0; w < 4; w++) { tot.merge(s_s[w], s_c[w]); df += s_df[w]; n += s_n[w]; }
  const bool adf = test == BNL_MI_ADF || test == BNL_X2_ADF;
  if (!adf) df = (double)(llx - 1) * (double)(lly - 1) * (double)llz;
  double stat = 0.0;
  if (n > 0 && !(adf && (double)n < 5.0 * (double)llx * (double)lly * (double)llz)) {
    if (test == BNL_MI || test == BNL_MI_ADF) stat = tot.value() / (double)n * (2.0 * (double)n);
    else stat = tot.value();
  }
  out[0] = stat; out[1] = df; out[2] = bnl::pchisq_upper(stat, df); out[3] = (double)n;
}

std::mutex g_setup_mutex;
int *g_table_ptr[NOALLOC_MAX_DEVICES];      // device address of g_noalloc_table, per device
int g_sm_count[NOALLOC_MAX_DEVICES];

int fail(char *errbuf, int errlen, int code, const char *fmt, ...) {
  if (errbuf && errlen > 0) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(errbuf, (size_t)errlen, fmt, ap);
    va_end(ap);
  }
  return code;
}

}  // namespace

extern "C" int bnl_ci_discrete_noalloc(int test_e, int k, const uint8_t *const *d_cols,
                                       const int *nlevels, int64_t nobs, double *d_out,
                                       void *stream, char *errbuf, int errlen) {
  if (test_e != BNL_MI && test_e != BNL_MI_ADF && test_e != BNL_X2 && test_e != BNL_X2_ADF)
    return fail(errbuf, errlen, BNL_EUNSUPPORTED, "test %d is not a discrete asymptotic test", test_e);
  if (k < 2 || k > BNL_MAXK || !d_cols || !nlevels || !d_out || nobs < 0)
    return fail(errbuf, errlen, BNL_EINVAL, "bad arguments (need 2..%d columns: x, y, z...)", BNL_MAXK);

  // the unit: cell = y + lly * (x + llx * z), z = first conditioning column fastest.
  NoallocUnit U;
  memset(&U, 0, sizeof U);
  long long mul = 1, llz = 1;
  const int order[2] = {1, 0};                   // y first, then x
  for (int c = 0; c < k; c++) {
    const int src = c < 2 ? order[c] : c;
    if (!d_cols[src] || ((uintptr_t)d_cols[src] & 15u))
      return fail(errbuf, errlen, BNL_EINVAL, "column %d: null or not 16-byte aligned", src);
    if (nlevels[src] < 1 || nlevels[src] > BNL_MAX_LEVELS)
      return fail(errbuf, errlen, BNL_EINVAL, "column %d: levels must be in 1..%d", src, BNL_MAX_LEVELS);
    U.col[c] = d_cols[src];
    U.mul[c] = (int)mul;
    mul *= nlevels[src];
    if (c >= 2) llz *= nlevels[src];
    if (mul > BNL_NOALLOC_MAX_CELLS)
      return fail(errbuf, errlen, BNL_EUNSUPPORTED,
                  "the table needs more than %d cells, the size of the built-in scratch array; "
                  "use bnl_ci_discrete() for this test", BNL_NOALLOC_MAX_CELLS);
  }
  U.ncols = k; U.cells = (int)mul; U.llx = nlevels[0]; U.lly = nlevels[1]; U.llz = (int)llz;

  // once per device, with that device current: where does the scratch array live?
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return fail(errbuf, errlen, BNL_ECUDA, "cudaGetDevice: %s", cudaGetErrorString(e));
  if (dev < 0 || dev >= NOALLOC_MAX_DEVICES)
    return fail(errbuf, errlen, BNL_EUNSUPPORTED, "device ordinal %d not supported", dev);
  int *table = nullptr;
  int sms = 0;
  {
    std::lock_guard<std::mutex> lock(g_setup_mutex);
    if (!g_table_ptr[dev]) {
      int *p = nullptr;
      e = cudaGetSymbolAddress((void **)&p, g_noalloc_table);
      if (e != cudaSuccess || !p)
        return fail(errbuf, errlen, BNL_ECUDA, "cudaGetSymbolAddress: %s", cudaGetErrorString(e));
      e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
      if (e != cudaSuccess)
        return fail(errbuf, errlen, BNL_ECUDA, "cudaDeviceGetAttribute: %s", cudaGetErrorString(e));
      g_sm_count[dev] = sms;
      g_table_ptr[dev] = p;
    }
    table = g_table_ptr[dev];
    sms = g_sm_count[dev];
  }

  cudaStream_t st = (cudaStream_t)stream;
  // the array keeps the previous call's counts: clear the part this call uses.
  e = cudaMemsetAsync(table, 0, (size_t)U.cells * sizeof(int), st);
  if (e != cudaSuccess) return fail(errbuf, errlen, BNL_ECUDA, "cudaMemsetAsync: %s", cudaGetErrorString(e));
  const long long groups = nobs / 16;
  long long grid = (groups + 255) / 256;
  if (grid > 4ll * sms) grid = 4ll * sms;
  if (grid < 1) grid = 1;
  noalloc_count_kernel<<<(unsigned)grid, 256, 0, st>>>(U, nobs, table);
  noalloc_epilogue_kernel<<<1, 128, 0, st>>>(U, test_e, d_out);
  e = cudaGetLastError();
  if (e != cudaSuccess) return fail(errbuf, errlen, BNL_ECUDA, "kernel launch: %s", cudaGetErrorString(e));
  return BNL_OK;
}
