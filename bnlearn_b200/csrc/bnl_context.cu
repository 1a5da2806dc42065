// bnl_context.cu -- context, device resources and data-set registration
// (packing R's int32 factor columns into the uint8 HBM layout).
#include <stdarg.h>
#include <stdlib.h>
#include <string.h>

#include <atomic>
#include <chrono>
#include <thread>

#include <math.h>

#include "bnl_internal.h"
#include "bnl_special.cuh"

int bnl_fail(bnl_ctx *ctx, int code, const char *fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof buf, fmt, ap);
  va_end(ap);
  if (ctx) {
    std::lock_guard<std::mutex> lock(ctx->err_mutex);
    ctx->err = buf;
  }
  return code;
}

// ---------------------------------------------------------------------------
// pack_factors: int32 1-based codes (NA = INT_MIN) -> uint8 (code - 1, NA = 0xFF).
// One thread converts 4 rows: one 128-bit load, one 32-bit store, both coalesced.
// Codes outside [1, nlevels] raise the error flag (the reference would index
// out of bounds there; we refuse the column instead).
__global__ void __launch_bounds__(256)
pack_factors_kernel(const int *__restrict__ src, uint8_t *__restrict__ dst, long long nobs,
                    int nlevels, int *__restrict__ bad, int *__restrict__ has_na) {
  const long long quads = (nobs + 3) / 4;
  for (long long q = blockIdx.x * (long long)blockDim.x + threadIdx.x; q < quads;
       q += (long long)gridDim.x * blockDim.x) {
    const long long r = q * 4;
    int v[4];
    if (r + 3 < nobs) {
      const int4 w = *reinterpret_cast<const int4 *>(src + r);
      v[0] = w.x; v[1] = w.y; v[2] = w.z; v[3] = w.w;
    } else {
      for (int j = 0; j < 4; j++) v[j] = (r + j < nobs) ? src[r + j] : INT_MIN;
    }
    unsigned out = 0;
    for (int j = 0; j < 4; j++) {
      unsigned b;
      if (v[j] == INT_MIN) { b = BNL_NA_U8; if (r + j < nobs) *has_na = 1; }
      else if (v[j] < 1 || v[j] > nlevels) { b = BNL_NA_U8; *bad = 1; }
      else b = (unsigned)(v[j] - 1);
      out |= b << (8 * j);
    }
    *reinterpret_cast<unsigned *>(dst + r) = out;   // stride is padded: always in bounds
  }
}

// Factors with 128..32767 levels: int32 codes -> uint16 (code - 1, NA -> 0xFFFF).
__global__ void __launch_bounds__(256)
pack_factors16_kernel(const int *__restrict__ src, uint16_t *__restrict__ dst, long long nobs, int nlevels,
                      int *__restrict__ bad, int *__restrict__ has_na) {
  for (long long r = blockIdx.x * (long long)blockDim.x + threadIdx.x; r < nobs;
       r += (long long)gridDim.x * blockDim.x) {
    const int v = src[r];
    unsigned b;
    if (v == INT_MIN) { b = BNL_NA_U16; *has_na = 1; }
    else if (v < 1 || v > nlevels) { b = BNL_NA_U16; *bad = 1; }
    else b = (unsigned)(v - 1);
    dst[r] = (uint16_t)b;
  }
}

// Row gather for bnl_data_resample: out[c][r] = in[c][rows[r]] for every packed
// factor column c (blockIdx.y). Reads are random bytes, writes coalesced. A row
// index outside the source raises flags[0]; an NA that lands in the replicate
// raises the column's flag so that the NA-free fast paths stay valid.
__global__ void __launch_bounds__(256)
gather_u8_kernel(const uint8_t *__restrict__ src, size_t src_stride, long long src_nobs,
                 uint8_t *__restrict__ dst, size_t dst_stride, const int *__restrict__ rows,
                 long long nrows, int *__restrict__ flags) {
  const int c = blockIdx.y;
  const uint8_t *in = src + src_stride * c;
  uint8_t *out = dst + dst_stride * c;
  int saw_na = 0, bad = 0;
  for (long long r = blockIdx.x * (long long)blockDim.x + threadIdx.x; r < nrows;
       r += (long long)gridDim.x * blockDim.x) {
    const int k = rows[r];
    if (k < 0 || k >= src_nobs) { bad = 1; continue; }
    const uint8_t v = in[k];
    saw_na |= v == BNL_NA_U8;
    out[r] = v;
  }
  if (bad) flags[0] = 1;
  if (saw_na) flags[1 + c] = 1;
}

__global__ void __launch_bounds__(256)
gather_u16_kernel(const uint16_t *__restrict__ src, size_t src_stride, long long src_nobs,
                  uint16_t *__restrict__ dst, size_t dst_stride, const int *__restrict__ rows,
                  long long nrows, int *__restrict__ flags, int flag_base) {
  const int c = blockIdx.y;
  int saw_na = 0;
  for (long long r = blockIdx.x * (long long)blockDim.x + threadIdx.x; r < nrows;
       r += (long long)gridDim.x * blockDim.x) {
    const int k = rows[r];
    if (k < 0 || k >= src_nobs) { flags[0] = 1; continue; }
    const uint16_t v = src[src_stride * c + k];
    saw_na |= v == BNL_NA_U16;
    dst[dst_stride * c + r] = v;
  }
  if (saw_na) flags[flag_base + c] = 1;
}

__global__ void __launch_bounds__(256)
gather_f64_kernel(const double *__restrict__ src, size_t src_stride, long long src_nobs,
                  double *__restrict__ dst, size_t dst_stride, const int *__restrict__ rows,
                  long long nrows, int *__restrict__ flags) {
  const int c = blockIdx.y;
  for (long long r = blockIdx.x * (long long)blockDim.x + threadIdx.x; r < nrows;
       r += (long long)gridDim.x * blockDim.x) {
    const int k = rows[r];
    if (k < 0 || k >= src_nobs) { flags[0] = 1; continue; }
    dst[dst_stride * c + r] = src[src_stride * c + k];
  }
}

// bnl_data_commit: the packed columns were written by somebody else (a broadcast
// from another GPU): find out which columns hold an NA (0xFF within the rows) and
// refuse codes that are neither NA nor below the column's number of levels.
__global__ void __launch_bounds__(256)
scan_packed_kernel(const uint8_t *__restrict__ u8, size_t stride, long long nobs,
                   const int *__restrict__ nlev_by_slot, int *__restrict__ flags) {
  const int c = blockIdx.y;
  const uint8_t *col = u8 + stride * c;
  const unsigned lv = (unsigned)nlev_by_slot[c];
  int na = 0, bad = 0;
  for (long long r = blockIdx.x * (long long)blockDim.x + threadIdx.x; r < nobs;
       r += (long long)gridDim.x * blockDim.x) {
    const unsigned b = col[r];
    na |= b == BNL_NA_U8;
    bad |= b != BNL_NA_U8 && b >= lv;
  }
  if (bad) flags[0] = 1;
  if (na) flags[1 + c] = 1;
}

// Host-packed nibbles -> the uint8 layout (0xF -> 0xFF). One thread: 8 bytes in, 16 out.
__global__ void __launch_bounds__(256)
unpack_nibbles_kernel(const uint8_t *__restrict__ src, uint8_t *__restrict__ dst, long long nbytes_in) {
  const long long words = (nbytes_in + 7) / 8;
  for (long long w = blockIdx.x * (long long)blockDim.x + threadIdx.x; w < words;
       w += (long long)gridDim.x * blockDim.x) {
    uint2 v = reinterpret_cast<const uint2 *>(src)[w];
    const long long left = nbytes_in - 8 * w;                  // bytes of this word that belong to the column
    if (left < 8) {                                            // the rest must unpack to the 0xFF row padding
      const unsigned long long keep = left <= 0 ? 0ull : (~0ull >> (8 * (8 - left)));
      unsigned long long x = ((unsigned long long)v.y << 32) | v.x;
      x = (x & keep) | ~keep;
      v.x = (unsigned)x; v.y = (unsigned)(x >> 32);
    }
    const unsigned in[2] = {v.x, v.y};
    unsigned out[4];
#pragma unroll
    for (int h = 0; h < 2; h++)
#pragma unroll
      for (int q = 0; q < 2; q++) {
        unsigned x = (in[h] >> (16 * q)) & 0xFFFFu;              // four nibbles
        x = (x | (x << 8)) & 0x00FF00FFu;
        x = (x | (x << 4)) & 0x0F0F0F0Fu;                        // one per byte
        const unsigned na = ((x + 0x01010101u) >> 4) & 0x01010101u;   // 0xF + 1 carries into bit 4
        out[2 * h + q] = x | (na * 0xF0u);
      }
    reinterpret_cast<uint4 *>(dst)[w] = make_uint4(out[0], out[1], out[2], out[3]);
  }
}

// ---------------------------------------------------------------------------

extern "C" int bnl_init(bnl_ctx **out, const int *devices, int ndev) {
  if (!out) return BNL_EINVAL;
  *out = nullptr;
  int count = 0;
  cudaError_t e = cudaGetDeviceCount(&count);
  bnl_ctx *ctx = new bnl_ctx();
  *out = ctx;   // returned even on failure so that bnl_last_error() can explain
  if (e != cudaSuccess || count == 0)
    return bnl_fail(ctx, BNL_ECUDA, "no CUDA device available (%s); this engine has no CPU fallback",
                    e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0");
  std::vector<int> ids;
  if (!devices || ndev <= 0) {
    int cur = 0;
    BNL_CUDA(ctx, cudaGetDevice(&cur));
    ids.push_back(cur);
  } else {
    for (int i = 0; i < ndev; i++) {
      if (devices[i] < 0 || devices[i] >= count)
        return bnl_fail(ctx, BNL_EINVAL, "device %d out of range (have %d)", devices[i], count);
      ids.push_back(devices[i]);
    }
  }
  for (int id : ids) {
    BnlDevice d;
    d.id = id;
    BNL_CUDA(ctx, cudaSetDevice(id));
    BNL_CUDA(ctx, cudaStreamCreateWithFlags(&d.stream, cudaStreamNonBlocking));
    BNL_CUDA(ctx, cudaStreamCreateWithFlags(&d.stream2, cudaStreamNonBlocking));
    BNL_CUDA(ctx, cudaEventCreateWithFlags(&d.ev_stage[0], cudaEventDisableTiming));
    BNL_CUDA(ctx, cudaEventCreateWithFlags(&d.ev_stage[1], cudaEventDisableTiming));
    BNL_CUDA(ctx, cudaEventCreate(&d.ev0));
    BNL_CUDA(ctx, cudaEventCreate(&d.ev1));
    BNL_CUDA(ctx, cudaEventCreate(&d.evc0));
    BNL_CUDA(ctx, cudaEventCreate(&d.evc1));
    BNL_CUDA(ctx, cudaEventCreate(&d.eve1));
    BNL_CUDA(ctx, cudaDeviceGetAttribute(&d.sm_count, cudaDevAttrMultiProcessorCount, id));
    BNL_CUDA(ctx, cudaDeviceGetAttribute(&d.max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, id));
    ctx->devs.push_back(d);
  }
  BNL_CUDA(ctx, cudaSetDevice(ids[0]));
  return BNL_OK;
}

// Released data-set blocks are kept (a few, bounded) and handed out again to the
// next registration of a similar size instead of going through cudaFree/cudaMalloc.
static const size_t kPoolBlocks = 6;
static const size_t kPoolBytes = 24ull << 30;

int bnl_pool_alloc(bnl_ctx *ctx, BnlDevice &dev, void **p, size_t bytes) {
  *p = nullptr;
  if (bytes == 0) return BNL_OK;
  int best = -1;
  for (size_t i = 0; i < dev.pool.size(); i++)
    if (dev.pool[i].second >= bytes && dev.pool[i].second <= bytes + bytes / 4 + 4096 &&
        (best < 0 || dev.pool[i].second < dev.pool[best].second))
      best = (int)i;
  if (best >= 0) {
    *p = dev.pool[best].first;
    dev.pool.erase(dev.pool.begin() + best);
    return BNL_OK;
  }
  cudaError_t e = cudaMalloc(p, bytes);
  if (e != cudaSuccess) {
    // give cached blocks back and retry once.
    for (auto &b : dev.pool) cudaFree(b.first);
    dev.pool.clear();
    e = cudaMalloc(p, bytes);
  }
  if (e != cudaSuccess)
    return bnl_fail(ctx, BNL_ENOMEM, "cudaMalloc of %zu bytes failed: %s", bytes, cudaGetErrorString(e));
  return BNL_OK;
}

void bnl_pool_free(BnlDevice &dev, void *p, size_t bytes) {
  if (!p) return;
  dev.pool.push_back(std::make_pair(p, bytes));
  size_t total = 0;
  for (auto &b : dev.pool) total += b.second;
  while (dev.pool.size() > kPoolBlocks || total > kPoolBytes) {
    total -= dev.pool.front().second;
    cudaFree(dev.pool.front().first);
    dev.pool.erase(dev.pool.begin());
  }
}

// log C(n, k) of the multinomial NML normaliser for n <= 1000 observations and
// k <= 100 categories (get_regret_table, src/scores/nml.regret.table.c:44-69):
// the k = 2 column by direct summation, the others by the linear recurrence
// C(n, k) = C(n, k - 1) + n / (k - 2) C(n, k - 2). Entry (k, n) at k + n (K + 1).
int bnl_regret_table(bnl_ctx *ctx, BnlDevice &dev) {
  if (dev.d_regret) return BNL_OK;
  const int N = BNL_REGRET_N, K = BNL_REGRET_K;
  std::vector<double> logs(N + 1, 0.0), logfacs(N + 1, 0.0), tab((size_t)(N + 1) * (K + 1), 0.0);
  for (int n = 1; n <= N; n++) { logs[n] = log((double)n); logfacs[n] = bnl::lgamma_pos((double)n + 1.0); }
  for (int n = 1; n <= N; n++) {
    double res = 0.0;
    for (int m = 0; m <= n; m++) {
      double term = logfacs[n] - logfacs[m] - logfacs[n - m];
      if (m) term += m * (logs[m] - logs[n]);
      if (n - m) term += (n - m) * (logs[n - m] - logs[n]);
      res += exp(term);
    }
    tab[2 + (size_t)n * (K + 1)] = log(res);
    double pp = 1.0, prev = exp(tab[2 + (size_t)n * (K + 1)]);      // read back, as fill_regrets_up_to_K does
    for (int k = 3; k <= K; k++) {
      const double rk = prev + pp / (double)(k - 2) * (double)n;
      tab[k + (size_t)n * (K + 1)] = log(rk);
      pp = prev; prev = rk;
    }
  }
  BNL_CUDA(ctx, cudaSetDevice(dev.id));
  BNL_CUDA(ctx, cudaMalloc((void **)&dev.d_regret, tab.size() * sizeof(double)));
  // on the device's own stream (the kernels run there and it does not wait for
  // the legacy default stream), and complete before the host vector goes away.
  BNL_CUDA(ctx, cudaMemcpyAsync(dev.d_regret, tab.data(), tab.size() * sizeof(double), cudaMemcpyHostToDevice,
                                dev.stream));
  BNL_CUDA(ctx, cudaStreamSynchronize(dev.stream));
  return BNL_OK;
}

static void free_dataset(bnl_ctx *ctx, BnlDataSet &ds) {
  for (size_t i = 0; i < ds.dev.size(); i++) {
    cudaSetDevice(ctx->devs[i].id);
    BnlDeviceData &d = ds.dev[i];
    bnl_pool_free(ctx->devs[i], d.u8, d.u8_bytes);
    bnl_pool_free(ctx->devs[i], d.u4, d.u4_bytes);
    d.u4 = nullptr; d.u4_bytes = 0;
    bnl_pool_free(ctx->devs[i], d.u16, d.u16_bytes);
    bnl_pool_free(ctx->devs[i], d.f64, d.f64_bytes);
    bnl_pool_free(ctx->devs[i], d.mean, d.mean_bytes);
    bnl_pool_free(ctx->devs[i], d.cov, d.cov_bytes);
    bnl_pool_free(ctx->devs[i], d.bge_nu, d.nu_bytes);
    d = BnlDeviceData();
  }
  ds.dev.clear();
}

BnlTraceSlot g_bnl_trace[32];
int g_bnl_trace_on = -1;
static double trace_now() {
  return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}
int bnl_trace_slot(const char *label) {
  for (int i = 0; i < 32; i++) {
    if (g_bnl_trace[i].label == label) return i;
    if (!g_bnl_trace[i].label) { g_bnl_trace[i].label = label; return i; }
  }
  return 31;
}
BnlTraceScope::BnlTraceScope(const char *label) : slot(-1), t0(0.0) {
  if (g_bnl_trace_on < 0) { const char *v = getenv("BNL_TRACE"); g_bnl_trace_on = (v && *v && atoi(v)) ? 1 : 0; }
  if (g_bnl_trace_on) { slot = bnl_trace_slot(label); t0 = trace_now(); }
}
BnlTraceScope::~BnlTraceScope() {
  if (slot >= 0) { g_bnl_trace[slot].seconds += trace_now() - t0; g_bnl_trace[slot].calls++; }
}

extern "C" void bnl_shutdown(bnl_ctx *ctx) {
  if (!ctx) return;
  if (g_bnl_trace_on > 0)
    for (int i = 0; i < 32 && g_bnl_trace[i].label; i++)
      fprintf(stderr, "[bnl trace] %-28s %10.3f ms  %8lld calls\n", g_bnl_trace[i].label,
              g_bnl_trace[i].seconds * 1e3, g_bnl_trace[i].calls);
  for (auto &kv : ctx->data) free_dataset(ctx, kv.second);
  for (auto &d : ctx->devs) {
    cudaSetDevice(d.id);
    for (auto &b : d.pool) cudaFree(b.first);
    d.pool.clear();
    cudaFree(d.stage); cudaFree(d.d_flags); cudaFree(d.d_regret);
    cudaFreeHost(d.h_pack);
    cudaFreeHost(d.h_nstage);
    for (auto &e : d.ev_nstage) cudaEventDestroy(e);
    if (d.ev_stage[0]) cudaEventDestroy(d.ev_stage[0]);
    if (d.ev_stage[1]) cudaEventDestroy(d.ev_stage[1]);
    if (d.stream2) cudaStreamDestroy(d.stream2);
    cudaFree(d.d_units); cudaFree(d.d_tables); cudaFree(d.d_out); cudaFree(d.d_gidx); cudaFree(d.d_order); cudaFree(d.d_tiles);
    cudaFree(d.d_basecols); cudaFree(d.d_runfirst); cudaFreeHost(d.h_runfirst);
    cudaFree(d.d_pairG); cudaFree(d.d_pairmeta);
    cudaFreeHost(d.h_units); cudaFreeHost(d.h_out); cudaFreeHost(d.h_tables); cudaFreeHost(d.h_pairmeta); cudaFreeHost(d.h_gidx); cudaFreeHost(d.h_order); cudaFreeHost(d.h_tiles);
    if (d.ev0) cudaEventDestroy(d.ev0);
    if (d.ev1) cudaEventDestroy(d.ev1);
    if (d.evc0) cudaEventDestroy(d.evc0);
    if (d.evc1) cudaEventDestroy(d.evc1);
    if (d.eve1) cudaEventDestroy(d.eve1);
    if (d.stream) cudaStreamDestroy(d.stream);
  }
  delete ctx;
}

extern "C" const char *bnl_last_error(const bnl_ctx *ctx) {
  return ctx ? ctx->err.c_str() : "null context";
}

extern "C" int bnl_last_call_profile(const bnl_ctx *ctx, double *total_ms, double *count_ms,
                                     double *epilogue_ms, int *kernel_launches) {
  if (!ctx) return BNL_EINVAL;
  if (total_ms) *total_ms = ctx->last_ms;
  if (count_ms) *count_ms = ctx->last_count_ms;
  if (epilogue_ms) *epilogue_ms = ctx->last_epi_ms;
  if (kernel_launches) *kernel_launches = ctx->last_launches;
  return BNL_OK;
}

static size_t round_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

extern "C" void bnl_host_pack_slice(const int *src, uint8_t *dst, long long n, int nlevels, int *bad,
                                    int *has_na);
extern "C" void bnl_host_pack_slice4(const int *src, uint8_t *dst, long long n, int nlevels, int *bad,
                                     int *has_na);

// Upload of the factor columns of a LARGE data frame. Two routes share the work
// and the PCIe link:
//   raw route   int32 column -> device staging (2 buffers) -> pack_factors_kernel
//   host route  host threads pack 1M-row blocks of a column into a pinned ring;
//               whoever finishes a column's last block sends its uint8 copy
// Columns are claimed dynamically: the host route walks the columns upwards, the
// raw route downwards, until they meet, so the split adapts to the box (PCIe
// rate, core count, pageable vs pinned source). 4x fewer bytes cross the link for
// host-packed columns. dcols[s] is the data-frame column of packed slot s.
static int upload_factors_hybrid(bnl_ctx *ctx, BnlDataSet &ds, BnlDevice &d0, BnlDeviceData &dd0,
                                 const void *const *col, const int *nlevels,
                                 const std::vector<int> &dcols, int *d_flags, std::vector<int> &host_flags) {
  const long long nobs = ds.nobs;
  const int nd = ds.nd;
  const long long kBlock = 1 << 20;
  const int nblocks = (int)((nobs + kBlock - 1) / kBlock);
  const int kSlots = 6;
  unsigned hw = std::thread::hardware_concurrency();
  int nthreads = (int)(hw ? hw : 8);
  if (const char *v = getenv("BNL_PACK_THREADS")) if (*v) nthreads = atoi(v);
  nthreads = std::max(0, std::min(nthreads, 32) - 1);           // the caller's thread feeds the raw route

  int rc;
  const size_t stage_ints = ((size_t)nobs + 4 + 63) & ~(size_t)63;     // keeps both buffers 16-byte aligned
  if ((rc = bnl_reserve_device(ctx, &d0.stage, &d0.stage_cap, 2 * stage_ints))) return rc;
  if (nthreads > 0 && (rc = bnl_reserve_pinned(ctx, &d0.h_pack, &d0.h_pack_cap, (size_t)kSlots * ds.stride_u8)))
    return rc;

  // the device buffer was filled with 0xFF on the main stream: order stream2 after it.
  BNL_CUDA(ctx, cudaEventRecord(d0.ev_stage[0], d0.stream));
  BNL_CUDA(ctx, cudaStreamWaitEvent(d0.stream2, d0.ev_stage[0], 0));

  // nibble route (BNL_PACK_NIBBLES=0 turns it off): device staging for kSlots half-size columns.
  // 1M-row blocks keep the nibble pairs of a column aligned (kBlock is even).
  bool use_nibbles = true;
  if (const char *v = getenv("BNL_PACK_NIBBLES")) if (*v) use_nibbles = atoi(v) != 0;
  const size_t nib_stride = ((size_t)(nobs + 1) / 2 + 255 + 16) & ~(size_t)255;
  uint8_t *d_nib = nullptr;
  if (use_nibbles && nthreads > 0) {
    if ((rc = bnl_pool_alloc(ctx, d0, (void **)&d_nib, (size_t)kSlots * nib_stride))) return rc;
  } else use_nibbles = false;
  std::mutex enqueue;
  // columns [0, cpu_next) belong to the host route, [gpu_next, nd) to the raw route.
  std::mutex claim;
  int cpu_next = 0, cpu_cur = -1, cpu_blk = 0, gpu_next = nd;
  std::vector<std::atomic<int>> done(nd), enq(nd);
  for (int i = 0; i < nd; i++) { done[i] = 0; enq[i] = 0; }
  std::vector<cudaEvent_t> ev(nd, nullptr);
  std::atomic<int> failed(0);
  host_flags.assign(1 + nd, 0);

  auto claim_cpu = [&](int *c, int *b) {
    std::lock_guard<std::mutex> lock(claim);
    if (cpu_cur < 0 || cpu_blk >= nblocks) {
      if (cpu_next >= gpu_next) return false;
      cpu_cur = cpu_next++;
      cpu_blk = 0;
    }
    *c = cpu_cur; *b = cpu_blk++;
    return true;
  };
  auto claim_raw = [&]() {
    std::lock_guard<std::mutex> lock(claim);
    return gpu_next - 1 < cpu_next ? -1 : --gpu_next;
  };

  auto packer = [&]() {
    cudaSetDevice(d0.id);
    int c, b;
    while (!failed.load() && claim_cpu(&c, &b)) {
      const int slot = c % kSlots;
      if (c >= kSlots) {                                        // the ring slot must have left the host
        while (!enq[c - kSlots].load() && !failed.load()) std::this_thread::yield();
        if (failed.load()) break;
        if (cudaEventSynchronize(ev[c - kSlots]) != cudaSuccess) { failed = 1; break; }
      }
      const long long r0 = (long long)b * kBlock, r1 = std::min(nobs, r0 + kBlock);
      const int dc = dcols[c];
      int bad = 0, na = 0;
      const bool nib = use_nibbles && nlevels[dc] <= 15;          // two codes per byte through the ring
      if (nib)
        bnl_host_pack_slice4(static_cast<const int *>(col[dc]) + r0, d0.h_pack + (size_t)slot * ds.stride_u8 + r0 / 2,
                             r1 - r0, nlevels[dc], &bad, &na);
      else
        bnl_host_pack_slice(static_cast<const int *>(col[dc]) + r0, d0.h_pack + (size_t)slot * ds.stride_u8 + r0,
                            r1 - r0, nlevels[dc], &bad, &na);
      if (bad) host_flags[0] = 1;
      if (na) host_flags[1 + c] = 1;
      if (done[c].fetch_add(1) + 1 == nblocks) {
        cudaError_t e = cudaEventCreateWithFlags(&ev[c], cudaEventDisableTiming);
        if (e == cudaSuccess && nib) {
          // nibbles go to a device staging slot; the ring slot is free once they have left the host
          // (event), the unpack kernel follows on the same stream before the slot's next copy
          const size_t nb = (size_t)((nobs + 1) / 2);
          uint8_t *dst4 = d_nib + (size_t)slot * nib_stride;
          std::lock_guard<std::mutex> lock(enqueue);             // copy + event + kernel stay together on stream2
          e = cudaMemcpyAsync(dst4, d0.h_pack + (size_t)slot * ds.stride_u8, nb, cudaMemcpyHostToDevice, d0.stream2);
          if (e == cudaSuccess) e = cudaEventRecord(ev[c], d0.stream2);
          if (e == cudaSuccess) {
            const long long words = ((long long)nb + 7) / 8;
            const int grid = (int)std::min<long long>((words + 255) / 256, (long long)d0.sm_count * 8);
            unpack_nibbles_kernel<<<grid, 256, 0, d0.stream2>>>(dst4, dd0.u8 + ds.stride_u8 * c, (long long)nb);
            e = cudaGetLastError();
          }
        } else if (e == cudaSuccess) {
          std::lock_guard<std::mutex> lock(enqueue);
          e = cudaMemcpyAsync(dd0.u8 + ds.stride_u8 * c, d0.h_pack + (size_t)slot * ds.stride_u8, (size_t)nobs,
                              cudaMemcpyHostToDevice, d0.stream2);
          if (e == cudaSuccess) e = cudaEventRecord(ev[c], d0.stream2);
        }
        if (e != cudaSuccess) failed = 1;
        enq[c] = 1;
      }
    }
  };
  std::vector<std::thread> workers;
  for (int t = 0; t < nthreads; t++) workers.emplace_back(packer);

  // raw route, on the caller's thread -- for sources in PINNED memory only. From pageable
  // memory (what R hands over) a 40 MB column crawls through the driver's bounce buffer at
  // 8-10 GB/s and holds the copy engine meanwhile (config B: 48 ms per registration, against
  // 16 ms from pinned memory); there the caller's thread packs like the others.
  bool raw_ok = true;
  {
    cudaPointerAttributes at;
    const void *probe = col[dcols[nd - 1]];
    raw_ok = cudaPointerGetAttributes(&at, probe) == cudaSuccess &&
             (at.type == cudaMemoryTypeHost || at.type == cudaMemoryTypeManaged);
    cudaGetLastError();                       // an unregistered pointer may leave a sticky-looking error code behind
    // With two codes per byte through the ring and eight or more packing threads the raw route only
    // takes memory bandwidth and PCIe time away from them (config B, 16 cores: 16.4 ms with it, 15.7
    // without); on a box with few cores the copy engine's share is worth having.
    int nib_cols = 0;
    for (int c = 0; c < nd; c++) nib_cols += nlevels[dcols[c]] <= 15;
    if (use_nibbles && nthreads >= 8 && 2 * nib_cols >= nd) raw_ok = false;
    if (const char *v = getenv("BNL_RAW_ROUTE")) if (*v) raw_ok = atoi(v) != 0;
    if (nthreads == 0) raw_ok = true;         // nobody else would move the data
  }
  if (!raw_ok) packer();
  int fed = 0, c;
  cudaError_t ge = cudaSuccess;
  while (raw_ok && ge == cudaSuccess && !failed.load() && (c = claim_raw()) >= 0) {
    const int k = fed & 1;
    if (fed >= 2) ge = cudaEventSynchronize(d0.ev_stage[k]);
    int *st = d0.stage + (size_t)k * stage_ints;
    const int dc = dcols[c];
    if (ge == cudaSuccess)
      ge = cudaMemcpyAsync(st, col[dc], (size_t)nobs * sizeof(int), cudaMemcpyHostToDevice, d0.stream);
    if (ge == cudaSuccess) {
      long long quads = (nobs + 3) / 4;
      int grid = (int)std::min<long long>((quads + 255) / 256, (long long)d0.sm_count * 16);
      pack_factors_kernel<<<grid, 256, 0, d0.stream>>>(st, dd0.u8 + ds.stride_u8 * c, nobs, nlevels[dc], d_flags,
                                                       d_flags + 1 + c);
      ge = cudaGetLastError();
    }
    if (ge == cudaSuccess) ge = cudaEventRecord(d0.ev_stage[k], d0.stream);
    fed++;
  }
  if (ge != cudaSuccess) failed = 1;
  for (auto &w : workers) w.join();
  cudaError_t e2 = cudaStreamSynchronize(d0.stream2);
  if (d_nib) bnl_pool_free(d0, d_nib, (size_t)kSlots * nib_stride);
  for (cudaEvent_t e : ev) if (e) cudaEventDestroy(e);
  if (failed.load() || e2 != cudaSuccess)
    return bnl_fail(ctx, BNL_ECUDA, "uploading the factor columns failed: %s",
                    cudaGetErrorString(ge != cudaSuccess ? ge : (e2 != cudaSuccess ? e2 : cudaGetLastError())));
  return BNL_OK;
}

// Numeric columns of a large data frame that sits in PAGEABLE host memory (what R
// hands over): cudaMemcpyAsync would bounce them through the driver's own small
// staging buffer at 8-10 GB/s. Here host threads copy 4 MiB chunks into pinned
// buffers of their own (two each, so the copy of one chunk overlaps the DMA of the
// previous) and queue the transfers on stream2. Columns already in pinned memory go
// straight through.
static int upload_numeric_staged(bnl_ctx *ctx, BnlDataSet &ds, BnlDevice &d0, BnlDeviceData &dd0,
                                 const void *const *col, const std::vector<int> &ncols_idx) {
  const size_t kChunk = 4u << 20;
  unsigned hw = std::thread::hardware_concurrency();
  int nthreads = (int)(hw ? hw : 8);
  if (const char *v = getenv("BNL_PACK_THREADS")) if (*v) nthreads = atoi(v);
  nthreads = std::max(1, std::min(nthreads, 16));
  const size_t col_bytes = (size_t)ds.nobs * sizeof(double);
  const long long chunks_per_col = (long long)((col_bytes + kChunk - 1) / kChunk);
  const long long total = chunks_per_col * (long long)ncols_idx.size();
  nthreads = (int)std::min<long long>(nthreads, total);
  int rc;
  if ((rc = bnl_reserve_pinned(ctx, &d0.h_nstage, &d0.h_nstage_cap, (size_t)2 * nthreads * kChunk))) return rc;
  while ((int)d0.ev_nstage.size() < 2 * nthreads) {
    cudaEvent_t e;
    BNL_CUDA(ctx, cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    d0.ev_nstage.push_back(e);
  }
  // stream2 must see the zero fill of the block queued on the main stream.
  BNL_CUDA(ctx, cudaEventRecord(d0.ev_stage[0], d0.stream));
  BNL_CUDA(ctx, cudaStreamWaitEvent(d0.stream2, d0.ev_stage[0], 0));
  std::atomic<long long> next(0);
  std::atomic<int> failed(0);
  auto worker = [&](int t) {
    cudaSetDevice(d0.id);
    long long k;
    int used = 0;
    while (!failed.load() && (k = next.fetch_add(1)) < total) {
      const int c = ncols_idx[(size_t)(k / chunks_per_col)];
      const size_t o = (size_t)(k % chunks_per_col) * kChunk;
      const size_t len = std::min(kChunk, col_bytes - o);
      const int slot = 2 * t + (used & 1);
      if (used >= 2 && cudaEventSynchronize(d0.ev_nstage[slot]) != cudaSuccess) { failed = 1; break; }
      uint8_t *buf = d0.h_nstage + (size_t)slot * kChunk;
      memcpy(buf, (const uint8_t *)col[c] + o, len);
      cudaError_t e = cudaMemcpyAsync((uint8_t *)(dd0.f64 + ds.stride_f64 * ds.slot[c]) + o, buf, len,
                                      cudaMemcpyHostToDevice, d0.stream2);
      if (e == cudaSuccess) e = cudaEventRecord(d0.ev_nstage[slot], d0.stream2);
      if (e != cudaSuccess) { failed = 1; break; }
      used++;
    }
  };
  std::vector<std::thread> th;
  for (int t = 1; t < nthreads; t++) th.emplace_back(worker, t);
  worker(0);
  for (auto &x : th) x.join();
  cudaSetDevice(d0.id);
  if (failed.load()) return bnl_fail(ctx, BNL_ECUDA, "staged upload of the numeric columns failed");
  // the main stream continues after everything queued on stream2.
  BNL_CUDA(ctx, cudaEventRecord(d0.ev_stage[1], d0.stream2));
  BNL_CUDA(ctx, cudaStreamWaitEvent(d0.stream, d0.ev_stage[1], 0));
  return BNL_OK;
}

extern "C" int bnl_data_register(bnl_ctx *ctx, int ncols, const void *const *col,
                                 const int *is_numeric, const int *nlevels, int64_t nobs,
                                 bnl_data *handle) {
  if (!ctx || ctx->devs.empty()) return BNL_EINVAL;
  if (ncols <= 0 || !col || !is_numeric || !nlevels || !handle || nobs < 0)
    return bnl_fail(ctx, BNL_EINVAL, "bnl_data_register: bad arguments");
  if (nobs >= (int64_t)INT32_MAX)
    return bnl_fail(ctx, BNL_EINVAL, "bnl_data_register: nobs must fit an int (R vectors on this path do)");

  BnlDataSet ds;
  ds.ncols = ncols;
  ds.nobs = nobs;
  ds.is_numeric.assign(is_numeric, is_numeric + ncols);
  ds.nlevels.assign(nlevels, nlevels + ncols);
  ds.slot.resize(ncols);
  ds.is_wide.assign(ncols, 0);
  for (int i = 0; i < ncols; i++) {
    if (is_numeric[i]) ds.slot[i] = ds.nc++;
    else {
      if (nlevels[i] < 1)
        return bnl_fail(ctx, BNL_EINVAL, "column %d: a factor needs at least one level", i);
      if (nlevels[i] > BNL_MAX_LEVELS_WIDE)
        return bnl_fail(ctx, BNL_EUNSUPPORTED, "column %d has %d levels; the uint16 layout holds at most %d",
                        i, nlevels[i], BNL_MAX_LEVELS_WIDE);
      if (nlevels[i] > BNL_MAX_LEVELS) { ds.is_wide[i] = 1; ds.slot[i] = ds.nw++; }
      else ds.slot[i] = ds.nd++;
    }
  }
  ds.stride_u8 = round_up((size_t)nobs + 1, BNL_COL_ALIGN);
  ds.stride_u16 = round_up((size_t)nobs + 1, BNL_COL_ALIGN);
  ds.stride_f64 = round_up((size_t)nobs + 1, 32);
  ds.dev.resize(ctx->devs.size());

  // device 0: upload (int32 staging -> pack kernel; doubles straight in).
  BnlDevice &d0 = ctx->devs[0];
  BnlDeviceData &dd0 = ds.dev[0];
  BNL_CUDA(ctx, cudaSetDevice(d0.id));
  int *stage = nullptr, *d_bad = nullptr;
  int rc = BNL_OK;
  auto cleanup = [&](int code) {
    if (code != BNL_OK) free_dataset(ctx, ds);
    return code;
  };
#define REG_CUDA(call)                                                                     \
  do {                                                                                     \
    cudaError_t e__ = (call);                                                              \
    if (e__ != cudaSuccess)                                                                \
      return cleanup(bnl_fail(ctx, e__ == cudaErrorMemoryAllocation ? BNL_ENOMEM : BNL_ECUDA, \
                              "%s failed: %s", #call, cudaGetErrorString(e__)));           \
  } while (0)

  if (ds.nd > 0) {
    dd0.u8_bytes = ds.stride_u8 * ds.nd;
    if ((rc = bnl_pool_alloc(ctx, d0, (void **)&dd0.u8, dd0.u8_bytes))) return cleanup(rc);
    REG_CUDA(cudaMemsetAsync(dd0.u8, BNL_NA_U8, ds.stride_u8 * ds.nd, d0.stream));
  }
  if (ds.nw > 0) {
    dd0.u16_bytes = ds.stride_u16 * ds.nw * sizeof(uint16_t);
    if ((rc = bnl_pool_alloc(ctx, d0, (void **)&dd0.u16, dd0.u16_bytes))) return cleanup(rc);
    REG_CUDA(cudaMemsetAsync(dd0.u16, 0xFF, dd0.u16_bytes, d0.stream));
  }
  const int nflags = 1 + ds.nd + ds.nw;             // bad code | NA per uint8 slot | NA per uint16 slot
  if (ds.nd + ds.nw > 0) {
    if ((rc = bnl_reserve_device(ctx, &d0.stage, &d0.stage_cap, 2 * ((size_t)nobs + 68)))) return cleanup(rc);
    if ((rc = bnl_reserve_device(ctx, &d0.d_flags, &d0.flags_cap, (size_t)nflags))) return cleanup(rc);
    stage = d0.stage;
    d_bad = d0.d_flags;
    REG_CUDA(cudaMemsetAsync(d_bad, 0, (size_t)nflags * sizeof(int), d0.stream));
  }
  if (ds.nc > 0) {
    dd0.f64_bytes = ds.stride_f64 * ds.nc * sizeof(double);
    if ((rc = bnl_pool_alloc(ctx, d0, (void **)&dd0.f64, dd0.f64_bytes))) return cleanup(rc);
    REG_CUDA(cudaMemsetAsync(dd0.f64, 0, ds.stride_f64 * ds.nc * sizeof(double), d0.stream));
  }
  for (int i = 0; i < ncols; i++)
    if (!col[i] && nobs > 0)
      return cleanup(bnl_fail(ctx, BNL_EINVAL, "column %d is a null pointer", i));
  // big data frames: host cores and the copy engine share the factor columns.
  std::vector<int> host_flags;
  const bool hybrid = ds.nd > 0 && (long long)ds.nd * nobs >= (8ll << 20) && !getenv("BNL_NO_HOST_PACK");
  if (hybrid) {
    std::vector<int> dcols(ds.nd);
    for (int i = 0; i < ncols; i++) if (!is_numeric[i] && !ds.is_wide[i]) dcols[ds.slot[i]] = i;
    if ((rc = upload_factors_hybrid(ctx, ds, d0, dd0, col, nlevels, dcols, d_bad, host_flags)))
      return cleanup(rc);
  }
  // numeric columns in pageable memory: chunked through pinned buffers by host threads.
  std::vector<char> staged(ncols, 0);
  if (ds.nc > 0 && (long long)ds.nc * nobs >= (8ll << 20) && !getenv("BNL_NO_HOST_PACK")) {
    std::vector<int> pageable;
    for (int i = 0; i < ncols; i++) {
      if (!is_numeric[i]) continue;
      cudaPointerAttributes at;
      const bool pinned = cudaPointerGetAttributes(&at, col[i]) == cudaSuccess &&
                          (at.type == cudaMemoryTypeHost || at.type == cudaMemoryTypeManaged);
      cudaGetLastError();                       // unregistered host memory is not an error here
      if (!pinned) { pageable.push_back(i); staged[i] = 1; }
    }
    if (!pageable.empty() && (rc = upload_numeric_staged(ctx, ds, d0, dd0, col, pageable))) return cleanup(rc);
  }
  for (int i = 0; i < ncols; i++) {
    if (nobs == 0) continue;
    if (staged[i]) continue;
    if (hybrid && !is_numeric[i] && !ds.is_wide[i]) continue;
    if (is_numeric[i]) {
      REG_CUDA(cudaMemcpyAsync(dd0.f64 + ds.stride_f64 * ds.slot[i], col[i],
                               (size_t)nobs * sizeof(double), cudaMemcpyHostToDevice, d0.stream));
    } else if (ds.is_wide[i]) {
      REG_CUDA(cudaMemcpyAsync(stage, col[i], (size_t)nobs * sizeof(int), cudaMemcpyHostToDevice,
                               d0.stream));
      int grid = (int)std::min<long long>((nobs + 255) / 256, (long long)d0.sm_count * 16);
      if (grid < 1) grid = 1;
      pack_factors16_kernel<<<grid, 256, 0, d0.stream>>>(stage, dd0.u16 + ds.stride_u16 * ds.slot[i], nobs,
                                                         nlevels[i], d_bad, d_bad + 1 + ds.nd + ds.slot[i]);
      REG_CUDA(cudaGetLastError());
    } else {
      REG_CUDA(cudaMemcpyAsync(stage, col[i], (size_t)nobs * sizeof(int), cudaMemcpyHostToDevice,
                               d0.stream));
      long long quads = (nobs + 3) / 4;
      int grid = (int)((quads + 255) / 256);
      if (grid > d0.sm_count * 16) grid = d0.sm_count * 16;
      pack_factors_kernel<<<grid, 256, 0, d0.stream>>>(stage, dd0.u8 + ds.stride_u8 * ds.slot[i],
                                                       nobs, nlevels[i], d_bad, d_bad + 1 + ds.slot[i]);
      REG_CUDA(cudaGetLastError());
    }
  }
  if (ds.nd + ds.nw > 0) {
    std::vector<int> flags(nflags, 0);
    REG_CUDA(cudaMemcpyAsync(flags.data(), d_bad, flags.size() * sizeof(int), cudaMemcpyDeviceToHost,
                             d0.stream));
    REG_CUDA(cudaStreamSynchronize(d0.stream));
    for (size_t i = 0; i < host_flags.size() && i < (size_t)(1 + ds.nd); i++) flags[i] |= host_flags[i];
    const int bad = flags[0];
    ds.slot_has_na.assign(flags.begin() + 1, flags.begin() + 1 + ds.nd);
    ds.wide_has_na.assign(flags.begin() + 1 + ds.nd, flags.end());
    if (bad)
      return cleanup(bnl_fail(ctx, BNL_EINVAL, "a factor column holds a code outside 1..nlevels"));
  } else {
    REG_CUDA(cudaStreamSynchronize(d0.stream));
  }

  // replicate on the other devices (peer copies; NVLink where available).
  for (size_t i = 1; i < ctx->devs.size(); i++) {
    BnlDevice &di = ctx->devs[i];
    BnlDeviceData &ddi = ds.dev[i];
    REG_CUDA(cudaSetDevice(di.id));
    if (ds.nd > 0) {
      ddi.u8_bytes = dd0.u8_bytes;
      if ((rc = bnl_pool_alloc(ctx, di, (void **)&ddi.u8, ddi.u8_bytes))) return cleanup(rc);
      REG_CUDA(cudaMemcpyPeerAsync(ddi.u8, di.id, dd0.u8, d0.id, ds.stride_u8 * ds.nd, di.stream));
    }
    if (ds.nw > 0) {
      ddi.u16_bytes = dd0.u16_bytes;
      if ((rc = bnl_pool_alloc(ctx, di, (void **)&ddi.u16, ddi.u16_bytes))) return cleanup(rc);
      REG_CUDA(cudaMemcpyPeerAsync(ddi.u16, di.id, dd0.u16, d0.id, dd0.u16_bytes, di.stream));
    }
    if (ds.nc > 0) {
      ddi.f64_bytes = dd0.f64_bytes;
      if ((rc = bnl_pool_alloc(ctx, di, (void **)&ddi.f64, ddi.f64_bytes))) return cleanup(rc);
      REG_CUDA(cudaMemcpyPeerAsync(ddi.f64, di.id, dd0.f64, d0.id,
                                   ds.stride_f64 * ds.nc * sizeof(double), di.stream));
    }
  }
  for (size_t i = 1; i < ctx->devs.size(); i++) {
    REG_CUDA(cudaSetDevice(ctx->devs[i].id));
    REG_CUDA(cudaStreamSynchronize(ctx->devs[i].stream));
  }
#undef REG_CUDA

  // Gaussian moments (means + covariance), once per data set and device.
  if (ds.nc > 0) {
    for (size_t i = 0; i < ctx->devs.size(); i++) {
      rc = bnl_gaussian_prepare(ctx, ds, (int)i);
      if (rc != BNL_OK) return cleanup(rc);
    }
  }
  cleanup(BNL_OK);
  cudaSetDevice(d0.id);
  *handle = ctx->next_handle++;
  ctx->data[*handle] = ds;
  return BNL_OK;
}

extern "C" int bnl_data_resample(bnl_ctx *ctx, bnl_data source, int64_t nrows, const int *rows,
                                 bnl_data *handle) {
  if (!ctx || ctx->devs.empty()) return BNL_EINVAL;
  if (!handle || nrows < 0 || (nrows > 0 && !rows) || nrows >= (int64_t)INT32_MAX)
    return bnl_fail(ctx, BNL_EINVAL, "bnl_data_resample: bad arguments");
  auto it = ctx->data.find(source);
  if (it == ctx->data.end()) return bnl_fail(ctx, BNL_EINVAL, "unknown data handle");
  const BnlDataSet &src = it->second;
  if (src.nobs == 0 && nrows > 0)
    return bnl_fail(ctx, BNL_EINVAL, "bnl_data_resample: the source data set has no rows");

  BnlDataSet ds;
  ds.ncols = src.ncols; ds.nd = src.nd; ds.nc = src.nc; ds.nw = src.nw;
  ds.is_wide = src.is_wide;
  ds.nobs = nrows;
  ds.is_numeric = src.is_numeric; ds.nlevels = src.nlevels; ds.slot = src.slot;
  ds.bge_mu = src.bge_mu; ds.bge_w = src.bge_w;      // nu falls back to the replicate's own means
  ds.ebic_gamma = src.ebic_gamma;
  ds.mask_col = src.mask_col;                         // intervention masks are columns: they follow the rows
  ds.stride_u8 = round_up((size_t)nrows + 1, BNL_COL_ALIGN);
  ds.stride_f64 = round_up((size_t)nrows + 1, 32);
  ds.stride_u16 = round_up((size_t)nrows + 1, BNL_COL_ALIGN);
  ds.dev.resize(ctx->devs.size());
  ds.slot_has_na.assign(ds.nd, 0);
  ds.wide_has_na.assign(ds.nw, 0);
  int rc = BNL_OK;
  auto fail = [&](int code) { free_dataset(ctx, ds); return code; };
#define RS_CUDA(call)                                                                      \
  do {                                                                                     \
    cudaError_t e__ = (call);                                                              \
    if (e__ != cudaSuccess)                                                                \
      return fail(bnl_fail(ctx, e__ == cudaErrorMemoryAllocation ? BNL_ENOMEM : BNL_ECUDA, \
                           "%s failed: %s", #call, cudaGetErrorString(e__)));              \
  } while (0)

  // every device gathers from its own replica; the row list goes to each of them.
  std::vector<std::vector<int>> flags(ctx->devs.size(), std::vector<int>(1 + ds.nd + ds.nw, 0));
  for (size_t d = 0; d < ctx->devs.size(); d++) {
    BnlDevice &dev = ctx->devs[d];
    BnlDeviceData &dd = ds.dev[d];
    const BnlDeviceData &sd = src.dev[d];
    RS_CUDA(cudaSetDevice(dev.id));
    if ((rc = bnl_reserve_device(ctx, &dev.stage, &dev.stage_cap, (size_t)nrows + 64))) return fail(rc);
    if ((rc = bnl_reserve_device(ctx, &dev.d_flags, &dev.flags_cap, (size_t)(1 + ds.nd + ds.nw)))) return fail(rc);
    RS_CUDA(cudaMemsetAsync(dev.d_flags, 0, (size_t)(1 + ds.nd + ds.nw) * sizeof(int), dev.stream));
    if (nrows > 0)
      RS_CUDA(cudaMemcpyAsync(dev.stage, rows, (size_t)nrows * sizeof(int), cudaMemcpyHostToDevice, dev.stream));
    int gx = (int)std::min<long long>((nrows + 255) / 256, (long long)dev.sm_count * 8);
    if (gx < 1) gx = 1;
    if (ds.nd > 0) {
      dd.u8_bytes = ds.stride_u8 * ds.nd;
      if ((rc = bnl_pool_alloc(ctx, dev, (void **)&dd.u8, dd.u8_bytes))) return fail(rc);
      RS_CUDA(cudaMemsetAsync(dd.u8, BNL_NA_U8, dd.u8_bytes, dev.stream));
      if (nrows > 0) {
        gather_u8_kernel<<<dim3(gx, ds.nd), 256, 0, dev.stream>>>(sd.u8, src.stride_u8, src.nobs, dd.u8,
                                                                 ds.stride_u8, dev.stage, nrows, dev.d_flags);
        RS_CUDA(cudaGetLastError());
      }
    }
    if (ds.nw > 0) {
      dd.u16_bytes = ds.stride_u16 * ds.nw * sizeof(uint16_t);
      if ((rc = bnl_pool_alloc(ctx, dev, (void **)&dd.u16, dd.u16_bytes))) return fail(rc);
      RS_CUDA(cudaMemsetAsync(dd.u16, 0xFF, dd.u16_bytes, dev.stream));
      if (nrows > 0) {
        gather_u16_kernel<<<dim3(gx, ds.nw), 256, 0, dev.stream>>>(sd.u16, src.stride_u16, src.nobs, dd.u16,
                                                                  ds.stride_u16, dev.stage, nrows, dev.d_flags,
                                                                  1 + ds.nd);
        RS_CUDA(cudaGetLastError());
      }
    }
    if (ds.nc > 0) {
      dd.f64_bytes = ds.stride_f64 * ds.nc * sizeof(double);
      if ((rc = bnl_pool_alloc(ctx, dev, (void **)&dd.f64, dd.f64_bytes))) return fail(rc);
      RS_CUDA(cudaMemsetAsync(dd.f64, 0, dd.f64_bytes, dev.stream));
      if (nrows > 0) {
        gather_f64_kernel<<<dim3(gx, ds.nc), 256, 0, dev.stream>>>(sd.f64, src.stride_f64, src.nobs, dd.f64,
                                                                  ds.stride_f64, dev.stage, nrows, dev.d_flags);
        RS_CUDA(cudaGetLastError());
      }
    }
    RS_CUDA(cudaMemcpyAsync(flags[d].data(), dev.d_flags, flags[d].size() * sizeof(int),
                            cudaMemcpyDeviceToHost, dev.stream));
  }
  for (size_t d = 0; d < ctx->devs.size(); d++) {
    RS_CUDA(cudaSetDevice(ctx->devs[d].id));
    RS_CUDA(cudaStreamSynchronize(ctx->devs[d].stream));
  }
#undef RS_CUDA
  if (flags[0][0]) return fail(bnl_fail(ctx, BNL_EINVAL, "bnl_data_resample: a row index is outside the source data set"));
  for (int c = 0; c < ds.nd; c++) ds.slot_has_na[c] = flags[0][1 + c];
  for (int c = 0; c < ds.nw; c++) ds.wide_has_na[c] = flags[0][1 + ds.nd + c];
  if (ds.nc > 0)
    for (size_t d = 0; d < ctx->devs.size(); d++)
      if ((rc = bnl_gaussian_prepare(ctx, ds, (int)d)) != BNL_OK) return fail(rc);
  cudaSetDevice(ctx->devs[0].id);
  *handle = ctx->next_handle++;
  ctx->data[*handle] = ds;
  return BNL_OK;
}

extern "C" int bnl_data_alloc(bnl_ctx *ctx, int ncols, const int *is_numeric, const int *nlevels,
                              int64_t nobs, bnl_data *handle) {
  if (!ctx || ctx->devs.empty()) return BNL_EINVAL;
  if (ncols <= 0 || !is_numeric || !nlevels || !handle || nobs < 0 || nobs >= (int64_t)INT32_MAX)
    return bnl_fail(ctx, BNL_EINVAL, "bnl_data_alloc: bad arguments");
  BnlDataSet ds;
  ds.ncols = ncols;
  ds.nobs = nobs;
  ds.is_numeric.assign(is_numeric, is_numeric + ncols);
  ds.nlevels.assign(nlevels, nlevels + ncols);
  ds.slot.resize(ncols);
  ds.is_wide.assign(ncols, 0);
  for (int i = 0; i < ncols; i++) {
    if (is_numeric[i]) ds.slot[i] = ds.nc++;
    else {
      if (nlevels[i] < 1 || nlevels[i] > BNL_MAX_LEVELS)
        return bnl_fail(ctx, BNL_EUNSUPPORTED, "column %d: 1..%d levels per factor", i, BNL_MAX_LEVELS);
      ds.slot[i] = ds.nd++;
    }
  }
  ds.stride_u8 = round_up((size_t)nobs + 1, BNL_COL_ALIGN);
  ds.stride_f64 = round_up((size_t)nobs + 1, 32);
  ds.dev.resize(ctx->devs.size());
  ds.slot_has_na.assign(ds.nd, 1);             // unknown until bnl_data_commit
  ds.committed = false;
  for (size_t d = 0; d < ctx->devs.size(); d++) {
    BnlDevice &dev = ctx->devs[d];
    BnlDeviceData &dd = ds.dev[d];
    int rc;
    if (cudaSetDevice(dev.id) != cudaSuccess) { free_dataset(ctx, ds); return bnl_fail(ctx, BNL_ECUDA, "cudaSetDevice failed"); }
    if (ds.nd > 0) {
      dd.u8_bytes = ds.stride_u8 * ds.nd;
      if ((rc = bnl_pool_alloc(ctx, dev, (void **)&dd.u8, dd.u8_bytes))) { free_dataset(ctx, ds); return rc; }
    }
    if (ds.nc > 0) {
      dd.f64_bytes = ds.stride_f64 * ds.nc * sizeof(double);
      if ((rc = bnl_pool_alloc(ctx, dev, (void **)&dd.f64, dd.f64_bytes))) { free_dataset(ctx, ds); return rc; }
    }
  }
  cudaSetDevice(ctx->devs[0].id);
  *handle = ctx->next_handle++;
  ctx->data[*handle] = ds;
  return BNL_OK;
}

extern "C" int bnl_data_device_view(bnl_ctx *ctx, bnl_data data, int dev_index, void **u8, size_t *u8_bytes,
                                    void **f64, size_t *f64_bytes) {
  if (!ctx) return BNL_EINVAL;
  auto it = ctx->data.find(data);
  if (it == ctx->data.end()) return bnl_fail(ctx, BNL_EINVAL, "unknown data handle");
  if (dev_index < 0 || dev_index >= (int)it->second.dev.size())
    return bnl_fail(ctx, BNL_EINVAL, "bnl_data_device_view: device index out of range");
  const BnlDeviceData &dd = it->second.dev[dev_index];
  if (u8) *u8 = dd.u8;
  if (u8_bytes) *u8_bytes = dd.u8_bytes;
  if (f64) *f64 = dd.f64;
  if (f64_bytes) *f64_bytes = dd.f64_bytes;
  return BNL_OK;
}

extern "C" int bnl_data_commit(bnl_ctx *ctx, bnl_data data) {
  if (!ctx) return BNL_EINVAL;
  auto it = ctx->data.find(data);
  if (it == ctx->data.end()) return bnl_fail(ctx, BNL_EINVAL, "unknown data handle");
  BnlDataSet &ds = it->second;
  int rc;
  if (ds.nd > 0) {
    std::vector<int> lv(ds.nd, 0), flags(1 + ds.nd, 0), any(1 + ds.nd, 0);
    for (int c = 0; c < ds.ncols; c++) if (!ds.is_numeric[c]) lv[ds.slot[c]] = ds.nlevels[c];
    // every replica is checked (each was filled by the caller), and the bytes between
    // nobs and the column stride are forced to the NA code: the counting kernels read
    // whole 16-row groups and rely on that padding to veto the rows past the end.
    for (size_t d = 0; d < ctx->devs.size(); d++) {
      BnlDevice &dev = ctx->devs[d];
      BNL_CUDA(ctx, cudaSetDevice(dev.id));
      // the caller may have rewritten the columns: a nibble copy made from the old ones is void
      BnlDeviceData &ddc = ds.dev[d];
      if (ddc.u4) { BNL_CUDA(ctx, cudaStreamSynchronize(dev.stream)); bnl_pool_free(dev, ddc.u4, ddc.u4_bytes); ddc.u4 = nullptr; ddc.u4_bytes = 0; }
      if ((rc = bnl_reserve_device(ctx, &dev.d_flags, &dev.flags_cap, (size_t)(2 + 2 * ds.nd)))) return rc;
      int *d_lv = dev.d_flags + 1 + ds.nd;
      BNL_CUDA(ctx, cudaMemsetAsync(dev.d_flags, 0, (size_t)(1 + ds.nd) * sizeof(int), dev.stream));
      BNL_CUDA(ctx, cudaMemcpyAsync(d_lv, lv.data(), (size_t)ds.nd * sizeof(int), cudaMemcpyHostToDevice, dev.stream));
      if (ds.stride_u8 > (size_t)ds.nobs)
        BNL_CUDA(ctx, cudaMemset2DAsync(ds.dev[d].u8 + ds.nobs, ds.stride_u8, BNL_NA_U8,
                                        ds.stride_u8 - (size_t)ds.nobs, (size_t)ds.nd, dev.stream));
      int gx = (int)std::min<long long>((ds.nobs + 255) / 256, (long long)dev.sm_count * 8);
      if (gx < 1) gx = 1;
      scan_packed_kernel<<<dim3(gx, ds.nd), 256, 0, dev.stream>>>(ds.dev[d].u8, ds.stride_u8, ds.nobs, d_lv,
                                                                  dev.d_flags);
      BNL_CUDA(ctx, cudaGetLastError());
      BNL_CUDA(ctx, cudaMemcpyAsync(flags.data(), dev.d_flags, flags.size() * sizeof(int), cudaMemcpyDeviceToHost,
                                    dev.stream));
      BNL_CUDA(ctx, cudaStreamSynchronize(dev.stream));
      for (size_t i = 0; i < flags.size(); i++) any[i] |= flags[i];
    }
    if (any[0]) return bnl_fail(ctx, BNL_EINVAL, "a packed factor column holds a code outside 0..nlevels-1");
    ds.slot_has_na.assign(any.begin() + 1, any.end());
  }
  if (ds.nc > 0)
    for (size_t d = 0; d < ctx->devs.size(); d++)
      if ((rc = bnl_gaussian_prepare(ctx, ds, (int)d)) != BNL_OK) return rc;
  ds.committed = true;
  cudaSetDevice(ctx->devs[0].id);
  return BNL_OK;
}

extern "C" int bnl_data_release(bnl_ctx *ctx, bnl_data handle) {
  if (!ctx) return BNL_EINVAL;
  auto it = ctx->data.find(handle);
  if (it == ctx->data.end()) return bnl_fail(ctx, BNL_EINVAL, "unknown data handle");
  free_dataset(ctx, it->second);
  ctx->data.erase(it);
  cudaSetDevice(ctx->devs[0].id);
  return BNL_OK;
}

extern "C" int bnl_data_flush(bnl_ctx *ctx) {
  if (!ctx) return BNL_EINVAL;
  for (auto &kv : ctx->data) free_dataset(ctx, kv.second);
  ctx->data.clear();
  if (!ctx->devs.empty()) cudaSetDevice(ctx->devs[0].id);
  return BNL_OK;
}
