// tma_probe.cu -- should the row tiles of the counting kernels be staged in shared
// memory by the bulk-copy engine (cp.async.bulk + mbarrier, "TMA-staged row tiles")
// instead of travelling through registers (LDG.128, software-pipelined)?
//
// Both kernels do the work of count_based_kernel<5, 1, 1024>: one CTA per unit, a
// uint16 base-index column shared by all units (L2-resident) and a uint8 candidate
// column of its own (from DRAM), cell = base + cmul * cand, one ATOMS per row into a
// table with 32 lane counters per cell. Only the way the rows arrive differs.
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o scripts/_build/tma_probe scripts/tma_probe.cu
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cuda_runtime.h>

#define CK(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { \
    printf("%s failed: %s\n", #call, cudaGetErrorString(e_)); exit(1); } } while (0)

constexpr int NT = 1024;
constexpr int LC = 5;

__device__ __forceinline__ uint4 ldg_stream(const uint4 *p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
  return r;
}

__device__ __forceinline__ void bump2(int *hist, unsigned v, unsigned slot4) {
  atomicAdd(reinterpret_cast<int *>(reinterpret_cast<char *>(hist) + (((v & 0xFFFFu) << (LC + 2)) + slot4)), 1);
  atomicAdd(reinterpret_cast<int *>(reinterpret_cast<char *>(hist) + (((v >> 16) << (LC + 2)) + slot4)), 1);
}

__device__ __forceinline__ void count16(int *hist, const unsigned *bw, const unsigned *cw, unsigned cmul, unsigned slot4) {
#pragma unroll
  for (int q = 0; q < 4; q++) {
    bump2(hist, bw[2 * q] + __byte_perm(cw[q], 0, 0x4140) * cmul, slot4);
    bump2(hist, bw[2 * q + 1] + __byte_perm(cw[q], 0, 0x4342) * cmul, slot4);
  }
}

// (A) registers: the loads of the next two 16-row groups in flight.
__global__ void __launch_bounds__(NT)
reg_kernel(const uint16_t *base, const uint8_t *cand, size_t cstride, long long rows, int cells, unsigned cmul,
           int *tables) {
  extern __shared__ __align__(256) int hist[];
  for (int c = threadIdx.x; c < cells << LC; c += NT) hist[c] = 0;
  __syncthreads();
  const unsigned slot4 = (threadIdx.x & 31u) << 2;
  const uint4 *bp = reinterpret_cast<const uint4 *>(base);
  const uint4 *cp = reinterpret_cast<const uint4 *>(cand + (size_t)blockIdx.x * cstride);
  const long long g1 = rows / 16;
  struct R { uint4 b0, b1, c; } r0, r1, r2;
  auto load = [&](R &r, long long g) { r.b0 = ldg_stream(bp + 2 * g); r.b1 = ldg_stream(bp + 2 * g + 1); r.c = ldg_stream(cp + g); };
  auto count = [&](const R &r) {
    const unsigned bw[8] = {r.b0.x, r.b0.y, r.b0.z, r.b0.w, r.b1.x, r.b1.y, r.b1.z, r.b1.w};
    const unsigned cw[4] = {r.c.x, r.c.y, r.c.z, r.c.w};
    count16(hist, bw, cw, cmul, slot4);
  };
  long long g = threadIdx.x;
  if (g < g1) load(r0, g);
  if (g + NT < g1) load(r1, g + NT);
#define STEP(A, C) if (g < g1) { if (g + 2 * NT < g1) load(C, g + 2 * NT); count(A); g += NT; }
  while (g < g1) { STEP(r0, r2) STEP(r1, r0) STEP(r2, r1) }
#undef STEP
  __syncthreads();
  int *gt = tables + (size_t)blockIdx.x * cells;
  for (int c = threadIdx.x; c < cells; c += NT) {
    int s = 0;
    for (int l = 0; l < 32; l++) s += hist[c * 32 + ((l + threadIdx.x) & 31)];
    gt[c] = s;
  }
}

// (B) bulk copies: STAGES tiles of TILE rows (base: 2 * TILE bytes, cand: TILE bytes)
// land in shared memory; an mbarrier per stage counts the bytes; every thread then
// reads its 16 rows with LDS.128 (TILE = 16 * NT) or 8 rows (TILE = 8 * NT).
__device__ __forceinline__ unsigned smem_u32(const void *p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(unsigned long long *bar, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect(unsigned long long *bar, unsigned bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void *dst, const void *src, unsigned bytes, unsigned long long *bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               :: "r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long *bar, unsigned parity) {
  asm volatile(
      "{\n .reg .pred p;\n WAIT_%=:\n"
      " mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      " @p bra DONE_%=;\n bra WAIT_%=;\n DONE_%=:\n}\n" :: "r"(smem_u32(bar)), "r"(parity) : "memory");
}

template <int RPT, int STAGES>       // rows per thread and tile: 16 or 8
__global__ void __launch_bounds__(NT)
tma_kernel(const uint16_t *base, const uint8_t *cand, size_t cstride, long long rows, int cells, unsigned cmul,
           int *tables) {
  constexpr int TILE = RPT * NT;
  extern __shared__ __align__(256) int hist[];
  __shared__ __align__(8) unsigned long long full[STAGES];
  char *stage0 = reinterpret_cast<char *>(hist) + (((size_t)cells << (LC + 2)) + 255) / 256 * 256;
  auto sbase = [&](int s) { return stage0 + (size_t)s * 3 * TILE; };
  auto scand = [&](int s) { return stage0 + (size_t)s * 3 * TILE + 2 * TILE; };
  for (int c = threadIdx.x; c < cells << LC; c += NT) hist[c] = 0;
  if (threadIdx.x == 0) {
    for (int s = 0; s < STAGES; s++) mbar_init(&full[s], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  const unsigned slot4 = (threadIdx.x & 31u) << 2;
  const uint8_t *cp = cand + (size_t)blockIdx.x * cstride;
  const long long ntiles = rows / TILE;
  auto issue = [&](long long t, int s) {
    mbar_expect(&full[s], 3 * TILE);
    bulk_g2s(sbase(s), base + t * TILE, 2 * TILE, &full[s]);
    bulk_g2s(scand(s), cp + t * TILE, TILE, &full[s]);
  };
  if (threadIdx.x == 0)
    for (int s = 0; s < STAGES && s < ntiles; s++) issue(s, s);
  for (long long t = 0; t < ntiles; t++) {
    const int s = (int)(t % STAGES);
    mbar_wait(&full[s], (unsigned)((t / STAGES) & 1));
    if (RPT == 16) {
      const uint4 b0 = reinterpret_cast<const uint4 *>(sbase(s))[2 * threadIdx.x];
      const uint4 b1 = reinterpret_cast<const uint4 *>(sbase(s))[2 * threadIdx.x + 1];
      const uint4 c = reinterpret_cast<const uint4 *>(scand(s))[threadIdx.x];
      const unsigned bw[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
      const unsigned cw[4] = {c.x, c.y, c.z, c.w};
      count16(hist, bw, cw, cmul, slot4);
    } else {
      const uint4 b0 = reinterpret_cast<const uint4 *>(sbase(s))[threadIdx.x];
      const uint2 c = reinterpret_cast<const uint2 *>(scand(s))[threadIdx.x];
      const unsigned bw[4] = {b0.x, b0.y, b0.z, b0.w};
      const unsigned cw[2] = {c.x, c.y};
#pragma unroll
      for (int q = 0; q < 2; q++) {
        bump2(hist, bw[2 * q] + __byte_perm(cw[q], 0, 0x4140) * cmul, slot4);
        bump2(hist, bw[2 * q + 1] + __byte_perm(cw[q], 0, 0x4342) * cmul, slot4);
      }
    }
    __syncthreads();                                   // every thread has read stage s
    if (threadIdx.x == 0 && t + STAGES < ntiles) issue(t + STAGES, s);
  }
  int *gt = tables + (size_t)blockIdx.x * cells;
  for (int c = threadIdx.x; c < cells; c += NT) {
    int s = 0;
    for (int l = 0; l < 32; l++) s += hist[c * 32 + ((l + threadIdx.x) & 31)];
    gt[c] = s;
  }
}

template <typename F>
float timed(F launch, int reps) {
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  launch();
  CK(cudaDeviceSynchronize());
  cudaEventRecord(e0);
  for (int i = 0; i < reps; i++) launch();
  cudaEventRecord(e1);
  CK(cudaEventSynchronize(e1));
  float ms = 0;
  cudaEventElapsedTime(&ms, e0, e1);
  return ms / reps;
}

int main() {
  const long long rows = 1 << 20;
  int sms = 0;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  const int units = sms * 4;
  for (int nbase : {96, 288}) {                         // 384 and 1152 cells: 48 KB and 144 KB of table
    const int cells = nbase * 4;
    std::vector<uint16_t> hb(rows);
    std::vector<uint8_t> hc((size_t)rows * 8);
    unsigned s = 12345u;
    for (auto &v : hb) { s = s * 1664525u + 1013904223u; v = (uint16_t)((s >> 8) % nbase); }
    for (auto &v : hc) { s = s * 1664525u + 1013904223u; v = (uint8_t)((s >> 24) & 3); }
    uint16_t *db; uint8_t *dc; int *t0, *t1;
    CK(cudaMalloc(&db, rows * 2));
    CK(cudaMalloc(&dc, (size_t)units * rows));
    CK(cudaMalloc(&t0, (size_t)units * cells * 4));
    CK(cudaMalloc(&t1, (size_t)units * cells * 4));
    CK(cudaMemcpy(db, hb.data(), rows * 2, cudaMemcpyHostToDevice));
    for (int u = 0; u < units; u++)                      // distinct columns (rotations of 8 random ones)
      CK(cudaMemcpy(dc + (size_t)u * rows, hc.data() + (size_t)(u % 8) * rows, rows, cudaMemcpyHostToDevice));
    const size_t tab = (size_t)cells << (LC + 2);
    CK(cudaFuncSetAttribute(reg_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)tab));
    const float a = timed([&] { reg_kernel<<<units, NT, tab>>>(db, dc, rows, rows, cells, nbase, t0); }, 10);
    CK(cudaGetLastError());
    printf("cells %4d (table %3zu KB)  registers, 2 groups ahead          : %.3f ms  %.2f T updates/s\n", cells,
           tab >> 10, a, (double)units * rows / a / 1e9);
    auto run_tma = [&](auto kernel, int rpt, int stages, const char *name) {
      const size_t smem = (tab + 255) / 256 * 256 + (size_t)stages * 3 * rpt * NT;
      if (smem > 227 * 1024) { printf("cells %4d  %s: does not fit (%zu KB)\n", cells, name, smem >> 10); return; }
      CK(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      CK(cudaMemset(t1, 0, (size_t)units * cells * 4));
      const float b = timed([&] { kernel<<<units, NT, smem>>>(db, dc, rows, rows, cells, nbase, t1); }, 10);
      CK(cudaGetLastError());
      std::vector<int> h0((size_t)units * cells), h1((size_t)units * cells);
      CK(cudaMemcpy(h0.data(), t0, h0.size() * 4, cudaMemcpyDeviceToHost));
      CK(cudaMemcpy(h1.data(), t1, h1.size() * 4, cudaMemcpyDeviceToHost));
      printf("cells %4d (table %3zu KB)  %-36s: %.3f ms  %.2f T updates/s  tables %s\n", cells, tab >> 10, name, b,
             (double)units * rows / b / 1e9, h0 == h1 ? "identical" : "DIFFER");
    };
    run_tma(tma_kernel<16, 2>, 16, 2, "bulk copy, 16384-row tiles x 2");
    run_tma(tma_kernel<16, 3>, 16, 3, "bulk copy, 16384-row tiles x 3");
    run_tma(tma_kernel<8, 3>, 8, 3, "bulk copy, 8192-row tiles x 3");
    run_tma(tma_kernel<8, 4>, 8, 4, "bulk copy, 8192-row tiles x 4");
    cudaFree(db); cudaFree(dc); cudaFree(t0); cudaFree(t1);
  }
  return 0;
}
