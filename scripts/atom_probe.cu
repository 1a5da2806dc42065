// atom_probe.cu -- throughput of three ways to bump per-lane counters in shared
// memory, without any global traffic: (A) shared-memory atomics on lane-private
// slots shared by 4 warps, (B) plain 32-bit load/add/store on warp-private slots,
// (C) the same with 16-bit counters. Decides whether a warp-private table layout
// could beat the atomic one in the counting kernels.
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o scripts/_build/atom_probe scripts/atom_probe.cu
#include <cstdio>
#include <cstdint>
#include <cstdlib>
#include <cuda_runtime.h>

#define CK(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { \
    printf("%s failed: %s\n", #call, cudaGetErrorString(e_)); exit(1); } } while (0)

template <int MODE>
__global__ void __launch_bounds__(256) probe(int cells, int iters, int words, unsigned *sink) {
  extern __shared__ unsigned tab[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int i = threadIdx.x; i < words; i += 256) tab[i] = 0;
  __syncthreads();
  unsigned s = threadIdx.x * 2654435761u + blockIdx.x;
  volatile unsigned *vt = tab;
  volatile unsigned short *vs = reinterpret_cast<volatile unsigned short *>(tab);
  for (int it = 0; it < iters; it++) {
#pragma unroll
    for (int j = 0; j < 16; j++) {
      s = s * 1664525u + 1013904223u;
      const int cell = (s >> 16) & (cells - 1);      // cells is a power of two here
      if (MODE == 0) {
        atomicAdd(&tab[cell * 64 + (warp & 1) * 32 + lane], 1u);
      } else if (MODE == 1) {
        const int idx = (warp * cells + cell) * 32 + lane;
        vt[idx] = vt[idx] + 1;
      } else {
        const int idx = (warp * cells + cell) * 32 + lane;
        vs[idx] = vs[idx] + 1;
      }
    }
  }
  __syncthreads();
  unsigned acc = 0;
  for (int i = threadIdx.x; i < words; i += 256) acc += tab[i];
  if (acc == 0xdeadbeef) sink[0] = acc;
}

template <int MODE>
void run(const char *name, int cells, int ctas_per_sm) {
  int sms = 0;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  const int iters = 4096;
  const size_t smem = (size_t)cells * 64 * 4 * (MODE == 0 ? 1 : (MODE == 1 ? 4 : 2));
  CK(cudaFuncSetAttribute(probe<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  unsigned *sink = nullptr;
  CK(cudaMalloc(&sink, 4));
  const int grid = sms * ctas_per_sm;
  const int words = (int)(smem / 4);
  probe<MODE><<<grid, 256, smem>>>(cells, 16, words, sink);
  CK(cudaDeviceSynchronize());
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  cudaEventRecord(e0);
  probe<MODE><<<grid, 256, smem>>>(cells, iters, words, sink);
  cudaEventRecord(e1);
  CK(cudaEventSynchronize(e1));
  float ms = 0;
  CK(cudaEventElapsedTime(&ms, e0, e1));
  cudaError_t err = cudaGetLastError();
  const double updates = (double)grid * 256 * iters * 16;
  printf("%-28s cells %4d ctas/sm %d smem %6zu B: %8.3f ms  %.2f T updates/s  (%s)\n", name, cells, ctas_per_sm,
         smem, ms, updates / ms / 1e9, cudaGetErrorString(err));
  cudaFree(sink);
}

// (D) address span: one CTA of NT threads per SM, a table of `cells` cells with
// (1 << LC) lane counters each, random cells. Same-bank collisions cannot happen
// for LC >= 5; what changes with `cells` is only how far apart the 32 addresses of
// one warp-wide atomic lie.
template <int LC, int NT>
__global__ void __launch_bounds__(NT) span_probe(int cells, int iters, int words, unsigned *sink) {
  extern __shared__ unsigned tab[];
  for (int i = threadIdx.x; i < words; i += NT) tab[i] = 0;
  __syncthreads();
  unsigned s = threadIdx.x * 2654435761u + blockIdx.x;
  const unsigned slot4 = (LC == 6 ? (threadIdx.x & 31u) + 32u * ((threadIdx.x >> 5) & 1u) : threadIdx.x & ((1u << LC) - 1u)) << 2;
  for (int it = 0; it < iters; it++) {
#pragma unroll
    for (int j = 0; j < 16; j++) {
      // full-rate instructions only (IMAD.HI would cap the loop at ~13 lanes/clk/SM)
      s = s * 1664525u + 1013904223u;
      const unsigned u = __byte_perm(s, 0, 0x4432) * (unsigned)cells;       // (s >> 16) * cells
      const unsigned off = ((u >> (14 - LC)) & (0xFFFFFFFFu << (LC + 2))) | slot4;
      atomicAdd(reinterpret_cast<unsigned *>(reinterpret_cast<char *>(tab) + off), 1u);
    }
  }
  __syncthreads();
  unsigned acc = 0;
  for (int i = threadIdx.x; i < words; i += NT) acc += tab[i];
  if (acc == 0xdeadbeef) sink[0] = acc;
}

template <int LC, int NT>
void run_span(int cells) {
  int sms = 0;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  const int iters = 2048;
  const size_t smem = ((size_t)cells << LC) * 4;
  if (smem > 226 * 1024) return;
  CK(cudaFuncSetAttribute(span_probe<LC, NT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  unsigned *sink = nullptr;
  CK(cudaMalloc(&sink, 4));
  const int words = (int)(smem / 4);
  span_probe<LC, NT><<<sms, NT, smem>>>(cells, 16, words, sink);
  CK(cudaDeviceSynchronize());
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  cudaEventRecord(e0);
  span_probe<LC, NT><<<sms, NT, smem>>>(cells, iters, words, sink);
  cudaEventRecord(e1);
  CK(cudaEventSynchronize(e1));
  float ms = 0;
  CK(cudaEventElapsedTime(&ms, e0, e1));
  const double updates = (double)sms * NT * iters * 16;
  printf("span: %2d counters/cell %4d threads cells %5d table %7zu B: %8.3f ms  %.2f T updates/s\n", 1 << LC, NT, cells,
         smem, ms, updates / ms / 1e9);
  cudaFree(sink);
}

int main(int argc, char **argv) {
  if (argc > 1) {
    for (int cells : {32, 64, 128, 256, 384, 512, 768, 880}) run_span<6, 1024>(cells);
    for (int cells : {64, 128, 256, 512, 768, 1024, 1400, 1750}) run_span<5, 1024>(cells);
    for (int cells : {256, 512, 1024, 2048, 3500}) run_span<4, 1024>(cells);
    for (int cells : {512, 1024, 2048, 4096, 7000}) run_span<3, 1024>(cells);
    for (int cells : {256, 1024, 1750}) run_span<5, 512>(cells);
    return 0;
  }
  for (int cells : {16, 64, 128}) {
    for (int c : {2, 4, 8}) {
      run<0>("atomics, 4 warps per slot", cells, c);
      if ((size_t)cells * 64 * 4 * 4 * c <= 200 * 1024) run<1>("private 32-bit load/add/store", cells, c);
      if ((size_t)cells * 64 * 4 * 2 * c <= 200 * 1024) run<2>("private 16-bit load/add/store", cells, c);
    }
  }
  return 0;
}
