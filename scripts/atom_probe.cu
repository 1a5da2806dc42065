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
#define CK(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { \
    printf("%s failed: %s\n", #call, cudaGetErrorString(e_)); exit(1); } } while (0)
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

int main() {
  for (int cells : {16, 64, 128}) {
    for (int c : {2, 4, 8}) {
      run<0>("atomics, 4 warps per slot", cells, c);
      if ((size_t)cells * 64 * 4 * 4 * c <= 200 * 1024) run<1>("private 32-bit load/add/store", cells, c);
      if ((size_t)cells * 64 * 4 * 2 * c <= 200 * 1024) run<2>("private 16-bit load/add/store", cells, c);
    }
  }
  return 0;
}
