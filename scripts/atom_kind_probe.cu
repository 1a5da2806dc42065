// atom_kind_probe.cu -- does the KIND of shared-memory update instruction matter? Lane-private
// counters (cell * 32 + lane: no bank conflicts, no address shared inside a warp), random cells,
// no global traffic; 1024 threads, one CTA per SM.
//   K0 atomicAdd(p, 1)           K1 atomicAdd(p, v) with v a register   K2 red.shared.add.u32 [p], 1
//   K3 red.shared.add.u32 [p], v K4 atom.shared.inc.u32                 K5 plain ld / add / st (lane-private: legal)
//   K6 red.shared.add.u64 on pairs of counters (one update per TWO rows when both fall in the same 8-byte slot -- bound only)
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
#define CK(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { printf("%s: %s\n", #call, cudaGetErrorString(e_)); exit(1); } } while (0)

template <int K>
__global__ void __launch_bounds__(1024) probe(int cells, int iters, int words, unsigned v, unsigned *sink) {
  extern __shared__ unsigned tab[];
  for (int i = threadIdx.x; i < words; i += 1024) tab[i] = 0;
  __syncthreads();
  unsigned s = threadIdx.x * 2654435761u + blockIdx.x;
  const unsigned slot4 = (threadIdx.x & 31u) << 2;
  const unsigned base = (unsigned)__cvta_generic_to_shared(tab);
  for (int it = 0; it < iters; it++) {
#pragma unroll
    for (int j = 0; j < 16; j++) {
      s = s * 1664525u + 1013904223u;
      const unsigned u = __byte_perm(s, 0, 0x4432) * (unsigned)cells;       // (s >> 16) * cells
      const unsigned off = ((u >> 9) & 0xFFFFFF80u) | slot4;                // cell * 128 + lane * 4
      unsigned *p = reinterpret_cast<unsigned *>(reinterpret_cast<char *>(tab) + off);
      if (K == 0) atomicAdd(p, 1u);
      else if (K == 1) atomicAdd(p, v);
      else if (K == 2) asm volatile("red.shared.add.u32 [%0], 1;" ::"r"(base + off) : "memory");
      else if (K == 3) asm volatile("red.shared.add.u32 [%0], %1;" ::"r"(base + off), "r"(v) : "memory");
      else if (K == 4) { unsigned o; asm volatile("atom.shared.inc.u32 %0, [%1], 0xffffffff;" : "=r"(o) : "r"(base + off) : "memory"); }
      else if (K == 5) { volatile unsigned *q = p; *q = *q + 1; }
      else if (K == 6) asm volatile("red.shared.add.u64 [%0], %1;" ::"r"(base + (off & ~7u)), "l"(0x100000001ull) : "memory");
    }
  }
  __syncthreads();
  unsigned acc = 0;
  for (int i = threadIdx.x; i < words; i += 1024) acc += tab[i];
  if (acc == 0xdeadbeef) sink[0] = acc;
}

template <int K>
void run(const char *name, int cells) {
  const size_t smem = (size_t)cells * 128;
  CK(cudaFuncSetAttribute(probe<K>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  unsigned *sink; CK(cudaMalloc(&sink, 4));
  const int iters = 2048, sms = 148;
  probe<K><<<sms, 1024, smem>>>(cells, 16, (int)(smem / 4), 1u, sink);
  CK(cudaDeviceSynchronize());
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  cudaEventRecord(e0);
  probe<K><<<sms, 1024, smem>>>(cells, iters, (int)(smem / 4), 1u, sink);
  cudaEventRecord(e1);
  CK(cudaEventSynchronize(e1));
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  const double upd = (double)sms * 1024 * iters * 16;
  printf("%-34s cells %5d: %8.3f ms  %.2f T updates/s  %.1f lanes/clk/SM\n", name, cells, ms, upd / ms / 1e9,
         upd / ms / 1e9 * 1e12 / 148 / 1.965e9 / 1e3 * 1e-9 * 1e3);
  cudaFree(sink);
}

int main() {
  for (int cells : {256, 1024}) {
    run<0>("atomicAdd(p, 1)", cells);
    run<1>("atomicAdd(p, v)", cells);
    run<2>("red.shared.add.u32 1", cells);
    run<3>("red.shared.add.u32 v", cells);
    run<4>("atom.shared.inc.u32", cells);
    run<5>("ld + add + st (lane-private)", cells);
    run<6>("red.shared.add.u64 (2 counters)", cells);
  }
  return 0;
}
