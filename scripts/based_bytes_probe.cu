// based_bytes_probe.cu -- how much would count_based_kernel gain from fewer loaded bytes per row?
// The kernel's inner loop (one CTA of 1024 threads per unit, 32 lane counters per cell, loads of the
// next two groups in flight) on 999 units x 1 M rows with
//   M0  uint16 base index + uint8 candidate (3 bytes per row: what ships)
//   M1  uint8 base index (base table <= 256 cells) + uint8 candidate (2 bytes)
//   M2  uint8 base index + nibble candidate (1.5 bytes; 32 rows per iteration)
//   M3  uint16 base index + nibble candidate (2.5 bytes; 32 rows per iteration, one group ahead)
//   M4  no loads at all, indices from a counter (the update rate alone)
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o scripts/_bin/based_bytes_probe scripts/based_bytes_probe.cu
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
#define CK(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { printf("%s: %s\n", #call, cudaGetErrorString(e_)); exit(1); } } while (0)
constexpr int NT = 1024, LC = 5;

__device__ __forceinline__ uint4 ldg_stream(const uint4 *p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
  return r;
}
__device__ __forceinline__ void bump2(int *hist, unsigned v, unsigned slot4) {     // two 16-bit cell indices
  atomicAdd(reinterpret_cast<int *>(reinterpret_cast<char *>(hist) + (((v & 0xFFFFu) << (LC + 2)) + slot4)), 1);
  atomicAdd(reinterpret_cast<int *>(reinterpret_cast<char *>(hist) + (((v >> 16) << (LC + 2)) + slot4)), 1);
}
// four 8-bit base indices b and four candidate codes c (bytes): cell = b + cmul * c as two 16-bit pairs
__device__ __forceinline__ void bump4(int *hist, unsigned b, unsigned c, unsigned cmul, unsigned slot4) {
  bump2(hist, __byte_perm(b, 0, 0x4140) + __byte_perm(c, 0, 0x4140) * cmul, slot4);
  bump2(hist, __byte_perm(b, 0, 0x4342) + __byte_perm(c, 0, 0x4342) * cmul, slot4);
}

// the same with the lane's slot folded into the table pointer once: cell * 128 + pointer is one multiply-add
__device__ __forceinline__ void bump4_lean(char *hp, unsigned b, unsigned c, unsigned cmul) {
  const unsigned lo = __byte_perm(b, 0, 0x4140) + __byte_perm(c, 0, 0x4140) * cmul;
  const unsigned hi = __byte_perm(b, 0, 0x4342) + __byte_perm(c, 0, 0x4342) * cmul;
  atomicAdd(reinterpret_cast<int *>(hp + ((lo & 0xFFFFu) << (LC + 2))), 1);
  atomicAdd(reinterpret_cast<int *>(hp + ((lo >> 16) << (LC + 2))), 1);
  atomicAdd(reinterpret_cast<int *>(hp + ((hi & 0xFFFFu) << (LC + 2))), 1);
  atomicAdd(reinterpret_cast<int *>(hp + ((hi >> 16) << (LC + 2))), 1);
}

// ... and with the address arithmetic spelled out: one multiply-add per update on a 32-bit shared address
__device__ __forceinline__ void red1(unsigned addr) { asm volatile("red.shared.add.u32 [%0], 1;" ::"r"(addr)); }
__device__ __forceinline__ void bump4_ptx(unsigned pre, unsigned b, unsigned c, unsigned cmul) {
  const unsigned lo = __byte_perm(b, 0, 0x4140) + __byte_perm(c, 0, 0x4140) * cmul;
  const unsigned hi = __byte_perm(b, 0, 0x4342) + __byte_perm(c, 0, 0x4342) * cmul;
  red1(__byte_perm(lo, 0, 0x4410) * 128u + pre);       // a byte-permute takes the 16-bit half: the compiler cannot
  red1(__byte_perm(lo, 0, 0x4432) * 128u + pre);       // fold it into a shift-and-mask, and * 128 + pre is one IMAD
  red1(__byte_perm(hi, 0, 0x4410) * 128u + pre);
  red1(__byte_perm(hi, 0, 0x4432) * 128u + pre);
}

template <int MODE>
__global__ void __launch_bounds__(NT)
probe(const uint16_t *base16, const uint8_t *base8, const uint8_t *cand, size_t cstride, long long rows, int cells,
      unsigned cmul, int *tables) {
  extern __shared__ __align__(256) int hist[];
  for (int c = threadIdx.x; c < cells << LC; c += NT) hist[c] = 0;
  __syncthreads();
  const unsigned slot4 = (threadIdx.x & 31u) << 2;
  const uint4 *b16 = reinterpret_cast<const uint4 *>(base16);
  const uint4 *b8 = reinterpret_cast<const uint4 *>(base8);
  const uint4 *cp = reinterpret_cast<const uint4 *>(cand + (size_t)blockIdx.x * cstride);
  const unsigned M = 0x0F0F0F0Fu;
  char *hp = reinterpret_cast<char *>(hist) + slot4;
  const unsigned pre = (unsigned)__cvta_generic_to_shared(hist) + slot4;
  if (MODE == 0 || MODE == 1 || MODE == 5 || MODE == 6) {
    const long long g1 = rows / 16;
    struct R { uint4 b0, b1, c; } r0, r1, r2;
    auto load = [&](R &r, long long g) {
      if (MODE == 0) { r.b0 = ldg_stream(b16 + 2 * g); r.b1 = ldg_stream(b16 + 2 * g + 1); } else r.b0 = ldg_stream(b8 + g);
      r.c = ldg_stream(cp + g);
    };
    auto count = [&](const R &r) {
      const unsigned cw[4] = {r.c.x, r.c.y, r.c.z, r.c.w};
      if (MODE == 0) {
        const unsigned bw[8] = {r.b0.x, r.b0.y, r.b0.z, r.b0.w, r.b1.x, r.b1.y, r.b1.z, r.b1.w};
#pragma unroll
        for (int q = 0; q < 4; q++) {
          bump2(hist, bw[2 * q] + __byte_perm(cw[q], 0, 0x4140) * cmul, slot4);
          bump2(hist, bw[2 * q + 1] + __byte_perm(cw[q], 0, 0x4342) * cmul, slot4);
        }
      } else if (MODE == 1) {
        const unsigned bw[4] = {r.b0.x, r.b0.y, r.b0.z, r.b0.w};
#pragma unroll
        for (int q = 0; q < 4; q++) bump4(hist, bw[q], cw[q], cmul, slot4);
      } else if (MODE == 5) {
        const unsigned bw[4] = {r.b0.x, r.b0.y, r.b0.z, r.b0.w};
#pragma unroll
        for (int q = 0; q < 4; q++) bump4_lean(hp, bw[q], cw[q], cmul);
      } else {
        const unsigned bw[4] = {r.b0.x, r.b0.y, r.b0.z, r.b0.w};
#pragma unroll
        for (int q = 0; q < 4; q++) bump4_ptx(pre, bw[q], cw[q], cmul);
      }
    };
    long long g = threadIdx.x;
    if (g < g1) load(r0, g);
    if (g + NT < g1) load(r1, g + NT);
#define STEP(A, C) if (g < g1) { if (g + 2 * NT < g1) load(C, g + 2 * NT); count(A); g += NT; }
    while (g < g1) { STEP(r0, r2) STEP(r1, r0) STEP(r2, r1) }
#undef STEP
  } else if (MODE == 2) {
    const long long g1 = rows / 32;                      // 32 rows per iteration: 2 x 16 base bytes, 16 candidate bytes
    struct R { uint4 b0, b1, c; } r0, r1, r2;
    auto load = [&](R &r, long long g) { r.b0 = ldg_stream(b8 + 2 * g); r.b1 = ldg_stream(b8 + 2 * g + 1); r.c = ldg_stream(cp + g); };
    auto count = [&](const R &r) {
      const unsigned bw[8] = {r.b0.x, r.b0.y, r.b0.z, r.b0.w, r.b1.x, r.b1.y, r.b1.z, r.b1.w};
      const unsigned cw[4] = {r.c.x, r.c.y, r.c.z, r.c.w};
#pragma unroll
      for (int q = 0; q < 4; q++) {
        // nibble word q holds rows 8q .. 8q + 7: even rows in the low nibbles, odd rows in the high ones; the base
        // bytes are laid out to match (the combine pass writes them in that order)
        bump4(hist, bw[2 * q], cw[q] & M, cmul, slot4);
        bump4(hist, bw[2 * q + 1], (cw[q] >> 4) & M, cmul, slot4);
      }
    };
    long long g = threadIdx.x;
    if (g < g1) load(r0, g);
    if (g + NT < g1) load(r1, g + NT);
#define STEP(A, C) if (g < g1) { if (g + 2 * NT < g1) load(C, g + 2 * NT); count(A); g += NT; }
    while (g < g1) { STEP(r0, r2) STEP(r1, r0) STEP(r2, r1) }
#undef STEP
  } else if (MODE == 3) {
    const long long g1 = rows / 32;
    struct R { uint4 b0, b1, b2, b3, c; } r0, r1;
    auto load = [&](R &r, long long g) {
      r.b0 = ldg_stream(b16 + 4 * g); r.b1 = ldg_stream(b16 + 4 * g + 1); r.b2 = ldg_stream(b16 + 4 * g + 2); r.b3 = ldg_stream(b16 + 4 * g + 3);
      r.c = ldg_stream(cp + g);
    };
    auto count = [&](const R &r) {
      const unsigned bw[16] = {r.b0.x, r.b0.y, r.b0.z, r.b0.w, r.b1.x, r.b1.y, r.b1.z, r.b1.w,
                               r.b2.x, r.b2.y, r.b2.z, r.b2.w, r.b3.x, r.b3.y, r.b3.z, r.b3.w};
      const unsigned cw[4] = {r.c.x, r.c.y, r.c.z, r.c.w};
#pragma unroll
      for (int q = 0; q < 4; q++) {
        const unsigned lo = cw[q] & M, hi = (cw[q] >> 4) & M;
        bump2(hist, bw[4 * q] + __byte_perm(lo, 0, 0x4140) * cmul, slot4);
        bump2(hist, bw[4 * q + 1] + __byte_perm(lo, 0, 0x4342) * cmul, slot4);
        bump2(hist, bw[4 * q + 2] + __byte_perm(hi, 0, 0x4140) * cmul, slot4);
        bump2(hist, bw[4 * q + 3] + __byte_perm(hi, 0, 0x4342) * cmul, slot4);
      }
    };
    long long g = threadIdx.x;
    if (g < g1) load(r0, g);
#define STEP(A, C) if (g < g1) { if (g + NT < g1) load(C, g + NT); count(A); g += NT; }
    while (g < g1) { STEP(r0, r1) STEP(r1, r0) }
#undef STEP
  } else {
    unsigned s = threadIdx.x * 2654435761u + blockIdx.x;
    const long long g1 = rows / 16;
    for (long long g = threadIdx.x; g < g1; g += NT) {
#pragma unroll
      for (int q = 0; q < 8; q++) {
        s = s * 1664525u + 1013904223u;
        const unsigned a = ((s >> 16) * (unsigned)cells) >> 16, b = ((s & 0xFFFFu) * (unsigned)cells) >> 16;
        bump2(hist, a | (b << 16), slot4);
      }
    }
  }
  __syncthreads();
  int *gt = tables + (size_t)blockIdx.x * cells;
  for (int c = threadIdx.x; c < cells; c += NT) {
    int s = 0;
    for (int l = 0; l < 32; l++) s += hist[c * 32 + ((l + threadIdx.x) & 31)];
    gt[c] = s;
  }
}

__global__ void fill(uint16_t *b16, uint8_t *b8, uint8_t *cand, size_t cstride, long long rows, int nunits, int basecells, int lev, int skew_pct) {
  const long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (i >= rows) return;
  unsigned long long z = (unsigned long long)i * 0x9e3779b97f4a7c15ull + 7;
  z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ull; z ^= z >> 27;
  const bool hot = (int)((z >> 40) % 100) < skew_pct;            // skewed tables: one cell holds skew_pct % of the rows
  b16[i] = hot ? 0 : (uint16_t)(z % (unsigned)basecells);
  b8[i] = hot ? 0 : (uint8_t)(z % (unsigned)(basecells > 256 ? 256 : basecells));
  for (int u = blockIdx.y; u < nunits; u += gridDim.y) {
    unsigned long long w = z * (2 * u + 3) + u;
    w = (w ^ (w >> 29)) * 0x94d049bb133111ebull; w ^= w >> 32;
    cand[(size_t)u * cstride + i] = hot && (w >> 50) % 4 ? 0 : (uint8_t)(w % (unsigned)lev);   // bytes < 16: valid as codes and as nibble pairs
  }
}

template <int MODE>
void run(const char *name, const uint16_t *b16, const uint8_t *b8, const uint8_t *cand, size_t cstride, long long rows,
         int nunits, int basecells, int lev, int *tables) {
  const int cells = basecells * lev;
  const size_t smem = (size_t)cells << (LC + 2);
  CK(cudaFuncSetAttribute(probe<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  float best = 1e30f;
  for (int rep = 0; rep < 4; rep++) {
    cudaEventRecord(e0);
    probe<MODE><<<nunits, NT, smem>>>(b16, b8, cand, cstride, rows, cells, (unsigned)basecells, tables);
    cudaEventRecord(e1);
    CK(cudaDeviceSynchronize());
    float ms; cudaEventElapsedTime(&ms, e0, e1); best = ms < best ? ms : best;
  }
  printf("base cells %4d x %d levels: %-52s %7.3f ms  %.2f T updates/s\n", basecells, lev, name, best,
         (double)nunits * rows / best / 1e9);
}

int main() {
  const long long rows = 1000000;
  const int nunits = 999;
  const size_t cstride = 1000064;
  uint16_t *b16; uint8_t *b8, *cand; int *tables;
  CK(cudaMalloc(&b16, rows * 2 + 256)); CK(cudaMalloc(&b8, rows + 256)); CK(cudaMalloc(&cand, cstride * nunits));
  CK(cudaMalloc(&tables, (size_t)nunits * 4096 * 4));
  for (int skew : {0, 30, 60})
  for (int basecells : {81, 243, 576}) {
    const int lev = 3;
    if (skew && basecells != 81) continue;
    printf("-- %d %% of the rows in one cell\n", skew);
    fill<<<dim3((unsigned)((rows + 255) / 256), 64), 256>>>(b16, b8, cand, cstride, rows, nunits, basecells, lev, skew);
    CK(cudaDeviceSynchronize());
    run<0>("M0 uint16 base + uint8 candidate (3 B/row)", b16, b8, cand, cstride, rows, nunits, basecells, lev, tables);
    if (basecells <= 256) run<1>("M1 uint8 base + uint8 candidate (2 B/row)", b16, b8, cand, cstride, rows, nunits, basecells, lev, tables);
    if (basecells <= 256) run<5>("M5 = M1 with the slot folded into the table pointer", b16, b8, cand, cstride, rows, nunits, basecells, lev, tables);
    if (basecells <= 256) run<6>("M6 = M1 with red.shared on a multiply-add address", b16, b8, cand, cstride, rows, nunits, basecells, lev, tables);
    if (basecells <= 256) run<2>("M2 uint8 base + nibble candidate (1.5 B/row)", b16, b8, cand, cstride, rows, nunits, basecells, lev, tables);
    run<3>("M3 uint16 base + nibble candidate (2.5 B/row)", b16, b8, cand, cstride, rows, nunits, basecells, lev, tables);
    run<4>("M4 no loads (update rate alone)", b16, b8, cand, cstride, rows, nunits, basecells, lev, tables);
  }
  return 0;
}
