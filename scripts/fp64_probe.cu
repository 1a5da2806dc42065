// fp64_probe.cu -- what the fp64 pipes of one B200 can do, register operands only:
// (a) DFMA chains (8 independent accumulators per thread), (b) DMMA m8n8k4 chains
// (8 independent accumulator fragments per warp), (c) both interleaved.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o fp64_probe fp64_probe.cu && ./fp64_probe
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ void dmma884(double &d0, double &d1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0, %1}, {%2}, {%3}, {%0, %1};"
               : "+d"(d0), "+d"(d1) : "d"(a), "d"(b));
}

template <int MODE>
__global__ void __launch_bounds__(256) probe(double *out, int iters, double a, double b) {
  double acc[16];
#pragma unroll
  for (int i = 0; i < 16; i++) acc[i] = threadIdx.x * 1e-9 + i;
  double x = a + threadIdx.x * 1e-12, y = b;
  for (int it = 0; it < iters; it++) {
    if (MODE == 0) {
#pragma unroll
      for (int i = 0; i < 16; i++) acc[i] = fma(acc[i], x, y);
    } else if (MODE == 1) {
#pragma unroll
      for (int i = 0; i < 8; i++) dmma884(acc[2 * i], acc[2 * i + 1], x, y);
    } else {
#pragma unroll
      for (int i = 0; i < 4; i++) {
        dmma884(acc[2 * i], acc[2 * i + 1], x, y);
        acc[8 + 2 * i] = fma(acc[8 + 2 * i], x, y);
        acc[9 + 2 * i] = fma(acc[9 + 2 * i], x, y);
      }
    }
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < 16; i++) s += acc[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// (d) the Gram kernel's register pattern: a 4 x 4 outer product of fragments, 16
// accumulator pairs, operands changing every k-step (ORDER 0: i-major, 1: j-major, 2: snake).
template <int ORDER, int WITH_DADD>
__global__ void __launch_bounds__(512, 1) outer(double *out, int iters, double a0, double b0) {
  double acc[4][4][2];
#pragma unroll
  for (int i = 0; i < 4; i++)
#pragma unroll
    for (int j = 0; j < 4; j++) acc[i][j][0] = acc[i][j][1] = 0.0;
  double a[4], b[4], ma[4], mb[4];
#pragma unroll
  for (int i = 0; i < 4; i++) { a[i] = a0 + i + threadIdx.x * 1e-9; b[i] = b0 - i; ma[i] = 1e-3 * i; mb[i] = 2e-3 * i; }
  for (int it = 0; it < iters; it++) {
    double x[4], y[4];
#pragma unroll
    for (int i = 0; i < 4; i++) { x[i] = WITH_DADD ? a[i] - ma[i] : a[i]; y[i] = WITH_DADD ? b[i] - mb[i] : b[i]; }
    if (ORDER == 0) {
#pragma unroll
      for (int i = 0; i < 4; i++)
#pragma unroll
        for (int j = 0; j < 4; j++) dmma884(acc[i][j][0], acc[i][j][1], x[i], y[j]);
    } else if (ORDER == 1) {
#pragma unroll
      for (int j = 0; j < 4; j++)
#pragma unroll
        for (int i = 0; i < 4; i++) dmma884(acc[i][j][0], acc[i][j][1], x[i], y[j]);
    } else {
#pragma unroll
      for (int i = 0; i < 4; i++)
#pragma unroll
        for (int jj = 0; jj < 4; jj++) { const int j = (i & 1) ? 3 - jj : jj; dmma884(acc[i][j][0], acc[i][j][1], x[i], y[j]); }
    }
    // new operands next step (cheap integer-free rotation keeps them live and distinct)
    const double t = a[0]; a[0] = a[1]; a[1] = a[2]; a[2] = a[3]; a[3] = t;
    const double u = b[0]; b[0] = b[1]; b[1] = b[2]; b[2] = b[3]; b[3] = u;
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < 4; i++)
#pragma unroll
    for (int j = 0; j < 4; j++) s += acc[i][j][0] + acc[i][j][1];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int ORDER, int WITH_DADD>
void run_outer(const char *name) {
  int sms = 148;
  double *out;
  cudaMalloc(&out, sizeof(double) * sms * 512);
  const int iters = 100000;
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  outer<ORDER, WITH_DADD><<<sms, 512>>>(out, 1000, 1.0000001, 1e-9);
  cudaDeviceSynchronize();
  cudaEventRecord(e0);
  outer<ORDER, WITH_DADD><<<sms, 512>>>(out, iters, 1.0000001, 1e-9);
  cudaEventRecord(e1);
  cudaDeviceSynchronize();
  float ms;
  cudaEventElapsedTime(&ms, e0, e1);
  const double fl = 16.0 * 16 * iters * 512.0 * sms;
  printf("%-40s 1 CTA/SM x 512 threads: %8.3f ms  %6.2f TFLOP/s (DMMA flops only)\n", name, ms, fl / ms / 1e9);
  cudaFree(out);
}

template <int MODE>
void run(const char *name, double flops_per_thread_iter, int ctas_per_sm) {
  int sms = 148;
  double *out;
  cudaMalloc(&out, sizeof(double) * sms * 8 * 256);
  const int iters = 200000;
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  probe<MODE><<<sms * ctas_per_sm, 256>>>(out, 1000, 1.0000001, 1e-9);
  cudaDeviceSynchronize();
  cudaEventRecord(e0);
  probe<MODE><<<sms * ctas_per_sm, 256>>>(out, iters, 1.0000001, 1e-9);
  cudaEventRecord(e1);
  cudaDeviceSynchronize();
  float ms;
  cudaEventElapsedTime(&ms, e0, e1);
  const double fl = flops_per_thread_iter * iters * 256.0 * sms * ctas_per_sm;
  printf("%-28s %d CTAs/SM x 256 threads: %8.3f ms  %6.2f TFLOP/s\n", name, ctas_per_sm, ms, fl / ms / 1e9);
  cudaFree(out);
}

int main() {
  for (int c : {1, 2, 4}) {
    if (c == 1) { run<0>("DFMA x16", 32, 1); run<1>("DMMA.884 x8", 8 * 2 * 8 * 8 * 4 / 32.0, 1); run<2>("DMMA x4 + DFMA x8", 4 * 16 + 16, 1); }
    if (c == 2) { run<0>("DFMA x16", 32, 2); run<1>("DMMA.884 x8", 8 * 2 * 8 * 8 * 4 / 32.0, 2); run<2>("DMMA x4 + DFMA x8", 4 * 16 + 16, 2); }
    if (c == 4) { run<0>("DFMA x16", 32, 4); run<1>("DMMA.884 x8", 8 * 2 * 8 * 8 * 4 / 32.0, 4); run<2>("DMMA x4 + DFMA x8", 4 * 16 + 16, 4); }
  }
  run_outer<0, 0>("4x4 outer product, i-major");
  run_outer<1, 0>("4x4 outer product, j-major");
  run_outer<2, 0>("4x4 outer product, snake");
  run_outer<0, 1>("4x4 outer product, i-major, 8 DADD/step");
  run_outer<2, 1>("4x4 outer product, snake, 8 DADD/step");
  return 0;
}
