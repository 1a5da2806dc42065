#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <cstring>
#include <thread>
#include <vector>
extern "C" void bnl_host_pack_slice(const int *src, uint8_t *dst, long long n, int nlevels, int *bad, int *has_na);
int main(int argc, char **argv) {
  int nt = argc > 1 ? atoi(argv[1]) : 1;
  long long n = 10000000; int ncol = 16;
  std::vector<int> src((size_t)n * ncol);
  for (size_t i = 0; i < src.size(); i++) src[i] = 1 + (int)(i * 2654435761u >> 30);
  std::vector<uint8_t> dst((size_t)n * ncol);
  memset(dst.data(), 1, dst.size());
  for (int rep = 0; rep < 3; rep++) {
    auto t0 = std::chrono::steady_clock::now();
    std::vector<std::thread> th;
    for (int t = 0; t < nt; t++) th.emplace_back([&, t] {
      int bad = 0, na = 0;
      for (int c = t; c < ncol; c += nt) bnl_host_pack_slice(src.data() + (size_t)c * n, dst.data() + (size_t)c * n, n, 4, &bad, &na);
    });
    for (auto &x : th) x.join();
    double s = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    printf("threads %d: %.1f ms, %.1f GB/s input\n", nt, s * 1e3, (double)n * ncol * 4 / s / 1e9);
  }
}
