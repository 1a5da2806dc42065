// bnl_hostpack.cpp -- host-side packing of R's int32 factor codes to the uint8
// device layout (code - 1, NA -> 0xFF). Compiled by g++ with vectorisation
// (-O3, function multi-versioning) apart from the CUDA sources.
//
// Why on the host: the caller's data frame is 4 bytes per value and usually sits
// in pageable memory. Packing it with the host cores into a pinned staging ring
// cuts the bytes that cross PCIe by 4x (bnl_data_register, bnl_context.cu).
#include <limits.h>
#include <stdint.h>

extern "C" {

// packs rows [0, n) of one column slice; ORs the flags (never clears them).
__attribute__((target_clones("arch=x86-64-v4", "avx2", "default")))
void bnl_host_pack_slice(const int *__restrict__ src, uint8_t *__restrict__ dst, long long n,
                         int nlevels, int *bad, int *has_na) {
  int any_na = 0, any_bad = 0;
  for (long long i = 0; i < n; i++) {
    const int v = src[i];
    const int na = v == INT_MIN;
    const int out_of_range = (v < 1) | (v > nlevels);
    any_na |= na;
    any_bad |= out_of_range & !na;
    dst[i] = (na | out_of_range) ? (uint8_t)0xFF : (uint8_t)(v - 1);
  }
  if (any_na) *has_na = 1;
  if (any_bad) *bad = 1;
}

// Two codes per byte (low nibble = the even row) for factors of at most 15 levels; NA and
// out-of-range -> 0xF. Halves what the packers write to the pinned ring and what the copy
// engine reads back from it: registration of a large data frame is bound by the host's
// memory traffic, not by PCIe. n even except for the last slice of a column.
__attribute__((target_clones("arch=x86-64-v4", "avx2", "default")))
void bnl_host_pack_slice4(const int *__restrict__ src, uint8_t *__restrict__ dst, long long n,
                          int nlevels, int *bad, int *has_na) {
  int any_na = 0, any_bad = 0;
  const long long pairs = n / 2;
  for (long long i = 0; i < pairs; i++) {
    const int a = src[2 * i], b = src[2 * i + 1];
    const int na_a = a == INT_MIN, na_b = b == INT_MIN;
    const int oor_a = (a < 1) | (a > nlevels), oor_b = (b < 1) | (b > nlevels);
    any_na |= na_a | na_b;
    any_bad |= (oor_a & !na_a) | (oor_b & !na_b);
    const int ca = (na_a | oor_a) ? 0xF : a - 1, cb = (na_b | oor_b) ? 0xF : b - 1;
    dst[i] = (uint8_t)(ca | (cb << 4));
  }
  if (n & 1) {
    const int a = src[n - 1];
    const int na_a = a == INT_MIN, oor_a = (a < 1) | (a > nlevels);
    any_na |= na_a;
    any_bad |= oor_a & !na_a;
    dst[pairs] = (uint8_t)(((na_a | oor_a) ? 0xF : a - 1) | 0xF0);
  }
  if (any_na) *has_na = 1;
  if (any_bad) *bad = 1;
}

}  // extern "C"
