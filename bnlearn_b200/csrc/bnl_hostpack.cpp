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

}  // extern "C"
