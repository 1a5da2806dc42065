// special_host.cpp -- host build of bnl_special.cuh behind a C ABI, so that the
// CPU test-suite can pin the exact code the kernels run (no GPU needed).
#include "bnl_special.cuh"

extern "C" {
double bnl_host_lgamma(double x) { return bnl::lgamma_pos(x); }
double bnl_host_pchisq_upper(double x, double df) { return bnl::pchisq_upper(x, df); }
double bnl_host_pt_two_sided(double t, double df) { return bnl::pt_two_sided(t, df); }
double bnl_host_pnorm_two_sided(double z) { return bnl::pnorm_two_sided(z); }
}
