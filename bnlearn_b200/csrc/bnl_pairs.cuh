// bnl_pairs.cuh -- all-pairs two-way tables as a tensor-core contraction (see bnl_pairs.cu).
#pragma once
#include "bnl_internal.h"
