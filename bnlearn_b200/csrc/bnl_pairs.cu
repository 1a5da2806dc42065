#include "bnl_pairs.cuh"
