// reference: include/thaDNN/thaDNN_rope.hpp:5-11, src/thaDNN/thaDNN_rope.cpp:25-63
#pragma once
#include "thaBLAS.hpp"

// Rotate the pairs (v[i], v[i+1]), i even < dim, of q -- and of k while i < kv_dim -- by the
// angle pos / 10000^((i % head_size) / head_size).  One slot per call, in place.
thablasStatus_t thaDNN_s_rope(thablasHandle_t* handle, int dim, int head_size, int kv_dim, int pos, float* q, float* k);
