// reference: include/thaDNN/thaDNN_rmsnorm.hpp:4-10, src/thaDNN/thaDNN_rmsnorm.cpp:35-83
#pragma once
#include "thaBLAS.hpp"

// o[b,i] = weight[i] * x[b,i] / sqrt(mean_i(x[b,:]^2) + 1e-5) for b < n_batches; rows are
// `dim` floats apart, the first `size` of each are normalised; o may alias x.
thablasStatus_t thaDNN_s_rmsnorm_v2_batch(thablasHandle_t* handle, int n_batches, float* o_batch, float* x_batch,
                                          float* weight, int size, int dim);
