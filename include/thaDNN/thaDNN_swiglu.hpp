// reference: include/thaDNN/thaDNN_swiglu.hpp:4-7, src/thaDNN/thaDNN_swiglu.cpp:5-35
#pragma once
#include "thaBLAS.hpp"

// hb[i] = hb[i] * (1 / (1 + expf(-hb[i]))) * hb2[i], in place
thablasStatus_t thaDNN_s_swiglu(thablasHandle_t* handle, float* hb, float* hb2, int hidden_dim);
