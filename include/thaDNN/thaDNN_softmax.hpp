// reference: include/thaDNN/thaDNN_softmax.hpp:5-7, src/thaDNN/thaDNN_softmax.cpp:62-119
#pragma once
#include "thaBLAS.hpp"

// In-place softmax of one device vector (max-subtract, expf, divide).  size <= 0 or a null
// pointer -> ALLOC_FAILED after "THABLAS ERROR: INVALID ARGUMENT" (:103-106).  The reference's
// header takes the handle by pointer and its .cpp by value: both exist.
thablasStatus_t thaDNN_s_softmax_v2(thablasHandle_t* handle, float* x, int size);
thablasStatus_t thaDNN_s_softmax_v2(thablasHandle_t handle, float* x, int size);
