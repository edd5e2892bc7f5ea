// thaBLAS.hpp -- the thaBLAS call surface of the reference (include/thaBLAS.hpp:1-143),
// kept name for name and argument for argument, on top of the B200 C-ABI (tha_b200.h).
//
// Host code that was written against the reference includes this header unchanged.  The
// implementation (hip_llama.cpp_b200/host/thaBLAS.cpp) contains no CUDA: every call forwards
// to an `extern "C"` launcher in libtha_b200.so.  The only type that changes is the stream:
// hipStream_t becomes cudaStream_t (declared below exactly as the CUDA runtime declares it,
// so including <cuda_runtime.h> next to this header is fine).
#pragma once

#include <stddef.h>

#ifndef __DRIVER_TYPES_H__
typedef struct CUstream_st* cudaStream_t;
#endif

// Status codes: same names and values as the reference (include/thaBLAS.hpp:5-19).
typedef enum {
  THABLAS_STATUS_SUCCESS = 0,
  THABLAS_STATUS_NOT_INITIALIZED = 1,
  THABLAS_STATUS_ALLOC_FAILED = 2,  // what the reference returns for an invalid argument (src/thaBLAS.cpp:80-84)
  THABLAS_STATUS_INVALID_VALUE = 3,
  THABLAS_STATUS_MAPPING_ERROR = 4,
  THABLAS_STATUS_EXECUTION_FAILED = 5,
  THABLAS_STATUS_INTERNAL_ERROR = 6,
  THABLAS_STATUS_NOT_SUPPORTED = 7,
  THABLAS_STATUS_ARCH_MISMATCH = 8,
  THABLAS_STATUS_HANDLE_IS_NULLPTR = 9,
  THABLAS_STATUS_INVALID_ENUM = 10,
  THABLAS_STATUS_UNKNOWN = 11,
} thablasStatus_t;

// include/thaBLAS.hpp:21-25
typedef struct {
  int current_gpu_id;
  cudaStream_t calc_stream;
  cudaStream_t copy_stream;
} thablasHandle_t;

// src/thaBLAS.cpp:49-64: current device + two streams; Destroy releases them (the reference's is a no-op).
thablasStatus_t thablasCreate(thablasHandle_t* handle);
thablasStatus_t thablasDestroy(thablasHandle_t handle);

// ---- level 1 ---------------------------------------------------------------------------
// B = A / val (src/thaBLAS.cpp:72-95); n <= 0 or a null pointer -> ALLOC_FAILED after
// printing "THABLAS ERROR: INVALID ARGUMENT", as the reference does.
thablasStatus_t thablas_Svds(thablasHandle_t handle, int n, float* A, float* B, float val);
// a += b (src/thaBLAS.cpp:104-126)
thablasStatus_t thaBLAS_s_vecaddvec(thablasHandle_t* handle, float* a, float* b, int size);

// ---- level 2 ---------------------------------------------------------------------------
// C[M] = A[M,K] . B[K] (src/thaBLAS.cpp:241-277)
thablasStatus_t thaBLAS_s_matmulvec(thablasHandle_t handle, float* C, float* B, float* A, int K, int M);
thablasStatus_t thaDNN_s_matmulvec_v2(thablasHandle_t handle, float* C, float* B, float* A, int K, int M);
// the header of the reference declares the by-value form and its .cpp defines a by-pointer
// overload; both exist here
thablasStatus_t thaDNN_s_matmulvec_v2(thablasHandle_t* handle, float* C, float* B, float* A, int K, int M);

// ---- level 3 ---------------------------------------------------------------------------
// C[m,n] = A[m,k] . B[k,n], row-major (src/thaBLAS.cpp:134-170)
thablasStatus_t thaBLAS_s_matmul(thablasHandle_t handle, int m, int n, int k, float* A, float* B, float* C);

// The weight matmul of the decode step (src/thaBLAS.cpp:191-228): for b < n_batches
//   C_b[0:M] = A[M,K] . B_b[K],  C_b = C_batch + Coff + has_pos*pos_d[b] + b*C_batch_size,
//   B_b = B_batch + b*B_batch_size.
thablasStatus_t thaBLAS_s_matmul_batch(thablasHandle_t* handle, int n_batches, float* C_batch, float* B_batch,
                                       float* A, int K, int M, int Coff, int has_pos, int pos_d[],
                                       int C_batch_size, int B_batch_size);

// Skinny GEMM with batch rows (src/thaBLAS.cpp:281-319): A [M,K] row-major, B holds N rows of K,
// C holds N rows of M.
thablasStatus_t thaBLAS_s_matmul_reduction(thablasHandle_t* handle, float* A, float* B, float* C, int M, int N, int K);
// src/thaBLAS.cpp:321-351 (the reference's unfinished MFMA kernel) and :353-359 (compile-time
// switch between the two): same contract as thaBLAS_s_matmul_reduction.
thablasStatus_t thaBLAS_s_sgemm_Mx16xK(thablasHandle_t* handle, float* d_A, float* d_B, float* d_D, int M, int N, int K);
thablasStatus_t thaBLAS_s_matmul_ifdef(thablasHandle_t* handle, float* d_A, float* d_B, float* d_D, int M, int N, int K);
