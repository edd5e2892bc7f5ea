// thaBLAS_s_matmul (src/thaBLAS.cpp:134-170): C[M,N] = A[M,K] . B[K,N], row-major fp32.
// Shapes of at least 32 x 64 x 64 with K % 4 == 0 run on the tensor cores (matmul_tc_route, gemm_tc.cu: tcgen05 3xTF32, fp32
// accuracy); everything else on this exact-fp32 shared-memory tiled kernel (64x64 tile, 4x4 micro-tile per thread).
#include "common.cuh"

namespace tha {

int matmul_tc_route(cudaStream_t st, int m, int n, int k, const float* A, const float* B, float* C);   // gemm_tc.cu

constexpr int TM = 64, TN = 64, TK = 16;

__global__ void __launch_bounds__(256) sgemm_tiled_kernel(int M, int N, int K, const float* __restrict__ A,
                                                          const float* __restrict__ B, float* __restrict__ C) {
  __shared__ float As[TK][TM + 4];
  __shared__ float Bs[TK][TN + 4];
  const int tx = threadIdx.x % 16, ty = threadIdx.x / 16;
  const int m0 = blockIdx.y * TM, n0 = blockIdx.x * TN;
  float acc[4][4] = {};
  for (int k0 = 0; k0 < K; k0 += TK) {
    for (int i = threadIdx.x; i < TM * TK; i += 256) {
      int r = i / TK, c = i % TK;
      As[c][r] = (m0 + r < M && k0 + c < K) ? A[(size_t)(m0 + r) * K + k0 + c] : 0.0f;
    }
    for (int i = threadIdx.x; i < TK * TN; i += 256) {
      int r = i / TN, c = i % TN;
      Bs[r][c] = (k0 + r < K && n0 + c < N) ? B[(size_t)(k0 + r) * N + n0 + c] : 0.0f;
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < TK; k++) {
      float a[4], b[4];
#pragma unroll
      for (int i = 0; i < 4; i++) a[i] = As[k][ty * 4 + i];
#pragma unroll
      for (int j = 0; j < 4; j++) b[j] = Bs[k][tx * 4 + j];
#pragma unroll
      for (int i = 0; i < 4; i++)
#pragma unroll
        for (int j = 0; j < 4; j++) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 4; i++)
#pragma unroll
    for (int j = 0; j < 4; j++) {
      int m = m0 + ty * 4 + i, n = n0 + tx * 4 + j;
      if (m < M && n < N) C[(size_t)m * N + n] = acc[i][j];
    }
}

}  // namespace tha

using namespace tha;

extern "C" int tha_matmul_f32(void* stream, int m, int n, int k, const float* A, const float* B, float* C) {
  if (m <= 0 || n <= 0 || k <= 0 || !A || !B || !C) {
    printf("THABLAS ERROR: INVALID ARGUMENT\n");
    fflush(stdout);
    return THA_STATUS_INVALID_ARG;
  }
  {  // shapes worth a tensor-core launch go to the tcgen05 3xTF32 kernel (gemm_tc.cu); the rest stay on CUDA cores
    const int st = matmul_tc_route((cudaStream_t)stream, m, n, k, A, B, C);
    if (st != THA_STATUS_NOT_SUPPORTED) return st;
  }
  dim3 grid(ceil_div(n, TN), ceil_div(m, TM));
  sgemm_tiled_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(m, n, k, A, B, C);
  count_launch();
  THA_LAUNCH_OK();
  return THA_STATUS_SUCCESS;
}
