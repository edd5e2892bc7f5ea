// Launchers + C-ABI entry points of the fp32 GEMV family (kernel in gemv.cuh).
#include "gemv.cuh"

namespace tha {

// Scalar fallback for shapes the vector kernel cannot take (K % 4 != 0 or unaligned
// pointers: the reference's own unit shapes 3x3x3, 100^3 ... go through here).
// One warp per output row, loops over the batch.
__global__ void gemv_scalar_kernel(const float* __restrict__ W, const float* __restrict__ X, float* out,
                                   int M, int K, int B, int x_stride, long long out_batch_stride,
                                   int Coff, int has_pos, const int* __restrict__ pos) {
  const int lane = threadIdx.x & 31;
  const int m = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (m >= M) return;
  const float* w = W + (size_t)m * K;
  for (int b = 0; b < B; b++) {
    const int pb = pos ? pos[b] : 0;
    if (pos && pb < 0) continue;
    const float* x = X + (size_t)b * x_stride;
    float acc = 0.0f;
    for (int k = lane; k < K; k += 32) acc = fmaf(w[k], x[k], acc);
    acc = warp_sum(acc);
    if (lane == 0) out[Coff + (size_t)has_pos * pb + (size_t)b * out_batch_stride + m] = acc;
  }
}

template <int BT>
static int launch_bt(cudaStream_t stream, int mode, const GemvP& p, bool pdl) {
  const int nk4 = p.K / 4;
  const int kc4 = GEMV_XS_FLOATS / BT / 4;
  const size_t smem = (size_t)BT * (nk4 < kc4 ? nk4 : kc4) * sizeof(float4);
  const int Mtot = (mode == GEMV_SWIGLU) ? p.M0 : (p.M0 + p.M1 + p.M2);
  dim3 grid(ceil_div(ceil_div(Mtot, 2), GEMV_WARPS), ceil_div(p.B, BT));
  dim3 block(GEMV_THREADS);
  void (*k)(const GemvP) = nullptr;
  switch (mode) {
    case GEMV_PLAIN: k = gemv_kernel<BT, GEMV_PLAIN>; break;
    case GEMV_RESID: k = gemv_kernel<BT, GEMV_RESID>; break;
    case GEMV_QKV: k = gemv_kernel<BT, GEMV_QKV>; break;
    case GEMV_SWIGLU: k = gemv_kernel<BT, GEMV_SWIGLU>; break;
    default: return THA_STATUS_INVALID_ARG;
  }
  if (smem > 48 * 1024) {
    // opt-in once per (kernel, device); cheap to repeat
    THA_CUDA_OK(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024));
  }
  THA_CUDA_OK(launch(k, grid, block, smem, stream, pdl, p));
  return THA_STATUS_SUCCESS;
}

int gemv_launch(cudaStream_t stream, int mode, const GemvP& p, bool pdl) {
  if (p.B <= 0 || p.K <= 0 || p.M0 <= 0) return THA_STATUS_INVALID_ARG;
  if (mode == GEMV_QKV && ((p.M0 | p.M1 | p.M2 | p.head_size) & 1)) return THA_STATUS_NOT_SUPPORTED;
  if (p.B == 1) return launch_bt<1>(stream, mode, p, pdl);
  if (p.B == 2) return launch_bt<2>(stream, mode, p, pdl);
  if (p.B <= 4) return launch_bt<4>(stream, mode, p, pdl);
  return launch_bt<8>(stream, mode, p, pdl);
}

static int plain_gemv(cudaStream_t stream, int B, float* C, const float* X, const float* A, int K, int M,
                      int Coff, int has_pos, const int* pos_d, long long C_bs, int X_bs) {
  if (B <= 0 || K <= 0 || M <= 0 || !A || !X || !C || (has_pos && !pos_d)) return THA_STATUS_INVALID_ARG;
  GemvP p = {};
  p.W0 = A; p.M0 = M; p.K = K; p.B = B; p.X = X; p.x_stride = X_bs;
  p.pos = has_pos ? pos_d : nullptr;
  p.out = C; p.out_batch_stride = C_bs; p.Coff = Coff; p.has_pos = has_pos;
  if (gemv_fast_ok(p)) return gemv_launch(stream, GEMV_PLAIN, p, false);
  gemv_scalar_kernel<<<ceil_div(M, 8), 256, 0, stream>>>(A, X, C, M, K, B, X_bs, C_bs, Coff, has_pos,
                                                        has_pos ? pos_d : nullptr);
  count_launch();
  THA_LAUNCH_OK();
  return THA_STATUS_SUCCESS;
}

}  // namespace tha

using namespace tha;

extern "C" int tha_matmul_batch_f32(void* stream, int n_batches, float* C_batch, const float* B_batch,
                                    const float* A, int K, int M, int Coff, int has_pos, const int* pos_d,
                                    int C_batch_size, int B_batch_size) {
  return plain_gemv((cudaStream_t)stream, n_batches, C_batch, B_batch, A, K, M, Coff, has_pos, pos_d,
                    C_batch_size, B_batch_size);
}

extern "C" int tha_matmulvec_f32(void* stream, float* C, const float* B, const float* A, int K, int M) {
  return plain_gemv((cudaStream_t)stream, 1, C, B, A, K, M, 0, 0, nullptr, M, K);
}

extern "C" int tha_matmul_reduction_f32(void* stream, const float* A, const float* B, float* C, int M, int N, int K) {
  // B: N rows of K (the batch rows), C: N rows of M  (src/thaBLAS.cpp:281-298)
  return plain_gemv((cudaStream_t)stream, N, C, B, A, K, M, 0, 0, nullptr, M, K);
}
