// Q8_0 batched GEMV (runq.c:317-342 semantics) for HBM-bound decode.
//
//   out[b, m] = sum_g  (float)(sum_{k in g} xq[b,k] * wq[m,k])  *  ws[m,g]  *  xs[b,g]
//
// * Activations are quantised ONCE per matmul input by q8_prep_kernel, bit-exactly as
//   runq.c:145-171 does (max|x|/127 scale, round-half-away), optionally after the
//   RMSNorm of run.cc:196-209 evaluated in the reference's operation order
//   w * (ss * x); the GEMV CTAs only copy the int8 rows + group scales into shared memory.
// * A warp owns a pair of consecutive rows (for w1|w3: the pair of each); lanes stride over
//   the rows with 128-bit loads (16 values) and dp4a.  A lane's 16-value partial is an exact
//   int32; it is scaled by the group's weight and activation scales and accumulated in fp32
//   (the GS/16 lanes of a group use the same scales, so this equals the reference's per-group
//   product up to fp32 rounding of the final sums).
// * Same epilogues and PDL weight prefetch as the fp32 kernel (gemv.cuh).
#include "gemv.cuh"

namespace tha {

constexpr int Q8_MAX_SMEM = 160 * 1024;

template <int BT, int MODE>
__global__ void __launch_bounds__(GEMV_THREADS) gemv_q8_kernel(const GemvP p) {
  constexpr int NR = (MODE == GEMV_SWIGLU) ? 4 : 2;  // weight rows per warp: a row pair (of w1 and of w3)
  constexpr int U = (BT <= 2) ? ((NR == 4) ? 4 : 8) : ((NR == 4) ? 2 : 4);   // 16-byte chunks per row in flight per lane
  extern __shared__ int4 smem_q8[];
  const int K = p.K, gs = p.gs, nchunk = K >> 4, ngrp = K / gs;
  const int lpg_shift = 31 - __clz(gs >> 4);         // chunk -> group: c >> lpg_shift (gs/16 is a power of two)
  int4* xq4 = smem_q8;                                             // [BT][nchunk]
  float* xsc = reinterpret_cast<float*>(smem_q8 + BT * nchunk);    // [BT][ngrp]

  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int b0 = blockIdx.y * BT;
  const int Mtot = (MODE == GEMV_SWIGLU) ? p.M0 : (p.M0 + p.M1 + p.M2);
  const int vr = (blockIdx.x * GEMV_WARPS + warp) * 2;
  const bool valid = vr < Mtot;
  int lr = valid ? vr : 0, src = 0, Msrc = p.M0;
  if (MODE == GEMV_QKV) {
    if (lr >= p.M0) { lr -= p.M0; src = 1; Msrc = p.M1;
      if (lr >= p.M1) { lr -= p.M1; src = 2; Msrc = p.M2; } }
  }
  const bool has2 = valid && (lr + 1 < Msrc);
  const int4* wrow[NR]; const float* srow[NR];
  {
    const int8_t* Q = (src == 0) ? p.Q0 : (src == 1 ? p.Q1 : p.Q2);
    const float* S = (src == 0) ? p.S0 : (src == 1 ? p.S1 : p.S2);
    const int lr1 = (lr + 1 < Msrc) ? lr + 1 : lr;
    wrow[0] = reinterpret_cast<const int4*>(Q + (size_t)lr * K);   srow[0] = S + (size_t)lr * ngrp;
    wrow[1] = reinterpret_cast<const int4*>(Q + (size_t)lr1 * K);  srow[1] = S + (size_t)lr1 * ngrp;
    if (NR == 4) {
      wrow[2] = reinterpret_cast<const int4*>(p.Q1 + (size_t)lr * K);   srow[2] = p.S1 + (size_t)lr * ngrp;
      wrow[3] = reinterpret_cast<const int4*>(p.Q1 + (size_t)lr1 * K);  srow[3] = p.S1 + (size_t)lr1 * ngrp;
    }
  }

  // ---- weight (and weight-scale) stream: U chunks per row in registers, the next U prefetched.
  // The first loads go out before the dependency wait: weights never depend on the previous kernel.
  // ---- weight (and weight-scale) stream: U chunks of every row in flight per lane (K = 4096: a whole
  // row pair per pass); enough warps are resident (2-3 CTAs per SM) to cover the compute phases.
  // The first pass goes out before the dependency wait: weights never depend on the previous kernel.
  const int T = (nchunk + 32 * U - 1) / (32 * U);
  int4 wcur[U][NR];
  float scur[U][NR];
  const int4 zi = make_int4(0, 0, 0, 0);
#pragma unroll
  for (int u = 0; u < U; u++) {
    const int c = u * 32 + lane;
#pragma unroll
    for (int r = 0; r < NR; r++) {
      wcur[u][r] = (c < nchunk) ? ldg_stream_i4(wrow[r] + c) : zi;
      scur[u][r] = (c < nchunk) ? __ldg(srow[r] + (c >> lpg_shift)) : 0.0f;
    }
  }
  pdl_wait();
  pdl_trigger();

  // ---- the BT quantised activation rows -> shared memory ----
  for (int bb = 0; bb < BT; bb++) {
    const int b = (b0 + bb < p.B) ? b0 + bb : 0;
    for (int i = threadIdx.x; i < nchunk; i += GEMV_THREADS)
      xq4[bb * nchunk + i] = reinterpret_cast<const int4*>(p.XQ + (size_t)b * K)[i];
    for (int i = threadIdx.x; i < ngrp; i += GEMV_THREADS) xsc[bb * ngrp + i] = p.XS[(size_t)b * ngrp + i];
  }
  __syncthreads();

  float acc[NR][BT];
#pragma unroll
  for (int r = 0; r < NR; r++)
#pragma unroll
    for (int b = 0; b < BT; b++) acc[r][b] = 0.0f;

  for (int t = 0; t < T; t++) {
#pragma unroll
    for (int u = 0; u < U; u++) {
      const int c = (t * U + u) * 32 + lane;
      if (c < nchunk) {
        const int g = c >> lpg_shift;
#pragma unroll
        for (int b = 0; b < BT; b++) {
          const int4 xv = xq4[b * nchunk + c];
          const float xsv = xsc[b * ngrp + g];
#pragma unroll
          for (int r = 0; r < NR; r++) {
            const int4 wv = wcur[u][r];
            int iv = __dp4a(wv.x, xv.x, 0);
            iv = __dp4a(wv.y, xv.y, iv);
            iv = __dp4a(wv.z, xv.z, iv);
            iv = __dp4a(wv.w, xv.w, iv);
            // this lane's 16 of the group's values: an exact int32; scaled in the reference's product
            // order (runq.c:334) and accumulated in fp32
            acc[r][b] += ((float)iv) * scur[u][r] * xsv;
          }
        }
      }
    }
    if (t + 1 < T) {
#pragma unroll
      for (int u = 0; u < U; u++) {
        const int c = ((t + 1) * U + u) * 32 + lane;
#pragma unroll
        for (int r = 0; r < NR; r++) {
          wcur[u][r] = (c < nchunk) ? ldg_stream_i4(wrow[r] + c) : zi;
          scur[u][r] = (c < nchunk) ? __ldg(srow[r] + (c >> lpg_shift)) : 0.0f;
        }
      }
    }
  }

  float mine[NR];
#pragma unroll
  for (int r = 0; r < NR; r++) mine[r] = 0.0f;
#pragma unroll
  for (int r = 0; r < NR; r++)
#pragma unroll
    for (int b = 0; b < BT; b++) {
      const float v = warp_sum(acc[r][b]);
      if (lane == b) mine[r] = v;
    }
  gemv_epilogue<MODE, NR>(p, mine, valid && lane < BT, b0 + lane, lr, src, has2, 1.0f);
}

// (RMSNorm +) quantise one activation row per block: x[b, 0:K] -> xq[b, 0:K] int8, xs[b, 0:K/gs].
// Bit-exact with runq.c:145-171 given the same fp32 input; with `rms_w` the input is
// w * (ss * x), ss = 1/sqrt(mean(x^2) + 1e-5), the reference's operation order (run.cc:196-209).
__global__ void __launch_bounds__(256) q8_prep_kernel(const float* __restrict__ X, int x_stride, const float* __restrict__ rms_w,
                                                      const int* __restrict__ pos, int K, int gs, int8_t* __restrict__ xq,
                                                      float* __restrict__ xs) {
  __shared__ float red[33];
  pdl_wait();
  pdl_trigger();
  const int b = blockIdx.x, lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
  const int ngrp = K / gs;
  if (pos && pos[b] < 0) {   // idle slot: zeros in, zeros out
    for (int i = threadIdx.x; i < K; i += blockDim.x) xq[(size_t)b * K + i] = 0;
    for (int i = threadIdx.x; i < ngrp; i += blockDim.x) xs[(size_t)b * ngrp + i] = 0.0f;
    return;
  }
  const float* x = X + (size_t)b * x_stride;
  float ssn = 1.0f;
  if (rms_w) {
    float ss = 0.0f;
    for (int i = threadIdx.x; i < K; i += blockDim.x) { float t = x[i]; ss = fmaf(t, t, ss); }
    ss = block_sum(ss, red);
    ss /= K; ss += 1e-5f;
    ssn = 1.0f / sqrtf(ss);
  }
  for (int g = warp; g < ngrp; g += nw) {
    float amax = 0.0f;
    for (int i = lane; i < gs; i += 32) {
      float v = x[g * gs + i];
      if (rms_w) v = rms_w[g * gs + i] * (ssn * v);
      amax = fmaxf(amax, fabsf(v));
    }
    amax = warp_max(amax);
    const float scale = __fdiv_rn(amax, 127.0f);
    if (lane == 0) xs[(size_t)b * ngrp + g] = scale;
    for (int i = lane; i < gs; i += 32) {
      float v = x[g * gs + i];
      if (rms_w) v = rms_w[g * gs + i] * (ssn * v);
      const float r = __fdiv_rn(v, scale);
      xq[(size_t)b * K + g * gs + i] = (r != r) ? (int8_t)0 : (int8_t)__float2int_rn(roundf(r));
    }
  }
}

int q8_prep_launch(cudaStream_t st, int B, const float* X, int x_stride, const float* rms_w, const int* pos, int K, int gs,
                   int8_t* xq, float* xs, bool pdl) {
  THA_CUDA_OK(launch(q8_prep_kernel, dim3(B), dim3(256), 0, st, pdl, X, x_stride, rms_w, pos, K, gs, xq, xs));
  return THA_STATUS_SUCCESS;
}

template <int BT>
static int launch_q8_bt(cudaStream_t stream, int mode, const GemvP& p, bool pdl) {
  const size_t smem = (size_t)BT * p.K + (size_t)BT * (p.K / p.gs) * sizeof(float);
  if (smem > (size_t)Q8_MAX_SMEM) return THA_STATUS_NOT_SUPPORTED;
  const int Mtot = (mode == GEMV_SWIGLU) ? p.M0 : (p.M0 + p.M1 + p.M2);
  dim3 grid(ceil_div(ceil_div(Mtot, 2), GEMV_WARPS), ceil_div(p.B, BT));
  void (*k)(const GemvP) = nullptr;
  switch (mode) {
    case GEMV_PLAIN: k = gemv_q8_kernel<BT, GEMV_PLAIN>; break;
    case GEMV_RESID: k = gemv_q8_kernel<BT, GEMV_RESID>; break;
    case GEMV_QKV: k = gemv_q8_kernel<BT, GEMV_QKV>; break;
    case GEMV_SWIGLU: k = gemv_q8_kernel<BT, GEMV_SWIGLU>; break;
    default: return THA_STATUS_INVALID_ARG;
  }
  if (smem > 40 * 1024) THA_CUDA_OK(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, Q8_MAX_SMEM));
  THA_CUDA_OK(launch(k, grid, dim3(GEMV_THREADS), smem, stream, pdl, p));
  return THA_STATUS_SUCCESS;
}

// Requirements: K multiple of 16 and of gs; gs in {16, 32, 64, 128, ...} with gs/16 a power of
// two <= 32; 16-byte aligned int8 rows.
int gemv_q8_launch(cudaStream_t stream, int mode, const GemvP& p, bool pdl) {
  if (p.B <= 0 || p.K <= 0 || p.M0 <= 0 || p.gs <= 0 || !p.Q0 || !p.S0 || !p.XQ || !p.XS || !aligned16(p.XQ)) return THA_STATUS_INVALID_ARG;
  const int lpg = p.gs / 16;
  if (p.K % 16 || p.K % p.gs || p.gs % 16 || (lpg & (lpg - 1)) || lpg > 32) return THA_STATUS_NOT_SUPPORTED;
  if (!aligned16(p.Q0) || (p.Q1 && !aligned16(p.Q1)) || (p.Q2 && !aligned16(p.Q2))) return THA_STATUS_NOT_SUPPORTED;
  if (mode == GEMV_QKV && ((p.M0 | p.M1 | p.M2 | p.head_size) & 1)) return THA_STATUS_NOT_SUPPORTED;
  if (p.B == 1) return launch_q8_bt<1>(stream, mode, p, pdl);
  if (p.B == 2) return launch_q8_bt<2>(stream, mode, p, pdl);
  if (p.B <= 4) return launch_q8_bt<4>(stream, mode, p, pdl);
  return launch_q8_bt<8>(stream, mode, p, pdl);
}

}  // namespace tha

using namespace tha;

extern "C" int tha_matmul_batch_q8(void* stream, int n_batches, float* out, int out_stride, const int8_t* xq,
                                   const float* xs, const int8_t* wq, const float* ws, int K, int M, int gs) {
  if (n_batches <= 0 || !out || !xq || !xs || !wq || !ws || K <= 0 || M <= 0 || gs <= 0) return THA_STATUS_INVALID_ARG;
  GemvP p = {};
  p.Q0 = wq; p.S0 = ws; p.M0 = M; p.K = K; p.B = n_batches; p.gs = gs; p.XQ = xq; p.XS = xs;
  p.out = out; p.out_batch_stride = out_stride;
  return gemv_q8_launch((cudaStream_t)stream, GEMV_PLAIN, p, false);
}
