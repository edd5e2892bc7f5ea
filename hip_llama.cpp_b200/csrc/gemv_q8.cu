// Q8_0 batched GEMV (runq.c:317-342 semantics) for HBM-bound decode.
//
//   out[b, m] = sum_g  (float)(sum_{k in g} xq[b,k] * wq[m,k])  *  ws[m,g]  *  xs[b,g]
//
// * Activations are quantised INSIDE the kernel prologue, bit-exactly as
//   runq.c:145-171 does (max|x|/127 scale, round-half-away), optionally after the
//   RMSNorm of run.cc:196-209 evaluated in the reference's operation order
//   w * (ss * x); int8 values + group scales live in shared memory only.
// * A warp owns two consecutive rows = one contiguous stream of 2*K int8; lanes
//   stride over it with 128-bit loads (16 values) and dp4a; the GS/16 lanes that
//   share a group add their int32 partials with shuffles, so every group sum is an
//   exact int32 before it meets the fp32 scales -- as in the reference.
// * Same epilogues and PDL weight prefetch as the fp32 kernel (gemv.cuh).
#include "gemv.cuh"

namespace tha {

constexpr int Q8_MAX_SMEM = 160 * 1024;

template <int BT, int MODE>
__global__ void __launch_bounds__(GEMV_THREADS) gemv_q8_kernel(const GemvP p) {
  constexpr int NS = (MODE == GEMV_SWIGLU) ? 2 : 1;  // row-pair streams per warp (w1 and w3)
  constexpr int U = 3;                               // int4 loads in flight per stream per lane
  extern __shared__ int4 smem_q8[];
  __shared__ float red[33];
  const int K = p.K, gs = p.gs, nchunk = K >> 4, ngrp = K / gs, lpg = gs >> 4;  // lanes per group
  int4* xq4 = smem_q8;                                             // [BT][nchunk]
  float* xsc = reinterpret_cast<float*>(smem_q8 + BT * nchunk);    // [BT][ngrp]
  int8_t* xq1 = reinterpret_cast<int8_t*>(xq4);

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
  const int nrow = has2 ? 2 : 1;
  const int tot = nrow * nchunk;                      // int4 chunks in this warp's stream
  const int8_t* qb[NS]; const float* sb[NS];
  {
    const int8_t* Q = (src == 0) ? p.Q0 : (src == 1 ? p.Q1 : p.Q2);
    const float* S = (src == 0) ? p.S0 : (src == 1 ? p.S1 : p.S2);
    qb[0] = Q + (size_t)lr * K;  sb[0] = S + (size_t)lr * ngrp;
    if (NS == 2) { qb[1] = p.Q1 + (size_t)lr * K; sb[1] = p.S1 + (size_t)lr * ngrp; }
  }

  const int T = (tot + 32 * U - 1) / (32 * U);
  int4 wcur[NS][U], wnext[NS][U];
  const int4 zi = make_int4(0, 0, 0, 0);
#pragma unroll
  for (int s = 0; s < NS; s++)
#pragma unroll
    for (int u = 0; u < U; u++) {
      const int c = u * 32 + lane;
      wcur[s][u] = (c < tot) ? ldg_stream_i4(reinterpret_cast<const int4*>(qb[s]) + c) : zi;
    }
  pdl_wait();
  pdl_trigger();

  // ---- prologue: (rmsnorm) + quantise the BT activation rows into shared memory ----
  if (p.XQ) {  // activations already quantised by the caller (tha_matmul_batch_q8)
    for (int bb = 0; bb < BT; bb++) {
      const int b = (b0 + bb < p.B) ? b0 + bb : 0;
      for (int i = threadIdx.x; i < nchunk; i += GEMV_THREADS)
        xq4[bb * nchunk + i] = reinterpret_cast<const int4*>(p.XQ + (size_t)b * K)[i];
      for (int i = threadIdx.x; i < ngrp; i += GEMV_THREADS) xsc[bb * ngrp + i] = p.XS[(size_t)b * ngrp + i];
    }
  } else
  for (int bb = 0; bb < BT; bb++) {
    const int b = b0 + bb;
    const float* x = p.X + (size_t)(b < p.B ? b : 0) * p.x_stride;
    float ssn = 1.0f;
    if (p.rms_w) {
      float ss = 0.0f;
      for (int i = threadIdx.x; i < K; i += GEMV_THREADS) { float t = x[i]; ss = fmaf(t, t, ss); }
      ss = block_sum(ss, red);
      ss /= K; ss += 1e-5f;
      ssn = 1.0f / sqrtf(ss);
    }
    for (int g = warp; g < ngrp; g += GEMV_WARPS) {
      float amax = 0.0f;
      for (int i = lane; i < gs; i += 32) {
        float v = x[g * gs + i];
        if (p.rms_w) v = p.rms_w[g * gs + i] * (ssn * v);
        amax = fmaxf(amax, fabsf(v));
      }
      amax = warp_max(amax);
      const float scale = __fdiv_rn(amax, 127.0f);
      if (lane == 0) xsc[bb * ngrp + g] = scale;
      for (int i = lane; i < gs; i += 32) {
        float v = x[g * gs + i];
        if (p.rms_w) v = p.rms_w[g * gs + i] * (ssn * v);
        const float r = __fdiv_rn(v, scale);
        xq1[(size_t)bb * K + g * gs + i] = (r != r) ? (int8_t)0 : (int8_t)__float2int_rn(roundf(r));
      }
    }
  }
  __syncthreads();

  float acc[NS][2][BT];
#pragma unroll
  for (int s = 0; s < NS; s++)
#pragma unroll
    for (int r = 0; r < 2; r++)
#pragma unroll
      for (int b = 0; b < BT; b++) acc[s][r][b] = 0.0f;

  for (int t = 0; t < T; t++) {
    if (t + 1 < T) {
#pragma unroll
      for (int s = 0; s < NS; s++)
#pragma unroll
        for (int u = 0; u < U; u++) {
          const int c = (t + 1) * 32 * U + u * 32 + lane;
          wnext[s][u] = (c < tot) ? ldg_stream_i4(reinterpret_cast<const int4*>(qb[s]) + c) : zi;
        }
    }
#pragma unroll
    for (int u = 0; u < U; u++) {
      const int c = t * 32 * U + u * 32 + lane;
      const bool in = c < tot;                       // warp-uniform per group of lpg lanes
      const int row = (in && c >= nchunk) ? 1 : 0;
      const int cr = in ? c - row * nchunk : 0;       // chunk within the row
      const int g = cr / lpg;
      float wsv[NS];
#pragma unroll
      for (int s = 0; s < NS; s++) wsv[s] = in ? sb[s][row * ngrp + g] : 0.0f;
#pragma unroll
      for (int b = 0; b < BT; b++) {
        const int4 xv = xq4[b * nchunk + cr];
        const float xsv = xsc[b * ngrp + g];
#pragma unroll
        for (int s = 0; s < NS; s++) {
          const int4 wv = wcur[s][u];
          int iv = __dp4a(wv.x, xv.x, 0);
          iv = __dp4a(wv.y, xv.y, iv);
          iv = __dp4a(wv.z, xv.z, iv);
          iv = __dp4a(wv.w, xv.w, iv);
          for (int o = lpg >> 1; o > 0; o >>= 1) iv += __shfl_xor_sync(THA_FULL_MASK, iv, o);
          if (in && (lane & (lpg - 1)) == 0) {
            const float v = ((float)iv) * wsv[s] * xsv;    // runq.c:334 product order
            if (row == 0) acc[s][0][b] += v; else acc[s][1][b] += v;
          }
        }
      }
    }
    if (t + 1 < T) {
#pragma unroll
      for (int s = 0; s < NS; s++)
#pragma unroll
        for (int u = 0; u < U; u++) wcur[s][u] = wnext[s][u];
    }
  }

  constexpr int NR = NS * 2;
  float mine[NR];
#pragma unroll
  for (int r = 0; r < NR; r++) mine[r] = 0.0f;
#pragma unroll
  for (int s = 0; s < NS; s++)
#pragma unroll
    for (int r = 0; r < 2; r++)
#pragma unroll
      for (int b = 0; b < BT; b++) {
        const float v = warp_sum(acc[s][r][b]);
        if (lane == b) mine[s * 2 + r] = v;
      }
  gemv_epilogue<MODE, NR>(p, mine, valid && lane < BT, b0 + lane, lr, src, has2, 1.0f);
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
  if (p.B <= 0 || p.K <= 0 || p.M0 <= 0 || p.gs <= 0 || !p.Q0 || !p.S0 || (!p.X && !p.XQ)) return THA_STATUS_INVALID_ARG;
  if (p.XQ && (!p.XS || !aligned16(p.XQ))) return THA_STATUS_INVALID_ARG;
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
