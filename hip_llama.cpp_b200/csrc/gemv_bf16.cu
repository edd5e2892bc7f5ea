// Batched GEMV over bf16 weights for HBM-bound decode (1-8 slots) of the bf16 model path: the kernel of gemv.cuh with
// 128-bit loads that carry EIGHT weights instead of four (13.2 GB per Llama-2-7B token instead of 26.4).
//
// Differences to the fp32 kernel, all forced by the rule "a vector entering a weight matmul is rounded to bf16"
// (oracle/llama2_oracle.c, orc_set_bf16_activations):
//   * with a folded RMSNorm the 1/rms factor must be known BEFORE the activations are staged (the value that is rounded is
//     g * (ss * x), run.cc:196-209 order), so the CTA first sums the squares of the rows it serves (the rows are L2-resident:
//     a few KB per slot), then stages bf16-rounded values; the finished dot product is not scaled afterwards;
//   * weights are widened in registers (a bf16 is the upper half of an fp32: one shift / mask per value).
// Epilogues, PDL behaviour and row mapping are those of gemv.cuh (gemv_epilogue is shared).
#include "gemv.cuh"

namespace tha {

__device__ __forceinline__ float gb_round_bf16(float f) {
  const uint32_t u = __float_as_uint(f);
  if ((u & 0x7FFFFFFFu) > 0x7F800000u) return f;
  return __uint_as_float((u + 0x7FFFu + ((u >> 16) & 1u)) & 0xFFFF0000u);
}
__device__ __forceinline__ float dot8_bf16(const uint4& w, const float4& x0, const float4& x1, float acc) {
  acc = fmaf(__uint_as_float(w.x << 16), x0.x, acc);
  acc = fmaf(__uint_as_float(w.x & 0xFFFF0000u), x0.y, acc);
  acc = fmaf(__uint_as_float(w.y << 16), x0.z, acc);
  acc = fmaf(__uint_as_float(w.y & 0xFFFF0000u), x0.w, acc);
  acc = fmaf(__uint_as_float(w.z << 16), x1.x, acc);
  acc = fmaf(__uint_as_float(w.z & 0xFFFF0000u), x1.y, acc);
  acc = fmaf(__uint_as_float(w.w << 16), x1.z, acc);
  acc = fmaf(__uint_as_float(w.w & 0xFFFF0000u), x1.w, acc);
  return acc;
}
__device__ __forceinline__ uint4 ldg_stream_u4(const uint4* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
  return r;
}

template <int BT, int MODE>
__global__ void __launch_bounds__(GEMV_THREADS) gemv_bf16_kernel(const GemvP p) {
  constexpr int NR = (MODE == GEMV_SWIGLU) ? 4 : 2;
  constexpr int U = (NR == 4) ? 2 : 4;                // 128-bit loads (8 weights) in flight per row per lane
  constexpr int KC4 = GEMV_XS_FLOATS / BT / 4;        // float4 per staged chunk per slot
  constexpr int IPC = KC4 / (2 * 32 * U);             // inner iterations per chunk (an iteration eats 2 float4 of x per load)
  extern __shared__ float4 xs4[];
  __shared__ float red[33];
  __shared__ float scale_s[BT];

  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int nk4 = p.K >> 2, nk8 = p.K >> 3;
  const int kc4_alloc = nk4 < KC4 ? nk4 : KC4;
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
  const uint4* wrow[NR];
  {
    const uint16_t* base = reinterpret_cast<const uint16_t*>((src == 0) ? p.W0 : (src == 1 ? p.W1 : p.W2));
    const int lr1 = (lr + 1 < Msrc) ? lr + 1 : lr;
    wrow[0] = reinterpret_cast<const uint4*>(base + (size_t)lr * p.K);
    wrow[1] = reinterpret_cast<const uint4*>(base + (size_t)lr1 * p.K);
    if (MODE == GEMV_SWIGLU) {
      const uint16_t* b1 = reinterpret_cast<const uint16_t*>(p.W1);
      wrow[2] = reinterpret_cast<const uint4*>(b1 + (size_t)lr * p.K);
      wrow[3] = reinterpret_cast<const uint4*>(b1 + (size_t)lr1 * p.K);
    }
  }

  float acc[NR][BT];
#pragma unroll
  for (int r = 0; r < NR; r++)
#pragma unroll
    for (int b = 0; b < BT; b++) acc[r][b] = 0.0f;

  const int T = (nk8 + 32 * U - 1) / (32 * U);
  uint4 wcur[NR][U], wnext[NR][U];
  const uint4 z4 = make_uint4(0u, 0u, 0u, 0u);
#pragma unroll
  for (int r = 0; r < NR; r++)
#pragma unroll
    for (int u = 0; u < U; u++) {
      const int i = u * 32 + lane;
      wcur[r][u] = (i < nk8) ? ldg_stream_u4(wrow[r] + i) : z4;
    }
  pdl_wait();
  pdl_trigger();

  // ---- 1/rms of every row served by this CTA (run.cc:196-205), needed before anything is rounded
  if (p.rms_w) {
#pragma unroll
    for (int b = 0; b < BT; b++) {
      float ss = 0.0f;
      if (b0 + b < p.B) {
        const float4* xr = reinterpret_cast<const float4*>(p.X + (size_t)(b0 + b) * p.x_stride);
        for (int i = threadIdx.x; i < nk4; i += GEMV_THREADS) { const float4 v = xr[i]; ss = dot4(v, v, ss); }
      }
      const float tot = block_sum(ss, red);
      if (threadIdx.x == 0) {
        float m = tot / p.K;
        m += 1e-5f;
        scale_s[b] = 1.0f / sqrtf(m);
      }
    }
    __syncthreads();
  }

  for (int t = 0; t < T; t++) {
    const int c0 = (t / IPC) * KC4;
    if (t % IPC == 0) {
      if (t > 0) __syncthreads();
      const int n4 = (nk4 - c0) < KC4 ? (nk4 - c0) : KC4;
      for (int i = threadIdx.x; i < n4; i += GEMV_THREADS) {
        const float4 g = p.rms_w ? reinterpret_cast<const float4*>(p.rms_w)[c0 + i] : make_float4(1.f, 1.f, 1.f, 1.f);
#pragma unroll
        for (int b = 0; b < BT; b++) {
          float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
          if (b0 + b < p.B) v = reinterpret_cast<const float4*>(p.X + (size_t)(b0 + b) * p.x_stride)[c0 + i];
          if (p.rms_w) {
            const float sc = scale_s[b];
            v.x = g.x * (sc * v.x); v.y = g.y * (sc * v.y); v.z = g.z * (sc * v.z); v.w = g.w * (sc * v.w);
          }
          v.x = gb_round_bf16(v.x); v.y = gb_round_bf16(v.y); v.z = gb_round_bf16(v.z); v.w = gb_round_bf16(v.w);
          xs4[b * kc4_alloc + i] = v;
        }
      }
      __syncthreads();
    }
    if (t + 1 < T) {
#pragma unroll
      for (int r = 0; r < NR; r++)
#pragma unroll
        for (int u = 0; u < U; u++) {
          const int i = (t + 1) * 32 * U + u * 32 + lane;
          wnext[r][u] = (i < nk8) ? ldg_stream_u4(wrow[r] + i) : z4;
        }
    }
#pragma unroll
    for (int u = 0; u < U; u++) {
      const int i = t * 32 * U + u * 32 + lane;      // index of the 8-value group
      if (i < nk8) {
#pragma unroll
        for (int b = 0; b < BT; b++) {
          const float4 x0 = xs4[b * kc4_alloc + (2 * i - c0)];
          const float4 x1 = xs4[b * kc4_alloc + (2 * i + 1 - c0)];
#pragma unroll
          for (int r = 0; r < NR; r++) acc[r][b] = dot8_bf16(wcur[r][u], x0, x1, acc[r][b]);
        }
      }
    }
    if (t + 1 < T) {
#pragma unroll
      for (int r = 0; r < NR; r++)
#pragma unroll
        for (int u = 0; u < U; u++) wcur[r][u] = wnext[r][u];
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

template <int BT>
static int launch_bt_bf16(cudaStream_t stream, int mode, const GemvP& p, bool pdl) {
  const int nk4 = p.K / 4;
  const int kc4 = GEMV_XS_FLOATS / BT / 4;
  const size_t smem = (size_t)BT * (nk4 < kc4 ? nk4 : kc4) * sizeof(float4);
  const int Mtot = (mode == GEMV_SWIGLU) ? p.M0 : (p.M0 + p.M1 + p.M2);
  dim3 grid(ceil_div(ceil_div(Mtot, 2), GEMV_WARPS), ceil_div(p.B, BT));
  void (*k)(const GemvP) = nullptr;
  switch (mode) {
    case GEMV_PLAIN: k = gemv_bf16_kernel<BT, GEMV_PLAIN>; break;
    case GEMV_RESID: k = gemv_bf16_kernel<BT, GEMV_RESID>; break;
    case GEMV_QKV: k = gemv_bf16_kernel<BT, GEMV_QKV>; break;
    case GEMV_SWIGLU: k = gemv_bf16_kernel<BT, GEMV_SWIGLU>; break;
    default: return THA_STATUS_INVALID_ARG;
  }
  if (smem > 48 * 1024) THA_CUDA_OK(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024));
  THA_CUDA_OK(launch(k, grid, dim3(GEMV_THREADS), smem, stream, pdl, p));
  return THA_STATUS_SUCCESS;
}

// p.W0 / W1 / W2 carry bf16 data (cast).  K % 8 == 0, 16-byte aligned rows; a staged chunk must hold whole iterations.
int gemv_bf16_launch(cudaStream_t stream, int mode, const GemvP& p, bool pdl) {
  if (p.B <= 0 || p.K <= 0 || p.M0 <= 0 || p.K % 8 || p.x_stride % 4) return THA_STATUS_INVALID_ARG;
  if (mode == GEMV_QKV && ((p.M0 | p.M1 | p.M2 | p.head_size) & 1)) return THA_STATUS_NOT_SUPPORTED;
  if (p.B == 1) return launch_bt_bf16<1>(stream, mode, p, pdl);
  if (p.B == 2) return launch_bt_bf16<2>(stream, mode, p, pdl);
  if (p.B <= 4) return launch_bt_bf16<4>(stream, mode, p, pdl);
  return launch_bt_bf16<8>(stream, mode, p, pdl);
}

}  // namespace tha
