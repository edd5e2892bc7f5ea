// Batched fp32 GEMV for HBM-bound decode: out[b, m] = sum_k W[m, k] * x[b, k].
//
// One pass over the weights serves every slot of a batch tile (the reference
// re-reads the weight row once per slot: grid (M, B), src/thaBLAS.cpp:191-228).
//
// Mapping.  A warp owns two consecutive output rows (a RoPE pair) for the whole
// kernel; its 32 lanes stride over K with 128-bit streaming loads, U loads per
// row in flight plus the next U prefetched.  The activations of the batch tile
// (BT slots) are staged in shared memory in chunks of XS_FLOATS/BT floats, with
// the RMSNorm weight folded in on the way (the 1/rms scale is applied to the
// finished dot product).  blockIdx.y walks batch tiles when B > BT.
//
// Programmatic dependent launch: the first weight loads are issued BEFORE
// griddepcontrol.wait -- weights never depend on the previous kernel -- so the
// HBM stream of kernel N+1 overlaps the tail of kernel N.
//
// Epilogues (MODE):
//   PLAIN   reference addressing  C + Coff + has_pos*pos[b] + b*C_batch_size   (thaBLAS_s_matmul_batch)
//   RESID   x[b, m] += dot                       (wo, w2 + the residual add, src/thaDNN.cpp:56-59,69-72)
//   QKV     three stacked matrices; RoPE on q and k rows from host-built cos/sin
//           tables; k and v rows land in the KV cache row of pos[b]   (src/thaDNN.cpp:44-49)
//   SWIGLU  two matrices (w1, w3): hb[b, m] = silu(w1.x) * (w3.x)      (src/thaDNN.cpp:63-67)
#pragma once
#include "common.cuh"

namespace tha {

constexpr int GEMV_WARPS = 8;
constexpr int GEMV_THREADS = GEMV_WARPS * 32;
constexpr int GEMV_XS_FLOATS = 16384;  // staged activations per CTA (64 KB)

enum GemvMode { GEMV_PLAIN = 0, GEMV_RESID = 1, GEMV_QKV = 2, GEMV_SWIGLU = 3 };

struct GemvP {
  const float* W0; const float* W1; const float* W2;  // stacked row sources (same K)
  int M0, M1, M2;
  int K, B;
  const float* X; int x_stride;     // activations [B, x_stride]
  const float* rms_w;               // nullable: RMSNorm weight folded into the staging
  const int* pos;                   // nullable: pos[b] < 0 marks an idle slot
  float* out; long long out_batch_stride; int Coff, has_pos;   // PLAIN / RESID / SWIGLU / q rows of QKV
  float* kc; float* vc; long long kv_batch_stride; int loff, kv_dim, head_size;  // QKV
  const float* rope_cos; const float* rope_sin;  // [seq_len, head_size/2]
  // Q8_0 variant (gemv_q8.cu): int8 rows + one fp32 scale per `gs` values (runq.c:34-37)
  const int8_t* Q0; const int8_t* Q1; const int8_t* Q2; const float* S0; const float* S1; const float* S2; int gs;
  const int8_t* XQ; const float* XS;   // optional pre-quantised activations [B,K] / [B,K/gs] (standalone matmul API)
};

#ifdef __CUDACC__

__device__ __forceinline__ float silu_mul(float a, float c) {
  float v = a;
  v *= (1.0f / (1.0f + expf(-v)));   // same operation order as run.cc:344-347
  v *= c;
  return v;
}

// Shared epilogue of the fp32 and Q8_0 GEMV kernels.  `mine[r]` holds this lane's
// finished dot products (row pair lr, lr+1 of source `src`) for slot b.
template <int MODE, int NR>
__device__ __forceinline__ void gemv_epilogue(const GemvP& p, float (&mine)[NR], bool active, int b, int lr,
                                              int src, bool has2, float scale) {
  if (!active || b >= p.B) return;
  const int pb = p.pos ? p.pos[b] : 0;
  if (pb < 0) return;  // idle slot
#pragma unroll
  for (int r = 0; r < NR; r++) mine[r] *= scale;

  if (MODE == GEMV_PLAIN) {
    float* o = p.out + p.Coff + (size_t)p.has_pos * pb + (size_t)b * p.out_batch_stride + lr;
    o[0] = mine[0];
    if (has2) o[1] = mine[1];
  } else if (MODE == GEMV_RESID) {
    float* o = p.out + (size_t)b * p.out_batch_stride + lr;
    o[0] += mine[0];
    if (has2) o[1] += mine[1];
  } else if (MODE == GEMV_SWIGLU) {
    float* o = p.out + (size_t)b * p.out_batch_stride + lr;
    o[0] = silu_mul(mine[0], mine[NR == 4 ? 2 : 0]);
    if (has2) o[1] = silu_mul(mine[1], mine[NR == 4 ? 3 : 1]);
  } else {  // GEMV_QKV: rows are pairs (lr, lr+1), lr even
    float v0 = mine[0], v1 = mine[1];
    if (src < 2) {  // q or k: rotate (run.cc:271-286)
      const int j = (lr % p.head_size) >> 1;
      const float c = p.rope_cos[(size_t)pb * (p.head_size >> 1) + j];
      const float s = p.rope_sin[(size_t)pb * (p.head_size >> 1) + j];
      const float r0 = v0 * c - v1 * s, r1 = v0 * s + v1 * c;
      v0 = r0; v1 = r1;
    }
    float* o;
    if (src == 0) o = p.out + (size_t)b * p.out_batch_stride + lr;
    else {
      float* cache = (src == 1) ? p.kc : p.vc;
      o = cache + (size_t)b * p.kv_batch_stride + p.loff + (size_t)pb * p.kv_dim + lr;
    }
    o[0] = v0;
    o[1] = v1;
  }
}

template <int BT, int MODE>
__global__ void __launch_bounds__(GEMV_THREADS) gemv_kernel(const GemvP p) {
  constexpr int NR = (MODE == GEMV_SWIGLU) ? 4 : 2;   // physical weight rows per warp
  constexpr int U = (NR == 4) ? 2 : 4;                // float4 loads in flight per row per lane
  constexpr int KC4 = GEMV_XS_FLOATS / BT / 4;        // float4 per staged chunk per slot
  constexpr int IPC = KC4 / (32 * U);                 // inner iterations per chunk
  extern __shared__ float4 xs4[];                     // [BT][kc4_alloc]
  __shared__ float red[33];
  __shared__ float scale_s[BT];

  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int nk4 = p.K >> 2;
  const int kc4_alloc = nk4 < KC4 ? nk4 : KC4;        // row pitch of xs4 (float4)
  const int b0 = blockIdx.y * BT;
  const int Mtot = (MODE == GEMV_SWIGLU) ? p.M0 : (p.M0 + p.M1 + p.M2);

  // ---- which rows does this warp own -------------------------------------------------
  const int vr = (blockIdx.x * GEMV_WARPS + warp) * 2;
  const bool valid = vr < Mtot;
  int lr = valid ? vr : 0, src = 0, Msrc = p.M0;
  if (MODE == GEMV_QKV) {
    if (lr >= p.M0) { lr -= p.M0; src = 1; Msrc = p.M1;
      if (lr >= p.M1) { lr -= p.M1; src = 2; Msrc = p.M2; } }
  }
  const bool has2 = valid && (lr + 1 < Msrc);
  const float4* wrow[NR];
  {
    const float* base = (src == 0) ? p.W0 : (src == 1 ? p.W1 : p.W2);
    const int lr1 = (lr + 1 < Msrc) ? lr + 1 : lr;
    wrow[0] = reinterpret_cast<const float4*>(base + (size_t)lr * p.K);
    wrow[1] = reinterpret_cast<const float4*>(base + (size_t)lr1 * p.K);
    if (MODE == GEMV_SWIGLU) {
      wrow[2] = reinterpret_cast<const float4*>(p.W1 + (size_t)lr * p.K);
      wrow[3] = reinterpret_cast<const float4*>(p.W1 + (size_t)lr1 * p.K);
    }
  }

  float acc[NR][BT];
#pragma unroll
  for (int r = 0; r < NR; r++)
#pragma unroll
    for (int b = 0; b < BT; b++) acc[r][b] = 0.0f;
  float ss[BT];
#pragma unroll
  for (int b = 0; b < BT; b++) ss[b] = 0.0f;

  const int T = (nk4 + 32 * U - 1) / (32 * U);
  float4 wcur[NR][U], wnext[NR][U];
  const float4 z4 = make_float4(0.f, 0.f, 0.f, 0.f);
  // first weight loads: independent of the previous kernel, issue before the dependency wait
#pragma unroll
  for (int r = 0; r < NR; r++)
#pragma unroll
    for (int u = 0; u < U; u++) {
      int i = u * 32 + lane;
      wcur[r][u] = (i < nk4) ? ldg_stream(wrow[r] + i) : z4;
    }
  pdl_wait();
  pdl_trigger();

  for (int t = 0; t < T; t++) {
    const int c0 = (t / IPC) * KC4;   // first float4 of the chunk holding iteration t
    if (t % IPC == 0) {
      if (t > 0) __syncthreads();
      // ---- stage x[b0 .. b0+BT)[c0*4 ..) -> shared, folding the RMSNorm weight in
      const int n4 = (nk4 - c0) < KC4 ? (nk4 - c0) : KC4;
      for (int i = threadIdx.x; i < n4; i += GEMV_THREADS) {
        float4 g = p.rms_w ? reinterpret_cast<const float4*>(p.rms_w)[c0 + i] : make_float4(1.f, 1.f, 1.f, 1.f);
#pragma unroll
        for (int b = 0; b < BT; b++) {
          float4 v = z4;
          if (b0 + b < p.B) v = reinterpret_cast<const float4*>(p.X + (size_t)(b0 + b) * p.x_stride)[c0 + i];
          if (p.rms_w) {
            ss[b] = dot4(v, v, ss[b]);
            v.x *= g.x; v.y *= g.y; v.z *= g.z; v.w *= g.w;
          }
          xs4[b * kc4_alloc + i] = v;
        }
      }
      __syncthreads();
    }
    // ---- prefetch the next iteration's weights
    if (t + 1 < T) {
#pragma unroll
      for (int r = 0; r < NR; r++)
#pragma unroll
        for (int u = 0; u < U; u++) {
          int i = (t + 1) * 32 * U + u * 32 + lane;
          wnext[r][u] = (i < nk4) ? ldg_stream(wrow[r] + i) : z4;
        }
    }
    // ---- FMA: every staged slot against the rows in registers
#pragma unroll
    for (int u = 0; u < U; u++) {
      const int i = t * 32 * U + u * 32 + lane;
      if (i < nk4) {
#pragma unroll
        for (int b = 0; b < BT; b++) {
          const float4 xv = xs4[b * kc4_alloc + (i - c0)];
#pragma unroll
          for (int r = 0; r < NR; r++) acc[r][b] = dot4(wcur[r][u], xv, acc[r][b]);
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

  // ---- RMSNorm scale per slot: 1/sqrt(mean(x^2) + 1e-5)   (run.cc:196-205)
  if (p.rms_w) {
#pragma unroll
    for (int b = 0; b < BT; b++) {
      float tot = block_sum(ss[b], red);
      if (threadIdx.x == 0) {
        float m = tot / p.K;
        m += 1e-5f;
        scale_s[b] = 1.0f / sqrtf(m);
      }
    }
    __syncthreads();
  }

  // ---- warp reduction; lane b finishes slot b0+b
  float mine[NR];
#pragma unroll
  for (int r = 0; r < NR; r++) mine[r] = 0.0f;
#pragma unroll
  for (int r = 0; r < NR; r++)
#pragma unroll
    for (int b = 0; b < BT; b++) {
      float v = warp_sum(acc[r][b]);
      if (lane == b) mine[r] = v;
    }
  gemv_epilogue<MODE, NR>(p, mine, valid && lane < BT, b0 + lane, lr, src, has2, p.rms_w ? scale_s[lane < BT ? lane : 0] : 1.0f);
}

#endif  // __CUDACC__

// Host-side launcher: picks the batch tile and the grid.  Returns a THA_STATUS_*.
int gemv_launch(cudaStream_t stream, int mode, const GemvP& p, bool pdl);
// Requirements of the vector kernel (otherwise callers use the scalar fallback).
inline bool gemv_fast_ok(const GemvP& p) {
  return (p.K % 4 == 0) && (p.x_stride % 4 == 0) && aligned16(p.W0) && aligned16(p.X) &&
         (!p.W1 || aligned16(p.W1)) && (!p.W2 || aligned16(p.W2)) && (!p.rms_w || aligned16(p.rms_w));
}

}  // namespace tha
