// Causal attention for batched prompt ingestion: many query rows of one sequence against its
// KV cache in one pass (the decode kernel of attention.cu reads the whole prefix once per row,
// which is quadratic in traffic for a long prompt).
//
// Same arithmetic as the reference's per-position loop (run.cc:288-321: score = q.k / sqrt(hs),
// softmax over t <= pos, weighted sum of V) evaluated tile by tile with an online softmax:
//   CTA = 64 consecutive query rows of one (sequence, head); loop over 64-key tiles of the cache
//   (which already holds this layer's K/V rows of the chunk -- tc_rope_scatter_kernel ran first);
//   S = Q K^T on a 4x4 register micro-tile per thread out of transposed shared-memory tiles
//   (conflict-free 128-bit reads), running max / sum per row, P staged in shared memory, O += P V.
// K/V tiles are read once per 64 queries; scores never leave the SM.
#include "common.cuh"
#include "workspace.h"

namespace tha {

struct AttnPrefillP {
  const float* q; float* out; float* out_hi; float* out_lo;   // [rows, dim]
  const float* kc; const float* vc;                           // cache base of the layer
  long long kv_batch_stride; int kv_dim, kv_mul, dim;
  const int* row_slot; const int* row_pos;                    // [rows]
  const int* tile_row0; const int* tile_n;                    // [tiles]: first row and row count (<= 64) of each query tile
  KvPaged pg;                                                 // pg.table != nullptr: kc / vc are block pools, rows found through the table
};

constexpr int PF_TQ = 64, PF_TK = 64;

__device__ __forceinline__ float pf_tf32(float v) { return __uint_as_float((__float_as_uint(v) + 0x1000u) & 0xFFFFE000u); }

template <int HS>
__global__ void __launch_bounds__(256) attn_prefill_kernel(const AttnPrefillP p) {
  constexpr int CW = HS / 16;              // output columns per thread
  constexpr int NV = (CW >= 4) ? CW / 4 : 1;
  constexpr int VW = (CW >= 4) ? 4 : CW;   // floats per V read
  extern __shared__ float4 pf_smem4[];
  float* Qt = reinterpret_cast<float*>(pf_smem4);   // [HS][64]   Qt[k][row]
  float* Kt = Qt + HS * PF_TQ;                       // [HS][64]   Kt[k][key]
  float* Vs = Kt + HS * PF_TK;                       // [64][HS]
  float* Ps = Vs + PF_TK * HS;                       // [64][64]   P[row][key]

  pdl_wait();
  pdl_trigger();
  const int tile = blockIdx.x, h = blockIdx.y;
  const int row0 = p.tile_row0[tile], n = p.tile_n[tile];
  const int slot = p.row_slot[row0], pos0 = p.row_pos[row0];
  if (pos0 < 0 || n <= 0) return;
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  const int kh = h / p.kv_mul;
  const float* kbase = (p.pg.table ? p.kc : p.kc + (size_t)slot * p.kv_batch_stride) + kh * HS;
  const float* vbase = (p.pg.table ? p.vc : p.vc + (size_t)slot * p.kv_batch_stride) + kh * HS;
  const int kmax = pos0 + n - 1;   // last key any row of the tile may see
  const float sqrt_hs = sqrtf((float)HS);

  // ---- Q tile, transposed: thread = (row, every 4th float4 column) ----
  {
    const int r = tid & 63, c0 = tid >> 6;
    const int rr = (r < n) ? r : n - 1;
    const float* qrow = p.q + (size_t)(row0 + rr) * p.dim + h * HS;
    for (int c = c0; c < HS / 4; c += 4) {
      const float4 v = *reinterpret_cast<const float4*>(qrow + 4 * c);
      Qt[(4 * c + 0) * PF_TQ + r] = v.x; Qt[(4 * c + 1) * PF_TQ + r] = v.y;
      Qt[(4 * c + 2) * PF_TQ + r] = v.z; Qt[(4 * c + 3) * PF_TQ + r] = v.w;
    }
  }

  float m[4], l[4], o[4][CW];
#pragma unroll
  for (int i = 0; i < 4; i++) {
    m[i] = -INFINITY; l[i] = 0.0f;
#pragma unroll
    for (int c = 0; c < CW; c++) o[i][c] = 0.0f;
  }
  int qpos[4];
#pragma unroll
  for (int i = 0; i < 4; i++) { const int r = ty * 4 + i; qpos[i] = pos0 + ((r < n) ? r : n - 1); }

  for (int tk0 = 0; tk0 <= kmax; tk0 += PF_TK) {
    __syncthreads();   // the previous tile's K/V/P are no longer read (also orders the Q stores the first time)
    {
      const int r = tid & 63, c0 = tid >> 6;
      const int t = tk0 + r;
      const bool live = t <= kmax;
      const int tr = live ? t : kmax;
      const long long roff = p.pg.table ? kv_paged_off(p.pg, slot, tr, p.kv_dim) : (long long)tr * p.kv_dim;
      const float* krow = kbase + roff;
      const float* vrow = vbase + roff;
      for (int c = c0; c < HS / 4; c += 4) {
        const float4 kv = *reinterpret_cast<const float4*>(krow + 4 * c);
        Kt[(4 * c + 0) * PF_TK + r] = kv.x; Kt[(4 * c + 1) * PF_TK + r] = kv.y;
        Kt[(4 * c + 2) * PF_TK + r] = kv.z; Kt[(4 * c + 3) * PF_TK + r] = kv.w;
        float4 vv = *reinterpret_cast<const float4*>(vrow + 4 * c);
        if (!live) vv = make_float4(0.f, 0.f, 0.f, 0.f);   // a masked key has weight 0; keep 0 * garbage out of the sum
        *reinterpret_cast<float4*>(Vs + r * HS + 4 * c) = vv;
      }
    }
    __syncthreads();

    // ---- S = Q K^T : rows ty*4.., keys tx*4.. ----
    float s[4][4];
#pragma unroll
    for (int i = 0; i < 4; i++)
#pragma unroll
      for (int j = 0; j < 4; j++) s[i][j] = 0.0f;
#pragma unroll 8
    for (int k = 0; k < HS; k++) {
      const float4 a = *reinterpret_cast<const float4*>(Qt + k * PF_TQ + ty * 4);
      const float4 b = *reinterpret_cast<const float4*>(Kt + k * PF_TK + tx * 4);
      const float av[4] = {a.x, a.y, a.z, a.w}, bv[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
      for (int i = 0; i < 4; i++)
#pragma unroll
        for (int j = 0; j < 4; j++) s[i][j] = fmaf(av[i], bv[j], s[i][j]);
    }
    // ---- scale, causal mask, online softmax (row statistics shared by the 16 threads of a row) ----
#pragma unroll
    for (int i = 0; i < 4; i++) {
      float rmax = -INFINITY;
#pragma unroll
      for (int j = 0; j < 4; j++) {
        const int t = tk0 + tx * 4 + j;
        s[i][j] = (t <= qpos[i]) ? s[i][j] / sqrt_hs : -INFINITY;
        rmax = fmaxf(rmax, s[i][j]);
      }
#pragma unroll
      for (int off = 8; off > 0; off >>= 1) rmax = fmaxf(rmax, __shfl_xor_sync(THA_FULL_MASK, rmax, off));
      const float mn = fmaxf(m[i], rmax);          // finite: key 0 is visible to every row in the first tile
      const float corr = expf(m[i] - mn);          // exp(-inf) = 0 on the first tile
      float4 pv;
      pv.x = expf(s[i][0] - mn); pv.y = expf(s[i][1] - mn); pv.z = expf(s[i][2] - mn); pv.w = expf(s[i][3] - mn);
      l[i] = l[i] * corr + ((pv.x + pv.y) + (pv.z + pv.w));   // per-thread partial of the row sum; reduced at the end
#pragma unroll
      for (int c = 0; c < CW; c++) o[i][c] *= corr;
      m[i] = mn;
      *reinterpret_cast<float4*>(Ps + (ty * 4 + i) * PF_TK + tx * 4) = pv;
    }
    __syncthreads();

    // ---- O += P V : rows ty*4.., columns tx*VW + 64*v ----
#pragma unroll 2
    for (int t = 0; t < PF_TK; t += 4) {
      float pr[4][4];
#pragma unroll
      for (int i = 0; i < 4; i++) {
        const float4 a = *reinterpret_cast<const float4*>(Ps + (ty * 4 + i) * PF_TK + t);
        pr[i][0] = a.x; pr[i][1] = a.y; pr[i][2] = a.z; pr[i][3] = a.w;
      }
#pragma unroll
      for (int tt = 0; tt < 4; tt++) {
        float vv[CW];
        const float* vr = Vs + (t + tt) * HS;
        if constexpr (VW == 4) {
#pragma unroll
          for (int v = 0; v < NV; v++) {
            const float4 a = *reinterpret_cast<const float4*>(vr + tx * 4 + 64 * v);
            vv[4 * v + 0] = a.x; vv[4 * v + 1] = a.y; vv[4 * v + 2] = a.z; vv[4 * v + 3] = a.w;
          }
        } else {
#pragma unroll
          for (int c = 0; c < CW; c++) vv[c] = vr[tx * CW + c];
        }
#pragma unroll
        for (int i = 0; i < 4; i++)
#pragma unroll
          for (int c = 0; c < CW; c++) o[i][c] = fmaf(pr[i][tt], vv[c], o[i][c]);
      }
    }
  }

  // ---- normalise and store (fp32 row and its tf32 hi/lo split for the wo matmul) ----
#pragma unroll
  for (int i = 0; i < 4; i++) {
    float L = l[i];
#pragma unroll
    for (int off = 8; off > 0; off >>= 1) L += __shfl_xor_sync(THA_FULL_MASK, L, off);
    const int r = ty * 4 + i;
    if (r >= n) continue;
    const size_t base = (size_t)(row0 + r) * p.dim + h * HS;
#pragma unroll
    for (int c = 0; c < CW; c++) {
      const int col = (VW == 4) ? (tx * 4 + 64 * (c >> 2) + (c & 3)) : (tx * CW + c);
      const float v = o[i][c] / L;
      p.out[base + col] = v;
      if (p.out_hi) {
        const float hi = pf_tf32(v);
        p.out_hi[base + col] = hi;
        p.out_lo[base + col] = v - hi;
      }
    }
  }
}

bool attn_prefill_supported(int head_size, int kv_dim) {
  return (head_size == 16 || head_size == 32 || head_size == 64 || head_size == 128) && kv_dim % 4 == 0;
}

int attn_prefill_launch(cudaStream_t stream, int n_tiles, int n_heads, int n_kv_heads, int head_size, const float* q,
                        float* out, float* out_hi, float* out_lo, const float* kc, const float* vc,
                        long long kv_batch_stride, const int* row_slot, const int* row_pos, const int* tile_row0,
                        const int* tile_n, bool pdl, const KvPaged* pg) {
  if (!attn_prefill_supported(head_size, n_kv_heads * head_size)) return THA_STATUS_NOT_SUPPORTED;
  AttnPrefillP p;
  p.q = q; p.out = out; p.out_hi = out_hi; p.out_lo = out_lo; p.kc = kc; p.vc = vc;
  if (pg) p.pg = *pg; else p.pg = KvPaged{};
  p.kv_batch_stride = kv_batch_stride; p.kv_dim = n_kv_heads * head_size; p.kv_mul = n_heads / n_kv_heads;
  p.dim = n_heads * head_size;
  p.row_slot = row_slot; p.row_pos = row_pos; p.tile_row0 = tile_row0; p.tile_n = tile_n;
  const size_t smem = ((size_t)3 * head_size * 64 + 64 * 64) * sizeof(float);
  void (*k)(const AttnPrefillP) = nullptr;
  switch (head_size) {
    case 16: k = attn_prefill_kernel<16>; break;
    case 32: k = attn_prefill_kernel<32>; break;
    case 64: k = attn_prefill_kernel<64>; break;
    default: k = attn_prefill_kernel<128>; break;
  }
  if (smem > 40 * 1024) THA_CUDA_OK(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  THA_CUDA_OK(launch(k, dim3(n_tiles, n_heads), dim3(256), smem, stream, pdl, p));
  return THA_STATUS_SUCCESS;
}

}  // namespace tha
