// KV-cache multi-head attention for decode.
//
// (1) The reference's three stages (scores / softmax / weighted V sum,
//     src/thaDNN/thaDNN_mha.cpp) as three generic kernels, for both KV layouts
//     (v1 batch-major [B][L][S][kvd], v2 time-major [t][B][kvd]).  They keep the
//     reference entry points alive and serve head sizes the fused kernel rejects.
// (2) One fused kernel: QK^T, online softmax and PV in a single pass over the KV
//     rows, split over the time axis (flash-decoding) so a single long sequence
//     still fills the GPU; the last CTA of a (slot, segment) merges the splits.
//     K/V rows are read once, fully coalesced (a warp reads whole contiguous row
//     segments with 128-bit loads); scores never touch HBM.
#include <stdio.h>
#include <stdlib.h>

#include <mutex>

#include "common.cuh"
#include "workspace.h"
#include "ktrace.cuh"

namespace tha {

// ------------------------------------------------------------------ three stages ----
// att[b,h,t] = q[b,h,:] . K[b,t,h/kv_mul,:] / sqrt(hs)   for t <= pos[b]
__global__ void mha_scores_kernel(const int* __restrict__ pos, const float* __restrict__ q, float* __restrict__ att,
                                  const float* __restrict__ kbase, long long k_batch_stride, long long k_t_stride,
                                  int head_size, int att_row, int n_heads, int dim, int kv_mul) {
  const int h = blockIdx.x, b = blockIdx.y;
  const int pb = pos[b];
  if (pb < 0) return;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
  const float* qh = q + (size_t)b * dim + h * head_size;
  const float* kh = kbase + (size_t)b * k_batch_stride + (h / kv_mul) * head_size;
  float* a = att + ((size_t)b * n_heads + h) * att_row;
  const float inv = sqrtf((float)head_size);
  for (int t = warp; t <= pb; t += nw) {
    const float* k = kh + (size_t)t * k_t_stride;
    float s = 0.0f;
    for (int i = lane; i < head_size; i += 32) s = fmaf(qh[i], k[i], s);
    s = warp_sum(s);
    if (lane == 0) a[t] = s / inv;
  }
}

// softmax over t in [0, pos[b]] in place (thaDNN_mha.cpp:306-352)
__global__ void mha_softmax_kernel(const int* __restrict__ pos, float* __restrict__ att, int att_row, int n_heads) {
  __shared__ float red[33];
  const int h = blockIdx.x, b = blockIdx.y;
  const int pb = pos[b];
  if (pb < 0) return;
  float* x = att + ((size_t)b * n_heads + h) * att_row;
  const int n = pb + 1;
  float m = -INFINITY;
  for (int i = threadIdx.x; i < n; i += blockDim.x) m = fmaxf(m, x[i]);
  m = block_max(m, red);
  float s = 0.0f;
  for (int i = threadIdx.x; i < n; i += blockDim.x) { float e = expf(x[i] - m); x[i] = e; s += e; }
  s = block_sum(s, red);
  for (int i = threadIdx.x; i < n; i += blockDim.x) x[i] /= s;
}

// xb[b,h,i] = sum_t att[b,h,t] * V[b,t,h/kv_mul,i]   (thaDNN_mha.cpp:376-406)
__global__ void mha_weighted_kernel(const int* __restrict__ pos, float* __restrict__ xb, const float* __restrict__ att,
                                    const float* __restrict__ vbase, long long v_batch_stride, long long v_t_stride,
                                    int head_size, int att_row, int n_heads, int dim, int kv_mul) {
  extern __shared__ float part[];  // [8][head_size]
  const int h = blockIdx.x, b = blockIdx.y;
  const int pb = pos[b];
  if (pb < 0) return;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5, ny = blockDim.x >> 5;
  const float* a = att + ((size_t)b * n_heads + h) * att_row;
  const float* vh = vbase + (size_t)b * v_batch_stride + (h / kv_mul) * head_size;
  for (int i = tx; i < head_size; i += 32) {
    float s = 0.0f;
    for (int t = ty; t <= pb; t += ny) s = fmaf(a[t], vh[(size_t)t * v_t_stride + i], s);
    part[ty * head_size + i] = s;
  }
  __syncthreads();
  for (int i = threadIdx.x; i < head_size; i += blockDim.x) {
    float s = 0.0f;
    for (int y = 0; y < ny; y++) s += part[y * head_size + i];
    xb[(size_t)b * dim + h * head_size + i] = s;
  }
}

// ------------------------------------------------------------------ fused ----------
// RoPE + KV scatter inside the attention kernel (one launch and one dependency less per layer).  The wq|wk|wv matmul leaves
// nz split-K partials of the row [q | k | v]; a CTA sums its own slice of them: q (rotated, run.cc:271-286 with the host-built
// tables) for its segment, and -- in the CTA that owns the end of the context -- the new position's k (rotated) and v for
// the same columns, which it writes to the cache row of pos[b] and folds into the softmax from registers.  The cache rows
// 0 .. pos-1 are read as before.  q_out (nullable) keeps RunState.q as the reference leaves it.
struct AttnRope {
  const float* part; int nz; long long zstride; int ld;     // partials: part[z * zstride + b * ld + column]
  const float* rc; const float* rs;                         // cos / sin tables [seq_len][head_size / 2]
  float* q_out; float* kc_w; float* vc_w;                   // kc_w / vc_w: the same rows as AttnP.kc / vc, writable
  uint16_t* out_b16;                                        // nullable (bf16 model path): the output row rounded to bf16 (nearest even) as well
};
struct AttnP {
  const int* pos; const float* q; float* out;
  const float* kc; const float* vc;         // already offset by loff
  long long kv_batch_stride; int kv_dim, head_size, kv_mul, dim;
  int nseg, nsplit;
  float4* part_acc; float2* part_ml; int* counters;
  float* out_hi; float* out_lo;             // nullable: the tf32 hi/lo split of `out` for the tensor-core wo matmul
  int8_t* out_q; float* out_sT; int ldxs, gs;   // nullable: `out` quantised per group (runq.c:145-171) for the int8 wo matmul
  KvPaged pg;                                   // pg.table != nullptr: kc / vc are block pools (BatchManager), rows found through the table
  AttnRope rp;                                  // rp.part != nullptr (ROPE kernels): q / k / v of the new position come from the matmul's partials
};

__device__ __forceinline__ float attn_tf32(float v) { return __uint_as_float((__float_as_uint(v) + 0x1000u) & 0xFFFFE000u); }
// called by whole warps (the group maxima of the quantiser are shuffles)
__device__ __forceinline__ void attn_store(const AttnP& p, float* outp, float4 r, int b) {
  *reinterpret_cast<float4*>(outp) = r;
  if (p.out_q) {
    const int i4 = (int)((outp - p.out) - (size_t)b * p.dim) >> 2;
    q8_quant4(r, p.gs, i4, p.out_q + (size_t)b * p.dim, p.out_sT, p.ldxs, b, true);
  }
  if (p.rp.out_b16) {   // the rounding of step_bf16.cu's bf_round_kernel (nearest even; NaN stays NaN)
    auto b16 = [](float f) -> uint32_t {
      const uint32_t u = __float_as_uint(f);
      if ((u & 0x7FFFFFFFu) > 0x7F800000u) return (u >> 16) | 0x40u;
      return (u + 0x7FFFu + ((u >> 16) & 1u)) >> 16;
    };
    *reinterpret_cast<uint2*>(p.rp.out_b16 + (outp - p.out)) = make_uint2(b16(r.x) | (b16(r.y) << 16), b16(r.z) | (b16(r.w) << 16));
  }
  if (p.out_hi) {
    const size_t off = outp - p.out;
    const float4 h = make_float4(attn_tf32(r.x), attn_tf32(r.y), attn_tf32(r.z), attn_tf32(r.w));
    *reinterpret_cast<float4*>(p.out_hi + off) = h;
    *reinterpret_cast<float4*>(p.out_lo + off) = make_float4(r.x - h.x, r.y - h.y, r.z - h.z, r.w - h.w);
  }
}

// J = float4 slots per lane; a CTA covers a segment of 128*J query floats (whole heads).
template <int J, bool ROPE = false>
__global__ void __launch_bounds__(256) attn_fused_kernel(const AttnP p) {
  constexpr int F = 32 * J;  // slots per segment
  extern __shared__ float4 sm_acc[];                 // [warps][F]
  float2* sm_ml = reinterpret_cast<float2*>(sm_acc + (blockDim.x >> 5) * F);  // [warps][F]
  __shared__ int is_last;

  const int kt = threadIdx.x == 0 ? kt_begin(KT_ATTN) : -1;
  pdl_trigger();   // the next kernel may start its launch (it waits for this grid to complete before touching memory)
  pdl_wait();
  kt_stamp(kt, 1);
  const int b = blockIdx.x, seg = blockIdx.y, split = blockIdx.z;
  const int pb = p.pos[b];
  if (pb < 0) {   // idle slot: the int8 matmul that follows gets zeros (as the stand-alone quantiser gave it)
    if (p.rp.out_b16 && split == 0 && threadIdx.x < F)   // bf16 path: zeros into the wo matmul, as bf_round_kernel wrote
      *reinterpret_cast<uint2*>(p.rp.out_b16 + (size_t)b * p.dim + seg * (128 * J) + 4 * threadIdx.x) = make_uint2(0u, 0u);
    if (p.out_q && split == 0 && threadIdx.x < F)
      q8_quant4(make_float4(0.f, 0.f, 0.f, 0.f), p.gs, (seg * (128 * J) >> 2) + threadIdx.x, p.out_q + (size_t)b * p.dim, p.out_sT,
                p.ldxs, b, true);
    return;
  }
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int hs = p.head_size, LH = hs >> 2;  // lanes per head
  const float sqrt_hs = sqrtf((float)hs);

  const int n = pb + 1;
  // Splits in use: the launch provides p.nsplit CTAs per (slot, segment) for the longest context, but a split is only worth
  // its merge (partials through L2, a fence, an atomic, a second round of loads: ~5 us) when it has a full iteration of rows
  // to itself.  Short contexts use fewer splits -- one below 33 positions, which then stores its result directly; the
  // surplus CTAs leave at once.  Every CTA of the (slot, segment) derives the same count from pos[b].
  int nsp = (n + 31) >> 5;
  if (nsp > p.nsplit) nsp = p.nsplit;
  if (split >= nsp) return;
  const int chunk = (n + nsp - 1) / nsp;
  const int t0 = split * chunk;
  const int t1f = (t0 + chunk < n) ? (t0 + chunk) : n;
  // ROPE: the row of position pb is not in the cache yet -- the cache rows end at pb, the new row is folded in afterwards by
  // warp 0 of the CTA whose range reaches the end
  const bool own_new = ROPE && t1f == n && warp == 0;
  const int t1 = (ROPE && t1f == n) ? n - 1 : t1f;

  float4 qv[J];
  int koff[J];
  float4 knew[ROPE ? J : 1], vnew[ROPE ? J : 1];
#pragma unroll
  for (int j = 0; j < J; j++) {
    const int qo = seg * (128 * J) + 4 * (lane + 32 * j);
    koff[j] = ((qo / hs) / p.kv_mul) * hs + (qo % hs);
    if (!ROPE) {
      qv[j] = *reinterpret_cast<const float4*>(p.q + (size_t)b * p.dim + qo);
    } else {
      const float* pr = p.rp.part + (size_t)b * p.rp.ld;
      float4 a = make_float4(0.f, 0.f, 0.f, 0.f), kk = a, vv = a;
      for (int z = 0; z < p.rp.nz; z++) {   // fixed order, as tc_rope_scatter_kernel sums them
        const float4 t = *reinterpret_cast<const float4*>(pr + (size_t)z * p.rp.zstride + qo);
        a.x += t.x; a.y += t.y; a.z += t.z; a.w += t.w;
        if (own_new) {
          const float4 tk = *reinterpret_cast<const float4*>(pr + (size_t)z * p.rp.zstride + p.dim + koff[j]);
          const float4 tv = *reinterpret_cast<const float4*>(pr + (size_t)z * p.rp.zstride + p.dim + p.kv_dim + koff[j]);
          kk.x += tk.x; kk.y += tk.y; kk.z += tk.z; kk.w += tk.w;
          vv.x += tv.x; vv.y += tv.y; vv.z += tv.z; vv.w += tv.w;
        }
      }
      const int jj = (qo % hs) >> 1;         // the float4 holds the pairs jj and jj + 1 of its head (q and k alike)
      const float2 c2 = *reinterpret_cast<const float2*>(p.rp.rc + (size_t)pb * (hs >> 1) + jj);
      const float2 s2 = *reinterpret_cast<const float2*>(p.rp.rs + (size_t)pb * (hs >> 1) + jj);
      // (x, y) -> (x c - y s, x s + y c), written as in tc_rope_scatter_kernel: one product rounded, the other fused
      auto rot = [](float x, float y, float c, float sn, float& o0, float& o1) {
        o0 = fmaf(x, c, -__fmul_rn(y, sn));
        o1 = fmaf(x, sn, __fmul_rn(y, c));
      };
      rot(a.x, a.y, c2.x, s2.x, qv[j].x, qv[j].y);
      rot(a.z, a.w, c2.y, s2.y, qv[j].z, qv[j].w);
      if (p.rp.q_out && split == 0 && warp == 0) *reinterpret_cast<float4*>(p.rp.q_out + (size_t)b * p.dim + qo) = qv[j];
      if (own_new) {
        rot(kk.x, kk.y, c2.x, s2.x, knew[j].x, knew[j].y);
        rot(kk.z, kk.w, c2.y, s2.y, knew[j].z, knew[j].w);
        vnew[j] = vv;
        const long long ro = p.pg.table ? kv_paged_off(p.pg, b, pb, p.kv_dim) : (long long)b * p.kv_batch_stride + (long long)pb * p.kv_dim;
        *reinterpret_cast<float4*>(p.rp.kc_w + ro + koff[j]) = knew[j];
        *reinterpret_cast<float4*>(p.rp.vc_w + ro + koff[j]) = vnew[j];
      }
    }
  }
  float m[J], l[J];
  float4 acc[J];
#pragma unroll
  for (int j = 0; j < J; j++) { m[j] = -INFINITY; l[j] = 0.0f; acc[j] = make_float4(0.f, 0.f, 0.f, 0.f); }

  const float* kb = p.pg.table ? p.kc : p.kc + (size_t)b * p.kv_batch_stride;
  const float* vb = p.pg.table ? p.vc : p.vc + (size_t)b * p.kv_batch_stride;
  // A warp takes rows t, t+8, .., t+8(RW-1) per iteration (RW x 2 x J independent 128-bit loads in flight per lane)
  // and folds them into the running softmax with ONE rescale: exp(m_old - m_new) once per RW rows.  Scores are scaled
  // by 1/sqrt(hs) (exact for head sizes 16 and 64; 1 ulp from the reference's division otherwise) and exponentials
  // are ex2.approx-based (__expf, 2 ulp): both far inside the 1e-4 tolerance, and a third of the instructions of the
  // previous form (this kernel was issue- as well as latency-limited: sm__throughput 50 % at 4.3 TB/s).
  constexpr int RW = (J <= 2) ? 4 : 2;
  const float inv_sqrt_hs = 1.0f / sqrt_hs;
  const int nw = blockDim.x >> 5;   // warps of this CTA: 8, or fewer when the grid would not fit one wave (attn_pick_warps)
  for (int t = t0 + warp; t < t1; t += nw * RW) {
    float4 kr[RW][J], vr[RW][J];
    bool live[RW];
    long long roff[RW];
#pragma unroll
    for (int r = 0; r < RW; r++) {
      live[r] = (t + nw * r) < t1;
      const int tr = live[r] ? t + nw * r : t;
      roff[r] = p.pg.table ? kv_paged_off(p.pg, b, tr, p.kv_dim) : (long long)tr * p.kv_dim;
      const float* ka = kb + roff[r];
#pragma unroll
      for (int j = 0; j < J; j++) kr[r][j] = ldg_stream(reinterpret_cast<const float4*>(ka + koff[j]));
    }
#pragma unroll
    for (int r = 0; r < RW; r++) {
      const float* va = vb + roff[r];
#pragma unroll
      for (int j = 0; j < J; j++) vr[r][j] = ldg_stream(reinterpret_cast<const float4*>(va + koff[j]));
    }
#pragma unroll
    for (int j = 0; j < J; j++) {
      float sc[RW];
#pragma unroll
      for (int r = 0; r < RW; r++) sc[r] = dot4(qv[j], kr[r][j], 0.0f);
      for (int o = LH >> 1; o > 0; o >>= 1) {
#pragma unroll
        for (int r = 0; r < RW; r++) sc[r] += __shfl_xor_sync(THA_FULL_MASK, sc[r], o);
      }
      float mn = m[j];
#pragma unroll
      for (int r = 0; r < RW; r++) {
        sc[r] = live[r] ? sc[r] * inv_sqrt_hs : -INFINITY;
        mn = fmaxf(mn, sc[r]);
      }
      const float corr = __expf(m[j] - mn);   // exp(-inf) = 0 on the first step
      float ls = l[j] * corr;
      float4 a = make_float4(acc[j].x * corr, acc[j].y * corr, acc[j].z * corr, acc[j].w * corr);
#pragma unroll
      for (int r = 0; r < RW; r++) {
        const float pr = __expf(sc[r] - mn);  // dead rows: exp(-inf) = 0
        ls += pr;
        a.x = fmaf(pr, vr[r][j].x, a.x); a.y = fmaf(pr, vr[r][j].y, a.y);
        a.z = fmaf(pr, vr[r][j].z, a.z); a.w = fmaf(pr, vr[r][j].w, a.w);
      }
      l[j] = ls; acc[j] = a; m[j] = mn;
    }
  }

  if (ROPE && own_new) {   // ---- the new position, from registers
#pragma unroll
    for (int j = 0; j < J; j++) {
      float sc = dot4(qv[j], knew[j], 0.0f);
      for (int o = LH >> 1; o > 0; o >>= 1) sc += __shfl_xor_sync(THA_FULL_MASK, sc, o);
      sc *= inv_sqrt_hs;
      const float mn = fmaxf(m[j], sc);
      const float corr = __expf(m[j] - mn);   // exp(-inf) = 0 when this is the only row
      const float pr = __expf(sc - mn);
      l[j] = l[j] * corr + pr;
      acc[j] = make_float4(fmaf(pr, vnew[j].x, acc[j].x * corr), fmaf(pr, vnew[j].y, acc[j].y * corr),
                           fmaf(pr, vnew[j].z, acc[j].z * corr), fmaf(pr, vnew[j].w, acc[j].w * corr));
      m[j] = mn;
    }
  }

  // ---- merge the warps of this CTA ----
#pragma unroll
  for (int j = 0; j < J; j++) {
    const int f = lane + 32 * j;
    sm_acc[warp * F + f] = acc[j];
    sm_ml[warp * F + f] = make_float2(m[j], l[j]);
  }
  __syncthreads();
  const int f = threadIdx.x;
  float M = -INFINITY, L = 0.0f;
  float4 A = make_float4(0.f, 0.f, 0.f, 0.f);
  if (f < F) {
    for (int w = 0; w < nw; w++) M = fmaxf(M, sm_ml[w * F + f].x);
    for (int w = 0; w < nw; w++) {
      const float2 ml = sm_ml[w * F + f];
      const float e = (ml.x == -INFINITY) ? 0.0f : expf(ml.x - M);
      const float4 a = sm_acc[w * F + f];
      L += ml.y * e;
      A.x += a.x * e; A.y += a.y * e; A.z += a.z * e; A.w += a.w * e;
    }
  }
  float* outp = p.out + (size_t)b * p.dim + seg * (128 * J) + 4 * f;
  if (nsp == 1) {
    if (f < F) attn_store(p, outp, make_float4(A.x / L, A.y / L, A.z / L, A.w / L), b);
    kt_stamp(kt, 6);
    return;
  }

  // ---- publish this split's partial; the last CTA of (slot, segment) merges all ----
  const size_t pbase = ((size_t)(b * p.nseg + seg) * p.nsplit) * F;   // stride of the launch's split count: the scratch was sized by it
  if (f < F) {
    p.part_acc[pbase + (size_t)split * F + f] = A;
    p.part_ml[pbase + (size_t)split * F + f] = make_float2(M, L);
  }
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) {
    const int old = atomicAdd(&p.counters[b * p.nseg + seg], 1);
    is_last = (old == nsp - 1);
  }
  __syncthreads();
  if (!is_last) return;
  __threadfence();
  if (f < F) {
    float GM = -INFINITY;
    for (int s = 0; s < nsp; s++) GM = fmaxf(GM, __ldcg(&p.part_ml[pbase + (size_t)s * F + f]).x);
    float GL = 0.0f;
    float4 GA = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int s = 0; s < nsp; s++) {   // fixed order: result independent of arrival order
      const float2 ml = __ldcg(&p.part_ml[pbase + (size_t)s * F + f]);
      const float e = (ml.x == -INFINITY) ? 0.0f : expf(ml.x - GM);
      const float4 a = __ldcg(&p.part_acc[pbase + (size_t)s * F + f]);
      GL += ml.y * e;
      GA.x += a.x * e; GA.y += a.y * e; GA.z += a.z * e; GA.w += a.w * e;
    }
    attn_store(p, outp, make_float4(GA.x / GL, GA.y / GL, GA.z / GL, GA.w / GL), b);
  }
  if (threadIdx.x == 0) p.counters[b * p.nseg + seg] = 0;  // self-reset for the next launch
}

}  // namespace tha

// ------------------------------------------------------------------ launchers -------
namespace tha {

int attn_fused_nsplit(int B, int nseg, int seq_len);

// Segment choice for the fused kernel: a CTA covers G whole heads with G*hs = 128*J floats, J in 1..8.
// Returns J (0 = shape not supported by the fused kernel).  The NARROWEST segment wins: J = 1 needs 48 registers per
// thread (five CTAs per SM whose load and math phases interleave) where J = 6 needs 188 (one CTA per SM, its 8 warps in
// lock step): measured on stories110M at 256 slots, position 128: 140 k -> 172 k tok/s; 64 slots: 63 k -> 79 k;
// Llama-2-7B Q8_0 at 64 slots, position 256: 5.0 k (J = 8) -> 6.4 k tok/s (tools/r01/attn_sweep.sh).  A warp still reads
// 512 contiguous bytes of each K / V row.  THA_ATTN_J forces a width (development aid).
// Grouped-query models: a segment covers whole groups of query heads that share a K/V head where one fits (each K/V
// row is then read by one CTA, not by kv_mul of them).
static int fused_J(int n_heads, int head_size, int dim, int kv_mul = 1) {
  (void)n_heads;
  if (head_size < 4 || head_size > 128 || (head_size & (head_size - 1))) return 0;  // power of two, <= 128
  static int forced = -1;
  if (forced < 0) { const char* v = getenv("THA_ATTN_J"); forced = (v && *v) ? atoi(v) : 0; }
  int best = 0;
  for (int J = 1; J <= 8; J++) {
    const int seg = 128 * J;
    if (seg % head_size || dim % seg) continue;
    if (forced > 0 && J == forced) return J;
    if (!best) best = J;
    if (kv_mul > 1 && seg % (kv_mul * head_size) == 0) return J;
  }
  return best;
}

bool attn_fused_supported(int n_heads, int head_size, int dim, int kv_dim) {
  return fused_J(n_heads, head_size, dim) != 0 && (kv_dim % 4 == 0);
}

// Segment width and time split of the launch for B slots; grows the split scratch (never inside a stream capture).
int attn_fused_prepare(Workspace* ws, int B, int n_heads, int n_kv_heads, int head_size, int seq_len) {
  const int dim = n_heads * head_size;
  const int J = fused_J(n_heads, head_size, dim, n_heads / n_kv_heads);
  if (!J) return THA_STATUS_NOT_SUPPORTED;
  const int nseg = dim / (128 * J);
  const int nsplit = attn_fused_nsplit(B, nseg, seq_len);
  if (nsplit > 1) return ws_ensure_attn(ws, (size_t)B * nseg * nsplit * 32 * J, B * nseg);
  return THA_STATUS_SUCCESS;
}

int attn_fused_nsplit(int B, int nseg, int seq_len) {
  int target = 296 / (B * nseg);   // about two CTAs per SM
  int cap = seq_len / 32 > 0 ? seq_len / 32 : 1;
  if (target > cap) target = cap;
  if (target > 32) target = 32;
  if (target < 1) target = 1;
  return target;
}

// Warps per CTA.  Eight, unless the grid (one CTA per slot and segment, unsplit) would then need more than one wave of
// the machine: 256 slots x 6 segments of stories110M are 1536 CTAs against 740 resident ones (five per SM at 48 registers), i.e.
// three rounds of a latency-bound kernel (12 us per layer at position 10).  Warps per SM are what the register file allows either
// way, so narrower CTAs cost no bandwidth at long contexts; they put every (slot, segment) on the machine at once.
// A CTA needs one thread per float4 of its segment for the merge: warps >= J.
static int attn_pick_warps(void (*k)(const AttnP), int J, long long ctas) {
  static int sms = 0;
  if (!sms) { int dev = 0; sms = 148; if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev); }
  static int forced = -1;
  if (forced < 0) { const char* v = getenv("THA_ATTN_WARPS"); forced = (v && *v) ? atoi(v) : 0; }
  if (forced >= J && forced <= 8 && forced > 0) return forced;
  int smallest = 8;
  for (int nw = 8; nw >= 2 && nw >= J; nw >>= 1) {
    smallest = nw;
    int per_sm = 0;
    const size_t smem = (size_t)nw * 32 * J * (sizeof(float4) + sizeof(float2));
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k, nw * 32, smem) != cudaSuccess) { cudaGetLastError(); return 8; }
    if (ctas <= (long long)per_sm * sms) return nw;
  }
  return smallest;   // more than one wave whatever the shape: the most even distribution
}

// kc / vc are the caches already offset to the layer (base + loff).
static int attn_fused_launch_impl(cudaStream_t stream, Workspace* ws, int B, const int* pos_d, int n_heads, int n_kv_heads,
                                  const float* q, float* out, const float* kc, const float* vc, long long kv_batch_stride,
                                  int head_size, int seq_len, bool pdl, float* out_hi, float* out_lo, int8_t* out_q, float* out_sT,
                                  int ldxs, int gs, const KvPaged* pg, const AttnRope* rp) {
  const int dim = n_heads * head_size, kv_dim = n_kv_heads * head_size;
  const int J = fused_J(n_heads, head_size, dim, n_heads / n_kv_heads);
  if (!J || kv_dim % 4) return THA_STATUS_NOT_SUPPORTED;
  AttnP p;
  p.pos = pos_d; p.q = q; p.out = out; p.kc = kc; p.vc = vc; p.out_hi = out_hi; p.out_lo = out_lo;
  p.out_q = out_q; p.out_sT = out_sT; p.ldxs = ldxs; p.gs = gs;
  if (pg) p.pg = *pg; else p.pg = KvPaged{};
  p.rp = rp ? *rp : AttnRope{};
  if (rp && (J > 2 || !rp->part || !rp->rc || !rp->rs || !rp->kc_w || !rp->vc_w || (rp->ld & 3) || (rp->zstride & 3) || (head_size & 3)))
    return THA_STATUS_NOT_SUPPORTED;
  if (out_q && (gs < 32 || gs > 128 || (gs & (gs - 1)) || 128 % gs)) return THA_STATUS_NOT_SUPPORTED;
  p.kv_batch_stride = kv_batch_stride; p.kv_dim = kv_dim; p.head_size = head_size;
  p.kv_mul = n_heads / n_kv_heads; p.dim = dim;
  p.nseg = dim / (128 * J);
  p.nsplit = attn_fused_nsplit(B, p.nseg, seq_len);
  const int F = 32 * J;
  if (p.nsplit > 1) {
    int st = ws_ensure_attn(ws, (size_t)B * p.nseg * p.nsplit * F, B * p.nseg);
    if (st) return st;
  }
  p.part_acc = ws->part_acc; p.part_ml = ws->part_ml; p.counters = ws->counters;
  void (*k)(const AttnP) = nullptr;
  switch (J) {
    case 1: k = rp ? attn_fused_kernel<1, true> : attn_fused_kernel<1>; break;
    case 2: k = rp ? attn_fused_kernel<2, true> : attn_fused_kernel<2>; break;
    case 3: k = attn_fused_kernel<3>; break; case 4: k = attn_fused_kernel<4>; break;
    case 5: k = attn_fused_kernel<5>; break; case 6: k = attn_fused_kernel<6>; break;
    case 7: k = attn_fused_kernel<7>; break; case 8: k = attn_fused_kernel<8>; break;
  }
  const int nw = attn_pick_warps(k, J, (long long)B * p.nseg * p.nsplit);
  dim3 grid(B, p.nseg, p.nsplit), block(32 * nw);
  const size_t smem = (size_t)nw * F * (sizeof(float4) + sizeof(float2));
  if (smem > 40 * 1024) THA_CUDA_OK(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024));
  THA_CUDA_OK(launch(k, grid, block, smem, stream, pdl, p));
  return THA_STATUS_SUCCESS;
}

int attn_fused_launch(cudaStream_t stream, Workspace* ws, int B, const int* pos_d, int n_heads, int n_kv_heads,
                      const float* q, float* out, const float* kc, const float* vc, long long kv_batch_stride,
                      int head_size, int seq_len, bool pdl, float* out_hi, float* out_lo, int8_t* out_q, float* out_sT,
                      int ldxs, int gs, const KvPaged* pg) {
  return attn_fused_launch_impl(stream, ws, B, pos_d, n_heads, n_kv_heads, q, out, kc, vc, kv_batch_stride, head_size, seq_len, pdl,
                                out_hi, out_lo, out_q, out_sT, ldxs, gs, pg, nullptr);
}
// Segments of at most two float4 per lane (every multi-head model; wide grouped-query segments keep the separate kernel: the
// new row's k / v would cost them 16 registers per float4)
bool attn_fused_rope_supported(int n_heads, int n_kv_heads, int head_size) {
  const int J = fused_J(n_heads, head_size, n_heads * head_size, n_heads / n_kv_heads);
  return J == 1 || J == 2;
}
// The same with RoPE + KV scatter of the new position inside the kernel: part / nz / zstride / ld describe the wq|wk|wv
// partials, q_out (nullable) receives the rotated queries, the new k / v rows are written through kc / vc.
int attn_fused_rope_launch(cudaStream_t stream, Workspace* ws, int B, const int* pos_d, int n_heads, int n_kv_heads, float* q_out,
                           float* out, float* kc, float* vc, long long kv_batch_stride, int head_size, int seq_len, bool pdl,
                           float* out_hi, float* out_lo, int8_t* out_q, float* out_sT, int ldxs, int gs, const KvPaged* pg,
                           const float* part, int nz, long long zstride, int ld, const float* rc, const float* rs, uint16_t* out_b16) {
  AttnRope rp = {part, nz, zstride, ld, rc, rs, q_out, kc, vc, out_b16};
  return attn_fused_launch_impl(stream, ws, B, pos_d, n_heads, n_kv_heads, q_out, out, kc, vc, kv_batch_stride, head_size, seq_len, pdl,
                                out_hi, out_lo, out_q, out_sT, ldxs, gs, pg, &rp);
}

}  // namespace tha

using namespace tha;

extern "C" int tha_mha_scores_v1_f32(void* stream, int n_batches, const int* pos_d, int n_heads, int n_layers,
                                     const float* q, float* att, const float* key_cache, int head_size,
                                     int seq_len, int loff, int kv_dim, int dim, int kv_mul) {
  if (n_batches <= 0 || !pos_d || !q || !att || !key_cache) return THA_STATUS_INVALID_ARG;
  mha_scores_kernel<<<dim3(n_heads, n_batches), 256, 0, (cudaStream_t)stream>>>(
      pos_d, q, att, key_cache + loff, (long long)n_layers * seq_len * kv_dim, kv_dim, head_size, seq_len,
      n_heads, dim, kv_mul);
  count_launch();
  THA_LAUNCH_OK();
  return THA_STATUS_SUCCESS;
}

extern "C" int tha_mha_softmax_f32(void* stream, int n_batches, float* att, const int* pos_d, int seq_len, int n_heads) {
  if (n_batches <= 0 || !pos_d || !att) return THA_STATUS_INVALID_ARG;
  mha_softmax_kernel<<<dim3(n_heads, n_batches), 256, 0, (cudaStream_t)stream>>>(pos_d, att, seq_len, n_heads);
  count_launch();
  THA_LAUNCH_OK();
  return THA_STATUS_SUCCESS;
}

extern "C" int tha_mha_weighted_v1_f32(void* stream, int n_batches, const int* pos_d, int n_heads, float* xb,
                                       const float* att, const float* value_cache, int head_size, int seq_len,
                                       int loff, int kv_dim, int kv_mul, int dim, int n_layers) {
  if (n_batches <= 0 || !pos_d || !xb || !att || !value_cache) return THA_STATUS_INVALID_ARG;
  mha_weighted_kernel<<<dim3(n_heads, n_batches), 256, 8 * head_size * sizeof(float), (cudaStream_t)stream>>>(
      pos_d, xb, att, value_cache + loff, (long long)n_layers * seq_len * kv_dim, kv_dim, head_size, seq_len,
      n_heads, dim, kv_mul);
  count_launch();
  THA_LAUNCH_OK();
  return THA_STATUS_SUCCESS;
}

extern "C" int tha_mha_scores_v2_f32(void* stream, int n_batches, const int* pos_d, int n_heads, const float* q,
                                     float* att, const float* key_cache, int head_size, int n_words, int kv_dim,
                                     int dim, int kv_mul) {
  if (n_batches <= 0 || !pos_d || !q || !att || !key_cache) return THA_STATUS_INVALID_ARG;
  // time-major: row of (t, b) at t*B*kv_dim + b*kv_dim  (thaDNN_mha.cpp:78)
  mha_scores_kernel<<<dim3(n_heads, n_batches), 256, 0, (cudaStream_t)stream>>>(
      pos_d, q, att, key_cache, (long long)kv_dim, (long long)n_batches * kv_dim, head_size, n_words, n_heads,
      dim, kv_mul);
  count_launch();
  THA_LAUNCH_OK();
  return THA_STATUS_SUCCESS;
}

extern "C" int tha_mha_weighted_v2_f32(void* stream, int n_batches, const int* pos_d, int n_heads, float* xb,
                                       const float* att, const float* value_cache, int head_size, int n_words,
                                       int kv_dim, int kv_mul, int dim) {
  if (n_batches <= 0 || !pos_d || !xb || !att || !value_cache) return THA_STATUS_INVALID_ARG;
  mha_weighted_kernel<<<dim3(n_heads, n_batches), 256, 8 * head_size * sizeof(float), (cudaStream_t)stream>>>(
      pos_d, xb, att, value_cache, (long long)kv_dim, (long long)n_batches * kv_dim, head_size, n_words, n_heads,
      dim, kv_mul);
  count_launch();
  THA_LAUNCH_OK();
  return THA_STATUS_SUCCESS;
}

extern "C" int tha_mha_fused_v1_f32(void* stream, int n_batches, const int* pos_d, int n_heads, int n_kv_heads,
                                    int n_layers, const float* q, float* xb, const float* key_cache,
                                    const float* value_cache, int head_size, int seq_len, int loff) {
  if (n_batches <= 0 || !pos_d || !q || !xb || !key_cache || !value_cache || n_kv_heads <= 0) return THA_STATUS_INVALID_ARG;
  Workspace* ws = workspace();
  if (!ws) return THA_STATUS_EXECUTION_FAILED;
  const int kv_dim = n_kv_heads * head_size;
  return attn_fused_launch((cudaStream_t)stream, ws, n_batches, pos_d, n_heads, n_kv_heads, q, xb, key_cache + loff,
                           value_cache + loff, (long long)n_layers * seq_len * kv_dim, head_size, seq_len, false, nullptr, nullptr, nullptr, nullptr, 0, 0, nullptr);
}

// ------------------------------------------------------------------ no-allocation variant -------
// tha_mha_fused_v1_f32_static: the fused attention above for callers that may not allocate (CUDA graph capture, serving
// loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
// The split scratch (per-split accumulators, maxima / sums, arrival counters) lives in ONE file-scope device array of a fixed
// size instead of the Workspace's growable buffers.  Every GPU that loads the library holds the array until exit; every
// call on a device shares it, so two calls that overlap on that GPU (different streams) race -- serialise them.
namespace tha {

constexpr int ATTN_STATIC_SLOTS = 32768;      // float4 + float2 per slot; a launch needs B * nseg * nsplit * 32 J slots
constexpr int ATTN_STATIC_COUNTERS = 256;     // one per (slot, segment) of a split launch (B * nseg <= 148 when nsplit > 1)
constexpr size_t ATTN_STATIC_ACC_OFF = 0;
constexpr size_t ATTN_STATIC_ML_OFF = ATTN_STATIC_ACC_OFF + (size_t)ATTN_STATIC_SLOTS * sizeof(float4);
constexpr size_t ATTN_STATIC_CNT_OFF = ATTN_STATIC_ML_OFF + (size_t)ATTN_STATIC_SLOTS * sizeof(float2);
constexpr size_t ATTN_STATIC_BYTES = ATTN_STATIC_CNT_OFF + (size_t)ATTN_STATIC_COUNTERS * sizeof(int);   // 787,456 bytes

static __device__ __align__(256) unsigned char attn_static_scratch[ATTN_STATIC_BYTES];

constexpr int ATTN_STATIC_MAX_DEV = 64;
static unsigned char* attn_static_ptr[ATTN_STATIC_MAX_DEV];   // device address of the array, per device (nullptr: not resolved yet)
static std::mutex attn_static_mu;

// The array's name is only a host-side handle here: its device address is asked for once per device, with that device current.
static int attn_static_base(unsigned char** base) {
  int dev = -1;
  THA_CUDA_OK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= ATTN_STATIC_MAX_DEV) return THA_STATUS_NOT_SUPPORTED;
  std::lock_guard<std::mutex> g(attn_static_mu);
  if (!attn_static_ptr[dev]) {
    void* ptr = nullptr;
    THA_CUDA_OK(cudaGetSymbolAddress(&ptr, attn_static_scratch));
    attn_static_ptr[dev] = static_cast<unsigned char*>(ptr);
  }
  *base = attn_static_ptr[dev];
  return THA_STATUS_SUCCESS;
}

}  // namespace tha

extern "C" int tha_mha_fused_v1_f32_static(void* stream, int n_batches, const int* pos_d, int n_heads, int n_kv_heads,
                                           int n_layers, const float* q, float* xb, const float* key_cache,
                                           const float* value_cache, int head_size, int seq_len, int loff) {
  if (n_batches <= 0 || !pos_d || !q || !xb || !key_cache || !value_cache || n_kv_heads <= 0 || n_heads <= 0 ||
      n_heads % n_kv_heads || seq_len <= 0)
    return THA_STATUS_INVALID_ARG;
  const int dim = n_heads * head_size, kv_dim = n_kv_heads * head_size;
  const int J = fused_J(n_heads, head_size, dim, n_heads / n_kv_heads);
  if (!J || kv_dim % 4) return THA_STATUS_NOT_SUPPORTED;
  AttnP p;
  p.pos = pos_d; p.q = q; p.out = xb; p.kc = key_cache + loff; p.vc = value_cache + loff;
  p.out_hi = nullptr; p.out_lo = nullptr; p.out_q = nullptr; p.out_sT = nullptr; p.ldxs = 0; p.gs = 0;
  p.pg = KvPaged{};
  p.rp = AttnRope{};
  p.kv_batch_stride = (long long)n_layers * seq_len * kv_dim; p.kv_dim = kv_dim; p.head_size = head_size;
  p.kv_mul = n_heads / n_kv_heads; p.dim = dim;
  p.nseg = dim / (128 * J);
  p.nsplit = attn_fused_nsplit(n_batches, p.nseg, seq_len);
  const int F = 32 * J;
  cudaStream_t st = (cudaStream_t)stream;
  p.part_acc = nullptr; p.part_ml = nullptr; p.counters = nullptr;
  if (p.nsplit > 1) {
    const size_t slots = (size_t)n_batches * p.nseg * p.nsplit * F;
    const int counters = n_batches * p.nseg;
    if (slots > (size_t)ATTN_STATIC_SLOTS || counters > ATTN_STATIC_COUNTERS) {
      char msg[256];
      snprintf(msg, sizeof(msg),
               "tha_mha_fused_v1_f32_static: launch needs %zu scratch slots and %d counters, the static scratch holds %d and %d "
               "(use tha_mha_fused_v1_f32)", slots, counters, ATTN_STATIC_SLOTS, ATTN_STATIC_COUNTERS);
      set_last_error(msg, cudaErrorInvalidValue);
      return THA_STATUS_INVALID_ARG;
    }
    unsigned char* base = nullptr;
    const int rc = attn_static_base(&base);
    if (rc) return rc;
    p.part_acc = reinterpret_cast<float4*>(base + ATTN_STATIC_ACC_OFF);
    p.part_ml = reinterpret_cast<float2*>(base + ATTN_STATIC_ML_OFF);
    p.counters = reinterpret_cast<int*>(base + ATTN_STATIC_CNT_OFF);
    // the arrival counters must start at zero: the kernel resets them when it finishes, but the array outlives every
    // call (an aborted launch, or a larger launch before this one) -- so clear the ones this launch reads, on its stream
    THA_CUDA_OK(cudaMemsetAsync(p.counters, 0, (size_t)counters * sizeof(int), st));
  }
  void (*k)(const AttnP) = nullptr;
  switch (J) {
    case 1: k = attn_fused_kernel<1>; break; case 2: k = attn_fused_kernel<2>; break;
    case 3: k = attn_fused_kernel<3>; break; case 4: k = attn_fused_kernel<4>; break;
    case 5: k = attn_fused_kernel<5>; break; case 6: k = attn_fused_kernel<6>; break;
    case 7: k = attn_fused_kernel<7>; break; case 8: k = attn_fused_kernel<8>; break;
  }
  const int nw = attn_pick_warps(k, J, (long long)n_batches * p.nseg * p.nsplit);   // as attn_fused_launch: bit-identical results
  dim3 grid(n_batches, p.nseg, p.nsplit), block(32 * nw);
  const size_t smem = (size_t)nw * F * (sizeof(float4) + sizeof(float2));
  if (smem > 40 * 1024) THA_CUDA_OK(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024));
  THA_CUDA_OK(launch(k, grid, block, smem, st, false, p));
  return THA_STATUS_SUCCESS;
}
