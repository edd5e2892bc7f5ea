// The decode step for MORE than 8 slots, and batched prompt ingestion (prefill): weight matmuls on
// the tensor cores (gemm_tc.cu), everything else fused into four small kernels.  The unit of work
// is a ROW = one (slot, position, token): a decode step has one row per slot, a prefill chunk one row
// per prompt position (rows of a sequence are contiguous, positions ascending).  Per layer:
//
//   prep          x += sum_z partial[z]  (the residual add of the previous wo / w2 matmul; layer 0: the
//                 embedding gather, src/thaDNN.cpp:31-36), then RMSNorm folded: xg = x * g split
//                 into tf32 hi/lo, 1/rms kept aside as the matmul's column scale (run.cc:196-209)
//   gemm          wq|wk|wv (one launch, split-K partials)                     (src/thaDNN.cpp:44-46)
//   rope_scatter  sum the partials, RoPE on q and k, K/V rows into the cache  (src/thaDNN.cpp:48-49)
//   attention     fused QK^T / online softmax / PV (attention.cu), output also written as hi/lo
//   gemm          wo  -> partials (the add happens in the next prep)          (src/thaDNN.cpp:56-59)
//   prep          x += partials, RMSNorm(ffn) folded, split
//   gemm          w1|w3 (one launch)                                          (src/thaDNN.cpp:63-64)
//   swiglu_split  silu(w1.x) * (w3.x) -> hi/lo                                (src/thaDNN.cpp:66-67)
//   gemm          w2  -> partials                                             (src/thaDNN.cpp:69-72)
// then prep(final norm) -> gemm(wcls) -> logits (-> argmax); prefill runs the classifier only on the last
// row of each sequence (gathered by the final prep), and its attention is attn_prefill.cu (one pass over the
// cache per 64 query rows) instead of the per-row decode kernel.
//
// Split-K partials are summed by their consumer in a fixed order, so the step is bit-reproducible
// run to run (no floating-point atomics anywhere).
#include <map>
#include <mutex>
#include <vector>

#include "chain_tc.h"
#include "common.cuh"
#include "gemm_tc.h"
#include "tp.h"
#include "workspace.h"
#include "ktrace.cuh"

namespace tha {

int argmax_launch(cudaStream_t st, int B, const float* logits, int vocab, int* next, bool pdl);
bool attn_fused_supported(int n_heads, int head_size, int dim, int kv_dim);
int attn_fused_prepare(Workspace* ws, int B, int n_heads, int n_kv_heads, int head_size, int seq_len);
int attn_fused_launch(cudaStream_t stream, Workspace* ws, int B, const int* pos_d, int n_heads, int n_kv_heads,
                      const float* q, float* out, const float* kc, const float* vc, long long kv_batch_stride,
                      int head_size, int seq_len, bool pdl, float* out_hi, float* out_lo, int8_t* out_q = nullptr,
                      float* out_sT = nullptr, int ldxs = 0, int gs = 0, const KvPaged* pg = nullptr);

bool attn_prefill_supported(int head_size, int kv_dim);
bool attn_fused_rope_supported(int n_heads, int n_kv_heads, int head_size);
int attn_fused_rope_launch(cudaStream_t stream, Workspace* ws, int B, const int* pos_d, int n_heads, int n_kv_heads, float* q_out,
                           float* out, float* kc, float* vc, long long kv_batch_stride, int head_size, int seq_len, bool pdl,
                           float* out_hi, float* out_lo, int8_t* out_q, float* out_sT, int ldxs, int gs, const KvPaged* pg,
                           const float* part, int nz, long long zstride, int ld, const float* rc, const float* rs,
                           uint16_t* out_b16 = nullptr);
int attn_prefill_launch(cudaStream_t stream, int n_tiles, int n_heads, int n_kv_heads, int head_size, const float* q,
                        float* out, float* out_hi, float* out_lo, const float* kc, const float* vc,
                        long long kv_batch_stride, const int* row_slot, const int* row_pos, const int* tile_row0,
                        const int* tile_n, bool pdl, const KvPaged* pg = nullptr);

__device__ __forceinline__ float st_tf32(float v) { return __uint_as_float((__float_as_uint(v) + 0x1000u) & 0xFFFFE000u); }

// ---- tensor-parallel exchange (tp.h) -------------------------------------------------------------------------------
// The consumer of a wo / w2 matmul reads every rank's partials.  Before it does, block 0 tells every rank "my partials
// of exchange e are in place" (the matmul before this kernel has completed: pdl_wait / stream order), and every block
// waits until all ranks have said so.  Flags only grow (e = step epoch * exchanges per step + k + 1), nothing is reset.
// Two exchange buffers alternate (wo -> 0, w2 -> 1): a rank can only overwrite buffer 0 of layer l+1 after it passed the
// w2 exchange of layer l, which every rank signals after it finished reading buffer 0 of layer l.
constexpr long long TP_TIMEOUT_CYCLES = 40000000000ll;   // ~20 s: a lost peer ends in an error flag, not a hang
__device__ __forceinline__ void tp_wait_peers(const TpPeers& t) {
  const unsigned e = *t.epoch * (unsigned)t.nx + (unsigned)t.k + 1u;
  if (blockIdx.x == 0 && (int)threadIdx.x < t.n) {
    __threadfence_system();
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(t.peer_flags[threadIdx.x] + t.rank), "r"(e) : "memory");
  }
  if ((int)threadIdx.x < t.n) {
    const unsigned* f = t.peer_flags[t.rank] + threadIdx.x;
    const long long t0 = clock64();
    for (;;) {
      unsigned v;
      asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(f) : "memory");
      if ((int)(v - e) >= 0) break;
      if (clock64() - t0 > TP_TIMEOUT_CYCLES) { *t.err = 1; break; }
    }
  }
  __syncthreads();
}
// partials may live in a peer's HBM: system-scope load, never served from a stale line
__device__ __forceinline__ float4 ld_peer4(const float* p) {
  float4 r;
  asm volatile("ld.relaxed.sys.global.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p) : "memory");
  return r;
}
// x += every rank's partials, ranks in ascending order, splits in ascending order (identical on every rank)
__device__ __forceinline__ void tp_add_partials(float4& v, const TpPeers& t, long long zstride, size_t off) {
  for (int r = 0; r < t.n; r++) {
#pragma unroll 4
    for (int z = 0; z < t.nz[r]; z++) {
      const float4 pz = ld_peer4(t.part[r] + (size_t)z * zstride + off);
      v.x += pz.x; v.y += pz.y; v.z += pz.z; v.w += pz.w;
    }
  }
}
__global__ void tp_epoch_kernel(unsigned* epoch) {
  pdl_trigger();   // the next kernel may start its launch (it waits for this grid to complete before touching memory)
  pdl_wait();
  if (threadIdx.x == 0) *epoch += 1u;
}

// One block per output row.  x row: embedding row (emb != nullptr) or x + sum of nz partials.  With `gather`
// output row r is computed from source row gather[r] (the classifier input of a prefill chunk: the last
// row of every sequence) and x is left untouched.
__global__ void __launch_bounds__(1024) tc_prep_kernel(float* __restrict__ x, const float* __restrict__ part, int nz,
                                                      long long zstride, const float* __restrict__ emb,
                                                      const int* __restrict__ token, const int* __restrict__ pos,
                                                      const float* __restrict__ gain, int d, float* __restrict__ xh,
                                                      float* __restrict__ xl, float* __restrict__ scale,
                                                      const int* __restrict__ gather, const TpPeers tp) {
  __shared__ float red[33];
  const int kt = threadIdx.x == 0 ? kt_begin(KT_PREP) : -1;
  pdl_trigger();   // the next kernel may start its launch (it waits for this grid to complete before touching memory)
  pdl_wait();
  kt_stamp(kt, 1);
  if (tp.n && !emb) tp_wait_peers(tp);
  const int ob = blockIdx.x;
  const int b = gather ? gather[ob] : ob;
  const int d4 = d >> 2;
  float4* xr = reinterpret_cast<float4*>(x + (size_t)b * d);
  float4* hr = reinterpret_cast<float4*>(xh + (size_t)ob * d);
  float4* lr = reinterpret_cast<float4*>(xl + (size_t)ob * d);
  if (pos[b] < 0) {   // idle slot: feed zeros to the matmul
    for (int i = threadIdx.x; i < d4; i += blockDim.x) { hr[i] = make_float4(0.f, 0.f, 0.f, 0.f); lr[i] = hr[i]; }
    if (threadIdx.x == 0) scale[ob] = 0.0f;
    return;
  }
  const float4* er = emb ? reinterpret_cast<const float4*>(emb + (size_t)token[b] * d) : nullptr;
  float ss = 0.0f;
  for (int i = threadIdx.x; i < d4; i += blockDim.x) {
    float4 v;
    if (er) v = er[i];
    else {
      v = xr[i];
      if (tp.n) tp_add_partials(v, tp, zstride, (size_t)b * d + 4 * i);
      else {
        // eight partials of this element are requested before the first one is added (one L2 round trip per eight splits,
        // not per four -- the block may have 1024 threads: 64 registers each); the sum keeps the fixed order z = 0, 1, ...
        const float* pp = part + (size_t)b * d + 4 * i;
        for (int z0 = 0; z0 < nz; z0 += 8) {
          float4 pz[8];
#pragma unroll
          for (int u = 0; u < 8; u++)
            if (z0 + u < nz) pz[u] = ldg_stream(reinterpret_cast<const float4*>(pp + (size_t)(z0 + u) * zstride));
#pragma unroll
          for (int u = 0; u < 8; u++)
            if (z0 + u < nz) { v.x += pz[u].x; v.y += pz[u].y; v.z += pz[u].z; v.w += pz[u].w; }
        }
      }
    }
    if ((er || nz) && !gather) xr[i] = v;
    ss = dot4(v, v, ss);
    const float4 g = reinterpret_cast<const float4*>(gain)[i];
    v.x *= g.x; v.y *= g.y; v.z *= g.z; v.w *= g.w;
    const float4 h = make_float4(st_tf32(v.x), st_tf32(v.y), st_tf32(v.z), st_tf32(v.w));
    hr[i] = h;
    lr[i] = make_float4(v.x - h.x, v.y - h.y, v.z - h.z, v.w - h.w);
  }
  const float tot = block_sum(ss, red);
  if (threadIdx.x == 0) {
    float m = tot / d;
    m += 1e-5f;
    scale[ob] = 1.0f / sqrtf(m);
    kt_stamp(kt, 6);
  }
}

// q|k|v rows of every slot (sum of nz partials): RoPE on q and k pairs (run.cc:271-286 with the host-built
// cos/sin tables), q -> state, k/v -> the cache row of pos[b].  A thread takes TWO pairs a block-width apart (two
// independent chains of loads).  grid (pairs / 512, B).
__global__ void __launch_bounds__(256) tc_rope_scatter_kernel(const float* __restrict__ part, int nz, long long zstride,
                                                              int ld, const int* __restrict__ pos, int d, int kvd, int hs,
                                                              const float* __restrict__ rc, const float* __restrict__ rs,
                                                              float* __restrict__ q, float* __restrict__ kc,
                                                              float* __restrict__ vc, long long kv_bs, int loff,
                                                              const int* __restrict__ slot, const KvPaged pg) {
  const int kt = threadIdx.x == 0 ? kt_begin(KT_ROPE) : -1;
  pdl_trigger();   // the next kernel may start its launch (it waits for this grid to complete before touching memory)
  pdl_wait();
  kt_stamp(kt, 1);
  const int b = blockIdx.y, pb = pos[b];
  if (pb < 0) return;
  const int sb = slot ? slot[b] : b;   // cache slab of this row
  // where the row of position pb lives: the slot's dense slab, or a block of the pool (BatchManager)
  const long long rowoff = pg.table ? kv_paged_off(pg, sb, pb, kvd) : (long long)sb * kv_bs + loff + (long long)pb * kvd;
  float2 v[2];
  int idx[2];
#pragma unroll
  for (int u = 0; u < 2; u++) {
    const int pair = (blockIdx.x * 2 + u) * blockDim.x + threadIdx.x;
    idx[u] = pair * 2;
    v[u] = make_float2(0.f, 0.f);
    if (idx[u] < d + 2 * kvd) {
#pragma unroll 4
      for (int z = 0; z < nz; z++) {
        const float2 pz = *reinterpret_cast<const float2*>(part + (size_t)z * zstride + (size_t)b * ld + idx[u]);
        v[u].x += pz.x; v[u].y += pz.y;
      }
    }
  }
#pragma unroll
  for (int u = 0; u < 2; u++) {
    const int i = idx[u];
    if (i >= d + 2 * kvd) continue;
    if (i < d + kvd) {
      const int li = (i < d) ? i : i - d;
      const int j = (li % hs) >> 1;
      const float c = rc[(size_t)pb * (hs >> 1) + j], s = rs[(size_t)pb * (hs >> 1) + j];
      // one product rounded, the other fused: the SAME expression as the attention kernel's in-kernel RoPE (attention.cu), so the
      // two forms of the step leave identical bits in the cache
      const float r0 = fmaf(v[u].x, c, -__fmul_rn(v[u].y, s)), r1 = fmaf(v[u].x, s, __fmul_rn(v[u].y, c));
      if (i < d) *reinterpret_cast<float2*>(q + (size_t)b * d + li) = make_float2(r0, r1);
      else *reinterpret_cast<float2*>(kc + rowoff + li) = make_float2(r0, r1);
    } else {
      *reinterpret_cast<float2*>(vc + rowoff + (i - d - kvd)) = v[u];
    }
  }
  kt_stamp(kt, 6);
}

int rope_scatter_launch(cudaStream_t st, bool pdl, int R, const float* part, int nz, long long zstride, int ld, const int* pos, int d,
                        int kvd, int hs, const float* rc, const float* rs, float* q, float* kc, float* vc, long long kv_bs, int loff,
                        const int* slot, const KvPaged* pg) {
  THA_CUDA_OK(launch(tc_rope_scatter_kernel, dim3(ceil_div(ld / 2, 512), R), dim3(256), 0, st, pdl, part, nz, zstride, ld, pos, d, kvd, hs,
                     rc, rs, q, kc, vc, kv_bs, loff, slot, pg ? *pg : KvPaged{}));
  return THA_STATUS_SUCCESS;
}

// hb = silu(a) * c with a = w1.x, c = w3.x (sum of nz partials each), written as tf32 hi/lo.  Two consecutive values per
// thread (64-bit loads / stores; h even): 256 slots x 2048 values are then 1024 blocks = one wave, not two.  grid (h / 512, B)
__global__ void __launch_bounds__(256) tc_swiglu_split_kernel(const float* __restrict__ part, int nz, long long zstride,
                                                              int ld, const int* __restrict__ pos, int h,
                                                              float* __restrict__ hh, float* __restrict__ hl) {
  const int kt = threadIdx.x == 0 ? kt_begin(KT_SWIGLU) : -1;
  pdl_trigger();   // the next kernel may start its launch (it waits for this grid to complete before touching memory)
  pdl_wait();
  kt_stamp(kt, 1);
  const int b = blockIdx.y;
  const int i = 2 * (blockIdx.x * blockDim.x + threadIdx.x);
  if (i >= h) return;
  float2 r = make_float2(0.f, 0.f);
  if (pos[b] >= 0) {
    float2 a = make_float2(0.f, 0.f), c = a;
#pragma unroll 4
    for (int z = 0; z < nz; z++) {
      const float2 az = *reinterpret_cast<const float2*>(part + (size_t)z * zstride + (size_t)b * ld + i);
      const float2 cz = *reinterpret_cast<const float2*>(part + (size_t)z * zstride + (size_t)b * ld + h + i);
      a.x += az.x; a.y += az.y; c.x += cz.x; c.y += cz.y;
    }
    r = a;
    r.x *= (1.0f / (1.0f + expf(-r.x)));   // run.cc:344-347 operation order
    r.y *= (1.0f / (1.0f + expf(-r.y)));
    r.x *= c.x; r.y *= c.y;
  }
  const float2 hi = make_float2(st_tf32(r.x), st_tf32(r.y));
  *reinterpret_cast<float2*>(hh + (size_t)b * h + i) = hi;
  *reinterpret_cast<float2*>(hl + (size_t)b * h + i) = make_float2(r.x - hi.x, r.y - hi.y);
  kt_stamp(kt, 6);
}


// ---- Q8_0 rows (runq.c:344-481 on the tensor cores) ---------------------------------------------------------------
// The activation that enters a matmul is quantised exactly as runq.c:145-171 does (scale = max|x| / 127 per group,
// round half away from zero) -- the int8 values and scales are bit-identical to q8_prep_kernel's for the same input --
// and handed to the int8 tensor-core matmul (gemm_q8i_kernel) as int8 rows + transposed group scales.

// q8_quant4 (common.cuh): four consecutive values per thread, a group = gs/4 consecutive lanes.
// One block per output row (up to 1024 threads, one or more float4 per thread, all loads of a thread independent):
// x row = embedding row or x + sum of partials (written back), RMSNorm in the reference's order w * (ss * x)
// (run.cc:196-209), group quantisation.  gs in {32, 64, 128}; d % gs == 0.
constexpr int PREP_MAX_V = 4;   // float4 per thread kept in registers: d <= 4 * 1024 * 4
__global__ void __launch_bounds__(1024) tc_prep_q8_kernel(float* __restrict__ x, const float* __restrict__ part, int nz,
                                                          long long zstride, const float* __restrict__ emb,
                                                          const int* __restrict__ token, const int* __restrict__ pos,
                                                          const float* __restrict__ gain, int d, int gs,
                                                          int8_t* __restrict__ xq, float* __restrict__ xsT, int ldxs,
                                                          const int* __restrict__ gather, const TpPeers tp) {
  __shared__ float red[33];
  const int kt = threadIdx.x == 0 ? kt_begin(KT_PREP) : -1;
  pdl_trigger();   // the next kernel may start its launch (it waits for this grid to complete before touching memory)
  pdl_wait();
  kt_stamp(kt, 1);
  if (tp.n && !emb) tp_wait_peers(tp);
  const int ob = blockIdx.x;
  const int b = gather ? gather[ob] : ob;
  const int d4 = d >> 2;
  float* xr = x + (size_t)b * d;
  int8_t* qr = xq + (size_t)ob * d;
  if (pos[b] < 0) {   // idle slot: zeros in, zeros out
    for (int i = threadIdx.x; i < d4; i += blockDim.x) reinterpret_cast<unsigned int*>(qr)[i] = 0u;
    for (int g = threadIdx.x; g < d / gs; g += blockDim.x) xsT[(size_t)g * ldxs + ob] = 0.0f;
    return;
  }
  const float* er = emb ? emb + (size_t)token[b] * d : nullptr;
  float4 v[PREP_MAX_V];
  float ss = 0.0f;
#pragma unroll
  for (int k = 0; k < PREP_MAX_V; k++) {
    const int i = threadIdx.x + k * blockDim.x;
    v[k] = make_float4(0.f, 0.f, 0.f, 0.f);
    if (i < d4) {
      if (er) v[k] = reinterpret_cast<const float4*>(er)[i];
      else {
        v[k] = reinterpret_cast<const float4*>(xr)[i];
        if (tp.n) tp_add_partials(v[k], tp, zstride, (size_t)b * d + 4 * i);
        else {
#pragma unroll 8
          for (int z = 0; z < nz; z++) {   // fixed order
            const float4 pz = *reinterpret_cast<const float4*>(part + (size_t)z * zstride + (size_t)b * d + 4 * i);
            v[k].x += pz.x; v[k].y += pz.y; v[k].z += pz.z; v[k].w += pz.w;
          }
        }
      }
      if ((er || nz) && !gather) reinterpret_cast<float4*>(xr)[i] = v[k];
      ss = dot4(v[k], v[k], ss);
    }
  }
  ss = block_sum(ss, red);
  ss /= d; ss += 1e-5f;
  const float ssn = 1.0f / sqrtf(ss);
#pragma unroll
  for (int k = 0; k < PREP_MAX_V; k++) {
    const int i = threadIdx.x + k * blockDim.x;
    if (k * (int)blockDim.x < d4) {               // block-uniform: whole warps take part in the group reductions
      const bool in = i < d4;
      float4 t = make_float4(0.f, 0.f, 0.f, 0.f);
      if (in) {
        const float4 g4 = reinterpret_cast<const float4*>(gain)[i];
        t.x = g4.x * (ssn * v[k].x); t.y = g4.y * (ssn * v[k].y); t.z = g4.z * (ssn * v[k].z); t.w = g4.w * (ssn * v[k].w);
      }
      q8_quant4(t, gs, i, qr, xsT, ldxs, ob, in);
    }
  }
  kt_stamp(kt, 6);
}

// One float4 per thread: src row (attention output), or with h > 0 the SwiGLU of the w1|w3 partials (run.cc:344-347
// order), quantised per group (int8 row + transposed scales).  grid (ceil(K / 4 / 256), rows)
__global__ void __launch_bounds__(256) tc_quant_split_kernel(const float* __restrict__ src, long long ld, int nz,
                                                             long long zstride, int h, const int* __restrict__ pos, int K,
                                                             int gs, int8_t* __restrict__ oq, float* __restrict__ osT, int ldxs) {
  const int kt = threadIdx.x == 0 ? kt_begin(KT_SWIGLU) : -1;
  pdl_trigger();   // the next kernel may start its launch (it waits for this grid to complete before touching memory)
  pdl_wait();
  kt_stamp(kt, 1);
  const int b = blockIdx.y;
  const int i4 = blockIdx.x * blockDim.x + threadIdx.x;
  const bool in = 4 * i4 < K;   // K % gs == 0: a group is never cut, but a warp may be
  float4 r = make_float4(0.f, 0.f, 0.f, 0.f);
  if (in && pos[b] >= 0) {
    if (h > 0) {
      float4 a = make_float4(0.f, 0.f, 0.f, 0.f), c = a;
#pragma unroll 8
      for (int z = 0; z < nz; z++) {
        const float4 pa = *reinterpret_cast<const float4*>(src + (size_t)z * zstride + (size_t)b * ld + 4 * i4);
        const float4 pc = *reinterpret_cast<const float4*>(src + (size_t)z * zstride + (size_t)b * ld + h + 4 * i4);
        a.x += pa.x; a.y += pa.y; a.z += pa.z; a.w += pa.w;
        c.x += pc.x; c.y += pc.y; c.z += pc.z; c.w += pc.w;
      }
      r.x = a.x * (1.0f / (1.0f + expf(-a.x))); r.x *= c.x;
      r.y = a.y * (1.0f / (1.0f + expf(-a.y))); r.y *= c.y;
      r.z = a.z * (1.0f / (1.0f + expf(-a.z))); r.z *= c.z;
      r.w = a.w * (1.0f / (1.0f + expf(-a.w))); r.w *= c.w;
    } else {
      r = *reinterpret_cast<const float4*>(src + (size_t)b * ld + 4 * i4);
    }
  }
  q8_quant4(r, gs, i4, oq + (size_t)b * K, osT, ldxs, b, in);
  kt_stamp(kt, 6);
}

// ---- scratch ----------------------------------------------------------------------------
struct TcScratch {
  float *xh = nullptr, *xl = nullptr, *hh = nullptr, *hl = nullptr, *scale = nullptr, *part = nullptr;
  float *xr = nullptr, *qr = nullptr, *xbr = nullptr;   // row-major activations of a prefill chunk (decode uses RunState's)
  int* tab = nullptr;                                   // prefill row / tile tables
  int8_t *xq = nullptr, *hq = nullptr;                  // Q8_0: quantised matmul inputs [R, d], [R, hidden]
  float *xsT = nullptr, *hsT = nullptr;                 // their group scales, transposed [groups, ldxs]
  size_t n_xq = 0, n_hq = 0, n_xsT = 0, n_hsT = 0;
  // chained matmuls (chain_tc.cu): attention output hi/lo, sums of squares per row tile, counter blocks (one per launch of a step)
  float *ah = nullptr, *al = nullptr, *ss = nullptr; int* ctr = nullptr;
  size_t n_ah = 0, n_al = 0, n_ss = 0; int ctr_stride = 0, n_ctr_blocks = 0;
  size_t n_x = 0, n_xl = 0, n_h = 0, n_hl = 0, n_part = 0, n_xr = 0, n_qr = 0, n_xbr = 0, n_tab = 0; int n_b = 0;
};
static std::map<Workspace*, TcScratch> g_tc_scratch;   // per workspace (= per device, or per tensor-parallel rank)
static std::mutex g_tc_scratch_mu;

void step_tc_release(Workspace* ws) {
  std::lock_guard<std::mutex> lk(g_tc_scratch_mu);
  auto it = g_tc_scratch.find(ws);
  if (it == g_tc_scratch.end()) return;
  TcScratch& s = it->second;
  void* ptrs[] = {s.xh, s.xl, s.hh, s.hl, s.scale, s.part, s.xr, s.qr, s.xbr, s.tab, s.xq, s.hq, s.xsT, s.hsT, s.ah, s.al, s.ss, s.ctr};
  for (void* p : ptrs) if (p) cudaFree(p);
  g_tc_scratch.erase(it);
}

static int grow(float** p, size_t* cap, size_t need, Workspace* ws) {
  if (need <= *cap) return THA_STATUS_SUCCESS;
  if (*p) THA_CUDA_OK(cudaFree(*p));
  THA_CUDA_OK(cudaMalloc(p, need * sizeof(float)));
  *cap = need;
  ws_drop_graphs(ws);   // cached graphs captured the old pointers
  gemm_tc_forget();
  return THA_STATUS_SUCCESS;
}

struct TcPlan { int z_qkv, z_wo, z_w13, z_w2; size_t part_floats; };

// Split-K factor.  Every CTA streams its share of the weights at the same rate, so the matmul runs at
// (CTAs / (waves * 148)) of the achievable bandwidth: pick the split count that fills whole waves of the 148 SMs
// (one CTA per SM), with at least 4 k-blocks per split and a charge per split for the partials the consumer
// has to sum (their bytes relative to the weight bytes: z * N / K).  E.g. 7B wq|wk|wv at 64 rows: 96 tiles ->
// z = 3 (288 CTAs, 2 waves at 97 %) instead of z = 2 (192 CTAs, 65 %).
static int pick_splits(int tiles, int K, int N, int align) {
  const int kb = ceil_div(K, 32);
  int best = 1;
  float best_score = -1.0f;
  const float charge = 0.012f + 0.25f * (float)N / (float)K / 16.0f;
  for (int z = 1; z <= 16; z++) {
    if (z > 1 && kb / z < 4) break;
    const int per = ceil_div(ceil_div(kb, z), align) * align;   // Q8_0: a split starts on a 4-k-block ring slot
    if (ceil_div(kb, per) != z) continue;                        // the kernel needs every split non-empty
    const int ctas = tiles * z;
    const int sms = sm_budget();
    const float eff = (float)ctas / (float)(ceil_div(ctas, sms) * sms);
    const float score = eff - charge * (z - 1);
    if (score > best_score) { best_score = score; best = z; }
  }
  return best;
}

// Split count for the persistent int8 matmul: 148 CTAs walk `items` = tiles x splits work items round-robin, an item
// costs its (padded) slots plus a few slot-times of flush / restart, every extra split a little for the partial sums.
static int pick_splits_q8(int tiles, int K, int N, int gs) {
  const int slots = ceil_div(K, 128);
  const int nt = gemm_q8i_pick_nt(N, gs), sb = gs / nt, ny = ceil_div(N, nt);
  int best = 1;
  float best_cost = 1e30f;
  for (int z = 1; z <= 16 && z <= slots; z++) {
    const int per = ceil_div(slots, z);
    if (ceil_div(slots, per) != z) continue;
    const int per_pad = ceil_div(per, sb) * sb;
    const int items = tiles * ny * z;
    const float cost = (float)ceil_div(items, sm_budget()) * (per_pad + 3.0f) + 0.5f * (z - 1);
    if (cost < best_cost) { best_cost = cost; best = z; }
  }
  return best;
}

// Local shapes of the pipeline: d = residual width; da = attention width (heads on this rank x head size), kvd, hd =
// this rank's share under tensor parallelism (da == d, hd == hidden_dim without it).
struct Dims { int d, da, kvd, hd, hs; };
static Dims dims_of(const tha_config_t& c, const TpCtx* tp) {
  Dims m;
  m.d = c.dim; m.hs = tp ? tp->head_size : c.dim / c.n_heads;
  m.da = m.hs * c.n_heads; m.kvd = m.hs * c.n_kv_heads; m.hd = c.hidden_dim;
  return m;
}

static int plan_z(int tiles, int K, int R, bool q8, int gs) {
  if (q8) return pick_splits_q8(tiles, K, R, gs);
  return pick_splits(tiles * ceil_div(R, gemm_tc_pick_nt(R)), K, R, 1);
}

static TcPlan make_plan(const Dims& m, int R, bool q8, int gs = 64) {
  const int d = m.d, da = m.da, hd = m.hd, kvd = m.kvd;
  TcPlan p;
  p.z_qkv = plan_z(ceil_div(da, 128) + 2 * ceil_div(kvd, 128), d, R, q8, gs);
  p.z_wo = plan_z(ceil_div(d, 128), da, R, q8, gs);
  p.z_w13 = plan_z(2 * ceil_div(hd, 128), d, R, q8, gs);
  p.z_w2 = plan_z(ceil_div(d, 128), hd, R, q8, gs);
  {   // development aid: THA_TC_Z="qkv,wo,w13,w2" overrides the planner (0 keeps its choice)
    static int ov[4] = {-1, 0, 0, 0};
    if (ov[0] < 0) {
      ov[0] = 0;
      const char* v = getenv("THA_TC_Z");
      if (v && *v) sscanf(v, "%d,%d,%d,%d", &ov[0], &ov[1], &ov[2], &ov[3]);
    }
    if (ov[0] > 0) p.z_qkv = ov[0];
    if (ov[1] > 0) p.z_wo = ov[1];
    if (ov[2] > 0) p.z_w13 = ov[2];
    if (ov[3] > 0) p.z_w2 = ov[3];
  }
  size_t m_ = (size_t)p.z_qkv * (da + 2 * kvd);
  if ((size_t)p.z_wo * d > m_) m_ = (size_t)p.z_wo * d;
  if ((size_t)p.z_w13 * 2 * hd > m_) m_ = (size_t)p.z_w13 * 2 * hd;
  if ((size_t)p.z_w2 * d > m_) m_ = (size_t)p.z_w2 * d;
  p.part_floats = m_ * R;
  return p;
}

static bool rows_supported(const tha_config_t* c, int B, const TpCtx* tp) {
  if (B < 1) return false;
  const int d = c->dim, hd = c->hidden_dim;
  if (c->n_heads < 1 || c->n_kv_heads < 1 || c->n_heads % c->n_kv_heads) return false;
  if (!tp && d % c->n_heads) return false;
  const int hs = tp ? tp->head_size : d / c->n_heads, kvd = hs * c->n_kv_heads, da = hs * c->n_heads;
  if (d % 4 || hd % 4 || kvd % 4 || da % 4 || (hs & 1) || d < 32 || hd < 32 || da < 32) return false;
  return attn_fused_supported(c->n_heads, hs, da, kvd);
}
bool step_tc_supported(const tha_config_t* c, int B) { return rows_supported(c, B, nullptr); }

bool prefill_tc_supported(const tha_config_t* c, const TpCtx* tp) {
  if (!rows_supported(c, 1, tp)) return false;
  const int hs = tp ? tp->head_size : c->dim / c->n_heads;
  return attn_prefill_supported(hs, hs * c->n_kv_heads);
}

static void chain_needs(const tha_config_t& c, int R, size_t* part_floats, size_t* ss_floats, int* ctr_ints);

// Allocate / grow everything R rows need (never inside a stream capture).
static int prepare_rows(Workspace* ws, const tha_config_t* c, int R, bool prefill, int n_tab, bool q8, int gs = 64,
                        const TpCtx* tp = nullptr) {
  THA_TRY_RC(gemm_tc_init(nullptr));
  std::lock_guard<std::mutex> lk(g_tc_scratch_mu);
  TcScratch& s = g_tc_scratch[ws];
  const Dims dm = dims_of(*c, tp);
  const TcPlan pl = make_plan(dm, R, q8, gs);
  if (q8) {   // int8 rows (sized in floats: bytes / 4) and transposed scales (rows padded to the widest column tile)
    const size_t ldxs = (size_t)ceil_div(R, 64) * 64;
    THA_TRY_RC(grow(reinterpret_cast<float**>(&s.xq), &s.n_xq, ((size_t)R * c->dim + 3) / 4, ws));
    THA_TRY_RC(grow(reinterpret_cast<float**>(&s.hq), &s.n_hq, ((size_t)R * c->hidden_dim + 3) / 4, ws));
    THA_TRY_RC(grow(&s.xsT, &s.n_xsT, ldxs * (c->dim / 32), ws));
    THA_TRY_RC(grow(&s.hsT, &s.n_hsT, ldxs * (c->hidden_dim / 32), ws));
  } else {
    THA_TRY_RC(grow(&s.xh, &s.n_x, (size_t)R * c->dim, ws));
    THA_TRY_RC(grow(&s.xl, &s.n_xl, (size_t)R * c->dim, ws));
    THA_TRY_RC(grow(&s.hh, &s.n_h, (size_t)R * c->hidden_dim, ws));
    THA_TRY_RC(grow(&s.hl, &s.n_hl, (size_t)R * c->hidden_dim, ws));
  }
  size_t part_need = pl.part_floats;
  if (!q8 && !prefill && !tp && chain_mode() != 0) {   // the chained step (chain_tc.cu)
    size_t pf = 0, sf = 0;
    int ci = 0;
    chain_needs(*c, R, &pf, &sf, &ci);
    if (pf > part_need) part_need = pf;
    THA_TRY_RC(grow(&s.ah, &s.n_ah, (size_t)R * dm.da, ws));
    THA_TRY_RC(grow(&s.al, &s.n_al, (size_t)R * dm.da, ws));
    THA_TRY_RC(grow(&s.ss, &s.n_ss, sf, ws));
    const int stride = ceil_div(ci, 32) * 32, blocks = c->n_layers + 1;
    if (stride > s.ctr_stride || blocks > s.n_ctr_blocks) {
      if (s.ctr) THA_CUDA_OK(cudaFree(s.ctr));
      s.ctr_stride = stride > s.ctr_stride ? stride : s.ctr_stride;
      s.n_ctr_blocks = blocks > s.n_ctr_blocks ? blocks : s.n_ctr_blocks;
      THA_CUDA_OK(cudaMalloc(&s.ctr, (size_t)s.ctr_stride * s.n_ctr_blocks * sizeof(int)));
      THA_CUDA_OK(cudaMemset(s.ctr, 0, (size_t)s.ctr_stride * s.n_ctr_blocks * sizeof(int)));   // every launch leaves its block zero again
      ws_drop_graphs(ws);
    }
  }
  THA_TRY_RC(grow(&s.part, &s.n_part, part_need, ws));
  if (R > s.n_b) {
    if (s.scale) cudaFree(s.scale);
    THA_CUDA_OK(cudaMalloc(&s.scale, (size_t)R * sizeof(float)));
    s.n_b = R;
    ws_drop_graphs(ws);
  }
  if (prefill) {
    THA_TRY_RC(grow(&s.xr, &s.n_xr, (size_t)R * c->dim, ws));
    THA_TRY_RC(grow(&s.qr, &s.n_qr, (size_t)R * c->dim, ws));
    THA_TRY_RC(grow(&s.xbr, &s.n_xbr, (size_t)R * c->dim, ws));
    if ((size_t)n_tab > s.n_tab) {
      if (s.tab) THA_CUDA_OK(cudaFree(s.tab));
      THA_CUDA_OK(cudaMalloc(&s.tab, (size_t)n_tab * sizeof(int)));
      s.n_tab = n_tab;
    }
    return THA_STATUS_SUCCESS;
  }
  return attn_fused_prepare(ws, R, c->n_heads, c->n_kv_heads, dm.hs, c->seq_len);
}

int step_tc_prepare(Workspace* ws, const tha_config_t* c, int B, const TpCtx* tp) { return prepare_rows(ws, c, B, false, 0, false, 64, tp); }
int step_tc_q8_prepare(Workspace* ws, const tha_config_t* c, int B, int gs, const TpCtx* tp) { return prepare_rows(ws, c, B, false, 0, true, gs, tp); }

// Q8_0 on the tensor cores: group size a multiple of the 32-value k-block (32 / 64 / 128), dims a multiple of it and >= 128.
bool step_tc_q8_supported(const tha_config_t* c, int B, int gs, const TpCtx* tp) {
  if (!rows_supported(c, B, tp)) return false;
  if (gs != 32 && gs != 64 && gs != 128) return false;
  const int da = tp ? tp->head_size * c->n_heads : c->dim;
  return c->dim % gs == 0 && c->hidden_dim % gs == 0 && da % gs == 0 && c->dim >= 128 && c->hidden_dim >= 128 && da >= 128 &&
         c->dim <= 16384;
}
bool step_tc_tp_supported(const tha_config_t* c, int B, const TpCtx* tp) { return rows_supported(c, B, tp); }

static int gemm(cudaStream_t st, bool pdl, int n_mats, const float* const* W, const int* M, const int* off, int K,
                const float* xh, const float* xl, int N, float* out, long long ldo, long long zstride, int splits,
                const float* colscale) {
  TcArgs a = {};
  a.n_mats = n_mats;
  for (int i = 0; i < n_mats; i++) { a.W[i] = W[i]; a.M[i] = M[i]; a.out_off[i] = off[i]; }
  a.K = K; a.Xhi = xh; a.Xlo = xl; a.N = N; a.ldx = K; a.out = out; a.ldo = ldo; a.zstride = zstride; a.splits = splits;
  a.colscale = colscale;
  return gemm_tc_launch(st, a, pdl);
}

static int gemm_q8(cudaStream_t st, bool pdl, int n_mats, const int8_t* const* Wq, const float* const* Ws, int gs, const int* M,
                   const int* off, int K, const int8_t* xq, const float* xsT, int ldxs, int N, float* out, long long ldo,
                   long long zstride, int splits) {
  Q8iArgs a = {};
  a.n_mats = n_mats;
  for (int i = 0; i < n_mats; i++) { a.Wq[i] = Wq[i]; a.Ws[i] = Ws[i]; a.M[i] = M[i]; a.out_off[i] = off[i]; }
  a.gs = gs; a.K = K; a.Xq = xq; a.xsT = xsT; a.ldxs = ldxs; a.N = N; a.out = out; a.ldo = ldo; a.zstride = zstride; a.splits = splits;
  return gemm_q8i_launch(st, a, pdl);
}

// Development aid (THA_SKIP bit mask): leave a class of kernels out of the step to read its LIVE share of the step time off
// the difference (results are garbage): 1 layer matmuls, 2 attention, 4 prep, 8 RoPE/scatter, 16 SwiGLU / quantise,
// 32 classifier, 64 argmax.
static int skip_mask() { return skip_mode(); }   // workspace.cu: THA_SKIP / tha_set_option("skip", mask)

// What differs between a decode step and a prefill chunk.
struct RowsIO {
  int R;                                   // rows
  float *x, *q, *xb;                       // [R, d] residual stream / rotated queries / attention output
  const int *token, *pos, *slot;           // [R]; slot == nullptr: row r lives in cache slab r (decode)
  int n_tiles; const int *tile_row0, *tile_n;   // prefill attention tiles (n_tiles == 0: per-row decode attention)
  int n_out; const int* out_rows;          // classifier rows (out_rows == nullptr: every row)
  float* logits; int* next;                // [n_out, vocab] (nullable: skip the classifier), [n_out] (nullable)
};

// ---- chained matmuls (chain_tc.cu) ---------------------------------------------------------------------------------
// Launch k of a decode step: k = 0: wq|wk|wv of layer 0 (its input comes from the embedding prep); k = 1..L: wo, w1|w3,
// w2 of layer k-1 and then wq|wk|wv of layer k, or the classifier after the last layer.  Attention sits between two
// launches.  fp32 weights, one GPU (the tensor-parallel exchange and the Q8_0 rows keep one launch per matmul).
static void chain_splits_override(int which, int* z) {   // development aid: THA_CHAIN_Z="wo,w13,w2,qkv,cls" (0 keeps the planner's)
  static int ov[5] = {-1, 0, 0, 0, 0};
  if (ov[0] < 0) {
    ov[0] = 0;
    const char* v = getenv("THA_CHAIN_Z");
    if (v && *v) sscanf(v, "%d,%d,%d,%d,%d", &ov[0], &ov[1], &ov[2], &ov[3], &ov[4]);
  }
  if (ov[which] > 0) *z = ov[which];
}

static ChainArgs chain_of_launch(int k, const tha_config_t& c, const tha_weights_f32_t& w, const tha_state_t& s, const TcScratch& sc,
                                 const RowsIO& io, const float* rc, const float* rs) {
  const Dims dm = dims_of(c, nullptr);
  const int d = dm.d, da = dm.da, hd = dm.hd, hs = dm.hs, kvd = dm.kvd, L = c.n_layers;
  ChainArgs a = {};
  a.N = io.R; a.pos = io.pos; a.slot = io.slot;
  auto qkv = [&](int l) {
    ChainMatmul m = {};
    m.mode = CH_ROPE; m.n_mats = 3;
    m.W[0] = w.wq + (size_t)l * da * d; m.W[1] = w.wk + (size_t)l * d * kvd; m.W[2] = w.wv + (size_t)l * d * kvd;
    m.M[0] = da; m.M[1] = kvd; m.M[2] = kvd; m.K = d; m.Xhi = sc.xh; m.Xlo = sc.xl; m.colscale = sc.scale;
    m.out = io.q; m.ldo = da; m.kc = s.key_cache; m.vc = s.value_cache; m.kv_bs = (long long)L * c.seq_len * kvd;
    m.loff = (long long)l * c.seq_len * kvd; m.kvd = kvd; m.hs = hs; m.rc = rc; m.rs = rs;
    chain_splits_override(3, &m.splits);
    return m;
  };
  if (k == 0) { a.n = 1; a.mm[0] = qkv(0); return a; }
  const int l = k - 1;
  {
    ChainMatmul& m = a.mm[0];   // wo: x += wo . attention output; next input x * rms_ffn
    m.mode = CH_RESID; m.n_mats = 1; m.W[0] = w.wo + (size_t)l * d * da; m.M[0] = d; m.K = da; m.Xhi = sc.ah; m.Xlo = sc.al;
    m.out = io.x; m.ldo = d; m.oh = sc.xh; m.ol = sc.xl; m.gain = w.rms_ffn_weight + (size_t)l * d; m.scale_out = sc.scale;
    chain_splits_override(0, &m.splits);
  }
  {
    ChainMatmul& m = a.mm[1];   // w1 | w3 + SwiGLU
    m.mode = CH_SWIGLU; m.n_mats = 2; m.W[0] = w.w1 + (size_t)l * d * hd; m.W[1] = w.w3 + (size_t)l * d * hd; m.M[0] = hd; m.M[1] = hd;
    m.K = d; m.Xhi = sc.xh; m.Xlo = sc.xl; m.colscale = sc.scale; m.oh = sc.hh; m.ol = sc.hl;
    chain_splits_override(1, &m.splits);
  }
  {
    ChainMatmul& m = a.mm[2];   // w2: x += w2 . hb; next input x * (rms_att of the next layer | rms_final)
    m.mode = CH_RESID; m.n_mats = 1; m.W[0] = w.w2 + (size_t)l * d * hd; m.M[0] = d; m.K = hd; m.Xhi = sc.hh; m.Xlo = sc.hl;
    m.out = io.x; m.ldo = d; m.oh = sc.xh; m.ol = sc.xl;
    m.gain = k < L ? w.rms_att_weight + (size_t)k * d : w.rms_final_weight; m.scale_out = sc.scale;
    chain_splits_override(2, &m.splits);
  }
  a.n = 3;
  if (k < L) { a.mm[3] = qkv(k); a.n = 4; }
  else if (io.logits) {
    ChainMatmul& m = a.mm[3];
    m.mode = CH_PLAIN; m.n_mats = 1; m.W[0] = w.wcls; m.M[0] = c.vocab_size; m.K = d; m.Xhi = sc.xh; m.Xlo = sc.xl; m.colscale = sc.scale;
    m.out = io.logits; m.ldo = c.vocab_size;
    chain_splits_override(4, &m.splits);
    a.n = 4;
  }
  return a;
}

// Shapes the chained step covers: fp32, one GPU, a decode step (one row per slot, every row gets logits).
static bool chain_supported(const tha_config_t& c, int R, const TcScratch& sc, const tha_weights_f32_t& w, const tha_state_t& s) {
  if (chain_mode() == 0) return false;
  const Dims dm = dims_of(c, nullptr);
  if (dm.hd % 16 || dm.d % 4 || dm.da != dm.d) return false;
  (void)R; (void)sc; (void)w; (void)s;
  return true;
}

static void chain_needs(const tha_config_t& c, int R, size_t* part_floats, size_t* ss_floats, int* ctr_ints) {
  const tha_weights_f32_t w0 = {};
  const tha_state_t s0 = {};
  const TcScratch sc0;
  RowsIO io = {};
  io.R = R;
  float dummy = 0.0f;
  io.logits = &dummy;   // shapes only: nothing is dereferenced
  *part_floats = 0; *ss_floats = 0; *ctr_ints = 0;
  const int ks[3] = {0, 1, c.n_layers};
  for (int i = 0; i < 3; i++) {
    const ChainArgs a = chain_of_launch(ks[i], c, w0, s0, sc0, io, nullptr, nullptr);
    size_t pf, sf; int ci;
    chain_tc_need(a, &pf, &sf, &ci);
    if (pf > *part_floats) *part_floats = pf;
    if (sf > *ss_floats) *ss_floats = sf;
    if (ci > *ctr_ints) *ctr_ints = ci;
  }
}

// The decode step as L + 1 matmul chains with attention in between (28 launches for 12 layers instead of 112).
static int issue_rows_chain(cudaStream_t st, Workspace* ws, const tha_config_t& c, const tha_weights_f32_t& w, const tha_state_t& s,
                            const TcScratch& sc, const RowsIO& io, const float* rc, const float* rs, bool pdl) {
  const Dims dm = dims_of(c, nullptr);
  const int R = io.R, d = dm.d, L = c.n_layers, kvd = dm.kvd, hs = dm.hs;
  const long long kv_bs = (long long)L * c.seq_len * kvd;
  int prep_threads = ceil_div(d / 4, 32) * 32;
  if (prep_threads > 1024) prep_threads = 1024;
  if (prep_threads < 128) prep_threads = 128;
  const int skip = skip_mask();
  const TpPeers none = {};
  const int* nogather = nullptr;
  // embedding gather + RMSNorm(att, layer 0) folded + split (src/thaDNN.cpp:31-36, run.cc:196-209)
  THA_CUDA_OK(launch(tc_prep_kernel, dim3(R), dim3(prep_threads), 0, st, pdl, io.x, (const float*)nullptr, 0, 0LL,
                     w.token_embedding_table, io.token, io.pos, w.rms_att_weight, d, sc.xh, sc.xl, sc.scale, nogather, none));
  for (int k = 0; k <= L; k++) {
    ChainArgs a = chain_of_launch(k, c, w, s, sc, io, rc, rs);
    a.part = sc.part; a.part_floats = sc.n_part; a.ss_part = sc.ss; a.ss_floats = sc.n_ss;
    a.ctr = sc.ctr + (size_t)k * sc.ctr_stride; a.ctr_ints = sc.ctr_stride;
    if (!(skip & 1)) THA_TRY_RC(chain_tc_launch(st, a, pdl));
    if (k == L) break;
    const int loff = k * c.seq_len * kvd;
    if (!(skip & 2))
      THA_TRY_RC(attn_fused_launch(st, ws, R, io.pos, c.n_heads, c.n_kv_heads, io.q, io.xb, s.key_cache + loff, s.value_cache + loff,
                                   kv_bs, hs, c.seq_len, pdl, sc.ah, sc.al));
  }
  if (io.next && !(skip & 64)) THA_TRY_RC(argmax_launch(st, io.n_out, io.logits, c.vocab_size, io.next, pdl));
  return THA_STATUS_SUCCESS;
}

// `wq` != nullptr: Q8_0 weights (then `w` only lends its embedding table and norm gains).
// `tp` != nullptr: this rank's share of a tensor-parallel model (c = the LOCAL config: n_heads, n_kv_heads, hidden_dim
// are this rank's; dim is the full residual width): the wo / w2 partials go to the exchange buffers and their consumer
// sums every rank's (tp.h).
static int issue_rows(cudaStream_t st, Workspace* ws, const tha_config_t& c, const tha_weights_f32_t& w,
                      const tha_weights_q8_t* wq, const tha_state_t& s, const RowsIO& io, const float* rc, const float* rs,
                      bool pdl, const TpCtx* tp = nullptr) {
  TcScratch sc;
  {
    std::lock_guard<std::mutex> lk(g_tc_scratch_mu);
    sc = g_tc_scratch[ws];
  }
  const int R = io.R;
  const bool q8 = wq != nullptr;
  if (!q8 && !tp && io.n_tiles == 0 && !io.out_rows && io.n_out == R && io.logits && sc.ctr && chain_supported(c, R, sc, w, s) &&
      !kv_pool_find(s.key_cache, nullptr))
    return issue_rows_chain(st, ws, c, w, s, sc, io, rc, rs, pdl);
  const int gs = wq ? wq->group_size : 0;
  const Dims dm = dims_of(c, tp);
  const TcPlan pl = make_plan(dm, R, q8, q8 ? gs : 64);
  const int d = dm.d, da = dm.da, hd = dm.hd, hs = dm.hs, kvd = dm.kvd, L = c.n_layers;
  const long long kv_bs = (long long)L * c.seq_len * kvd;
  const int ld_qkv = da + 2 * kvd, ld_h = 2 * hd;
  const int* nogather = nullptr;
  const int ldxs = ceil_div(R, 64) * 64;
  int prep_threads = ceil_div(d / 4, 32) * 32;   // one float4 per thread where the row allows it
  if (prep_threads > 1024) prep_threads = 1024;
  if (prep_threads < 128) prep_threads = 128;
  // tensor parallelism: split counts of every rank's w2 (local hidden sizes may differ), exchange bookkeeping
  int z_w2_of[TP_MAX] = {0};
  const bool tpx = tp && tp->tp > 1;
  if (tpx) {
    size_t need = (size_t)pl.z_wo * R * d;
    for (int r = 0; r < tp->tp; r++) {
      z_w2_of[r] = plan_z(ceil_div(d, 128), tp->hidden_of[r], R, q8, q8 ? gs : 64);
      if ((size_t)z_w2_of[r] * R * d > need) need = (size_t)z_w2_of[r] * R * d;
    }
    if (need > tp->xchg_floats || !tp->connected || tp->hidden_of[tp->rank] != hd) return THA_STATUS_INVALID_ARG;
  }
  const int skip = skip_mask();
  KvPool pool;
  const bool paged = kv_pool_find(s.key_cache, &pool);
  if (paged && (tp || s.value_cache != pool.v)) return THA_STATUS_INVALID_ARG;
  int xk = 0;   // exchanges issued so far in this step
  auto peers_of = [&](int which, bool w2) -> TpPeers {
    TpPeers t = {};
    if (!tpx) return t;
    t.n = tp->tp; t.rank = tp->rank; t.k = xk++; t.nx = 2 * L;
    for (int r = 0; r < tp->tp; r++) {
      t.part[r] = tp->xchg(r, which);
      t.nz[r] = w2 ? z_w2_of[r] : pl.z_wo;
      t.peer_flags[r] = tp->flags(r);
    }
    t.epoch = tp->epoch(); t.err = tp->err();
    return t;
  };
  float* out_wo = tpx ? tp->xchg(tp->rank, 0) : sc.part;
  float* out_w2 = tpx ? tp->xchg(tp->rank, 1) : sc.part;
  // the (fused RMSNorm +) split that feeds a matmul whose input is the residual stream
  auto prep = [&](int rows, const float* part, int nz, long long zs, const float* emb, const float* gain, const int* gather,
                  const TpPeers& tpp) -> int {
    if (skip & 4) return THA_STATUS_SUCCESS;
    if (wq) {
      THA_CUDA_OK(launch(tc_prep_q8_kernel, dim3(rows), dim3(prep_threads), 0, st, pdl, io.x, part, nz, zs, emb, io.token,
                         io.pos, gain, d, gs, sc.xq, sc.xsT, ldxs, gather, tpp));
    } else {
      THA_CUDA_OK(launch(tc_prep_kernel, dim3(rows), dim3(prep_threads), 0, st, pdl, io.x, part, nz, zs, emb, io.token,
                         io.pos, gain, d, sc.xh, sc.xl, sc.scale, gather, tpp));
    }
    return THA_STATUS_SUCCESS;
  };
  bool mm_hidden = false;   // Q8_0: the next matmul reads the quantised hidden activations (w2) instead of the dim-wide ones
  // n stacked matrices of layer l: fp32 pointers or (int8, scales) pairs
  auto mm = [&](int n, const float* const* W, const tha_qtensor_t* const* Q, int l, const int* M, const int* off,
                int K, const float* xh, const float* xl, int N, float* out, long long ldo, long long zs, int z,
                const float* colscale) -> int {
    if ((skip & 1) && out != io.logits) return THA_STATUS_SUCCESS;
    if ((skip & 32) && out == io.logits) return THA_STATUS_SUCCESS;
    if (!wq) return gemm(st, pdl, n, W, M, off, K, xh, xl, N, out, ldo, zs, z, colscale);
    const int8_t* q[3]; const float* sc3[3];
    for (int i = 0; i < n; i++) {
      const size_t elems = (size_t)M[i] * K;
      q[i] = Q[i]->q + (size_t)l * elems;
      sc3[i] = Q[i]->s + (size_t)l * (elems / gs);
    }
    const bool hidden_in = mm_hidden;       // which quantised buffer feeds this matmul
    return gemm_q8(st, pdl, n, q, sc3, gs, M, off, K, hidden_in ? sc.hq : sc.xq, hidden_in ? sc.hsT : sc.xsT, ldxs, N, out, ldo, zs, z);
  };
  int prev_z = 0;                 // partials waiting to be added into x
  long long prev_zs = 0;
  TpPeers none = {};
  for (int l = 0; l < L; l++) {
    const int loff = l * c.seq_len * kvd;
    THA_TRY_RC(prep(R, out_w2, prev_z, prev_zs, l == 0 ? w.token_embedding_table : (const float*)nullptr,
                    w.rms_att_weight + (size_t)l * d, nogather, l == 0 ? none : peers_of(1, true)));
    {
      const float* W[3] = {nullptr, nullptr, nullptr};
      if (!wq) { W[0] = w.wq + (size_t)l * da * d; W[1] = w.wk + (size_t)l * d * kvd; W[2] = w.wv + (size_t)l * d * kvd; }
      const tha_qtensor_t* Q[3] = {wq ? &wq->wq : nullptr, wq ? &wq->wk : nullptr, wq ? &wq->wv : nullptr};
      const int M[3] = {da, kvd, kvd}, off[3] = {0, da, da + kvd};
      THA_TRY_RC(mm(3, W, Q, l, M, off, d, sc.xh, sc.xl, R, sc.part, ld_qkv, (long long)R * ld_qkv, pl.z_qkv, sc.scale));
    }
    const KvPaged pg = paged ? kv_paged_of(pool, L, kvd, l) : KvPaged{};
    const float* kcl = paged ? s.key_cache : s.key_cache + loff;      // layer's cache rows: pool base, or slab + layer offset
    const float* vcl = paged ? s.value_cache : s.value_cache + loff;
    // decode rows: RoPE + KV scatter ride in the attention kernel (one launch and one dependency less per layer)
    const bool rope_in_attn = io.n_tiles == 0 && !(!io.logits && l == L - 1) && !(skip & 2) && rope_fuse_mode() &&
                              attn_fused_rope_supported(c.n_heads, c.n_kv_heads, hs);
    if (!(skip & 8) && !rope_in_attn)
    THA_CUDA_OK(launch(tc_rope_scatter_kernel, dim3(ceil_div(ld_qkv / 2, 512), R), dim3(256), 0, st, pdl, (const float*)sc.part,
                       pl.z_qkv, (long long)R * ld_qkv, ld_qkv, io.pos, da, kvd, hs, rc, rs, io.q, s.key_cache, s.value_cache,
                       kv_bs, loff, io.slot, pg));
    if (!io.logits && l == L - 1) {   // cache-only ingestion: the last layer's K/V rows are in place
      if (tpx) THA_CUDA_OK(launch(tp_epoch_kernel, dim3(1), dim3(32), 0, st, pdl, tp->epoch()));
      return THA_STATUS_SUCCESS;
    }
    if (skip & 2) {
    } else if (io.n_tiles > 0) {
      THA_TRY_RC(attn_prefill_launch(st, io.n_tiles, c.n_heads, c.n_kv_heads, hs, io.q, io.xb, wq ? nullptr : sc.xh,
                                     wq ? nullptr : sc.xl, kcl, vcl, kv_bs, io.slot, io.pos,
                                     io.tile_row0, io.tile_n, pdl, &pg));
    } else if (rope_in_attn) {
      THA_TRY_RC(attn_fused_rope_launch(st, ws, R, io.pos, c.n_heads, c.n_kv_heads, io.q, io.xb, const_cast<float*>(kcl),
                                        const_cast<float*>(vcl), kv_bs, hs, c.seq_len, pdl, wq ? nullptr : sc.xh, wq ? nullptr : sc.xl,
                                        wq ? sc.xq : nullptr, wq ? sc.xsT : nullptr, ldxs, gs, &pg, sc.part, pl.z_qkv,
                                        (long long)R * ld_qkv, ld_qkv, rc, rs));
    } else {
      // Q8_0: the quantiser of the attention output (runq.c:437) runs in the attention kernel's final store
      THA_TRY_RC(attn_fused_launch(st, ws, R, io.pos, c.n_heads, c.n_kv_heads, io.q, io.xb, kcl,
                                   vcl, kv_bs, hs, c.seq_len, pdl, wq ? nullptr : sc.xh, wq ? nullptr : sc.xl,
                                   wq ? sc.xq : nullptr, wq ? sc.xsT : nullptr, ldxs, gs, &pg));
    }
    if ((skip & 16) || (wq && io.n_tiles == 0 && !(skip & 2))) {
    } else if (wq)   // prompt rows: runq.c:437 as its own launch
      THA_CUDA_OK(launch(tc_quant_split_kernel, dim3(ceil_div(da / 4, 256), R), dim3(256), 0, st, pdl, (const float*)io.xb, (long long)da,
                         0, 0LL, 0, io.pos, da, gs, sc.xq, sc.xsT, ldxs));
    {
      const float* W[1] = {wq ? nullptr : w.wo + (size_t)l * d * da};
      const tha_qtensor_t* Q[1] = {wq ? &wq->wo : nullptr};
      const int M[1] = {d}, off[1] = {0};
      THA_TRY_RC(mm(1, W, Q, l, M, off, da, sc.xh, sc.xl, R, out_wo, d, (long long)R * d, pl.z_wo, nullptr));
    }
    THA_TRY_RC(prep(R, out_wo, pl.z_wo, (long long)R * d, nullptr, w.rms_ffn_weight + (size_t)l * d, nogather, peers_of(0, false)));
    {
      const float* W[2] = {wq ? nullptr : w.w1 + (size_t)l * d * hd, wq ? nullptr : w.w3 + (size_t)l * d * hd};
      const tha_qtensor_t* Q[2] = {wq ? &wq->w1 : nullptr, wq ? &wq->w3 : nullptr};
      const int M[2] = {hd, hd}, off[2] = {0, hd};
      THA_TRY_RC(mm(2, W, Q, l, M, off, d, sc.xh, sc.xl, R, sc.part, ld_h, (long long)R * ld_h, pl.z_w13, sc.scale));
    }
    if (skip & 16) {
    } else if (wq)   // runq.c:457-465: SwiGLU, then quantise hb
      THA_CUDA_OK(launch(tc_quant_split_kernel, dim3(ceil_div(hd / 4, 256), R), dim3(256), 0, st, pdl, (const float*)sc.part,
                         (long long)ld_h, pl.z_w13, (long long)R * ld_h, hd, io.pos, hd, gs, sc.hq, sc.hsT, ldxs));
    else
      THA_CUDA_OK(launch(tc_swiglu_split_kernel, dim3(ceil_div(hd, 512), R), dim3(256), 0, st, pdl, (const float*)sc.part, pl.z_w13,
                         (long long)R * ld_h, ld_h, io.pos, hd, sc.hh, sc.hl));
    {
      const float* W[1] = {wq ? nullptr : w.w2 + (size_t)l * d * hd};
      const tha_qtensor_t* Q[1] = {wq ? &wq->w2 : nullptr};
      const int M[1] = {d}, off[1] = {0};
      mm_hidden = true;
      const int rc_w2 = mm(1, W, Q, l, M, off, hd, sc.hh, sc.hl, R, out_w2, d, (long long)R * d, pl.z_w2, nullptr);
      mm_hidden = false;
      THA_TRY_RC(rc_w2);
    }
    prev_z = pl.z_w2;
    prev_zs = (long long)R * d;
  }
  THA_TRY_RC(prep(io.n_out, out_w2, prev_z, prev_zs, nullptr, w.rms_final_weight, io.out_rows, peers_of(1, true)));
  {
    const float* W[1] = {wq ? nullptr : w.wcls};
    const tha_qtensor_t* Q[1] = {wq ? &wq->wcls : nullptr};
    const int M[1] = {c.vocab_size}, off[1] = {0};
    THA_TRY_RC(mm(1, W, Q, 0, M, off, d, sc.xh, sc.xl, io.n_out, io.logits, c.vocab_size, 0, 1, sc.scale));
  }
  if (io.next && !(skip & 64)) THA_TRY_RC(argmax_launch(st, io.n_out, io.logits, c.vocab_size, io.next, pdl));
  if (tpx) THA_CUDA_OK(launch(tp_epoch_kernel, dim3(1), dim3(32), 0, st, pdl, tp->epoch()));
  return THA_STATUS_SUCCESS;
}

int step_tc_issue(cudaStream_t st, Workspace* ws, int B, const tha_config_t& c, const tha_weights_f32_t& w,
                  const tha_state_t& s, const int* token_d, const int* pos_d, float* logits, int* next_d,
                  const float* rc, const float* rs, bool pdl, const TpCtx* tp) {
  RowsIO io = {};
  io.R = B; io.x = s.x; io.q = s.q; io.xb = s.xb; io.token = token_d; io.pos = pos_d; io.slot = nullptr;
  io.n_tiles = 0; io.n_out = B; io.out_rows = nullptr; io.logits = logits; io.next = next_d;
  return issue_rows(st, ws, c, w, nullptr, s, io, rc, rs, pdl, tp);
}

static tha_weights_f32_t q8_norms_view(const tha_weights_q8_t& wq) {
  tha_weights_f32_t w = {};
  w.token_embedding_table = wq.token_embedding_table;
  w.rms_att_weight = wq.rms_att_weight; w.rms_ffn_weight = wq.rms_ffn_weight; w.rms_final_weight = wq.rms_final_weight;
  return w;
}

int step_tc_q8_issue(cudaStream_t st, Workspace* ws, int B, const tha_config_t& c, const tha_weights_q8_t& wq,
                     const tha_state_t& s, const int* token_d, const int* pos_d, float* logits, int* next_d,
                     const float* rc, const float* rs, bool pdl, const TpCtx* tp) {
  RowsIO io = {};
  io.R = B; io.x = s.x; io.q = s.q; io.xb = s.xb; io.token = token_d; io.pos = pos_d; io.slot = nullptr;
  io.n_tiles = 0; io.n_out = B; io.out_rows = nullptr; io.logits = logits; io.next = next_d;
  return issue_rows(st, ws, c, q8_norms_view(wq), &wq, s, io, rc, rs, pdl, tp);
}

// Batched prompt ingestion.  Host arrays: sequence i occupies cache slab slot[i], its len[i] tokens (consecutive
// in `tokens`) go to positions pos0[i] .. pos0[i]+len[i]-1.  Every row's K/V lands in the cache; with `logits`
// the classifier runs on the last row of each sequence ([n_seqs, vocab], and argmax -> next[n_seqs]).
int step_tc_prefill(cudaStream_t st, Workspace* ws, const tha_config_t& c, const tha_weights_f32_t* w32,
                    const tha_weights_q8_t* wq, const tha_state_t& s, int n_seqs, const int* slot, const int* pos0,
                    const int* len, const int* tokens, float* logits, int* next_d, const float* rc, const float* rs,
                    const TpCtx* tp) {
  if (!prefill_tc_supported(&c, tp)) return THA_STATUS_NOT_SUPPORTED;
  if (wq && !step_tc_q8_supported(&c, 1, wq->group_size, tp)) return THA_STATUS_NOT_SUPPORTED;
  const tha_weights_f32_t w = wq ? q8_norms_view(*wq) : *w32;
  long long R = 0, T = 0;
  for (int i = 0; i < n_seqs; i++) {
    if (len[i] < 1 || pos0[i] < 0 || slot[i] < 0 || (long long)pos0[i] + len[i] > c.seq_len) return THA_STATUS_INVALID_ARG;
    R += len[i];
    T += ceil_div(len[i], 64);
  }
  if (R < 1 || R > (1 << 20)) return THA_STATUS_INVALID_ARG;
  // tables: [slot R | pos R | token R | tile_row0 T | tile_n T | last n_seqs]
  std::vector<int> tab((size_t)3 * R + 2 * T + n_seqs);
  int *h_slot = tab.data(), *h_pos = h_slot + R, *h_tok = h_pos + R, *h_t0 = h_tok + R, *h_tn = h_t0 + T, *h_last = h_tn + T;
  int r = 0, t = 0;
  for (int i = 0; i < n_seqs; i++) {
    for (int j = 0; j < len[i]; j += 64) { h_t0[t] = r + j; h_tn[t] = (len[i] - j < 64) ? len[i] - j : 64; t++; }
    for (int j = 0; j < len[i]; j++, r++) {
      h_slot[r] = slot[i]; h_pos[r] = pos0[i] + j; h_tok[r] = tokens[r];
      if (tokens[r] < 0 || tokens[r] >= c.vocab_size) return THA_STATUS_INVALID_ARG;
    }
    h_last[i] = r - 1;
  }
  THA_TRY_RC(prepare_rows(ws, &c, (int)R, true, (int)tab.size(), wq != nullptr, wq ? wq->group_size : 64, tp));
  TcScratch sc;
  {
    std::lock_guard<std::mutex> lk(g_tc_scratch_mu);
    sc = g_tc_scratch[ws];
  }
  // pageable source: the copy is staged before the call returns, so `tab` may go out of scope
  THA_CUDA_OK(cudaMemcpyAsync(sc.tab, tab.data(), tab.size() * sizeof(int), cudaMemcpyHostToDevice, st));
  RowsIO io = {};
  io.R = (int)R; io.x = sc.xr; io.q = sc.qr; io.xb = sc.xbr;
  io.slot = sc.tab; io.pos = sc.tab + R; io.token = sc.tab + 2 * R;
  io.n_tiles = (int)T; io.tile_row0 = sc.tab + 3 * R; io.tile_n = sc.tab + 3 * R + T;
  io.n_out = n_seqs; io.out_rows = sc.tab + 3 * R + 2 * T; io.logits = logits; io.next = logits ? next_d : nullptr;
  return issue_rows(st, ws, c, w, wq, s, io, rc, rs, pdl_enabled() != 0, tp);
}

}  // namespace tha
