// The bf16 model path (BASELINE.json configs[3] "fp32/bf16"; the reference declares 16-bit weight structs at
// include/models.hpp:77-117 and never uses them): weights stored as bf16, the residual stream, RMSNorm, RoPE, attention,
// SwiGLU and the KV cache in fp32 exactly as run.cc:244-374, and every vector that ENTERS a weight matmul rounded to bf16
// (round to nearest even) -- which is what the tensor cores consume.  The oracle restates the same rule on the CPU
// (oracle/llama2_oracle.c, orc_set_bf16_activations), so parity is a statement about fp32 accumulation order, not about
// bf16 rounding.
//
// One code path for a decode step (row = slot) and for prompt ingestion (row = prompt position), the row pipeline of
// step_tc.cu with gemm_bf16_kernel as the matmul:
//   prep  x += partials | embedding row; RMSNorm in the reference's order g * (ss * x); round -> bf16 row
//   gemm  wq|wk|wv -> partials;  rope_scatter (step_tc.cu);  attention (fp32);  round -> bf16
//   gemm  wo -> partials;  prep(ffn);  gemm w1|w3;  SwiGLU + round;  gemm w2 -> partials;  ... prep(final); gemm wcls
#include <stdlib.h>

#include <map>
#include <mutex>
#include <vector>

#include "common.cuh"
#include "gemm_tc.h"
#include "tp.h"
#include "workspace.h"

namespace tha {

int argmax_launch(cudaStream_t st, int B, const float* logits, int vocab, int* next, bool pdl);
bool attn_fused_supported(int n_heads, int head_size, int dim, int kv_dim);
int attn_fused_prepare(Workspace* ws, int B, int n_heads, int n_kv_heads, int head_size, int seq_len);
int attn_fused_launch(cudaStream_t stream, Workspace* ws, int B, const int* pos_d, int n_heads, int n_kv_heads,
                      const float* q, float* out, const float* kc, const float* vc, long long kv_batch_stride,
                      int head_size, int seq_len, bool pdl, float* out_hi, float* out_lo, int8_t* out_q = nullptr,
                      float* out_sT = nullptr, int ldxs = 0, int gs = 0, const KvPaged* pg = nullptr);
bool attn_prefill_supported(int head_size, int kv_dim);
int attn_prefill_launch(cudaStream_t stream, int n_tiles, int n_heads, int n_kv_heads, int head_size, const float* q,
                        float* out, float* out_hi, float* out_lo, const float* kc, const float* vc,
                        long long kv_batch_stride, const int* row_slot, const int* row_pos, const int* tile_row0,
                        const int* tile_n, bool pdl, const KvPaged* pg = nullptr);
bool attn_fused_rope_supported(int n_heads, int n_kv_heads, int head_size);
int attn_fused_rope_launch(cudaStream_t stream, Workspace* ws, int B, const int* pos_d, int n_heads, int n_kv_heads, float* q_out,
                           float* out, float* kc, float* vc, long long kv_batch_stride, int head_size, int seq_len, bool pdl,
                           float* out_hi, float* out_lo, int8_t* out_q, float* out_sT, int ldxs, int gs, const KvPaged* pg,
                           const float* part, int nz, long long zstride, int ld, const float* rc, const float* rs,
                           uint16_t* out_b16 = nullptr);
bool attn_prefill_tc_supported(int head_size, int kv_dim);
int attn_prefill_tc_launch(cudaStream_t stream, int n_tiles, int n_heads, int n_kv_heads, int head_size, const float* q, float* out,
                           uint16_t* out_b16, const float* kc, const float* vc, long long kv_batch_stride, const int* row_slot,
                           const int* row_pos, const int* tile_row0, const int* tile_n, bool pdl, const KvPaged* pg,
                           const uint16_t* k16, const uint16_t* vt16, int staged_rows, int staged_ld, const int* row_stage);
int rope_scatter_launch(cudaStream_t st, bool pdl, int R, const float* part, int nz, long long zstride, int ld, const int* pos, int d,
                        int kvd, int hs, const float* rc, const float* rs, float* q, float* kc, float* vc, long long kv_bs, int loff,
                        const int* slot, const KvPaged* pg = nullptr);

__device__ __forceinline__ float bf16_to_f32(uint16_t v) { return __uint_as_float((uint32_t)v << 16); }
// round to nearest even onto bf16 (finite inputs; NaN stays NaN)
__device__ __forceinline__ uint16_t f32_to_bf16(float f) {
  const uint32_t u = __float_as_uint(f);
  if ((u & 0x7FFFFFFFu) > 0x7F800000u) return (uint16_t)((u >> 16) | 0x40u);
  return (uint16_t)((u + 0x7FFFu + ((u >> 16) & 1u)) >> 16);
}
__device__ __forceinline__ uint2 pack4_bf16(float4 v) {
  uint2 r;
  r.x = (uint32_t)f32_to_bf16(v.x) | ((uint32_t)f32_to_bf16(v.y) << 16);
  r.y = (uint32_t)f32_to_bf16(v.z) | ((uint32_t)f32_to_bf16(v.w) << 16);
  return r;
}

constexpr int BPREP_MAX_V = 4;   // float4 per thread kept in registers: d <= 4 * 1024 * 4
// One block per output row: x row = embedding row (bf16 table) or x + sum of nz partials (written back), RMSNorm
// g * (ss * x) (run.cc:196-209 operation order), rounded to bf16.  With `gather`, output row r comes from source row
// gather[r] and x is left untouched (classifier input of a prefill chunk).
__global__ void __launch_bounds__(1024) bf_prep_kernel(float* __restrict__ x, const float* __restrict__ part, int nz,
                                                       long long zstride, const uint16_t* __restrict__ emb,
                                                       const int* __restrict__ token, const int* __restrict__ pos,
                                                       const float* __restrict__ gain, int d, uint16_t* __restrict__ xb,
                                                       const int* __restrict__ gather) {
  __shared__ float red[33];
  pdl_wait();
  pdl_trigger();
  const int ob = blockIdx.x;
  const int b = gather ? gather[ob] : ob;
  const int d4 = d >> 2;
  float* xr = x + (size_t)b * d;
  uint2* orow = reinterpret_cast<uint2*>(xb + (size_t)ob * d);
  if (pos[b] < 0) {   // idle slot: zeros into the matmul
    for (int i = threadIdx.x; i < d4; i += blockDim.x) orow[i] = make_uint2(0u, 0u);
    return;
  }
  const uint16_t* er = emb ? emb + (size_t)token[b] * d : nullptr;
  float4 v[BPREP_MAX_V];
  float ss = 0.0f;
#pragma unroll
  for (int k = 0; k < BPREP_MAX_V; k++) {
    const int i = threadIdx.x + k * blockDim.x;
    v[k] = make_float4(0.f, 0.f, 0.f, 0.f);
    if (i < d4) {
      if (er) {
        const uint2 e = reinterpret_cast<const uint2*>(er)[i];
        v[k] = make_float4(bf16_to_f32((uint16_t)(e.x & 0xFFFFu)), bf16_to_f32((uint16_t)(e.x >> 16)),
                           bf16_to_f32((uint16_t)(e.y & 0xFFFFu)), bf16_to_f32((uint16_t)(e.y >> 16)));
      } else {
        v[k] = reinterpret_cast<const float4*>(xr)[i];
#pragma unroll 8
        for (int z = 0; z < nz; z++) {   // fixed order
          const float4 pz = *reinterpret_cast<const float4*>(part + (size_t)z * zstride + (size_t)b * d + 4 * i);
          v[k].x += pz.x; v[k].y += pz.y; v[k].z += pz.z; v[k].w += pz.w;
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
  for (int k = 0; k < BPREP_MAX_V; k++) {
    const int i = threadIdx.x + k * blockDim.x;
    if (i < d4) {
      const float4 g4 = reinterpret_cast<const float4*>(gain)[i];
      orow[i] = pack4_bf16(make_float4(g4.x * (ssn * v[k].x), g4.y * (ssn * v[k].y), g4.z * (ssn * v[k].z), g4.w * (ssn * v[k].w)));
    }
  }
}

// One float4 per thread: src row (attention output), or with h > 0 the SwiGLU of the w1|w3 partials (run.cc:344-347 order),
// rounded to bf16.  grid (ceil(K / 4 / 256), rows)
__global__ void __launch_bounds__(256) bf_round_kernel(const float* __restrict__ src, long long ld, int nz, long long zstride, int h,
                                                       const int* __restrict__ pos, int K, uint16_t* __restrict__ out) {
  pdl_wait();
  pdl_trigger();
  const int b = blockIdx.y;
  const bool idle = pos[b] < 0;
  // two float4 per thread, a block-width apart: both (pairs of) loads are in flight before the first exponential
#pragma unroll
  for (int u = 0; u < 2; u++) {
    const int i4 = (blockIdx.x * 2 + u) * blockDim.x + threadIdx.x;
    if (4 * i4 >= K) continue;
    float4 r = make_float4(0.f, 0.f, 0.f, 0.f);
    if (!idle) {
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
    reinterpret_cast<uint2*>(out + (size_t)b * K)[i4] = pack4_bf16(r);
  }
}

// x[b] = embedding row of token[b] (bf16 table -> fp32 residual stream); idle slots untouched.  grid (B), the GEMV step's first kernel
__global__ void __launch_bounds__(256) bf_embed_kernel(float* __restrict__ x, const uint16_t* __restrict__ table,
                                                       const int* __restrict__ token, const int* __restrict__ pos, int d) {
  pdl_wait();
  pdl_trigger();
  const int b = blockIdx.x;
  if (pos[b] < 0) return;
  const uint2* er = reinterpret_cast<const uint2*>(table + (size_t)token[b] * d);
  float4* xr = reinterpret_cast<float4*>(x + (size_t)b * d);
  for (int i = threadIdx.x; i < (d >> 2); i += blockDim.x) {
    const uint2 e = er[i];
    xr[i] = make_float4(bf16_to_f32((uint16_t)(e.x & 0xFFFFu)), bf16_to_f32((uint16_t)(e.x >> 16)),
                        bf16_to_f32((uint16_t)(e.y & 0xFFFFu)), bf16_to_f32((uint16_t)(e.y >> 16)));
  }
}
int embed_bf16_launch(cudaStream_t st, int B, float* x, const uint16_t* table, const int* token, const int* pos, int d, bool pdl) {
  THA_CUDA_OK(launch(bf_embed_kernel, dim3(B), dim3(256), 0, st, pdl, x, table, token, pos, d));
  return THA_STATUS_SUCCESS;
}

__global__ void bf_convert_kernel(const float* __restrict__ src, uint16_t* __restrict__ dst, size_t n) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) dst[i] = f32_to_bf16(src[i]);
}

// K rows and V rows of the prompt positions just written to the cache -> bf16 (nearest even): k16[srow[r]][c] row-major and
// vt16[c][srow[r]] TRANSPOSED, r = row of the call, srow[r] = its staging row (a sequence starts at a multiple of 64), c = kv column.  32 x 32 tiles through shared memory: both sides coalesced.
// grid (kvd / 32, R / 32), 256 threads (32 x 8)
__global__ void __launch_bounds__(256) bf_kv_stage_kernel(const float* __restrict__ kc, const float* __restrict__ vc, long long kv_bs,
                                                          const int* __restrict__ slot, const int* __restrict__ pos, const KvPaged pg,
                                                          const int* __restrict__ srow, int R, int kvd, int ld,
                                                          uint16_t* __restrict__ k16, uint16_t* __restrict__ vt16) {
  __shared__ uint16_t t[32][34];
  pdl_wait();
  pdl_trigger();
  const int c0 = blockIdx.x * 32, r0 = blockIdx.y * 32;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
#pragma unroll
  for (int i = 0; i < 32; i += 8) {
    const int r = r0 + ty + i, c = c0 + tx;
    uint16_t kb = 0, vb = 0;
    if (r < R && c < kvd && pos[r] >= 0) {
      const long long ro = pg.table ? kv_paged_off(pg, slot[r], pos[r], kvd) : (long long)slot[r] * kv_bs + (long long)pos[r] * kvd;
      kb = f32_to_bf16(kc[ro + c]);
      vb = f32_to_bf16(vc[ro + c]);
    }
    if (r < R && c < kvd) k16[(size_t)srow[r] * kvd + c] = kb;
    t[ty + i][tx] = vb;
  }
  __syncthreads();
#pragma unroll
  for (int i = 0; i < 32; i += 8) {
    const int c = c0 + ty + i, r = r0 + tx;
    if (c < kvd && r < R) vt16[(size_t)c * ld + srow[r]] = t[tx][ty + i];
  }
}

// ---- scratch (per workspace) ---------------------------------------------------------------------------------------
struct BfScratch {
  uint16_t *xb = nullptr, *hb = nullptr;             // matmul inputs [R, d], [R, hidden]
  float* part = nullptr;                             // split-K partials
  float *xr = nullptr, *qr = nullptr, *xbr = nullptr; int* tab = nullptr;   // prefill rows
  uint16_t *k16 = nullptr, *vt16 = nullptr;          // prompt attention on the tensor cores: this layer's K rows [R, kvd] and V rows TRANSPOSED [kvd, ld] in bf16
  size_t n_xb = 0, n_hb = 0, n_part = 0, n_xr = 0, n_qr = 0, n_xbr = 0, n_tab = 0, n_k16 = 0, n_vt16 = 0;
};
static std::map<Workspace*, BfScratch> g_bf_scratch;
static std::mutex g_bf_mu;

template <typename T>
static int bf_grow(T** p, size_t* cap, size_t need, Workspace* ws) {
  if (need <= *cap) return THA_STATUS_SUCCESS;
  if (*p) THA_CUDA_OK(cudaFree(*p));
  THA_CUDA_OK(cudaMalloc(p, need * sizeof(T)));
  *cap = need;
  ws_drop_graphs(ws);
  gemm_tc_forget();
  return THA_STATUS_SUCCESS;
}

// Split-K factor of one matmul: whole waves of the 148 persistent CTAs, at least two 64-value k-blocks per split, a
// charge per split for the partials its consumer sums (as step_tc.cu's planner).  Prompt-sized row counts (>= 512: 256-row
// work items, gemm_bf16.cu bf_mt) are compute-bound, and there a split is expensive: it writes and re-reads R x M fp32
// partials, 4 F / (K BW) of the matmul's own time (F ~ 1.3 PFLOP/s, BW ~ 6.5 TB/s: 0.2 at K = 4096) -- at 2048 rows the old
// charge picked three splits for wq|wk|wv and the RoPE kernel read 303 MB per layer.
static int bf_splits(int tiles, int K, int R) {
  const int kb = ceil_div(K, 64);
  const bool big = R >= 512;
  const int items1 = (big ? ceil_div(tiles, 2) : tiles) * ceil_div(R, gemm_bf16_pick_nt(R));
  int best = 1;
  float best_score = -1.0f;
  const float charge = big ? 800.0f / (float)K : 0.012f + 0.25f * (float)R / (float)K / 16.0f;
  for (int z = 1; z <= 16; z++) {
    if (z > 1 && kb / z < 2) break;
    if (ceil_div(kb, ceil_div(kb, z)) != z) continue;
    const int ctas = items1 * z;
    const float eff = (float)ctas / (float)(ceil_div(ctas, 148) * 148);
    const float score = eff - charge * (z - 1);
    if (score > best_score) { best_score = score; best = z; }
  }
  return best;
}

struct BfPlan { int z_qkv, z_wo, z_w13, z_w2; size_t part_floats; };
static BfPlan bf_plan(const tha_config_t& c, int R) {
  const int d = c.dim, hd = c.hidden_dim, hs = d / c.n_heads, kvd = hs * c.n_kv_heads;
  BfPlan p;
  p.z_qkv = bf_splits(ceil_div(d, 128) + 2 * ceil_div(kvd, 128), d, R);
  p.z_wo = bf_splits(ceil_div(d, 128), d, R);
  p.z_w13 = bf_splits(2 * ceil_div(hd, 128), d, R);
  p.z_w2 = bf_splits(ceil_div(d, 128), hd, R);
  size_t m = (size_t)p.z_qkv * (d + 2 * kvd);
  if ((size_t)p.z_wo * d > m) m = (size_t)p.z_wo * d;
  if ((size_t)p.z_w13 * 2 * hd > m) m = (size_t)p.z_w13 * 2 * hd;
  if ((size_t)p.z_w2 * d > m) m = (size_t)p.z_w2 * d;
  p.part_floats = m * R;
  return p;
}

bool step_bf16_supported(const tha_config_t* c) {
  if (!c || c->dim <= 0 || c->n_heads <= 0 || c->n_kv_heads <= 0 || c->dim % c->n_heads || c->n_heads % c->n_kv_heads) return false;
  const int d = c->dim, hd = c->hidden_dim, hs = d / c->n_heads, kvd = hs * c->n_kv_heads;
  if (d % 8 || hd % 8 || kvd % 4 || (hs & 1) || d < 64 || hd < 64 || d > 4 * 1024 * BPREP_MAX_V) return false;
  return attn_fused_supported(c->n_heads, hs, d, kvd);
}

static int bf_prepare(Workspace* ws, const tha_config_t& c, int R, bool prefill, int n_tab, int Rs = 0) {
  THA_TRY_RC(gemm_tc_init(nullptr));
  std::lock_guard<std::mutex> lk(g_bf_mu);
  BfScratch& s = g_bf_scratch[ws];
  const BfPlan pl = bf_plan(c, R);
  THA_TRY_RC(bf_grow(&s.xb, &s.n_xb, (size_t)R * c.dim, ws));
  THA_TRY_RC(bf_grow(&s.hb, &s.n_hb, (size_t)R * c.hidden_dim, ws));
  THA_TRY_RC(bf_grow(&s.part, &s.n_part, pl.part_floats, ws));
  if (prefill) {
    THA_TRY_RC(bf_grow(&s.xr, &s.n_xr, (size_t)R * c.dim, ws));
    THA_TRY_RC(bf_grow(&s.qr, &s.n_qr, (size_t)R * c.dim, ws));
    THA_TRY_RC(bf_grow(&s.xbr, &s.n_xbr, (size_t)R * c.dim, ws));
    THA_TRY_RC(bf_grow(&s.tab, &s.n_tab, (size_t)n_tab, ws));
    const int kvd_ = c.dim / c.n_heads * c.n_kv_heads;
    const size_t before = s.n_k16 + s.n_vt16;
    THA_TRY_RC(bf_grow(&s.k16, &s.n_k16, (size_t)Rs * kvd_, ws));
    THA_TRY_RC(bf_grow(&s.vt16, &s.n_vt16, (size_t)kvd_ * Rs, ws));
    if (s.n_k16 + s.n_vt16 != before) {   // the padding columns of V^T are read (and masked): they must hold finite values
      THA_CUDA_OK(cudaMemset(s.k16, 0, s.n_k16 * sizeof(uint16_t)));
      THA_CUDA_OK(cudaMemset(s.vt16, 0, s.n_vt16 * sizeof(uint16_t)));
    }
    return THA_STATUS_SUCCESS;
  }
  const int hs = c.dim / c.n_heads;
  return attn_fused_prepare(ws, R, c.n_heads, c.n_kv_heads, hs, c.seq_len);
}
int step_bf16_prepare(Workspace* ws, const tha_config_t* c, int B) { return bf_prepare(ws, *c, B, false, 0); }

struct BfRows {
  int R;
  float *x, *q, *xb;
  const int *token, *pos, *slot;
  int n_tiles; const int *tile_row0, *tile_n; int tile_rows;   // prefill attention tiles (rows per tile: 64 CUDA cores, 128 tensor cores)
  bool staged;                                                 // every sequence starts at position 0: K / V^T staged in bf16 for the tensor-core attention
  const int* srow; int Rs;                                     // staging row of every row, staging rows in all
  int n_out; const int* out_rows;
  float* logits; int* next;
};

static int bf_attn_tc = -1;   // THA_ATTN_TC=0: keep the CUDA-core prefill attention (development aid)
static int bf_swiglu_mode() {   // THA_BF16_SWIGLU=0: SwiGLU stays its own kernel behind the w1|w3 matmul
  static int v = -1;
  if (v < 0) { const char* e = getenv("THA_BF16_SWIGLU"); v = (e && *e) ? atoi(e) : 1; }
  return v;
}
static int bf_stage_mode() {   // THA_ATTN_STAGE=0: the tensor-core attention converts the fp32 cache rows itself (the first form)
  static int v = -1;
  if (v < 0) { const char* e = getenv("THA_ATTN_STAGE"); v = (e && *e) ? atoi(e) : 1; }
  return v;
}

static int bf_issue_rows(cudaStream_t st, Workspace* ws, const tha_config_t& c, const tha_weights_bf16_t& w, const tha_state_t& s,
                         const BfRows& io, const float* rc, const float* rs, bool pdl) {
  BfScratch sc;
  {
    std::lock_guard<std::mutex> lk(g_bf_mu);
    sc = g_bf_scratch[ws];
  }
  const int R = io.R, d = c.dim, hd = c.hidden_dim, hs = d / c.n_heads, kvd = hs * c.n_kv_heads, L = c.n_layers;
  const BfPlan pl = bf_plan(c, R);
  const long long kv_bs = (long long)L * c.seq_len * kvd;
  const int ld_qkv = d + 2 * kvd, ld_h = 2 * hd;
  int prep_threads = ceil_div(d / 4, 32) * 32;
  if (prep_threads > 1024) prep_threads = 1024;
  if (prep_threads < 128) prep_threads = 128;
  // many rows (prompt ingestion): 256 threads with up to four float4 each -- eight CTAs per SM and four independent loads per
  // thread hide the L2 latency that a 1024-thread CTA per row (lowest latency for a handful of decode rows) leaves exposed
  if (R >= 512 && d / 4 <= 256 * BPREP_MAX_V && prep_threads > 256) prep_threads = 256;
  const int* nogather = nullptr;
  auto prep = [&](int rows, int nz, long long zs, const uint16_t* emb, const float* gain, const int* gather) -> int {
    THA_CUDA_OK(launch(bf_prep_kernel, dim3(rows), dim3(prep_threads), 0, st, pdl, io.x, (const float*)sc.part, nz, zs, emb, io.token,
                       io.pos, gain, d, sc.xb, gather));
    return THA_STATUS_SUCCESS;
  };
  auto mm = [&](int n, const uint16_t* const* W, const int* M, const int* off, int K, const uint16_t* X, int N, float* out,
                long long ldo, long long zs, int z) -> int {
    BfArgs a = {};
    a.n_mats = n;
    for (int i = 0; i < n; i++) { a.W[i] = W[i]; a.M[i] = M[i]; a.out_off[i] = off[i]; }
    a.K = K; a.X = X; a.N = N; a.ldx = K; a.out = out; a.ldo = ldo; a.zstride = zs; a.splits = z; a.colscale = nullptr;
    return gemm_bf16_launch(st, a, pdl);
  };
  int prev_z = 0;
  long long prev_zs = 0;
  KvPool pool;
  const bool paged = kv_pool_find(s.key_cache, &pool);
  if (paged && s.value_cache != pool.v) return THA_STATUS_INVALID_ARG;
  for (int l = 0; l < L; l++) {
    const int loff = l * c.seq_len * kvd;
    const KvPaged pg = paged ? kv_paged_of(pool, L, kvd, l) : KvPaged{};
    const float* kcl = paged ? s.key_cache : s.key_cache + loff;
    const float* vcl = paged ? s.value_cache : s.value_cache + loff;
    THA_TRY_RC(prep(R, prev_z, prev_zs, l == 0 ? w.token_embedding_table : (const uint16_t*)nullptr, w.rms_att_weight + (size_t)l * d, nogather));
    {
      const uint16_t* W[3] = {w.wq + (size_t)l * d * d, w.wk + (size_t)l * d * kvd, w.wv + (size_t)l * d * kvd};
      const int M[3] = {d, kvd, kvd}, off[3] = {0, d, d + kvd};
      THA_TRY_RC(mm(3, W, M, off, d, sc.xb, R, sc.part, ld_qkv, (long long)R * ld_qkv, pl.z_qkv));
    }
    // decode rows: RoPE + KV scatter and the bf16 rounding of the output ride in the attention kernel (7 launches per layer, not 10)
    const bool rope_in_attn = io.n_tiles == 0 && !(!io.logits && l == L - 1) && rope_fuse_mode() &&
                              attn_fused_rope_supported(c.n_heads, c.n_kv_heads, hs);
    if (!rope_in_attn)
    THA_TRY_RC(rope_scatter_launch(st, pdl, R, sc.part, pl.z_qkv, (long long)R * ld_qkv, ld_qkv, io.pos, d, kvd, hs, rc, rs, io.q,
                                   s.key_cache, s.value_cache, kv_bs, loff, io.slot, &pg));
    if (!io.logits && l == L - 1) return THA_STATUS_SUCCESS;   // cache-only ingestion: the last layer's K/V rows are in place
    bool rounded = false;   // the attention kernel already left its output as bf16 rows in sc.xb
    if (io.n_tiles > 0 && io.tile_rows == 128) {
      // prompt rows, head size 64 / 128: Q K^T and P V on the tensor cores (attn_prefill_tc.cu)
      if (io.staged) {   // this layer's K rows and V rows (transposed) in bf16: the attention tiles load them with TMA
        THA_CUDA_OK(launch(bf_kv_stage_kernel, dim3(ceil_div(kvd, 32), ceil_div(R, 32)), dim3(256), 0, st, pdl, kcl, vcl, kv_bs, io.slot,
                           io.pos, pg, io.srow, R, kvd, io.Rs, sc.k16, sc.vt16));
      }
      THA_TRY_RC(attn_prefill_tc_launch(st, io.n_tiles, c.n_heads, c.n_kv_heads, hs, io.q, io.xb, sc.xb, kcl, vcl, kv_bs, io.slot, io.pos,
                                        io.tile_row0, io.tile_n, pdl, &pg, io.staged ? sc.k16 : nullptr, io.staged ? sc.vt16 : nullptr,
                                        io.Rs, io.Rs, io.srow));
      rounded = true;
    } else if (io.n_tiles > 0)
      THA_TRY_RC(attn_prefill_launch(st, io.n_tiles, c.n_heads, c.n_kv_heads, hs, io.q, io.xb, nullptr, nullptr, kcl,
                                     vcl, kv_bs, io.slot, io.pos, io.tile_row0, io.tile_n, pdl, &pg));
    else if (rope_in_attn) {
      THA_TRY_RC(attn_fused_rope_launch(st, ws, R, io.pos, c.n_heads, c.n_kv_heads, io.q, io.xb, const_cast<float*>(kcl),
                                        const_cast<float*>(vcl), kv_bs, hs, c.seq_len, pdl, nullptr, nullptr, nullptr, nullptr, 0, 0, &pg,
                                        sc.part, pl.z_qkv, (long long)R * ld_qkv, ld_qkv, rc, rs, sc.xb));
      rounded = true;
    } else
      THA_TRY_RC(attn_fused_launch(st, ws, R, io.pos, c.n_heads, c.n_kv_heads, io.q, io.xb, kcl, vcl,
                                   kv_bs, hs, c.seq_len, pdl, nullptr, nullptr, nullptr, nullptr, 0, 0, &pg));
    if (!rounded)
      THA_CUDA_OK(launch(bf_round_kernel, dim3(ceil_div(d / 4, 512), R), dim3(256), 0, st, pdl, (const float*)io.xb, (long long)d, 0, 0LL, 0,
                         io.pos, d, sc.xb));
    {
      const uint16_t* W[1] = {w.wo + (size_t)l * d * d};
      const int M[1] = {d}, off[1] = {0};
      THA_TRY_RC(mm(1, W, M, off, d, sc.xb, R, sc.part, d, (long long)R * d, pl.z_wo));
    }
    THA_TRY_RC(prep(R, pl.z_wo, (long long)R * d, nullptr, w.rms_ffn_weight + (size_t)l * d, nogather));
    bool swiglu_fused = false;
    {
      const uint16_t* W[2] = {w.w1 + (size_t)l * d * hd, w.w3 + (size_t)l * d * hd};
      const int M[2] = {hd, hd}, off[2] = {0, hd};
      // prompt-sized row counts without split-K: SwiGLU in the matmul's read-out -- the fp32 w1|w3 outputs (180 MB per 7B layer at
      // 2048 rows) are never written, the bf16 input of w2 comes straight out of TMEM
      BfArgs a = {};
      a.n_mats = 2; a.W[0] = W[0]; a.W[1] = W[1]; a.M[0] = hd; a.M[1] = hd; a.out_off[0] = 0; a.out_off[1] = hd;
      a.K = d; a.X = sc.xb; a.N = R; a.ldx = d; a.out = sc.part; a.ldo = ld_h; a.zstride = (long long)R * ld_h; a.splits = pl.z_w13;
      a.swiglu_out = sc.hb; a.ld_swiglu = hd;
      swiglu_fused = bf_swiglu_mode() && gemm_bf16_swiglu_ok(a);
      if (swiglu_fused) THA_TRY_RC(gemm_bf16_launch(st, a, pdl));
      else THA_TRY_RC(mm(2, W, M, off, d, sc.xb, R, sc.part, ld_h, (long long)R * ld_h, pl.z_w13));
    }
    if (!swiglu_fused)
    THA_CUDA_OK(launch(bf_round_kernel, dim3(ceil_div(hd / 4, 512), R), dim3(256), 0, st, pdl, (const float*)sc.part, (long long)ld_h, pl.z_w13,
                       (long long)R * ld_h, hd, io.pos, hd, sc.hb));
    {
      const uint16_t* W[1] = {w.w2 + (size_t)l * d * hd};
      const int M[1] = {d}, off[1] = {0};
      THA_TRY_RC(mm(1, W, M, off, hd, sc.hb, R, sc.part, d, (long long)R * d, pl.z_w2));
    }
    prev_z = pl.z_w2;
    prev_zs = (long long)R * d;
  }
  THA_TRY_RC(prep(io.n_out, prev_z, prev_zs, nullptr, w.rms_final_weight, io.out_rows));
  {
    const uint16_t* W[1] = {w.wcls};
    const int M[1] = {c.vocab_size}, off[1] = {0};
    THA_TRY_RC(mm(1, W, M, off, d, sc.xb, io.n_out, io.logits, c.vocab_size, 0, 1));
  }
  if (io.next) THA_TRY_RC(argmax_launch(st, io.n_out, io.logits, c.vocab_size, io.next, pdl));
  return THA_STATUS_SUCCESS;
}

int step_bf16_issue(cudaStream_t st, Workspace* ws, int B, const tha_config_t& c, const tha_weights_bf16_t& w, const tha_state_t& s,
                    const int* token_d, const int* pos_d, float* logits, int* next_d, const float* rc, const float* rs, bool pdl) {
  BfRows io = {};
  io.R = B; io.x = s.x; io.q = s.q; io.xb = s.xb; io.token = token_d; io.pos = pos_d; io.slot = nullptr;
  io.n_tiles = 0; io.n_out = B; io.out_rows = nullptr; io.logits = logits; io.next = next_d;
  return bf_issue_rows(st, ws, c, w, s, io, rc, rs, pdl);
}

int step_bf16_prefill(cudaStream_t st, Workspace* ws, const tha_config_t& c, const tha_weights_bf16_t& w, const tha_state_t& s,
                      int n_seqs, const int* slot, const int* pos0, const int* len, const int* tokens, float* logits, int* next_d,
                      const float* rc, const float* rs) {
  const int hs = c.dim / c.n_heads;
  if (!step_bf16_supported(&c) || !attn_prefill_supported(hs, hs * c.n_kv_heads)) return THA_STATUS_NOT_SUPPORTED;
  if (bf_attn_tc < 0) { const char* v = getenv("THA_ATTN_TC"); bf_attn_tc = (v && *v) ? atoi(v) : 1; }
  const int TQ = (bf_attn_tc && attn_prefill_tc_supported(hs, hs * c.n_kv_heads)) ? 128 : 64;   // query rows per attention tile
  long long R = 0, T = 0;
  for (int i = 0; i < n_seqs; i++) {
    if (len[i] < 1 || pos0[i] < 0 || slot[i] < 0 || (long long)pos0[i] + len[i] > c.seq_len) return THA_STATUS_INVALID_ARG;
    R += len[i];
    T += ceil_div(len[i], TQ);
  }
  if (R < 1 || R > (1 << 20)) return THA_STATUS_INVALID_ARG;
  std::vector<int> tab((size_t)3 * R + 2 * T + n_seqs + R);
  int *h_slot = tab.data(), *h_pos = h_slot + R, *h_tok = h_pos + R, *h_t0 = h_tok + R, *h_tn = h_t0 + T, *h_last = h_tn + T;
  int* h_srow = h_last + n_seqs;   // staging row of every row: sequences start at multiples of 64
  int r = 0, t = 0, Rs = 0;
  for (int i = 0; i < n_seqs; i++) {
    for (int j = 0; j < len[i]; j++) h_srow[r + j] = Rs + j;
    Rs += ceil_div(len[i], 64) * 64;
    for (int j = 0; j < len[i]; j += TQ) { h_t0[t] = r + j; h_tn[t] = (len[i] - j < TQ) ? len[i] - j : TQ; t++; }
    for (int j = 0; j < len[i]; j++, r++) {
      h_slot[r] = slot[i]; h_pos[r] = pos0[i] + j; h_tok[r] = tokens[r];
      if (tokens[r] < 0 || tokens[r] >= c.vocab_size) return THA_STATUS_INVALID_ARG;
    }
    h_last[i] = r - 1;
  }
  THA_TRY_RC(bf_prepare(ws, c, (int)R, true, (int)tab.size(), Rs));
  BfScratch sc;
  {
    std::lock_guard<std::mutex> lk(g_bf_mu);
    sc = g_bf_scratch[ws];
  }
  THA_CUDA_OK(cudaMemcpyAsync(sc.tab, tab.data(), tab.size() * sizeof(int), cudaMemcpyHostToDevice, st));
  BfRows io = {};
  io.R = (int)R; io.x = sc.xr; io.q = sc.qr; io.xb = sc.xbr;
  io.slot = sc.tab; io.pos = sc.tab + R; io.token = sc.tab + 2 * R;
  io.n_tiles = (int)T; io.tile_row0 = sc.tab + 3 * R; io.tile_n = sc.tab + 3 * R + T; io.tile_rows = TQ;
  io.staged = TQ == 128 && bf_attn_tc >= 1 && bf_stage_mode();
  for (int i = 0; i < n_seqs; i++) io.staged = io.staged && pos0[i] == 0;   // staged keys are addressed by prompt row
  io.srow = sc.tab + 3 * R + 2 * T + n_seqs; io.Rs = Rs;
  io.n_out = n_seqs; io.out_rows = sc.tab + 3 * R + 2 * T; io.logits = logits; io.next = logits ? next_d : nullptr;
  return bf_issue_rows(st, ws, c, w, s, io, rc, rs, pdl_enabled() != 0);
}

int bf16_convert_launch(cudaStream_t st, const float* src, uint16_t* dst, size_t n) {
  int blocks = (int)((n + 255) / 256);
  if (blocks > 148 * 16) blocks = 148 * 16;
  if (blocks < 1) blocks = 1;
  bf_convert_kernel<<<blocks, 256, 0, st>>>(src, dst, n);
  count_launch();
  THA_LAUNCH_OK();
  return THA_STATUS_SUCCESS;
}

}  // namespace tha
