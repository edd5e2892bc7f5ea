// The whole decode step (thaDNN_s_forward_batch, src/thaDNN.cpp:13-81) as ONE persistent
// kernel: one CTA per SM, every layer, grid-wide barriers between the dependent phases.
//
// Why: stories110M moves 438 MB per step = 67 us at HBM speed, i.e. ~1 us per matmul.  A kernel
// launch per matmul (even graph-replayed, even with programmatic dependent launch) costs several
// us of launch + ramp + drain, so the multi-kernel step sat at 9-16 % of the roofline
// (profiles/r01a_*).  Here the weights never wait for anything:
//
//   * warp 16 of every CTA is a PRODUCER: it walks the CTA's static list of weight row-pairs for
//     the whole step (its share of wq|wk|wv, wo, w1&w3, w2 of every layer, then of wcls) and
//     streams them HBM -> shared memory with 1-D bulk async copies (cp.async.bulk, the TMA
//     engine; SASS UBLKCP) into a ring of 10 x 16 KB slots, gated only by slot reuse
//     (mbarrier full/empty pairs) -- never by a grid barrier.  ~24 MB are in flight chip-wide.
//   * warps 0-15 are CONSUMERS: per phase they wait for the grid barrier (activations of the
//     previous phase are complete), stage the B activation rows in shared memory (RMSNorm
//     folded in), take the chunks the producer has already landed, do the fp32 FMAs out of
//     shared memory, run the fused epilogue (RoPE + KV-cache write / residual add / SwiGLU /
//     logits + running argmax) and arrive at the next barrier.
//   * attention is a phase of the same kernel: (slot, head, time-split) items spread over all
//     2368 warps, online softmax per warp, the last warp of a (slot, head) merges the splits.
//
// Phases per layer: QKV | ATTN | WO(+residual) | W1&W3(+SwiGLU) | W2(+residual); then CLS and a
// tiny ARGMAX/advance tail.  5 barriers per layer.
//
// Limits of this kernel (anything else takes the multi-kernel path in forward.cu):
//   B <= 8, dim <= 1024, hidden_dim <= 2048, B_tile * max(dim, hidden) <= 16384 floats,
//   head_size in {32, 64, 128}, even row counts.  Covers stories15M/42M/110M; Llama-2-7B
//   matmuls are 100x larger and launch-latency is noise there.
#include <stdlib.h>
#include <string.h>

#include <map>

#include "common.cuh"
#include "workspace.h"

namespace tha {

constexpr int MG_NW = 16;                      // consumer warps
constexpr int MG_CT = MG_NW * 32;              // consumer threads
constexpr int MG_THREADS = MG_CT + 32;         // + producer warp
constexpr int MG_NSLOT = MG_NW;                // one ring slot per consumer warp (see the ring notes below)
constexpr int MG_SMEM_BUDGET = 227 * 1024;     // opt-in maximum of a CTA on sm_100
constexpr int MG_XS_MAX_FLOATS = 16384;        // staged activations: at most 64 KB
constexpr int MG_MAX_B = 8;
constexpr unsigned long long MG_WAIT_NS = 100ull * 1000 * 1000;   // a wait longer than this (a step takes < 1 ms) raises the abort flag

struct MegaP {
  tha_config_t cfg;
  int B, advance;
  const float *tok_emb, *rms_att, *rms_ffn, *rms_final, *wq, *wk, *wv, *wo, *w1, *w2, *w3, *wcls;
  float *x, *q, *xb, *hb, *kc, *vc, *logits;
  int* next; int* token; int* pos;
  const float *rope_cos, *rope_sin;
  unsigned* bar;            // [0] arrivals, [1] exits, [2] abort flag
  float* att_acc; float2* att_ml; int* att_cnt; int smax;
  float* amax_val; int* amax_idx;   // [grid][MG_MAX_B]
  int slot_bytes, xs_floats;        // shared-memory carve-up chosen by the host (ring = 16 slots)
  unsigned long long* trace;        // nullable: [grid][phases][4] globaltimer stamps (THA_MEGA_TRACE=1)
};

// ------------------------------------------------------------------ PTX helpers ----
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, unsigned count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, unsigned bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, unsigned parity) {
  unsigned ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
  return ok != 0;
}
// 1-D bulk copy global -> shared through the TMA engine, completion on an mbarrier; the L2
// policy marks the weights evict-first (read once per step) so KV rows and activations stay.
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, unsigned bytes, uint64_t* bar, uint64_t policy) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;"
      ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)), "l"(policy) : "memory");
}
__device__ __forceinline__ uint64_t policy_evict_first() {
  uint64_t p;
  asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p));
  return p;
}
__device__ __forceinline__ void bar_consumers() { asm volatile("bar.sync 1, %0;" ::"n"(MG_CT) : "memory"); }
__device__ __forceinline__ unsigned ld_acquire(const unsigned* p) {
  unsigned v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void red_release_add(unsigned* p, unsigned v) {
  asm volatile("red.release.gpu.global.add.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

struct MegaShared {
  uint64_t full[MG_NSLOT];
  uint64_t empty[MG_NSLOT];
  float red[MG_NW * MG_MAX_B];
  float scale[MG_MAX_B];
  int pos[MG_MAX_B];
  int tok[MG_MAX_B];
  float amax_v[MG_NW * MG_MAX_B];
  int amax_i[MG_NW * MG_MAX_B];
  float rope_c[MG_MAX_B][64];   // cos/sin row of pos[b] (head_size/2 <= 64 entries)
  float rope_s[MG_MAX_B][64];
  int aborted;
};

__device__ __forceinline__ unsigned long long gtime_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}
// First waiter to time out records where: bar[4] = code, bar[5] = blockIdx, bar[6] = warp, bar[7] = detail.
__device__ __noinline__ void raise_abort(unsigned* gbar, MegaShared* sh, unsigned code, unsigned detail) {
  if (atomicExch(&gbar[2], 1u) == 0u) {
    gbar[4] = code; gbar[5] = blockIdx.x; gbar[6] = threadIdx.x >> 5; gbar[7] = detail;
  }
  sh->aborted = 1;
}
// Wait on an mbarrier, bounded in time; on timeout raise the global abort flag so that every
// other wait in the grid falls through too and the kernel ends instead of hanging the GPU.
__device__ __forceinline__ void mbar_wait(uint64_t* bar, unsigned parity, unsigned* gbar, MegaShared* sh, unsigned code) {
  if (mbar_try_wait(bar, parity)) return;
  const unsigned long long t0 = gtime_ns();
  unsigned n = 0;
  while (!mbar_try_wait(bar, parity)) {
    if ((++n & 0x3f) == 0) {
      if (*((volatile int*)&sh->aborted) || *((volatile unsigned*)&gbar[2])) { sh->aborted = 1; return; }
      if (gtime_ns() - t0 > MG_WAIT_NS) { raise_abort(gbar, sh, code, parity); return; }
    }
  }
}

// ------------------------------------------------------------------ work partition ----
// A phase's weight rows are dealt out in UNITS: a RoPE row pair (QKV), one w1 row + the same
// w3 row (W13), or a single row (WO, W2, CLS).  CTA c owns units [c*P/grid, (c+1)*P/grid), a
// contiguous block of rows, and walks it in CHUNKS of 1, 2 or 4 units (as many as fit a ring
// slot, at most 4 rows).  Chunk i of the CTA's whole-step sequence lives in ring slot i % 16
// and is consumed by warp i % 16: every slot has exactly one consumer, which takes its chunks in
// order, so a full/empty mbarrier pair is never asked about a phase more than one ahead.
__device__ __forceinline__ void cta_units(int P, int& u0, int& u1) {
  u0 = (int)((long long)blockIdx.x * P / gridDim.x);
  u1 = (int)((long long)(blockIdx.x + 1) * P / gridDim.x);
}

enum { PH_QKV = 0, PH_ATTN = 1, PH_WO = 2, PH_W13 = 3, PH_W2 = 4, PH_CLS = 5 };

struct PhaseW {      // the weight side of a GEMV phase
  int P;             // units in the phase
  int K;             // row length (floats)
  int upr;           // rows per unit (2: QKV pair, W13 w1+w3; 1: others)
  int nmax;          // units per full chunk (power of two; nmax * upr <= 4 rows)
};

__device__ __forceinline__ PhaseW phase_weights(const tha_config_t& c, int kind, int slot_bytes) {
  const int d = c.dim, hd = c.hidden_dim, kvd = (d / c.n_heads) * c.n_kv_heads;
  PhaseW w;
  switch (kind) {
    case PH_QKV: w.P = (d + 2 * kvd) >> 1; w.K = d; w.upr = 2; break;
    case PH_WO:  w.P = d; w.K = d; w.upr = 1; break;
    case PH_W13: w.P = hd; w.K = d; w.upr = 2; break;
    case PH_W2:  w.P = d; w.K = hd; w.upr = 1; break;
    default:     w.P = c.vocab_size; w.K = d; w.upr = 1; break;
  }
  const int fit = slot_bytes / (w.upr * w.K * (int)sizeof(float));   // host guarantees >= 1
  const int cap = 4 / w.upr;
  const int n = fit < cap ? fit : cap;
  w.nmax = n >= 4 ? 4 : (n >= 2 ? 2 : 1);
  return w;
}
// Units in the chunk that starts with `remaining` units left.
__device__ __forceinline__ int chunk_units(int remaining, int nmax) {
  int n = remaining < nmax ? remaining : nmax;
  return n >= 4 ? 4 : (n >= 2 ? 2 : 1);
}

// Global address of the first row of unit `u` (for W13: of the w1 row, or the w3 row).
__device__ __forceinline__ const float* unit_src(const MegaP& p, int kind, int layer, int u, int which13) {
  const tha_config_t& c = p.cfg;
  const int d = c.dim, hd = c.hidden_dim, kvd = (d / c.n_heads) * c.n_kv_heads;
  switch (kind) {
    case PH_QKV: {
      int r = u * 2;
      if (r < d) return p.wq + ((size_t)layer * d + r) * d;
      r -= d;
      if (r < kvd) return p.wk + ((size_t)layer * kvd + r) * d;
      r -= kvd;
      return p.wv + ((size_t)layer * kvd + r) * d;
    }
    case PH_WO:  return p.wo + ((size_t)layer * d + u) * d;
    case PH_W13: return (which13 ? p.w3 : p.w1) + ((size_t)layer * hd + u) * d;
    case PH_W2:  return p.w2 + ((size_t)layer * d + u) * hd;
    default:     return p.wcls + (size_t)u * d;
  }
}

// ------------------------------------------------------------------ producer ----
// One lane streams this CTA's weight chunks of the whole step, in consumption order.
__device__ void producer_loop(const MegaP& p, MegaShared* sh, unsigned char* ring) {
  const tha_config_t& c = p.cfg;
  const uint64_t pol = policy_evict_first();
  unsigned ci = 0;
  const int nph = c.n_layers * 5 + 1;
  for (int ph = 0; ph < nph; ph++) {
    const int kind = (ph == nph - 1) ? PH_CLS : ph % 5;
    if (kind == PH_ATTN) continue;
    const int layer = ph / 5;
    const PhaseW w = phase_weights(c, kind, p.slot_bytes);
    int u0, u1;
    cta_units(w.P, u0, u1);
    const unsigned row_bytes = (unsigned)w.K * sizeof(float);
    for (int u = u0; u < u1; ci++) {
      const int n = chunk_units(u1 - u, w.nmax);
      const int slot = ci % MG_NSLOT;
      const unsigned par = (ci / MG_NSLOT) & 1u;
      mbar_wait(&sh->empty[slot], par ^ 1u, p.bar, sh, 100u + ph);
      if (*((volatile int*)&sh->aborted)) return;
      unsigned char* dst = ring + (size_t)slot * p.slot_bytes;
      mbar_expect_tx(&sh->full[slot], n * w.upr * row_bytes);
      if (kind == PH_W13) {          // [n w1 rows][n w3 rows]
        bulk_g2s(dst, unit_src(p, kind, layer, u, 0), n * row_bytes, &sh->full[slot], pol);
        bulk_g2s(dst + (size_t)n * row_bytes, unit_src(p, kind, layer, u, 1), n * row_bytes, &sh->full[slot], pol);
      } else if (kind == PH_QKV) {   // one copy per pair: a chunk may straddle wq|wk|wv
        for (int i = 0; i < n; i++)
          bulk_g2s(dst + (size_t)i * 2 * row_bytes, unit_src(p, kind, layer, u + i, 0), 2 * row_bytes, &sh->full[slot], pol);
      } else {
        bulk_g2s(dst, unit_src(p, kind, layer, u, 0), n * row_bytes, &sh->full[slot], pol);
      }
      u += n;
    }
  }
}

// ------------------------------------------------------------------ consumer pieces ----
// Stage the B activation rows (row b at rows[b], K floats) into shared memory.  512/BT threads
// per row; with `gain` the RMSNorm weight is folded in and the function returns, in lane l < 16
// of every warp, scale = 1/sqrt(mean(x^2)+1e-5) of row l / (16/BT) (run.cc:196-209; it is applied
// to the finished dot products).  One block barrier.  `copy_out`: also write the raw rows to p.x
// (layer 0: the embedding gather, src/thaDNN.cpp:31-36).
template <int BT>
__device__ __forceinline__ float stage_x(const MegaP& p, MegaShared* sh, float4* xs4, const float* const (&rows)[BT],
                                         int K, const float* gain, bool copy_out) {
  constexpr int TPR = MG_CT / BT;          // threads per row (>= 64: a warp never spans two rows)
  constexpr int WPR = MG_NW / BT;          // warps per row
  const int tid = threadIdx.x, K4 = K >> 2, lane = tid & 31, warp = tid >> 5;
  const int b = tid / TPR, j = tid % TPR;
  const float* row = nullptr;
#pragma unroll
  for (int bb = 0; bb < BT; bb++) if (bb == b) row = rows[bb];
  float ss = 0.0f;
  for (int i = j; i < K4; i += TPR) {
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (row) v = __ldcg(reinterpret_cast<const float4*>(row) + i);
    if (copy_out && row) reinterpret_cast<float4*>(p.x + (size_t)b * K)[i] = v;
    if (gain) {
      const float4 g = __ldg(reinterpret_cast<const float4*>(gain) + i);
      ss = dot4(v, v, ss);
      v.x *= g.x; v.y *= g.y; v.z *= g.z; v.w *= g.w;
    }
    xs4[b * K4 + i] = v;
  }
  if (gain) {
    ss = warp_sum(ss);
    if (lane == 0) sh->red[warp] = ss;
  }
  bar_consumers();
  float scale = 1.0f;
  if (gain) {
    float t = (lane < MG_NW) ? sh->red[lane] : 0.0f;
#pragma unroll
    for (int o = WPR >> 1; o > 0; o >>= 1) t += __shfl_xor_sync(THA_FULL_MASK, t, o);
    float m = t / K;
    m += 1e-5f;
    scale = 1.0f / sqrtf(m);
  }
  return scale;
}

// Reduce N = 2^k per-lane values across the warp with N-1 + (5-k) shuffles; afterwards value
// index (lane >> (5-k)) is complete in every lane of its group, in v[0].
template <int N>
__device__ __forceinline__ float warp_reduce_many(float (&v)[N], int lane) {
  int o = 16;
#pragma unroll
  for (int cnt = N; cnt > 1; cnt >>= 1, o >>= 1) {
    const int half = cnt >> 1;
    const bool upper = (lane & o) != 0;
#pragma unroll
    for (int i = 0; i < half; i++) {
      const float send = upper ? v[i] : v[i + half];
      const float keep = upper ? v[i + half] : v[i];
      v[i] = keep + __shfl_xor_sync(THA_FULL_MASK, send, o);
    }
  }
  float r = v[0];
  for (; o > 0; o >>= 1) r += __shfl_xor_sync(THA_FULL_MASK, r, o);
  return r;
}

__device__ __forceinline__ float silu_mul_mg(float a, float c) {
  float v = a;
  v *= (1.0f / (1.0f + expf(-v)));   // run.cc:344-347 operation order
  v *= c;
  return v;
}

// One chunk: NR weight rows (in the slot) against the BT staged activation rows, then the
// phase epilogue.  u0 = first unit of the chunk.
template <int BT, int NR>
__device__ __forceinline__ void consume_chunk(const MegaP& p, MegaShared* sh, int kind, int layer, int u0,
                                              const float4* ws4, const float4* xs4, int K, float scale_l) {
  constexpr int N = NR * BT;
  constexpr int LOGN = (N == 1) ? 0 : (N == 2) ? 1 : (N == 4) ? 2 : (N == 8) ? 3 : (N == 16) ? 4 : 5;
  constexpr int SH = 5 - LOGN;
  const int lane = threadIdx.x & 31, K4 = K >> 2;
  const int idx = lane >> SH;              // which (r, b) this lane finishes
  const int r = idx / BT, b = idx % BT;
  const bool writer = (lane & ((1 << SH) - 1)) == 0;
  const int pb = sh->pos[b];
  const tha_config_t& c = p.cfg;
  const int d = c.dim, hs = d / c.n_heads, kvd = hs * c.n_kv_heads;
  // residual phases: fetch the old x element now, its L2 round trip hides behind the dot product
  float xold = 0.0f;
  float* xdst = nullptr;
  if ((kind == PH_WO || kind == PH_W2) && writer && pb >= 0) {
    xdst = p.x + (size_t)b * d + (u0 + r);
    xold = __ldcg(xdst);
  }
  const float scale = __shfl_sync(THA_FULL_MASK, scale_l, b * (MG_NW / BT));
  float acc[N];
#pragma unroll
  for (int i = 0; i < N; i++) acc[i] = 0.0f;
#pragma unroll 2
  for (int k4 = lane; k4 < K4; k4 += 32) {
    float4 w[NR];
#pragma unroll
    for (int r = 0; r < NR; r++) w[r] = ws4[r * K4 + k4];
#pragma unroll
    for (int b = 0; b < BT; b++) {
      const float4 xv = xs4[b * K4 + k4];
#pragma unroll
      for (int r = 0; r < NR; r++) acc[r * BT + b] = dot4(w[r], xv, acc[r * BT + b]);
    }
  }
  float v = warp_reduce_many<N>(acc, lane);

  if (kind == PH_QKV) {
    v *= scale;
    const float v1 = __shfl_sync(THA_FULL_MASK, v, ((idx ^ BT) << SH) & 31);   // partner row of the RoPE pair
    if (NR >= 2 && writer && !(r & 1) && pb >= 0) {
      int row = (u0 + (r >> 1)) * 2;       // virtual row over wq|wk|wv
      float o0 = v, o1 = v1;
      float* dst;
      if (row < d + kvd) {                 // q or k: rotate (run.cc:271-286)
        const int lr = (row < d) ? row : row - d;
        const int j = (lr % hs) >> 1;
        const float fc = sh->rope_c[b][j], fs = sh->rope_s[b][j];
        o0 = v * fc - v1 * fs;
        o1 = v * fs + v1 * fc;
        dst = (row < d) ? p.q + (size_t)b * d + lr
                        : p.kc + ((size_t)b * c.n_layers + layer) * c.seq_len * kvd + (size_t)pb * kvd + lr;
      } else {
        dst = p.vc + ((size_t)b * c.n_layers + layer) * c.seq_len * kvd + (size_t)pb * kvd + (row - d - kvd);
      }
      *reinterpret_cast<float2*>(dst) = make_float2(o0, o1);
    }
  } else if (kind == PH_WO || kind == PH_W2) {
    if (xdst) *xdst = xold + v;
  } else if (kind == PH_W13) {
    v *= scale;
    constexpr int NU = NR / 2;             // slot rows: [NU w1 rows][NU w3 rows]
    const float cpart = __shfl_sync(THA_FULL_MASK, v, ((idx ^ (NU * BT)) << SH) & 31);   // the w3 row
    if (NR >= 2 && writer && r < NU && pb >= 0) p.hb[(size_t)b * c.hidden_dim + u0 + r] = silu_mul_mg(v, cpart);
  } else {  // PH_CLS
    v *= scale;
    const int row = u0 + r;
    if (writer && pb >= 0 && p.logits) p.logits[(size_t)b * c.vocab_size + row] = v;
    // running greedy candidate of (this warp, slot b): best of the chunk's NR rows first
    float cv = (pb >= 0) ? v : -INFINITY;
    int cidx = row;
#pragma unroll
    for (int st = BT; st < N; st <<= 1) {
      const float ov = __shfl_xor_sync(THA_FULL_MASK, cv, st << SH);
      const int oi = __shfl_xor_sync(THA_FULL_MASK, cidx, st << SH);
      if (ov > cv || (ov == cv && oi < cidx)) { cv = ov; cidx = oi; }
    }
    if (writer && r == 0) {
      const int w = threadIdx.x >> 5;
      const float bv = sh->amax_v[w * MG_MAX_B + b];
      const int bi = sh->amax_i[w * MG_MAX_B + b];
      if (cv > bv || (cv == bv && cidx < bi)) { sh->amax_v[w * MG_MAX_B + b] = cv; sh->amax_i[w * MG_MAX_B + b] = cidx; }
    }
  }
}

// ------------------------------------------------------------------ grid barrier ----
// Monotonic arrival counter: barrier k completes when the counter reaches k * gridDim.x.
// All consumer threads call both; writes made before grid_arrive are visible after grid_wait.
__device__ __forceinline__ void grid_arrive(const MegaP& p) {
  bar_consumers();                                          // every consumer's writes happen-before thread 0 here
  if (threadIdx.x == 0) red_release_add(&p.bar[0], 1u);     // release at gpu scope, cumulative over them
}
__device__ __forceinline__ void grid_wait(const MegaP& p, MegaShared* sh, unsigned target) {
  if (threadIdx.x == 0 && ld_acquire(&p.bar[0]) < target) {
    const unsigned long long t0 = gtime_ns();
    unsigned n = 0;
    while (ld_acquire(&p.bar[0]) < target) {
      if ((++n & 0x3f) == 0) {
        if (*((volatile int*)&sh->aborted) || *((volatile unsigned*)&p.bar[2])) { sh->aborted = 1; break; }
        if (gtime_ns() - t0 > MG_WAIT_NS) { raise_abort(p.bar, sh, 10000u, target); break; }
      }
    }
  }
  bar_consumers();
}

// ------------------------------------------------------------------ attention phase ----
// Items (slot b, head h, time split s), one CTA per item.  Inside the CTA warp w takes rows
// w, w+16, ... of the split (32/LH rows per warp pass: LH = head_size/4 lanes hold one K/V row of
// the head as float4), U passes in flight, online softmax per lane group; lane groups merge by
// shuffles, the 16 warps merge through shared memory.  A context shorter than one CTA iteration
// (16 * 32/LH * U rows: 128 at head_size 64) is never split; longer ones are split over up to
// `smax` CTAs and the last CTA to arrive merges the partials in split order (result independent
// of arrival order).  `scr` = the activation staging area, free during this phase.
template <int HS>
__device__ void attn_phase(const MegaP& p, MegaShared* sh, int layer, float* scr) {
  constexpr int LH = HS / 4, TPW = 32 / LH, U = 4, RPI = MG_NW * TPW * U;   // rows per CTA iteration
  const tha_config_t& c = p.cfg;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int g = lane / LH, li = lane % LH;
  const int H = c.n_heads, d = c.dim, kvd = HS * c.n_kv_heads, kv_mul = H / c.n_kv_heads;
  const int items = p.B * H * p.smax;
  const float sqrt_hs = sqrtf((float)HS);
  float4* sm_acc = reinterpret_cast<float4*>(scr);                 // [16][LH]
  float2* sm_ml = reinterpret_cast<float2*>(scr + MG_NW * HS);     // [16]
  int* sm_flag = reinterpret_cast<int*>(scr + MG_NW * HS + 2 * MG_NW);
  for (int it = blockIdx.x; it < items; it += gridDim.x) {
    const int s = it % p.smax, bh = it / p.smax, h = bh % H, b = bh / H;
    const int pb = sh->pos[b];
    if (pb < 0) continue;
    const int n = pb + 1;
    int S = (n + RPI - 1) / RPI;
    if (S > p.smax) S = p.smax;
    const int chunk = (n + S - 1) / S;
    S = (n + chunk - 1) / chunk;
    if (s >= S) continue;
    const int t0 = s * chunk, t1 = (t0 + chunk < n) ? t0 + chunk : n;

    const float4 q4 = __ldcg(reinterpret_cast<const float4*>(p.q + (size_t)b * d + h * HS) + li);
    const size_t base = ((size_t)b * c.n_layers + layer) * c.seq_len * kvd + (size_t)(h / kv_mul) * HS + 4 * li;
    const float* kb = p.kc + base;
    const float* vb = p.vc + base;
    float m = -INFINITY, l = 0.0f;
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int tb = t0; tb < t1; tb += RPI) {     // warp-uniform trip count: the shuffles below need all lanes
      float4 k4[U], v4[U];
      bool in[U];
#pragma unroll
      for (int u = 0; u < U; u++) {
        const int t = tb + (u * MG_NW + warp) * TPW + g;
        in[u] = t < t1;
        const size_t off = (size_t)(in[u] ? t : t0) * kvd;
        k4[u] = __ldcg(reinterpret_cast<const float4*>(kb + off));
        v4[u] = __ldcg(reinterpret_cast<const float4*>(vb + off));
      }
      float sc[U];
#pragma unroll
      for (int u = 0; u < U; u++) {
        float x = dot4(q4, k4[u], 0.0f);
#pragma unroll
        for (int o = LH >> 1; o > 0; o >>= 1) x += __shfl_xor_sync(THA_FULL_MASK, x, o);
        sc[u] = in[u] ? x / sqrt_hs : -INFINITY;
      }
      float mn = m;
#pragma unroll
      for (int u = 0; u < U; u++) mn = fmaxf(mn, sc[u]);
      const float corr = (mn == -INFINITY) ? 1.0f : expf(m - mn);   // a lane group may have no row in this pass
      l *= corr;
      acc.x *= corr; acc.y *= corr; acc.z *= corr; acc.w *= corr;
#pragma unroll
      for (int u = 0; u < U; u++) {
        const float pe = in[u] ? expf(sc[u] - mn) : 0.0f;
        l += pe;
        acc.x = fmaf(pe, v4[u].x, acc.x); acc.y = fmaf(pe, v4[u].y, acc.y);
        acc.z = fmaf(pe, v4[u].z, acc.z); acc.w = fmaf(pe, v4[u].w, acc.w);
      }
      m = mn;
    }
    // merge the TPW lane groups of the warp
#pragma unroll
    for (int o = LH; o < 32; o <<= 1) {
      const float m2 = __shfl_xor_sync(THA_FULL_MASK, m, o), l2 = __shfl_xor_sync(THA_FULL_MASK, l, o);
      float4 a2;
      a2.x = __shfl_xor_sync(THA_FULL_MASK, acc.x, o); a2.y = __shfl_xor_sync(THA_FULL_MASK, acc.y, o);
      a2.z = __shfl_xor_sync(THA_FULL_MASK, acc.z, o); a2.w = __shfl_xor_sync(THA_FULL_MASK, acc.w, o);
      const float mn = fmaxf(m, m2);
      const float e1 = (m == -INFINITY) ? 0.0f : expf(m - mn), e2 = (m2 == -INFINITY) ? 0.0f : expf(m2 - mn);
      l = l * e1 + l2 * e2;    // symmetric form: both partners compute bit-identical results
      acc.x = acc.x * e1 + a2.x * e2; acc.y = acc.y * e1 + a2.y * e2;
      acc.z = acc.z * e1 + a2.z * e2; acc.w = acc.w * e1 + a2.w * e2;
      m = mn;
    }
    // merge the 16 warps through shared memory
    if (g == 0) sm_acc[warp * LH + li] = acc;
    if (lane == 0) sm_ml[warp] = make_float2(m, l);
    bar_consumers();
    float4* outp = reinterpret_cast<float4*>(p.xb + (size_t)b * d + h * HS) + li;
    const size_t pbase = (size_t)bh * p.smax;
    if (warp == 0) {
      float M = -INFINITY, L = 0.0f;
      float4 A = make_float4(0.f, 0.f, 0.f, 0.f);
      if (g == 0) {
#pragma unroll
        for (int w2 = 0; w2 < MG_NW; w2++) M = fmaxf(M, sm_ml[w2].x);
#pragma unroll
        for (int w2 = 0; w2 < MG_NW; w2++) {
          const float2 ml = sm_ml[w2];
          const float e = (ml.x == -INFINITY) ? 0.0f : expf(ml.x - M);
          const float4 a = sm_acc[w2 * LH + li];
          L += ml.y * e;
          A.x += a.x * e; A.y += a.y * e; A.z += a.z * e; A.w += a.w * e;
        }
      }
      if (S == 1) {
        if (g == 0) *outp = make_float4(A.x / L, A.y / L, A.z / L, A.w / L);
      } else {
        if (g == 0) reinterpret_cast<float4*>(p.att_acc + (pbase + s) * HS)[li] = A;
        if (lane == 0) p.att_ml[pbase + s] = make_float2(M, L);
        __threadfence();
        __syncwarp();
        int last = 0;
        if (lane == 0) last = (atomicAdd(&p.att_cnt[bh], 1) == S - 1);
        last = __shfl_sync(THA_FULL_MASK, last, 0);
        if (last) {
          __threadfence();
          if (g == 0) {
            float GM = -INFINITY;
            for (int s2 = 0; s2 < S; s2++) GM = fmaxf(GM, __ldcg(&p.att_ml[pbase + s2]).x);
            float GL = 0.0f;
            float4 GA = make_float4(0.f, 0.f, 0.f, 0.f);
            for (int s2 = 0; s2 < S; s2++) {
              const float2 ml = __ldcg(&p.att_ml[pbase + s2]);
              const float e = (ml.x == -INFINITY) ? 0.0f : expf(ml.x - GM);
              const float4 a = __ldcg(reinterpret_cast<const float4*>(p.att_acc + (pbase + s2) * HS) + li);
              GL += ml.y * e;
              GA.x += a.x * e; GA.y += a.y * e; GA.z += a.z * e; GA.w += a.w * e;
            }
            *outp = make_float4(GA.x / GL, GA.y / GL, GA.z / GL, GA.w / GL);
          }
          if (lane == 0) p.att_cnt[bh] = 0;   // self-reset for the next layer / step
        }
      }
    }
    if (it + (int)gridDim.x < items) bar_consumers();   // the scratch is reused by this CTA's next item
  }
}

// ------------------------------------------------------------------ the kernel ----
template <int BT>
__global__ void __launch_bounds__(MG_THREADS, 1) mega_step_kernel(const MegaP p) {
  extern __shared__ __align__(128) unsigned char smem_mg[];
  unsigned char* ring = smem_mg;
  float4* xs4 = reinterpret_cast<float4*>(smem_mg + (size_t)MG_NSLOT * p.slot_bytes);
  MegaShared* sh = reinterpret_cast<MegaShared*>(smem_mg + (size_t)MG_NSLOT * p.slot_bytes + (size_t)p.xs_floats * sizeof(float));
  const tha_config_t& c = p.cfg;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

  if (tid == 0) {
    for (int i = 0; i < MG_NSLOT; i++) { mbar_init(&sh->full[i], 1); mbar_init(&sh->empty[i], 1); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    sh->aborted = 0;
  }
  if (tid < MG_MAX_B) {
    sh->pos[tid] = (tid < p.B) ? p.pos[tid] : -1;
    sh->tok[tid] = (tid < p.B) ? p.token[tid] : 0;
    sh->scale[tid] = 1.0f;
  }
  if (tid < MG_NW * MG_MAX_B) { sh->amax_v[tid] = -INFINITY; sh->amax_i[tid] = 0x7fffffff; }
  {
    // RoPE row of every slot's position -> shared (asynchronously; first use is the QKV epilogue)
    const int half = (c.dim / c.n_heads) >> 1;
    if (tid < p.B * half) {
      const int b = tid / half, j = tid % half, pb = p.pos[b];
      if (pb >= 0) {
        asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(smem_u32(&sh->rope_c[b][j])), "l"(p.rope_cos + (size_t)pb * half + j));
        asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(smem_u32(&sh->rope_s[b][j])), "l"(p.rope_sin + (size_t)pb * half + j));
      }
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
    // the RMSNorm weights of the whole step -> L2 now (75 KB), so staging never waits on HBM for them
    const int lines_l = c.n_layers * c.dim * 4 / 128;
    for (int i = blockIdx.x * MG_THREADS + tid; i < 2 * lines_l + c.dim * 4 / 128; i += gridDim.x * MG_THREADS) {
      const char* a = (i < lines_l) ? (const char*)p.rms_att + (size_t)i * 128
                    : (i < 2 * lines_l) ? (const char*)p.rms_ffn + (size_t)(i - lines_l) * 128
                                        : (const char*)p.rms_final + (size_t)(i - 2 * lines_l) * 128;
      asm volatile("prefetch.global.L2 [%0];" ::"l"(a));
    }
  }
  __syncthreads();

  if (warp == MG_NW) {   // ---- producer warp
    if (lane == 0) producer_loop(p, sh, ring);
    return;
  }

  // ---- consumer warps
  const int d = c.dim, hd = c.hidden_dim, hs = d / c.n_heads;
  const int nph = c.n_layers * 5 + 1;
  unsigned ci = 0, nbar = 0;
  for (int ph = 0; ph < nph; ph++) {
    const int kind = (ph == nph - 1) ? PH_CLS : ph % 5;
    const int layer = ph / 5;
    unsigned long long* tr = (p.trace && tid == 0) ? p.trace + ((size_t)blockIdx.x * nph + ph) * 4 : nullptr;
    if (tr) tr[0] = gtime_ns();
    if (ph > 0) grid_wait(p, sh, ++nbar * gridDim.x);
    if (tr) tr[1] = gtime_ns();
    if (kind == PH_ATTN) {
      if (hs == 64) attn_phase<64>(p, sh, layer, reinterpret_cast<float*>(xs4));
      else if (hs == 32) attn_phase<32>(p, sh, layer, reinterpret_cast<float*>(xs4));
      else attn_phase<128>(p, sh, layer, reinterpret_cast<float*>(xs4));
    } else {
      const PhaseW w = phase_weights(c, kind, p.slot_bytes);
      // ---- stage the activations of this phase
      const float* rows[BT];
      const float* gain = nullptr;
#pragma unroll
      for (int b = 0; b < BT; b++) {
        const bool on = sh->pos[b] >= 0;
        const float* r = nullptr;
        if (on) {
          if (kind == PH_QKV) r = (layer == 0) ? p.tok_emb + (size_t)sh->tok[b] * d : p.x + (size_t)b * d;
          else if (kind == PH_WO) r = p.xb + (size_t)b * d;
          else if (kind == PH_W2) r = p.hb + (size_t)b * hd;
          else r = p.x + (size_t)b * d;
        }
        rows[b] = r;
      }
      if (kind == PH_QKV) gain = p.rms_att + (size_t)layer * d;
      else if (kind == PH_W13) gain = p.rms_ffn + (size_t)layer * d;
      else if (kind == PH_CLS) gain = p.rms_final;
      if (ph == 0) asm volatile("cp.async.wait_all;" ::: "memory");   // RoPE rows: visible to all after stage_x's barrier
      const float scale_l = stage_x<BT>(p, sh, xs4, rows, w.K, gain, kind == PH_QKV && layer == 0 && blockIdx.x == 0);
      if (tr) tr[2] = gtime_ns();
      // ---- consume this CTA's chunks of the phase (chunk idx -> warp idx % 16, slot idx % 16)
      int u0, u1;
      cta_units(w.P, u0, u1);
      for (int u = u0; u < u1; ci++) {
        const int n = chunk_units(u1 - u, w.nmax);
        if ((int)(ci % MG_NW) == warp) {
          const int slot = ci % MG_NSLOT;
          mbar_wait(&sh->full[slot], (ci / MG_NSLOT) & 1u, p.bar, sh, 1000u + ph);
          const float4* ws4 = reinterpret_cast<const float4*>(ring + (size_t)slot * p.slot_bytes);
          const int nr = n * w.upr;
          if (nr == 4) consume_chunk<BT, 4>(p, sh, kind, layer, u, ws4, xs4, w.K, scale_l);
          else if (nr == 2) consume_chunk<BT, 2>(p, sh, kind, layer, u, ws4, xs4, w.K, scale_l);
          else consume_chunk<BT, 1>(p, sh, kind, layer, u, ws4, xs4, w.K, scale_l);
          __syncwarp();
          if (lane == 0) mbar_arrive(&sh->empty[slot]);
        }
        u += n;
      }
    }
    if (kind == PH_CLS) {   // this CTA's argmax candidates (first maximum wins, src/llama.cpp:275-286)
      bar_consumers();
      if (tid < BT) {
        float v = -INFINITY; int i = 0x7fffffff;
        for (int w2 = 0; w2 < MG_NW; w2++) {
          const float ov = sh->amax_v[w2 * MG_MAX_B + tid]; const int oi = sh->amax_i[w2 * MG_MAX_B + tid];
          if (ov > v || (ov == v && oi < i)) { v = ov; i = oi; }
        }
        p.amax_val[blockIdx.x * MG_MAX_B + tid] = v;
        p.amax_idx[blockIdx.x * MG_MAX_B + tid] = i;
      }
    }
    if (p.trace) { bar_consumers(); if (tr) tr[3] = gtime_ns(); }
    grid_arrive(p);
  }

  // ---- tail: CTA 0 picks the greedy token per slot and advances the slot table
  if (blockIdx.x == 0) {
    grid_wait(p, sh, ++nbar * gridDim.x);
    if (warp < p.B && p.next) {
      float v = -INFINITY; int i = 0x7fffffff;
      for (int cta = lane; cta < (int)gridDim.x; cta += 32) {
        const float ov = __ldcg(&p.amax_val[cta * MG_MAX_B + warp]); const int oi = __ldcg(&p.amax_idx[cta * MG_MAX_B + warp]);
        if (ov > v || (ov == v && oi < i)) { v = ov; i = oi; }
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        const float ov = __shfl_xor_sync(THA_FULL_MASK, v, o);
        const int oi = __shfl_xor_sync(THA_FULL_MASK, i, o);
        if (ov > v || (ov == v && oi < i)) { v = ov; i = oi; }
      }
      const int pb = sh->pos[warp];
      if (lane == 0 && pb >= 0) {
        const int nx = (i == 0x7fffffff) ? 0 : i;
        p.next[warp] = nx;
        if (p.advance) { p.token[warp] = nx; p.pos[warp] = (pb + 1 < c.seq_len) ? pb + 1 : -1; }
      }
    }
  }
  // ---- exit protocol: the last CTA out resets the barrier words for the next launch
  bar_consumers();
  if (tid == 0) {
    __threadfence();
    if (atomicAdd(&p.bar[1], 1u) == gridDim.x - 1) {
      p.bar[0] = 0u;
      p.bar[1] = 0u;
      __threadfence();
    }
  }
}

// ------------------------------------------------------------------ host side ----
static int g_mega_enabled = -1;
int mega_enabled() {
  if (g_mega_enabled < 0) {
    const char* v = getenv("THA_MEGA");
    g_mega_enabled = (v && *v) ? (atoi(v) != 0) : 1;
  }
  return g_mega_enabled;
}
void mega_set_enabled(int v) { g_mega_enabled = v != 0; }

static int mega_bt(int B) { return B <= 1 ? 1 : (B <= 2 ? 2 : (B <= 4 ? 4 : 8)); }

// Shared-memory carve-up: activations first (B_tile rows of max(dim, hidden)), the rest is the ring.
static void mega_smem_plan(const tha_config_t* c, int B, int* slot_bytes, int* xs_floats) {
  const int bt = mega_bt(B), kmax = c->dim > c->hidden_dim ? c->dim : c->hidden_dim;
  *xs_floats = bt * kmax;
  if (*xs_floats < MG_NW * 128 + 2 * MG_NW + 8) *xs_floats = MG_NW * 128 + 2 * MG_NW + 8;   // attention merge scratch
  const int left = MG_SMEM_BUDGET - (int)sizeof(MegaShared) - 256 - *xs_floats * (int)sizeof(float);
  int sb = left / MG_NSLOT;
  sb -= sb % 128;
  *slot_bytes = sb;
}

bool mega_supported(const tha_config_t* c, int B) {
  if (B < 1 || B > MG_MAX_B) return false;
  const int d = c->dim, hd = c->hidden_dim;
  if (d <= 0 || hd <= 0 || d % c->n_heads || c->n_heads % c->n_kv_heads) return false;
  const int hs = d / c->n_heads, kvd = hs * c->n_kv_heads;
  if (hs != 32 && hs != 64 && hs != 128) return false;
  if (d % 4 || hd % 4 || (kvd & 1)) return false;
  int slot_bytes, xs_floats;
  mega_smem_plan(c, B, &slot_bytes, &xs_floats);
  if (xs_floats > MG_XS_MAX_FLOATS) return false;
  if (2 * d * (int)sizeof(float) > slot_bytes) return false;   // a RoPE pair / a w1+w3 row couple
  if (hd * (int)sizeof(float) > slot_bytes) return false;      // one w2 row
  return true;
}

struct MegaScratch {
  unsigned* bar = nullptr;
  float* att_acc = nullptr; float2* att_ml = nullptr; int* att_cnt = nullptr;
  float* amax_val = nullptr; int* amax_idx = nullptr;
  size_t att_items = 0, att_floats = 0; int cnt_cap = 0, grid = 0;
  unsigned long long* trace = nullptr;
};
static std::map<int, MegaScratch> g_scratch;   // per device

// Launch the persistent step kernel.  All pointer arguments as in decode_step_f32.
int mega_step_launch(cudaStream_t st, Workspace* ws, int B, const tha_config_t* cfg, const tha_weights_f32_t* w,
                     const tha_state_t* s, int* token_d, int* pos_d, float* logits_d, int* next_d, int advance,
                     const float* rope_cos, const float* rope_sin, bool prepare_only) {
  int dev = ws->device;
  MegaScratch& sc = g_scratch[dev];
  if (!sc.grid) {
    cudaDeviceProp prop;
    THA_CUDA_OK(cudaGetDeviceProperties(&prop, dev));
    sc.grid = prop.multiProcessorCount;
    THA_CUDA_OK(cudaMalloc(&sc.bar, 16 * sizeof(unsigned)));
    THA_CUDA_OK(cudaMemset(sc.bar, 0, 16 * sizeof(unsigned)));
    THA_CUDA_OK(cudaMalloc(&sc.amax_val, (size_t)sc.grid * MG_MAX_B * sizeof(float)));
    THA_CUDA_OK(cudaMalloc(&sc.amax_idx, (size_t)sc.grid * MG_MAX_B * sizeof(int)));
    const char* tv = getenv("THA_MEGA_TRACE");
    if (tv && atoi(tv)) {
      THA_CUDA_OK(cudaMalloc(&sc.trace, (size_t)sc.grid * 1024 * 4 * sizeof(unsigned long long)));
      THA_CUDA_OK(cudaMemset(sc.trace, 0, (size_t)sc.grid * 1024 * 4 * sizeof(unsigned long long)));
    }
    THA_CUDA_OK(cudaDeviceSynchronize());   // the memset runs on the legacy stream; callers use non-blocking streams
  }
  const int hs = cfg->dim / cfg->n_heads;
  int smax = sc.grid / (B * cfg->n_heads);   // time splits (CTAs) per (slot, head)
  if (smax > 16) smax = 16;
  if (smax < 1) smax = 1;
  const size_t items = (size_t)B * cfg->n_heads * smax;
  if (items * hs > sc.att_floats || items > sc.att_items || B * cfg->n_heads > sc.cnt_cap) {
    if (sc.att_acc) { cudaFree(sc.att_acc); cudaFree(sc.att_ml); cudaFree(sc.att_cnt); }
    sc.att_floats = items * hs; sc.att_items = items; sc.cnt_cap = B * cfg->n_heads;
    THA_CUDA_OK(cudaMalloc(&sc.att_acc, sc.att_floats * sizeof(float)));
    THA_CUDA_OK(cudaMalloc(&sc.att_ml, sc.att_items * sizeof(float2)));
    THA_CUDA_OK(cudaMalloc(&sc.att_cnt, sc.cnt_cap * sizeof(int)));
    THA_CUDA_OK(cudaMemset(sc.att_cnt, 0, sc.cnt_cap * sizeof(int)));
    THA_CUDA_OK(cudaDeviceSynchronize());
    ws_drop_graphs(ws);
  }
  if (prepare_only) return THA_STATUS_SUCCESS;

  MegaP p;
  memset(&p, 0, sizeof(p));
  p.cfg = *cfg; p.B = B; p.advance = advance;
  p.tok_emb = w->token_embedding_table; p.rms_att = w->rms_att_weight; p.rms_ffn = w->rms_ffn_weight;
  p.rms_final = w->rms_final_weight; p.wq = w->wq; p.wk = w->wk; p.wv = w->wv; p.wo = w->wo;
  p.w1 = w->w1; p.w2 = w->w2; p.w3 = w->w3; p.wcls = w->wcls;
  p.x = s->x; p.q = s->q; p.xb = s->xb; p.hb = s->hb; p.kc = s->key_cache; p.vc = s->value_cache;
  p.logits = logits_d; p.next = next_d; p.token = token_d; p.pos = pos_d;
  p.rope_cos = rope_cos; p.rope_sin = rope_sin;
  p.bar = sc.bar; p.att_acc = sc.att_acc; p.att_ml = sc.att_ml; p.att_cnt = sc.att_cnt; p.smax = smax;
  p.amax_val = sc.amax_val; p.amax_idx = sc.amax_idx;
  mega_smem_plan(cfg, B, &p.slot_bytes, &p.xs_floats);
  p.trace = sc.trace;

  void (*k)(const MegaP) = nullptr;
  switch (mega_bt(B)) {
    case 1: k = mega_step_kernel<1>; break;
    case 2: k = mega_step_kernel<2>; break;
    case 4: k = mega_step_kernel<4>; break;
    default: k = mega_step_kernel<8>; break;
  }
  const size_t smem = (size_t)MG_NSLOT * p.slot_bytes + (size_t)p.xs_floats * sizeof(float) + sizeof(MegaShared);
  THA_CUDA_OK(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  cudaLaunchConfig_t lc = {};
  lc.gridDim = dim3(sc.grid); lc.blockDim = dim3(MG_THREADS); lc.dynamicSmemBytes = smem; lc.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeCooperative;   // co-residency of all CTAs is required by the grid barrier
  attr[0].val.cooperative = 1;
  lc.attrs = attr; lc.numAttrs = 1;
  count_launch();
  THA_CUDA_OK(cudaLaunchKernelEx(&lc, k, p));
  return THA_STATUS_SUCCESS;
}

// Development aid: copy out the phase time stamps of the last step (THA_MEGA_TRACE=1).
extern "C" int tha_mega_trace(unsigned long long* out_host, int max_entries, int* grid_out) {
  int dev = 0;
  cudaGetDevice(&dev);
  auto it = g_scratch.find(dev);
  if (it == g_scratch.end() || !it->second.trace) return THA_STATUS_NOT_SUPPORTED;
  size_t n = (size_t)it->second.grid * 1024 * 4;
  if ((size_t)max_entries < n) n = max_entries;
  THA_CUDA_OK(cudaMemcpy(out_host, it->second.trace, n * sizeof(unsigned long long), cudaMemcpyDeviceToHost));
  if (grid_out) *grid_out = it->second.grid;
  return THA_STATUS_SUCCESS;
}

// Abort flag of the persistent kernel (a wait timed out).  Reads and clears it; call after a sync.
int mega_check_abort(int dev) {
  auto it = g_scratch.find(dev);
  if (it == g_scratch.end() || !it->second.bar) return 0;
  unsigned h[4] = {0, 0, 0, 0};
  if (cudaMemcpy(h, it->second.bar, sizeof(h), cudaMemcpyDeviceToHost) != cudaSuccess) return 1;
  if (h[2]) {
    unsigned r[8];
    cudaMemcpy(r, it->second.bar, sizeof(r), cudaMemcpyDeviceToHost);
    fprintf(stderr, "tha: persistent step kernel aborted: code %u (100+ph producer/empty, 1000+ph consumer/full, 10000 grid) "
                    "cta %u warp %u detail %u arrivals %u\n", r[4], r[5], r[6], r[7], r[0]);
    cudaMemset(it->second.bar, 0, 16 * sizeof(unsigned));
    return 1;
  }
  return 0;
}

}  // namespace tha
