// Causal attention for batched prompt ingestion ON THE TENSOR CORES (bf16 model path): Q K^T and P V as
// tcgen05.mma.kind::f16 with fp32 accumulators in TMEM, softmax in fp32 registers between them.
//
// attn_prefill.cu does both contractions on CUDA cores: 1.66 ms per Llama-2-7B layer at 2048 positions with the tensor
// pipe idle (profiles/r02_prefill_7b_2048_ncu_full_summary.txt) -- more than the layer's four bf16 matmuls together.
//
// Rule (restated by the oracle, oracle/llama2_oracle.c orc_attention_bf16_tiled; reference arithmetic: run.cc:288-321):
// keys are visited in tiles of 128 starting at key 0; q, k, v are rounded to bf16 (nearest even); scores, the running
// maximum m, the running sum l and the output accumulator are fp32; per tile  p = exp(s - m_new)  is rounded to bf16 for
// the P V product, l accumulates the UNROUNDED p;  out = acc / l.
//
// CTA = 128 consecutive query rows of one (sequence, head), 256 threads: two threads per query row (= TMEM lane), each with
// half of the row's key columns and half of its head-dimension accumulator:
//   Q tile -> shared memory once (bf16, K-major, 128-byte swizzled rows: the layout TMA would write);
//   per key tile: K rows -> shared (same layout, rows = keys);  V rows -> shared TRANSPOSED (rows = head dimension, K = keys:
//   the B operand of P V must be K-major);  one thread issues hs/16 MMAs: S[128 x 128] into TMEM;  every thread reads its row
//   of S (two passes: maximum, then exponentials), writes its row of P (bf16) into shared memory;  hs/... 8 MMAs: P V into
//   TMEM;  every thread folds its row of P V into its register accumulator (rescaled by exp(m_old - m_new)).
// Nothing overlaps inside a CTA; there are tiles x heads CTAs (512 per 7B layer at 2048 positions) to fill the machine.
#include <string.h>

#include "common.cuh"
#include "gemm_tc.h"
#include "tc_common.cuh"
#include "workspace.h"

namespace tha {

struct AttnPfTcP {
  const float* q; float* out; uint16_t* out_b16;              // [rows, dim]; out_b16 nullable: the row rounded to bf16 (wo's input)
  const float* kc; const float* vc;                           // cache base of the layer (dense: + loff applied; paged: pool base)
  long long kv_batch_stride; int kv_dim, kv_mul, dim;
  const int* row_slot; const int* row_pos;                    // [rows]
  const int* tile_row0; const int* tile_n;                    // [tiles]: first row and row count (<= 128)
  KvPaged pg;
  int* error_flag;
  int staged_rows;     // STAGED kernels: rows of the bf16 staging buffers (K row-major, V transposed) the tensor maps describe
  const int* row_stage; // STAGED: staging row of every row of the call (a sequence starts at a multiple of 64: TMA wants the
                        // inner coordinate of the transposed V tile 16-byte aligned)
};

constexpr int PTC_TQ = 128, PTC_TK = 128;
#ifndef PTC_CTAS_PER_SM
#define PTC_CTAS_PER_SM 1
#endif

// two fp32 -> packed bf16x2, round to nearest even (one instruction; a in the low half)
__device__ __forceinline__ uint32_t ptc_bf16x2(float a, float b) {
  uint32_t r;
  asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(b), "f"(a));
  return r;
}
__device__ __forceinline__ float ptc_round(float a) {
  const uint32_t u = __float_as_uint(a);
  return __uint_as_float((u + 0x7FFFu + ((u >> 16) & 1u)) & 0xFFFF0000u);
}
__device__ __forceinline__ void ptc_tmem_ld32(uint32_t taddr, uint32_t (&v)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, %17, %18, %19, "
      "%20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]),
        "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]),
        "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]),
        "=r"(v[30]), "=r"(v[31])
      : "r"(taddr));
}
// kind::f16, bf16 x bf16 -> fp32, both K-major, M = 128, N = n
__host__ __device__ constexpr uint32_t ptc_idesc(int n) {
  return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
}
__device__ __forceinline__ void ptc_mma(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
// byte offset of element (row r, column c) of a K-major bf16 tile whose rows are 128 bytes (64 values), SWIZZLE_128B
__device__ __forceinline__ int ptc_off(int r, int c) { return r * 128 + ((((c >> 3) ^ (r & 7))) << 4) + ((c & 7) << 1); }

// STAGED: the prompt's K rows (bf16, row-major) and V rows (bf16, TRANSPOSED: [kv column][prompt row]) were written once per
// layer by bf_kv_stage_kernel (step_bf16.cu); a key tile then is two TMA loads (mk: [128 keys x HS], mv: [HS x 128 keys]) into
// the swizzled layout the MMAs read, double-buffered so that the next tile lands while this one is multiplied and soft-maxed.
// (Unstaged, every CTA re-reads the fp32 cache rows, converts them and transposes V with 2-byte shared-memory stores: at 2048
// positions that was 363 us per 7B layer with the tensor pipe 4.6 % busy.)  Staged keys are addressed by prompt row: the
// sequence's row of position 0 is row_stage[row0] - pos0, so every sequence of the call must start at position 0.
template <int HS, bool STAGED>
__global__ void __launch_bounds__(256, PTC_CTAS_PER_SM)
attn_prefill_tc_kernel(const __grid_constant__ CUtensorMap mk, const __grid_constant__ CUtensorMap mv, const AttnPfTcP p) {
  constexpr int KBH = HS / 64;                       // 64-value k-blocks along the head dimension
  constexpr int Q_BYTES = KBH * PTC_TQ * 128;        // Q tile  [128 rows][HS]
  constexpr int K_BYTES = KBH * PTC_TK * 128;        // K tile  [128 keys][HS]
  constexpr int V_BYTES = 2 * HS * 128;              // V^T tile [HS rows][128 keys] as two k-blocks of 64 keys
  constexpr int P_BYTES = 2 * PTC_TQ * 128;          // P tile  [128 rows][128 keys] as two k-blocks of 64 keys
  constexpr int TMEM_COLS = 256;                     // S: 128 columns, P V: HS columns
  constexpr int HH = HS / 2;                         // head-dimension columns per thread
  extern __shared__ unsigned char ptc_raw[];
  __shared__ float s_max[2][PTC_TQ];                 // row maxima / sums of the two column halves
  unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(ptc_raw) + 1023) & ~(uintptr_t)1023);
  // P overwrites K: K is dead once S = Q K^T has completed, P is dead once P V has completed (both waited for), so the two
  // share one region (96 KB per CTA at head size 128)
  constexpr int KP_BYTES = K_BYTES > P_BYTES ? K_BYTES : P_BYTES;
  unsigned char* Qs = smem;
  unsigned char* Ks = Qs + Q_BYTES;                  // STAGED: two K buffers, then two V^T buffers
  unsigned char* Ps = Ks;
  unsigned char* Vt = Ks + (STAGED ? 2 : 1) * KP_BYTES;
  unsigned char* const K0 = Ks;
  unsigned char* const V0 = Vt;
  uint64_t* bar = reinterpret_cast<uint64_t*>(Vt + (STAGED ? 2 : 1) * V_BYTES);
  uint64_t* full = bar + 1;                          // STAGED: tile landed in buffer 0 / 1
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bar + 3);
  int* abort_flag = reinterpret_cast<int*>(tmem_slot + 1);

  // 8 warps: warp w serves TMEM lanes 32 (w & 3) .. (query rows) and column half w >> 2 -- two threads per query row, each
  // with half of the row's scores and half of its output accumulator (half the serial work per thread, twice the loaders)
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int qd = warp & 3, hf = warp >> 2;
  const int row = qd * 32 + lane;
  const int tile = blockIdx.x, h = blockIdx.y;
  if (tid == 0) {
    tc_mbar_init(bar, 1);
    tc_mbar_init(&full[0], 1); tc_mbar_init(&full[1], 1);
    *abort_flag = 0;
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    if (STAGED) {
      asm volatile("prefetch.tensormap [%0];" ::"l"(&mk) : "memory");
      asm volatile("prefetch.tensormap [%0];" ::"l"(&mv) : "memory");
    }
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tc_smem_u32(tmem_slot)), "r"(TMEM_COLS));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::);
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem_base = *tmem_slot;
  const uint32_t tmem_s = tmem_base, tmem_pv = tmem_base + 128;
  pdl_wait();
  pdl_trigger();

  const int row0 = p.tile_row0[tile], n = p.tile_n[tile];
  const int slot = p.row_slot[row0], pos0 = p.row_pos[row0];
  const bool live_tile = pos0 >= 0 && n > 0;
  const int kh = h / p.kv_mul;
  const float* kbase = (p.pg.table ? p.kc : p.kc + (size_t)slot * p.kv_batch_stride) + kh * HS;
  const float* vbase = (p.pg.table ? p.vc : p.vc + (size_t)slot * p.kv_batch_stride) + kh * HS;
  const int kmax = pos0 + n - 1;
  const float scale_log2 = 1.4426950408889634f / sqrtf((float)HS);   // scores in units of log2(e): exp(x) = exp2(x * log2 e)
  constexpr int CPL = (HS + 127) / 128;              // float4 per lane per row

  if (live_tile) {
    // ---- Q tile: warp w takes rows w, w+8, ...; a lane converts 4 consecutive values
    for (int r = warp; r < PTC_TQ; r += 8) {
      const int rr = r < n ? r : n - 1;
      const float* qrow = p.q + (size_t)(row0 + rr) * p.dim + h * HS;
#pragma unroll
      for (int cc = 0; cc < CPL; cc++) {
        const int c = 4 * lane + 128 * cc;
        if (c < HS) {
          const float4 v = *reinterpret_cast<const float4*>(qrow + c);
          uint2 w;
          w.x = ptc_bf16x2(v.x, v.y); w.y = ptc_bf16x2(v.z, v.w);
          *reinterpret_cast<uint2*>(Qs + (c >> 6) * (PTC_TQ * 128) + ptc_off(r, c & 63)) = w;
        }
      }
    }
  }
  float m_run = -INFINITY, l_run = 0.0f;             // l_run: this thread's half of the row sum (both halves share m)
  float acc[HH];
#pragma unroll
  for (int c = 0; c < HH; c++) acc[c] = 0.0f;
  const int qpos = pos0 + (row < n ? row : n - 1);
  unsigned phase = 0;
  bool ok = true;

  // STAGED: a tile = KBH boxes of K [128 keys x 64 values] + two boxes of V^T [HS x 64 keys], by prompt row
  const int key_row0 = STAGED ? p.row_stage[row0] - pos0 : 0;
  auto stage_tile = [&](int t0, int buf) {
    tc_mbar_expect_tx(&full[buf], K_BYTES + V_BYTES);
#pragma unroll
    for (int kb = 0; kb < KBH; kb++) tc_tma_2d(K0 + buf * KP_BYTES + kb * (PTC_TK * 128), &mk, kh * HS + kb * 64, key_row0 + t0, &full[buf]);
#pragma unroll
    for (int j = 0; j < 2; j++) tc_tma_2d(V0 + buf * V_BYTES + j * (HS * 128), &mv, key_row0 + t0 + 64 * j, kh * HS, &full[buf]);
  };
  if (STAGED && live_tile && tid == 0) stage_tile(0, 0);
  int it = 0;
  for (int t0 = 0; live_tile && t0 <= kmax; t0 += PTC_TK, it++) {
    __syncthreads();     // the previous tile's K / V^T / P are no longer read by an MMA (its completion was waited for)
    if (STAGED) {
      const int buf = it & 1;
      if (tid == 0 && t0 + PTC_TK <= kmax) stage_tile(t0 + PTC_TK, buf ^ 1);   // its buffer: the tile before this one, finished
      Ks = K0 + buf * KP_BYTES; Ps = Ks; Vt = V0 + buf * V_BYTES;
      if (ok) ok = tc_mbar_wait(&full[buf], (unsigned)(it >> 1) & 1u, abort_flag);
    } else {
    // ---- K tile (rows = keys): warp w takes rows w, w+8, ..; eight rows' loads are issued before the first is converted
#pragma unroll 1
    for (int rb = 0; rb < PTC_TK / 8; rb += 8) {
      float4 v[8][CPL];
#pragma unroll
      for (int u = 0; u < 8; u++) {
        const int t = t0 + warp + 8 * (rb + u);
        const int tr = t <= kmax ? t : kmax;
        const long long roff = p.pg.table ? kv_paged_off(p.pg, slot, tr, p.kv_dim) : (long long)tr * p.kv_dim;
#pragma unroll
        for (int cc = 0; cc < CPL; cc++) {
          const int c = 4 * lane + 128 * cc;
          v[u][cc] = (c < HS) ? ldg_stream(reinterpret_cast<const float4*>(kbase + roff + c)) : make_float4(0.f, 0.f, 0.f, 0.f);
        }
      }
#pragma unroll
      for (int u = 0; u < 8; u++) {
        const int r = warp + 8 * (rb + u);
        const bool live = t0 + r <= kmax;
#pragma unroll
        for (int cc = 0; cc < CPL; cc++) {
          const int c = 4 * lane + 128 * cc;
          if (c < HS) {
            uint2 w;
            w.x = live ? ptc_bf16x2(v[u][cc].x, v[u][cc].y) : 0u; w.y = live ? ptc_bf16x2(v[u][cc].z, v[u][cc].w) : 0u;
            *reinterpret_cast<uint2*>(Ks + (c >> 6) * (PTC_TK * 128) + ptc_off(r, c & 63)) = w;
          }
        }
      }
    }
    // ---- V tile, transposed: lane = key (consecutive lanes write consecutive bf16 of one V^T row), a warp takes every
    // eighth group of 4 head-dimension columns; two key blocks' loads are in flight together
    constexpr int VG = HS / 32;                      // column groups per warp
#pragma unroll 1
    for (int kb = 0; kb < PTC_TK / 32; kb += 2) {
      float4 v[2][VG];
      bool live[2];
#pragma unroll
      for (int u = 0; u < 2; u++) {
        const int t = t0 + (kb + u) * 32 + lane;
        live[u] = t <= kmax;
        const int tr = live[u] ? t : kmax;
        const long long roff = p.pg.table ? kv_paged_off(p.pg, slot, tr, p.kv_dim) : (long long)tr * p.kv_dim;
#pragma unroll
        for (int g = 0; g < VG; g++) v[u][g] = ldg_stream(reinterpret_cast<const float4*>(vbase + roff + 4 * warp + 32 * g));
      }
#pragma unroll
      for (int u = 0; u < 2; u++) {
        const int key = (kb + u) * 32 + lane;
        unsigned char* vtk = Vt + (key >> 6) * (HS * 128);
        const int kc = key & 63;
#pragma unroll
        for (int g = 0; g < VG; g++) {
          const int c = 4 * warp + 32 * g;
          const uint32_t w01 = live[u] ? ptc_bf16x2(v[u][g].x, v[u][g].y) : 0u, w23 = live[u] ? ptc_bf16x2(v[u][g].z, v[u][g].w) : 0u;
          *reinterpret_cast<uint16_t*>(vtk + ptc_off(c + 0, kc)) = (uint16_t)(w01 & 0xFFFFu);
          *reinterpret_cast<uint16_t*>(vtk + ptc_off(c + 1, kc)) = (uint16_t)(w01 >> 16);
          *reinterpret_cast<uint16_t*>(vtk + ptc_off(c + 2, kc)) = (uint16_t)(w23 & 0xFFFFu);
          *reinterpret_cast<uint16_t*>(vtk + ptc_off(c + 3, kc)) = (uint16_t)(w23 >> 16);
        }
      }
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // generic stores -> visible to the tensor core's reads
    }
    __syncthreads();
    // ---- S = Q K^T
    if (tid == 0) {
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      constexpr uint32_t idesc = ptc_idesc(PTC_TK);
#pragma unroll
      for (int kb = 0; kb < KBH; kb++) {
        const uint64_t a = tc_smem_desc(tc_smem_u32(Qs + kb * (PTC_TQ * 128))), b = tc_smem_desc(tc_smem_u32(Ks + kb * (PTC_TK * 128)));
#pragma unroll
        for (int k = 0; k < 4; k++) ptc_mma(tmem_s, a + (uint64_t)(2 * k), b + (uint64_t)(2 * k), idesc, (uint32_t)((kb | k) != 0));
      }
      tc_commit(bar);
    }
    if (ok) ok = tc_mbar_wait(bar, phase & 1u, abort_flag);
    phase++;
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    // ---- softmax of this thread's half row (keys 64 hf .. 64 hf + 63 of the tile): the maximum first (exchanged with the
    // other half through shared memory), then exponentials -> P (bf16) in shared memory
    const uint32_t ts = tmem_s + ((uint32_t)(qd * 32) << 16) + (uint32_t)(64 * hf);
    const int tk = t0 + 64 * hf;                     // first key of this thread's columns
    uint32_t s0[32], s1[32];
    ptc_tmem_ld32(ts, s0);
    ptc_tmem_ld32(ts + 32u, s1);
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
    float mt = -INFINITY;
#pragma unroll
    for (int j = 0; j < 32; j++) {
      if (tk + j <= qpos) mt = fmaxf(mt, __uint_as_float(s0[j]) * scale_log2);
      if (tk + 32 + j <= qpos) mt = fmaxf(mt, __uint_as_float(s1[j]) * scale_log2);
    }
    s_max[hf][row] = mt;
    __syncthreads();
    const float m_new = fmaxf(m_run, fmaxf(s_max[0][row], s_max[1][row]));   // finite: key 0 is visible to every row
    const float corr = exp2f(m_run - m_new);         // exp2(-inf) = 0 on the first tile
    float lsum = 0.0f;
    unsigned char* prow = Ps + hf * (PTC_TQ * 128);  // this half's 64 keys = one k-block of P
#pragma unroll
    for (int half = 0; half < 2; half++) {
      float pr[32];
#pragma unroll
      for (int j = 0; j < 32; j++) {
        const float sv = __uint_as_float(half ? s1[j] : s0[j]);
        pr[j] = (tk + 32 * half + j <= qpos) ? exp2f(sv * scale_log2 - m_new) : 0.0f;
        lsum += pr[j];
      }
#pragma unroll
      for (int j = 0; j < 32; j += 8) {
        uint4 w;
        w.x = ptc_bf16x2(pr[j], pr[j + 1]); w.y = ptc_bf16x2(pr[j + 2], pr[j + 3]);
        w.z = ptc_bf16x2(pr[j + 4], pr[j + 5]); w.w = ptc_bf16x2(pr[j + 6], pr[j + 7]);
        *reinterpret_cast<uint4*>(prow + ptc_off(row, 32 * half + j)) = w;
      }
    }
    l_run = l_run * corr + lsum;
    m_run = m_new;
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncthreads();
    // ---- P V
    if (tid == 0) {
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      constexpr uint32_t idesc = ptc_idesc(HS);
#pragma unroll
      for (int kb = 0; kb < 2; kb++) {
        const uint64_t a = tc_smem_desc(tc_smem_u32(Ps + kb * (PTC_TQ * 128))), b = tc_smem_desc(tc_smem_u32(Vt + kb * (HS * 128)));
#pragma unroll
        for (int k = 0; k < 4; k++) ptc_mma(tmem_pv, a + (uint64_t)(2 * k), b + (uint64_t)(2 * k), idesc, (uint32_t)((kb | k) != 0));
      }
      tc_commit(bar);
    }
    if (ok) ok = tc_mbar_wait(bar, phase & 1u, abort_flag);
    phase++;
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tpv = tmem_pv + ((uint32_t)(qd * 32) << 16) + (uint32_t)(HH * hf);
#pragma unroll
    for (int c0 = 0; c0 < HH; c0 += 32) {
      uint32_t v[32];
      ptc_tmem_ld32(tpv + (uint32_t)c0, v);
      asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
      for (int j = 0; j < 32; j++) acc[c0 + j] = acc[c0 + j] * corr + __uint_as_float(v[j]);
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  }

  // ---- the row sum is the sum of its two halves; normalise and store this thread's half row
  __syncthreads();
  s_max[hf][row] = l_run;
  __syncthreads();
  if (live_tile && row < n && ok) {
    const float inv = 1.0f / (s_max[0][row] + s_max[1][row]);
    float* orow = p.out + (size_t)(row0 + row) * p.dim + h * HS + HH * hf;
#pragma unroll
    for (int c = 0; c < HH; c += 4) {
      const float4 o = make_float4(acc[c] * inv, acc[c + 1] * inv, acc[c + 2] * inv, acc[c + 3] * inv);
      *reinterpret_cast<float4*>(orow + c) = o;
      if (p.out_b16) {
        uint2 w;
        w.x = ptc_bf16x2(o.x, o.y); w.y = ptc_bf16x2(o.z, o.w);
        *reinterpret_cast<uint2*>(p.out_b16 + (size_t)(row0 + row) * p.dim + h * HS + HH * hf + c) = w;
      }
    }
  }
  __syncthreads();
  if (warp == 0) {
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(TMEM_COLS));
  }
  if (tid == 0 && *abort_flag && p.error_flag) *((volatile int*)p.error_flag) = 1;
}

bool attn_prefill_tc_supported(int head_size, int kv_dim) { return (head_size == 64 || head_size == 128) && kv_dim % 4 == 0; }

// tiles of up to 128 query rows (tile_row0 / tile_n), never across sequences
// k16 / vt16 (nullable): the staged bf16 copies of this call's K rows [staged_rows][kv_dim] and V rows, transposed
// [kv_dim][staged_ld] (every sequence of the call starts at position 0; step_bf16.cu decides)
int attn_prefill_tc_launch(cudaStream_t stream, int n_tiles, int n_heads, int n_kv_heads, int head_size, const float* q, float* out,
                           uint16_t* out_b16, const float* kc, const float* vc, long long kv_batch_stride, const int* row_slot,
                           const int* row_pos, const int* tile_row0, const int* tile_n, bool pdl, const KvPaged* pg,
                           const uint16_t* k16, const uint16_t* vt16, int staged_rows, int staged_ld, const int* row_stage) {
  if (!attn_prefill_tc_supported(head_size, n_kv_heads * head_size)) return THA_STATUS_NOT_SUPPORTED;
  AttnPfTcP p;
  p.staged_rows = staged_rows; p.row_stage = row_stage;
  p.q = q; p.out = out; p.out_b16 = out_b16; p.kc = kc; p.vc = vc;
  if (pg) p.pg = *pg; else p.pg = KvPaged{};
  p.kv_batch_stride = kv_batch_stride; p.kv_dim = n_kv_heads * head_size; p.kv_mul = n_heads / n_kv_heads;
  p.dim = n_heads * head_size;
  p.row_slot = row_slot; p.row_pos = row_pos; p.tile_row0 = tile_row0; p.tile_n = tile_n;
  THA_TRY_RC(gemm_tc_init(&p.error_flag));
  const int kbh = head_size / 64;
  const size_t kp = (size_t)kbh * 128 * 128 > (size_t)2 * 128 * 128 ? (size_t)kbh * 128 * 128 : (size_t)2 * 128 * 128;
  const bool staged = k16 && vt16 && staged_rows > 0 && row_stage;
  const int nb = staged ? 2 : 1;
  const size_t smem = (size_t)kbh * 128 * 128 + nb * kp + (size_t)nb * 2 * head_size * 128 + 1024 + 64;
  CUtensorMap mk, mv;
  if (staged) {
    const int kvd = n_kv_heads * head_size;
    THA_TRY_RC(tc_map(k16, staged_rows, kvd, kvd, 128, &mk, 2));
    THA_TRY_RC(tc_map(vt16, kvd, staged_ld, staged_ld, head_size, &mv, 2));
  } else {
    memset(&mk, 0, sizeof(mk)); memset(&mv, 0, sizeof(mv));
  }
  void (*k)(const CUtensorMap, const CUtensorMap, const AttnPfTcP) =
      staged ? (head_size == 64 ? attn_prefill_tc_kernel<64, true> : attn_prefill_tc_kernel<128, true>)
             : (head_size == 64 ? attn_prefill_tc_kernel<64, false> : attn_prefill_tc_kernel<128, false>);
  THA_CUDA_OK(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  THA_CUDA_OK(launch(k, dim3(n_tiles, n_heads), dim3(256), smem, stream, pdl, mk, mv, p));
  return THA_STATUS_SUCCESS;
}

}  // namespace tha
