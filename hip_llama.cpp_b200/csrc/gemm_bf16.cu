// bf16 weight matmul on the tensor cores: tcgen05.mma.kind::f16 (bf16 x bf16 -> fp32 in TMEM) with BOTH operands read
// straight from TMA-written shared memory -- no conversion warps, the weights cross HBM and shared memory as the two
// bytes per value they are stored in.
//
//   out[n, row] = sum_k W[row, k] * X[n, k]      W: [M, K] row-major bf16, X: [N, K] row-major bf16 (rows = slots or
//                                                prompt positions, already rounded to bf16 by the producing kernel)
//
// This is the matmul of the bf16 model path (BASELINE.json configs[3] "fp32/bf16"; the reference declares the 16-bit
// weight structs at include/models.hpp:77-117 and never fills them) for more than 8 rows and for prompt ingestion, where
// the fp32 path's 3xTF32 scheme spends three tensor-core passes and a converter stage per useful one.
//
// Structure (persistent: 148 CTAs walk (weight tile, column tile, K split) items with running ring counters):
//   warp 0   TMA producer: per k-block of 64 values (one 128-byte swizzle row) the W tile [128 x 64] and the X tile
//            [NT x 64] into a ring of up to 8 stages; the first ring of W tiles goes out before griddepcontrol.wait
//   warp 1   MMA issuer: 4 x (M 128, N NT, K 16) per stage; tcgen05.commit frees the stage; the accumulator is DOUBLE
//            buffered in TMEM (2 x NT columns), so the next item's MMAs run while the epilogue drains the previous one
//   warp 2   TMEM allocation
//   warps 4-7  epilogue: thread = weight row (TMEM lane), 16 columns per tcgen05.ld, optional column scale, coalesced
//            stores (a warp writes 32 consecutive rows of one column); split-K partials to out + z * zstride
// Up to three matrices with the same K are stacked in one launch (wq|wk|wv, w1|w3).
#include <stdlib.h>

#include "gemm_tc.h"
#include "tc_common.cuh"

namespace tha {

constexpr int BF_THREADS = 256;
constexpr int BF_BK = 64;             // bf16 values per k-block: 128 bytes

struct BfP {
  int M[3]; int tiles[3]; int out_off[3];
  int K, N, ny, kblocks_total, kblocks_per_split;
  float* out; long long ldo; long long zstride;
  const float* colscale;
  int* error_flag;
  int splits, n_items;
  uint16_t* swiglu_out; long long ld_swiglu;   // SWIGLU kernels
};

// MT = weight tiles of 128 rows per work item.  MT = 2 (prompt-sized N only): a CTA multiplies 256 weight rows by the
// same 256 activation columns -- per 64-value k-block it pulls 32 + 32 KB through L2 for twice the flops of the 16 + 32 KB
// of a 128-row tile.  With 128 x 256 tiles the prompt matmuls were bound by L2 -> shared-memory traffic (768 MB per
// 4096 x 4096 x 2048 matmul, ~48 % of the sustained tensor peak); the two accumulators then fill all 512 TMEM columns, so
// the read-out of an item is no longer hidden behind the next item's MMAs (eight warps keep it at a few per cent).
template <int NT, int MT = 1>
struct BfSmem {
  static constexpr int W_BYTES = MT * TC_BM * 128, X_BYTES = NT * 128, STAGE = W_BYTES + X_BYTES;
  static constexpr int NST_RAW = (TC_SMEM_LIMIT - 2048) / STAGE;
  static constexpr int NST = NST_RAW > 8 ? 8 : NST_RAW;
  static constexpr int NBUF = (2 * MT * NT <= 512) ? 2 : 1;           // accumulator buffers of MT x NT columns
  static constexpr int TMEM_COLS = NBUF * MT * NT < 32 ? 32 : NBUF * MT * NT;   // powers of two: NT in {16, 32, 64, 128, 256}
  static constexpr int TOTAL = NST * STAGE + 1024 + (2 * NST + 4) * 8 + 64;
  static_assert(NST >= 3, "ring too short");
  static_assert(TOTAL <= TC_SMEM_LIMIT, "shared memory overflow");
};

// kind::f16 with bf16 operands, fp32 accumulate, both K-major, M = 128, N = nt (cute::UMMA::InstrDescriptor)
__host__ __device__ constexpr uint32_t bf_idesc(int nt) {
  return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(nt >> 3) << 17) | ((uint32_t)(TC_BM >> 4) << 24);
}
__device__ __forceinline__ void bf_mma(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}

// SWIGLU (MT = 2): the two 128-row halves of a work item are the SAME rows of two matrices (mw0 = w1, mw1 = w3); the read-out
// combines the two accumulators of a row into silu(a) * c (run.cc:344-347 order) and stores it as bf16.
template <int NT, int MT = 1, bool SWIGLU = false>
__global__ void __launch_bounds__(128 + 128 * MT, 1)
gemm_bf16_kernel(const __grid_constant__ CUtensorMap mw0, const __grid_constant__ CUtensorMap mw1,
                 const __grid_constant__ CUtensorMap mw2, const __grid_constant__ CUtensorMap mx, const BfP p) {
  using S = BfSmem<NT, MT>;
  constexpr int NBUF = S::NBUF;
  extern __shared__ unsigned char tc_smem_raw[];
  unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(tc_smem_raw) + 1023) & ~(uintptr_t)1023);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + S::NST * S::STAGE);
  uint64_t* full = bars;                 // stage landed (W + X bytes)
  uint64_t* empty = full + S::NST;       // the MMAs that read the stage have completed
  uint64_t* accum = empty + S::NST;      // [2] accumulator buffer complete
  uint64_t* accfree = accum + 2;         // [2] ... and read out by the four epilogue warps
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(accfree + 2);
  int* abort_flag = reinterpret_cast<int*>(tmem_slot + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int tiles01 = p.tiles[0] + p.tiles[1];
  struct Item { int mat, row0, n0, kb0, nkb; };
  auto item_of = [&](int it) -> Item {
    Item w;
    const int z = it % p.splits;
    const int t = it / p.splits;
    int tile = t / p.ny;
    w.n0 = (t - tile * p.ny) * NT;
    w.mat = 0;
    if (tile >= p.tiles[0]) { w.mat = 1; if (tile >= tiles01) { w.mat = 2; tile -= tiles01; } else tile -= p.tiles[0]; }
    w.row0 = tile * (SWIGLU ? TC_BM : TC_BM * MT);
    w.kb0 = z * p.kblocks_per_split;
    w.nkb = p.kblocks_total - w.kb0;
    if (w.nkb > p.kblocks_per_split) w.nkb = p.kblocks_per_split;
    return w;
  };

  if (threadIdx.x == 0) {
    for (int s = 0; s < S::NST; s++) { tc_mbar_init(&full[s], 1); tc_mbar_init(&empty[s], 1); }
    for (int b = 0; b < 2; b++) { tc_mbar_init(&accum[b], 1); tc_mbar_init(&accfree[b], 4 * MT); }
    *abort_flag = 0;
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&mw0) : "memory");
    if (p.tiles[1]) asm volatile("prefetch.tensormap [%0];" ::"l"(&mw1) : "memory");
    if (p.tiles[2]) asm volatile("prefetch.tensormap [%0];" ::"l"(&mw2) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&mx) : "memory");
  }
  if (warp == 2) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tc_smem_u32(tmem_slot)), "r"(S::TMEM_COLS));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::);
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ===== TMA producer
    if (lane == 0) {
      uint64_t pol_w;
      asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol_w));
      const bool stream_w = p.ny == 1;            // every weight tile is read once: do not let it displace the activations in L2
      int gb = 0;
      bool ok = true, first = true;
      auto load_w = [&](unsigned char* dst, const CUtensorMap* mw, int kcol, int row0, uint64_t* bar) {
        if (SWIGLU) {   // 128 rows of w1, then the same 128 rows of w3
          tc_tma_2d(dst, &mw0, kcol, row0, bar);
          tc_tma_2d(dst + TC_BM * 128, &mw1, kcol, row0, bar);
        } else if (stream_w) tc_tma_2d_hint(dst, mw, kcol, row0, bar, pol_w);
        else tc_tma_2d(dst, mw, kcol, row0, bar);
      };
      for (int it = blockIdx.x; it < p.n_items && ok; it += gridDim.x) {
        const Item w = item_of(it);
        const CUtensorMap* mw = (w.mat == 0) ? &mw0 : (w.mat == 1 ? &mw1 : &mw2);
        int i = 0;
        if (first) {
          // weights never depend on the previous kernel: the first ring goes out before the dependency wait (the bytes may
          // land before the matching expect_tx -- the phase cannot complete until that arrival)
          const int pre = w.nkb < S::NST ? w.nkb : S::NST;
          for (; i < pre; i++) load_w(smem + i * S::STAGE, mw, (w.kb0 + i) * BF_BK, w.row0, &full[i]);
          pdl_wait();
          for (int j = 0; j < pre; j++) {
            tc_mbar_expect_tx(&full[j], S::STAGE);
            tc_tma_2d(smem + j * S::STAGE + S::W_BYTES, &mx, (w.kb0 + j) * BF_BK, w.n0, &full[j]);
          }
          first = false;
        }
        for (; i < w.nkb; i++) {
          const int g = gb + i, s = g % S::NST;
          if (g >= S::NST && !tc_mbar_wait(&empty[s], ((g / S::NST) & 1u) ^ 1u, abort_flag)) { ok = false; break; }
          tc_mbar_expect_tx(&full[s], S::STAGE);
          load_w(smem + s * S::STAGE, mw, (w.kb0 + i) * BF_BK, w.row0, &full[s]);
          tc_tma_2d(smem + s * S::STAGE + S::W_BYTES, &mx, (w.kb0 + i) * BF_BK, w.n0, &full[s]);
        }
        gb += w.nkb;
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer
    if (lane == 0) {
      constexpr uint32_t idesc = bf_idesc(NT);
      int gb = 0, n_it = 0;
      bool ok = true;
      for (int it = blockIdx.x; it < p.n_items && ok; it += gridDim.x, n_it++) {
        const Item w = item_of(it);
        const int buf = n_it % NBUF, use = n_it / NBUF;   // use-th time this buffer is filled
        if (use >= 1) {                            // the buffer's previous accumulator must have been read out
          if (!tc_mbar_wait(&accfree[buf], (unsigned)(use & 1) ^ 1u, abort_flag)) break;
          asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        }
        const uint32_t d = tmem_base + (uint32_t)(buf * MT * NT);
        for (int i = 0; i < w.nkb; i++) {
          const int g = gb + i, s = g % S::NST;
          if (!tc_mbar_wait(&full[s], (g / S::NST) & 1u, abort_flag)) { ok = false; break; }
          asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
          const uint64_t adesc = tc_smem_desc(tc_smem_u32(smem + s * S::STAGE));
          const uint64_t bdesc = tc_smem_desc(tc_smem_u32(smem + s * S::STAGE + S::W_BYTES));
#pragma unroll
          for (int k = 0; k < BF_BK / 16; k++)     // UMMA K = 16 bf16 = 32 bytes = 2 descriptor units
#pragma unroll
            for (int m = 0; m < MT; m++)           // the m-th 128 rows of the W stage: 16 KB = 1024 descriptor units further
              bf_mma(d + (uint32_t)(m * NT), adesc + (uint64_t)(1024 * m + 2 * k), bdesc + (uint64_t)(2 * k), idesc, (uint32_t)((i | k) != 0));
          tc_commit(&empty[s]);
        }
        if (ok) tc_commit(&accum[buf]);
        gb += w.nkb;
      }
    }
  } else if (warp >= 4) {
    // ===== epilogue: thread = weight row of the tile (TMEM lane); warps 4-7 read the first 128 rows, 8-11 the second (MT = 2)
    const int q = warp & 3, mt = (warp - 4) >> 2;
    const int r = q * 32 + lane;
    pdl_trigger();
    pdl_wait();                                    // the column scales come from the previous kernel
    int n_it = 0;
    bool ok = true;
    for (int it = blockIdx.x; it < p.n_items; it += gridDim.x, n_it++) {
      const Item w = item_of(it);
      const int buf = n_it % NBUF, use = n_it / NBUF;
      if (ok) ok = tc_mbar_wait(&accum[buf], (unsigned)(use & 1), abort_flag);
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      if (SWIGLU) {
        // both accumulators of this thread's row; warps 4-7 take the first half of the columns, 8-11 the second
        const int srow = w.row0 + r;
        const uint32_t ta = tmem_base + (uint32_t)(buf * MT * NT) + ((uint32_t)(q * 32) << 16);
        if (ok && srow < p.M[0]) {
          uint16_t* o = p.swiglu_out + (size_t)(w.n0 + mt * (NT / 2)) * p.ld_swiglu + srow;
#pragma unroll 1
          for (int c = mt * (NT / 2); c < (mt + 1) * (NT / 2); c += 16) {
            uint32_t va[16], vc[16];
            tc_tmem_ld16(ta + (uint32_t)c, va);
            tc_tmem_ld16(ta + (uint32_t)(NT + c), vc);
            asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
            for (int j = 0; j < 16; j++) {
              if (w.n0 + c + j < p.N) {
                float h = __uint_as_float(va[j]);
                h *= (1.0f / (1.0f + expf(-h)));
                h *= __uint_as_float(vc[j]);
                const uint32_t u = __float_as_uint(h);   // round to nearest even, NaN stays NaN (step_bf16.cu f32_to_bf16)
                *o = ((u & 0x7FFFFFFFu) > 0x7F800000u) ? (uint16_t)((u >> 16) | 0x40u) : (uint16_t)((u + 0x7FFFu + ((u >> 16) & 1u)) >> 16);
              }
              o += p.ld_swiglu;
            }
          }
        }
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
        __syncwarp();
        if (lane == 0) tc_mbar_arrive(&accfree[buf]);
        continue;
      }
      const int row = w.row0 + mt * TC_BM + r;
      const bool row_ok = row < p.M[w.mat];
      float* outz = p.out + (size_t)(it % p.splits) * p.zstride + p.out_off[w.mat] + row;
      const uint32_t tq = tmem_base + (uint32_t)(buf * MT * NT + mt * NT) + ((uint32_t)(q * 32) << 16);
      // whole tile inside the output: no per-element bounds checks, the store address advanced by one add per column
      // (the checked form costs ~25 instructions per element: gemm_tc.cu has the measurement)
      const bool full = row_ok && (w.n0 + NT <= p.N);
      const bool cs_vec = p.colscale && ((reinterpret_cast<uintptr_t>(p.colscale) & 15) == 0);
      if (ok) {
#pragma unroll 1
        for (int c = 0; c < NT; c += 32) {          // two 16-column loads in flight
          uint32_t v[32];
          tc_tmem_ld16(tq + (uint32_t)c, *reinterpret_cast<uint32_t(*)[16]>(v));
          if (NT > 16) tc_tmem_ld16(tq + (uint32_t)(c + 16), *reinterpret_cast<uint32_t(*)[16]>(v + 16));
          constexpr int NC = NT > 16 ? 32 : 16;
          float sc[NC];
          if (full && p.colscale) {
#pragma unroll
            for (int j = 0; j < NC; j += 4) {
              if (cs_vec) {
                const float4 t = __ldg(reinterpret_cast<const float4*>(p.colscale + w.n0 + c + j));
                sc[j] = t.x; sc[j + 1] = t.y; sc[j + 2] = t.z; sc[j + 3] = t.w;
              } else {
#pragma unroll
                for (int e = 0; e < 4; e++) sc[j + e] = __ldg(p.colscale + w.n0 + c + j + e);
              }
            }
          }
          asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
          if (full) {
            float* o = outz + (size_t)(w.n0 + c) * p.ldo;
#pragma unroll
            for (int j = 0; j < NC; j++) {
              float x = __uint_as_float(v[j]);
              if (p.colscale) x *= sc[j];
              *o = x;
              o += p.ldo;
            }
          } else {
#pragma unroll
            for (int j = 0; j < NC; j++) {
              const int n = w.n0 + c + j;
              if (n < p.N && row_ok) {
                float x = __uint_as_float(v[j]);
                if (p.colscale) x *= p.colscale[n];
                outz[(size_t)n * p.ldo] = x;
              }
            }
          }
        }
      }
      asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
      __syncwarp();
      if (lane == 0) tc_mbar_arrive(&accfree[buf]);
    }
  }
  __syncthreads();
  if (warp == 2) {
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(S::TMEM_COLS));
  }
  if (threadIdx.x == 0 && *abort_flag && p.error_flag) *((volatile int*)p.error_flag) = 1;
}

// ------------------------------------------------------------------ host side ----
bool gemm_bf16_ok(const BfArgs& a) {
  if (a.K % 8 || a.ldx % 8 || a.K < BF_BK || a.N < 1 || a.n_mats < 1 || a.n_mats > 3 || !a.X || !a.out) return false;
  if (reinterpret_cast<uintptr_t>(a.X) & 15u) return false;
  for (int i = 0; i < a.n_mats; i++)
    if (!a.W[i] || (reinterpret_cast<uintptr_t>(a.W[i]) & 15u) || a.M[i] < 1) return false;
  return true;
}

int gemm_bf16_pick_nt(int N) { return N <= 16 ? 16 : (N <= 32 ? 32 : (N <= 64 ? 64 : (N <= 128 ? 128 : 256))); }

template <int NT, int MT = 1, bool SWIGLU = false>
static int bf_launch_nt(cudaStream_t st, const CUtensorMap* mw, const CUtensorMap& mx, const BfP& p, dim3 grid, bool pdl) {
  using S = BfSmem<NT, MT>;
  auto k = gemm_bf16_kernel<NT, MT, SWIGLU>;
  THA_CUDA_OK(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, S::TOTAL));
  THA_CUDA_OK(launch(k, grid, dim3(128 + 128 * MT), (size_t)S::TOTAL, st, pdl, mw[0], mw[1], mw[2], mx, p));
  return THA_STATUS_SUCCESS;
}
// 256-row work items for prompt-sized N (THA_BF16_MT=1 keeps 128-row items)
static int bf_mt(int NT, int N) {
  static int forced = -1;
  if (forced < 0) { const char* v = getenv("THA_BF16_MT"); forced = (v && *v) ? atoi(v) : 0; }
  if (forced == 1 || forced == 2) return NT == 256 ? forced : 1;
  return (NT == 256 && N >= 512) ? 2 : 1;
}

bool gemm_bf16_swiglu_ok(const BfArgs& a) {
  const int NT = gemm_bf16_pick_nt(a.N);
  return a.swiglu_out && a.n_mats == 2 && a.M[0] == a.M[1] && a.splits <= 1 && NT == 256 && bf_mt(NT, a.N) == 2 && !a.colscale &&
         (a.ld_swiglu % 1) == 0;
}

int gemm_bf16_launch(cudaStream_t st, const BfArgs& a, bool pdl) {
  if (!gemm_bf16_ok(a)) return THA_STATUS_INVALID_ARG;
  const int NT = gemm_bf16_pick_nt(a.N);
  const bool swiglu = a.swiglu_out != nullptr;
  if (swiglu && !gemm_bf16_swiglu_ok(a)) return THA_STATUS_NOT_SUPPORTED;
  const int MT = swiglu ? 1 : bf_mt(NT, a.N);   // SwiGLU: 128-row boxes of each of the two matrices (the kernel itself runs MT = 2)
  CUtensorMap mw[3], mx;
  BfP p = {};
  int tiles = 0;
  for (int i = 0; i < 3; i++) {
    if (i < a.n_mats) {
      THA_TRY_RC(tc_map(a.W[i], a.M[i], a.K, a.K, TC_BM * MT, &mw[i], 2));
      p.M[i] = a.M[i]; p.tiles[i] = ceil_div(a.M[i], TC_BM * MT); p.out_off[i] = a.out_off[i];
    } else {
      mw[i] = mw[0];
    }
    tiles += p.tiles[i];
  }
  if (swiglu) { tiles = p.tiles[0]; p.tiles[1] = 0; p.swiglu_out = a.swiglu_out; p.ld_swiglu = a.ld_swiglu; }   // an item = a pair of tiles
  THA_TRY_RC(tc_map(a.X, a.N, a.K, a.ldx, NT, &mx, 2));
  p.K = a.K; p.N = a.N;
  p.kblocks_total = ceil_div(a.K, BF_BK);
  const int splits = a.splits < 1 ? 1 : a.splits;
  p.kblocks_per_split = ceil_div(p.kblocks_total, splits);
  if (ceil_div(p.kblocks_total, p.kblocks_per_split) != splits) return THA_STATUS_INVALID_ARG;
  p.out = a.out; p.ldo = a.ldo; p.zstride = a.zstride; p.colscale = a.colscale;
  THA_TRY_RC(gemm_tc_init(&p.error_flag));
  p.ny = ceil_div(a.N, NT);
  p.splits = splits;
  p.n_items = tiles * p.ny * splits;
  int dev = 0, sms = 148;
  if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  dim3 grid(p.n_items < sms ? p.n_items : sms, 1, 1);
  switch (NT) {
    case 16: return bf_launch_nt<16>(st, mw, mx, p, grid, pdl);
    case 32: return bf_launch_nt<32>(st, mw, mx, p, grid, pdl);
    case 64: return bf_launch_nt<64>(st, mw, mx, p, grid, pdl);
    case 128: return bf_launch_nt<128>(st, mw, mx, p, grid, pdl);
    default: return swiglu ? bf_launch_nt<256, 2, true>(st, mw, mx, p, grid, pdl)
                           : (MT == 2 ? bf_launch_nt<256, 2>(st, mw, mx, p, grid, pdl) : bf_launch_nt<256>(st, mw, mx, p, grid, pdl));
  }
}

}  // namespace tha

using namespace tha;

// C-ABI: out[n*ldo + row] = sum_k W[row,k] X[n,k], W and X bf16 (raw 16-bit patterns), fp32 out
extern "C" int tha_gemm_bf16(void* stream, float* out, long long ldo, const uint16_t* W, int M, int K, const uint16_t* X, int N,
                             long long ldx) {
  if (!out || !W || !X || M < 1 || N < 1 || K < 1) return THA_STATUS_INVALID_ARG;
  BfArgs a = {};
  a.n_mats = 1; a.W[0] = W; a.M[0] = M; a.out_off[0] = 0; a.K = K; a.X = X; a.N = N; a.ldx = ldx;
  a.out = out; a.ldo = ldo; a.zstride = 0; a.splits = 1; a.colscale = nullptr;
  if (!gemm_bf16_ok(a)) return THA_STATUS_NOT_SUPPORTED;
  return gemm_bf16_launch((cudaStream_t)stream, a, false);
}
