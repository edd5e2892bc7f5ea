// Weight matmul on the 5th-generation tensor cores (tcgen05 + TMEM + TMA), fp32 in / fp32 out.
//
//   out[n, row] = sum_k W[row, k] * X[n, k]          W: [M, K] row-major fp32 (the checkpoint layout)
//                                                    X: [N, K] row-major fp32 (N = batch slots or prompt tokens)
//
// Used where the GEMV family stops being HBM-bound on CUDA cores: multi-sequence decode with more
// than 8 slots and batched prefill (replaces thaBLAS_s_matmul_batch,
// src/thaBLAS.cpp:191-228, whose grid re-reads every weight row once per slot).
//
// Precision: tcgen05 has no fp32 kind.  kind::tf32 reads 32-bit operands and uses 10 mantissa
// bits, which alone would put logits ~1e-3 off the CPU reference.  The kernel therefore runs the
// error-compensated 3xTF32 scheme: every operand v is split in shared memory into
// hi = tf32(v) (round to nearest) and lo = v - hi, and each k-block issues
//   D += Whi.Xhi + Wlo.Xhi + Whi.Xlo        (the lo.lo term is below fp32 rounding)
// with fp32 accumulation in TMEM.  Results agree with an fp32 FMA loop to ~1e-6 relative.
//
// Structure (one 128-row weight tile x NT columns per CTA, optional split-K over blockIdx.z):
//   warp 0      TMA producer: per k-block (32 floats = one 128-byte swizzle row) one
//               cp.async.bulk.tensor for the W tile [128 x 32] and one for the X tile [NT x 32],
//               SWIZZLE_128B, into a ring of stages; completion on the stage's `full` mbarrier.
//   warps 4-15  converters (3 groups of 4 warps, k-blocks round-robin): thread = weight row; read the
//               landed fp32 row (swizzle-aware, conflict-free), split into hi = tf32(w), lo = w - hi
//               and store both into the TMEM operand ring (tcgen05.st); arrive on `ready`.
//   warp 1      MMA issuer: one elected lane issues 4 (K=8 each) x 3 tcgen05.mma.kind::tf32 per
//               k-block, accumulator [128 lanes x NT columns] fp32 in TMEM; tcgen05.commit frees
//               the stage (`empty`) and finally signals the epilogue (`accum`).
//   warp 2      TMEM allocation / release.
//   warps 4-7   epilogue: tcgen05.ld (32 lanes x 32-bit x 16 columns per instruction), optional
//               per-column scale (the 1/rms factor of a folded RMSNorm), coalesced stores
//               out[n * ldo + off + row] (a warp writes 32 consecutive rows of one column).
// Up to three weight matrices with the same K are stacked in one launch (wq|wk|wv, w1|w3); a tile
// never straddles two matrices.
#include <cuda.h>
#include <stdlib.h>

#include <map>
#include <mutex>
#include <tuple>

#include "common.cuh"
#include "gemm_tc.h"
#include "tc_common.cuh"
#include "ktrace.cuh"
#include "workspace.h"

namespace tha {

// Persistent: a CTA walks the work items (weight tile, column tile, K split) blockIdx.x, blockIdx.x + gridDim.x, ...
// with running stage counters, so the weight ring keeps streaming while the previous item's accumulator is read out.
template <int NT>
__global__ void __launch_bounds__(TC_THREADS, 1)
gemm_tf32x3_kernel(const __grid_constant__ CUtensorMap mw0, const __grid_constant__ CUtensorMap mw1,
                   const __grid_constant__ CUtensorMap mw2, const __grid_constant__ CUtensorMap mxh,
                   const __grid_constant__ CUtensorMap mxl, const TcP p) {
  using S = TcSmem<NT>;
  extern __shared__ unsigned char tc_smem_raw[];
  unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(tc_smem_raw) + 1023) & ~(uintptr_t)1023);
  unsigned char* wring = smem;
  unsigned char* xring = smem + S::SW * S::W_STAGE;
  uint64_t* bars = reinterpret_cast<uint64_t*>(xring + S::SX * S::XS_STAGE);
  uint64_t* wfull = bars;                     // W stage landed
  uint64_t* wfree = wfull + S::SW;            // its rows are in registers: the slot may be refilled
  uint64_t* ready = wfree + S::SW;            // operand slot full: W hi/lo in TMEM (4 warps) + X hi/lo landed (TMA bytes)
  uint64_t* aempty = ready + S::SA;           // MMAs that read the operand slot have completed
  uint64_t* accum = aempty + S::SA;           // accumulator of the current item complete
  uint64_t* accfree = accum + 1;              // ... and read out by the 12 epilogue warps: the next item may overwrite it
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(accfree + 1);
  int* abort_flag = reinterpret_cast<int*>(tmem_slot + 1);
  __shared__ int kt_s;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int tiles01 = p.tiles[0] + p.tiles[1];
  struct Item { int mat, row0, n0, kb0, nkb, nst; };
  auto item_of = [&](int it) -> Item {
    Item w;
    const int z = it % p.splits;
    const int t = it / p.splits;
    int tile = t / p.ny;
    w.n0 = (t - tile * p.ny) * NT;
    w.mat = 0;
    if (tile >= p.tiles[0]) { w.mat = 1; if (tile >= tiles01) { w.mat = 2; tile -= tiles01; } else tile -= p.tiles[0]; }
    w.row0 = tile * TC_BM;
    w.kb0 = z * p.kblocks_per_split;
    w.nkb = p.kblocks_total - w.kb0;
    if (w.nkb > p.kblocks_per_split) w.nkb = p.kblocks_per_split;
    w.nst = (w.nkb + S::KG - 1) / S::KG;       // pipeline stages of KG k-blocks (the last one may be short)
    return w;
  };

  if (threadIdx.x == 0) {
    for (int s = 0; s < S::SW; s++) { tc_mbar_init(&wfull[s], 1); tc_mbar_init(&wfree[s], 4); }
    for (int s = 0; s < S::SA; s++) { tc_mbar_init(&ready[s], 5); tc_mbar_init(&aempty[s], 1); }
    tc_mbar_init(accum, 1);
    tc_mbar_init(accfree, 4 * TC_CGROUPS);
    *abort_flag = 0;
    kt_s = kt_begin((unsigned long long)KT_GEMM | ((unsigned long long)p.splits << 16) | ((unsigned long long)p.n_items << 32));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&mw0) : "memory");
    if (p.tiles[1]) asm volatile("prefetch.tensormap [%0];" ::"l"(&mw1) : "memory");
    if (p.tiles[2]) asm volatile("prefetch.tensormap [%0];" ::"l"(&mw2) : "memory");
  }
  if (warp == 3 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&mxh) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&mxl) : "memory");
  }
  if (warp == 2) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tc_smem_u32(tmem_slot)), "r"(S::TMEM_COLS));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::);
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem_base = *tmem_slot;
  const uint32_t tmem_d = tmem_base;                               // main accumulator (hi.hi terms)
  const uint32_t tmem_dl = tmem_base + (S::SPLIT_ACC ? NT : 0);    // lo-term accumulator
  const uint32_t tmem_a = tmem_base + S::ACC_COLS;                 // W operand ring
  const int kt = kt_s;

  if (warp == 0) {
    // ===== W producer: weights never depend on the previous kernel, so no programmatic-dependency wait here --
    // with PDL the whole ring is already in flight while the predecessor drains.
    if (lane == 0) {
      uint64_t pol_w;
      asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol_w));
      int gb = 0;                                 // running stage counter at the start of the item
      bool ok = true;
      for (int it = blockIdx.x; it < p.n_items && ok; it += gridDim.x) {
        const Item w = item_of(it);
        const CUtensorMap* mw = (w.mat == 0) ? &mw0 : (w.mat == 1 ? &mw1 : &mw2);
        for (int i = 0; i < w.nst; i++) {
          const int g = gb + i, s = g % S::SW;
          if (g >= S::SW && !tc_mbar_wait(&wfree[s], ((g / S::SW) & 1u) ^ 1u, abort_flag)) { ok = false; break; }
          const int nk = (w.nkb - i * S::KG) < S::KG ? (w.nkb - i * S::KG) : S::KG;
          tc_mbar_expect_tx(&wfull[s], nk * S::W_BYTES);
          TC_TRACE(0, i);
          for (int kk = 0; kk < nk; kk++)
            tc_tma_2d_hint(wring + s * S::W_STAGE + kk * S::W_BYTES, mw, (w.kb0 + i * S::KG + kk) * TC_BK, w.row0, &wfull[s], pol_w);
        }
        gb += w.nst;
      }
    }
  } else if (warp == 3) {
    // ===== X producer (activations: written by the previous kernel)
    if (lane == 0) {
      pdl_wait();
      kt_stamp(kt, 1);
      int gb = 0;
      bool ok = true;
      for (int it = blockIdx.x; it < p.n_items && ok; it += gridDim.x) {
        const Item w = item_of(it);
        for (int i = 0; i < w.nst; i++) {
          const int g = gb + i, s = g % S::SX;
          if (g >= S::SX && !tc_mbar_wait(&aempty[s], ((g / S::SX) & 1u) ^ 1u, abort_flag)) { ok = false; break; }
          const int nk = (w.nkb - i * S::KG) < S::KG ? (w.nkb - i * S::KG) : S::KG;
          TC_TRACE(1, i);
          tc_mbar_expect_tx(&ready[s], nk * S::XS_BYTES);
          for (int kk = 0; kk < nk; kk++) {
            unsigned char* st = xring + s * S::XS_STAGE + kk * S::XS_BYTES;
            tc_tma_2d(st, &mxh, (w.kb0 + i * S::KG + kk) * TC_BK, w.n0, &ready[s]);
            tc_tma_2d(st + S::X_BYTES, &mxl, (w.kb0 + i * S::KG + kk) * TC_BK, w.n0, &ready[s]);
          }
        }
        gb += w.nst;
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer: A (weights, hi / lo) from TMEM, B (activations, hi / lo) from shared memory
    if (lane == 0) {
      constexpr uint32_t idesc = tc_idesc(NT), idesc2 = tc_idesc(2 * NT);
      int gb = 0, n_it = 0;
      bool ok = true;
      for (int it = blockIdx.x; it < p.n_items && ok; it += gridDim.x, n_it++) {
        const Item w = item_of(it);
        if (n_it > 0) {                           // the previous item's accumulator must have been read out
          if (!tc_mbar_wait(accfree, (n_it - 1) & 1u, abort_flag)) break;
          asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        }
        for (int i = 0; i < w.nst; i++) {
          const int g = gb + i, sa = g % S::SA;
          TC_TRACE(8, i);
          if (!tc_mbar_wait(&ready[sa], (g / S::SA) & 1u, abort_flag)) { ok = false; break; }
          TC_TRACE(9, i);
          if (n_it == 0 && i == 0) kt_stamp(kt, 2);
          asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
          TC_TRACE(5, i);
          const int nk = (w.nkb - i * S::KG) < S::KG ? (w.nkb - i * S::KG) : S::KG;
          // Per K = 8 step TWO MMAs instead of three: the X tile is stored [Xhi rows | Xlo rows], so one
          // descriptor over 2*NT rows gives D[:, 0:NT] += Whi.Xhi and D[:, NT:2NT] += Whi.Xlo in a single
          // instruction (an MMA of this shape costs ~60 cycles for any N <= 128); Wlo.Xhi then adds into the
          // second half.  Columns [NT, 2NT) hold the small correction terms, [0, NT) the main sum.
#pragma unroll
          for (int kk = 0; kk < S::KG; kk++) {
            if (kk < nk) {
              const uint64_t xhi = tc_smem_desc(tc_smem_u32(xring + sa * S::XS_STAGE + kk * S::XS_BYTES));
              const uint32_t ahi = tmem_a + sa * S::A_STAGE + kk * S::A_COLS, alo = ahi + TC_BK;
#pragma unroll
              for (int k = 0; k < TC_BK / 8; k++) {       // UMMA K = 8 tf32: 8 TMEM columns of A, 32 bytes (2 descriptor units) of B
                const uint64_t adv = (uint64_t)(k * 2);
                tc_mma_tf32_ts(tmem_d, ahi + 8 * k, xhi + adv, idesc2, (uint32_t)((i | kk | k) != 0));
                tc_mma_tf32_ts(tmem_dl, alo + 8 * k, xhi + adv, idesc, 1u);
              }
            }
          }
          TC_TRACE(6, i);
          tc_commit(&aempty[sa]);
          TC_TRACE(7, i);
        }
        if (ok) tc_commit(accum);
        if (n_it == 0) kt_stamp(kt, 3);
        gb += w.nst;
      }
    }
  } else if (warp >= 4) {
    // ===== converters (shared -> hi/lo -> TMEM), then epilogue.  Thread = weight row of the tile.
    const int q = warp & 3;                             // this warp's 32-lane quarter of TMEM
    const int r = q * 32 + lane;
    const int grp = (warp >> 2) - 1;                    // converter group 0..TC_CGROUPS-1
    bool ok = true;
    int gb = 0, n_it = 0;
    for (int it = blockIdx.x; it < p.n_items && ok; it += gridDim.x, n_it++) {
      const Item w = item_of(it);
      int i0 = (grp - gb % TC_CGROUPS + TC_CGROUPS) % TC_CGROUPS;   // first stage of this item with running index = grp (mod groups)
      for (int i = i0; i < w.nst && ok; i += TC_CGROUPS) {
        const int g = gb + i, sw = g % S::SW, sa = g % S::SA;
        const int nk = (w.nkb - i * S::KG) < S::KG ? (w.nkb - i * S::KG) : S::KG;
        ok = tc_mbar_wait(&wfull[sw], (g / S::SW) & 1u, abort_flag);
        if (threadIdx.x == 128 + 128 * grp) TC_TRACE(2, i);
#pragma unroll
        for (int kk = 0; kk < S::KG; kk++) {
          if (kk < nk) {
            uint32_t hi[32], lo[32];
            const unsigned char* wrow = wring + sw * S::W_STAGE + kk * S::W_BYTES + r * 128;
#pragma unroll
            for (int j = 0; j < 8; j++) {                  // logical 16-byte chunk j lives at j ^ (row & 7) (SWIZZLE_128B)
              const float4 v = *reinterpret_cast<const float4*>(wrow + ((j ^ (r & 7)) << 4));
              const float h0 = tc_tf32(v.x), h1 = tc_tf32(v.y), h2 = tc_tf32(v.z), h3 = tc_tf32(v.w);
              hi[4 * j + 0] = __float_as_uint(h0); hi[4 * j + 1] = __float_as_uint(h1);
              hi[4 * j + 2] = __float_as_uint(h2); hi[4 * j + 3] = __float_as_uint(h3);
              lo[4 * j + 0] = __float_as_uint(v.x - h0); lo[4 * j + 1] = __float_as_uint(v.y - h1);
              lo[4 * j + 2] = __float_as_uint(v.z - h2); lo[4 * j + 3] = __float_as_uint(v.w - h3);
            }
            if (kk == 0 && g >= S::SA && ok) ok = tc_mbar_wait(&aempty[sa], ((g / S::SA) & 1u) ^ 1u, abort_flag);
            if (kk == 0 && threadIdx.x == 128 + 128 * grp) TC_TRACE(3, i);
            const uint32_t ta = tmem_a + sa * S::A_STAGE + kk * S::A_COLS + ((uint32_t)(q * 32) << 16);
            tc_tmem_st32(ta, hi);                          // consumes every loaded value: the shared-memory reads are complete
            tc_tmem_st32(ta + TC_BK, lo);
          }
        }
        __syncwarp();
        if (lane == 0) tc_mbar_arrive(&wfree[sw]);         // the W slot can be refilled while the stores drain
        asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
        __syncwarp();
        if (threadIdx.x == 128 + 128 * grp) TC_TRACE(4, i);
        if (lane == 0) tc_mbar_arrive(&ready[sa]);
      }
      // ---- epilogue of the item: lane = weight row within this warp's 32-lane TMEM quarter
      if (n_it == 0) {
        pdl_trigger();
        pdl_wait();   // returns at once (the X producer already waited); makes the predecessor's writes visible to these threads
      }
      if (ok) ok = tc_mbar_wait(accum, n_it & 1u, abort_flag);
      if (n_it == 0 && threadIdx.x == 128) kt_stamp(kt, 4);
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      const int row = w.row0 + r;
      const bool row_ok = row < p.M[w.mat];
      float* outz = p.out + (size_t)(it % p.splits) * p.zstride + p.out_off[w.mat] + row;
      if (ok) {
        // Whole tile inside the output (every tile of the model shapes): no per-element bounds checks, the 16 column
        // scales as four 128-bit read-only loads, the store address advanced by one 64-bit add per column.  (The
        // checked form below costs ~25 instructions per element -- 64-bit multiplies, two predicates, a branch -- and made
        // this read-out issue-bound: ~1 us per 16 columns with 12 warps on 4 schedulers.)
        const bool full = row_ok && (w.n0 + NT <= p.N);
        const bool cs_vec = p.colscale && ((reinterpret_cast<uintptr_t>(p.colscale) & 15) == 0);
#pragma unroll 1
        for (int c = 16 * grp; c < NT; c += 16 * TC_CGROUPS) {
          uint32_t v[16], w2[16];
          const uint32_t taddr = tmem_d + ((uint32_t)(q * 32) << 16) + (uint32_t)c;
          tc_tmem_ld16(taddr, v);
          if (S::SPLIT_ACC) tc_tmem_ld16(taddr + NT, w2);
          float sc[16];
          if (full && p.colscale) {
            if (cs_vec) {
#pragma unroll
              for (int j4 = 0; j4 < 4; j4++) {
                const float4 t = __ldg(reinterpret_cast<const float4*>(p.colscale + w.n0 + c) + j4);
                sc[4 * j4] = t.x; sc[4 * j4 + 1] = t.y; sc[4 * j4 + 2] = t.z; sc[4 * j4 + 3] = t.w;
              }
            } else {
#pragma unroll
              for (int j = 0; j < 16; j++) sc[j] = __ldg(p.colscale + w.n0 + c + j);
            }
          }
          asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
          if (n_it == 0 && threadIdx.x == 128) kt_stamp(kt, 7 + 2 * (c / (16 * TC_CGROUPS)));
          if (full) {
            float* o = outz + (size_t)(w.n0 + c) * p.ldo;
#pragma unroll
            for (int j = 0; j < 16; j++) {
              float x = __uint_as_float(v[j]);
              if (S::SPLIT_ACC) x += __uint_as_float(w2[j]);
              if (p.colscale) x *= sc[j];
              *o = x;
              o += p.ldo;
            }
          } else {
#pragma unroll
            for (int j = 0; j < 16; j++) {
              const int n = w.n0 + c + j;
              if (n < p.N && row_ok) {
                float x = __uint_as_float(v[j]);
                if (S::SPLIT_ACC) x += __uint_as_float(w2[j]);
                if (p.colscale) x *= p.colscale[n];
                outz[(size_t)n * p.ldo] = x;
              }
            }
          }
          if (n_it == 0 && threadIdx.x == 128) kt_stamp(kt, 8 + 2 * (c / (16 * TC_CGROUPS)));
        }
      }
      if (n_it == 0 && threadIdx.x == 128) kt_stamp(kt, 5);
      asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
      __syncwarp();
      if (lane == 0) tc_mbar_arrive(accfree);
      gb += w.nst;
    }
  }
  __syncthreads();
  if (threadIdx.x == 0) kt_stamp(kt, 6);
  if (warp == 2) {
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(S::TMEM_COLS));
  }
  if (threadIdx.x == 0 && *abort_flag && p.error_flag) *((volatile int*)p.error_flag) = 1;
}

__global__ void split_hilo_kernel(const float* __restrict__ x, long long ldx, float* __restrict__ hi, float* __restrict__ lo, int K) {
  const int n = blockIdx.y;
  for (int k = blockIdx.x * blockDim.x + threadIdx.x; k < K; k += gridDim.x * blockDim.x) {
    const float v = x[(size_t)n * ldx + k];
    const float h = tc_tf32(v);
    hi[(size_t)n * K + k] = h;
    lo[(size_t)n * K + k] = v - h;
  }
}

// ------------------------------------------------------------------ host side ----
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn g_encode = nullptr;
static long long* g_trace = nullptr;   // development aid: tha_gemm_trace()
static int tc_env(const char* n, int d) { const char* v = getenv(n); return (v && *v) ? atoi(v) : d; }
static std::mutex g_tc_mu;
static std::map<std::tuple<const void*, long long, int, long long, int, int>, CUtensorMap> g_maps;
struct TcErrFlag { int* host; int* dev; };   // one mapped pinned word per device: kernels store to it, the host reads it without a copy
static std::map<int, TcErrFlag> g_err_flags;

static int tc_get_encode() {
  if (g_encode) return THA_STATUS_SUCCESS;
  void* fn = nullptr;
  cudaDriverEntryPointQueryResult qres;
  THA_CUDA_OK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres));
  if (!fn || qres != cudaDriverEntryPointSuccess) { set_last_error("cuTensorMapEncodeTiled unavailable", cudaErrorNotSupported); return THA_STATUS_NOT_SUPPORTED; }
  g_encode = (EncodeTiledFn)fn;
  return THA_STATUS_SUCCESS;
}

// fp32 row-major [rows, K] with row pitch `ld` floats; box = [box_rows x 32 floats], 128-byte swizzle.
// (elem_bytes = 1: int8 row-major [rows, K], box = [box_rows x 128 values] -- the same 128-byte swizzled rows.)
int tc_map(const void* ptr, long long rows, int K, long long ld, int box_rows, CUtensorMap* out, int elem_bytes) {
  std::lock_guard<std::mutex> lk(g_tc_mu);
  THA_TRY_RC(tc_get_encode());
  auto key = std::make_tuple((const void*)ptr, rows, K, ld, box_rows, elem_bytes);
  auto it = g_maps.find(key);
  if (it != g_maps.end()) { *out = it->second; return THA_STATUS_SUCCESS; }
  CUtensorMap m;
  // int8 rows are described as 32-bit words (same bytes, same swizzle): coordinates along K are then in words;
  // bf16 rows (elem_bytes = 2) as 64 elements per 128-byte box row
  const cuuint64_t dims[2] = {(cuuint64_t)(elem_bytes == 1 ? K / 4 : K), (cuuint64_t)rows};
  const cuuint64_t strides[1] = {(cuuint64_t)ld * elem_bytes};
  const cuuint32_t box[2] = {(cuuint32_t)(elem_bytes == 2 ? 64 : 32), (cuuint32_t)box_rows};
  const cuuint32_t estr[2] = {1, 1};
  const CUtensorMapDataType dt = elem_bytes == 4 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32
                                 : (elem_bytes == 2 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_UINT32);
  CUresult r = g_encode(&m, dt, 2,
                        const_cast<void*>(ptr), dims, strides, box, estr,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                        (CUtensorMapL2promotion)tc_env("THA_TC_PROMO", (int)CU_TENSOR_MAP_L2_PROMOTION_L2_256B),
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { set_last_error("cuTensorMapEncodeTiled", cudaErrorInvalidValue); return THA_STATUS_EXECUTION_FAILED; }
  if (g_maps.size() > 4096) g_maps.clear();
  g_maps[key] = m;
  *out = m;
  return THA_STATUS_SUCCESS;
}

// Per-device error flag (allocated on first use; callers that capture graphs call this beforehand).
int gemm_tc_init(int** flag_out) {
  int dev = 0;
  THA_CUDA_OK(cudaGetDevice(&dev));
  std::lock_guard<std::mutex> lk(g_tc_mu);
  auto it = g_err_flags.find(dev);
  if (it == g_err_flags.end()) {
    TcErrFlag f = {nullptr, nullptr};
    THA_CUDA_OK(cudaHostAlloc((void**)&f.host, sizeof(int), cudaHostAllocMapped | cudaHostAllocPortable));
    *f.host = 0;
    THA_CUDA_OK(cudaHostGetDevicePointer((void**)&f.dev, f.host, 0));
    it = g_err_flags.emplace(dev, f).first;
  }
  if (flag_out) *flag_out = it->second.dev;
  return THA_STATUS_SUCCESS;
}

bool gemm_tc_ok(const TcArgs& a) {
  if (a.K % 4 || a.ldx % 4 || a.K < TC_BK || a.N < 1) return false;
  for (int i = 0; i < a.n_mats; i++)
    if (!a.W[i] || (reinterpret_cast<uintptr_t>(a.W[i]) & 15u) || a.M[i] < 1) return false;
  return a.Xhi && a.Xlo && (reinterpret_cast<uintptr_t>(a.Xhi) & 15u) == 0 && (reinterpret_cast<uintptr_t>(a.Xlo) & 15u) == 0;
}

int gemm_tc_pick_nt(int N) { return N <= 16 ? 16 : (N <= 32 ? 32 : (N <= 64 ? 64 : 128)); }   // wider N: more column tiles

template <int NT>
static int tc_launch_nt(cudaStream_t st, const CUtensorMap* mw, const CUtensorMap& mxh, const CUtensorMap& mxl, const TcP& p, dim3 grid, bool pdl) {
  using S = TcSmem<NT>;
  auto k = gemm_tf32x3_kernel<NT>;
  THA_CUDA_OK(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, S::TOTAL));
  THA_CUDA_OK(launch(k, grid, dim3(TC_THREADS), (size_t)S::TOTAL, st, pdl, mw[0], mw[1], mw[2], mxh, mxl, p));
  return THA_STATUS_SUCCESS;
}

int gemm_tc_launch(cudaStream_t st, const TcArgs& a, bool pdl) {
  if (!gemm_tc_ok(a) || a.n_mats < 1 || a.n_mats > 3 || !a.out) return THA_STATUS_INVALID_ARG;
  const int NT = gemm_tc_pick_nt(a.N);
  CUtensorMap mw[3], mxh, mxl;
  TcP p = {};
  int tiles = 0;
  for (int i = 0; i < 3; i++) {
    if (i < a.n_mats) {
      THA_TRY_RC(tc_map(a.W[i], a.M[i], a.K, a.K, TC_BM, &mw[i]));
      p.M[i] = a.M[i]; p.tiles[i] = ceil_div(a.M[i], TC_BM); p.out_off[i] = a.out_off[i];
    } else {
      mw[i] = mw[0]; p.M[i] = 0; p.tiles[i] = 0; p.out_off[i] = 0;
    }
    tiles += p.tiles[i];
  }
  THA_TRY_RC(tc_map(a.Xhi, a.N, a.K, a.ldx, NT, &mxh));
  THA_TRY_RC(tc_map(a.Xlo, a.N, a.K, a.ldx, NT, &mxl));
  p.K = a.K; p.N = a.N;
  p.kblocks_total = ceil_div(a.K, TC_BK);
  const int splits = a.splits < 1 ? 1 : a.splits;
  p.kblocks_per_split = ceil_div(p.kblocks_total, splits);
  if (ceil_div(p.kblocks_total, p.kblocks_per_split) != splits) return THA_STATUS_INVALID_ARG;   // caller sizes its partial buffers by `splits`
  p.out = a.out; p.ldo = a.ldo; p.zstride = a.zstride; p.colscale = a.colscale;
  p.trace = g_trace;
  THA_TRY_RC(gemm_tc_init(&p.error_flag));
  p.ny = ceil_div(a.N, NT);
  p.splits = splits;
  p.n_items = tiles * p.ny * splits;
  const int sms = sm_budget();
  dim3 grid(p.n_items < sms ? p.n_items : sms, 1, 1);
  switch (NT) {
    case 16: return tc_launch_nt<16>(st, mw, mxh, mxl, p, grid, pdl);
    case 32: return tc_launch_nt<32>(st, mw, mxh, mxl, p, grid, pdl);
    case 64: return tc_launch_nt<64>(st, mw, mxh, mxl, p, grid, pdl);
    default: return tc_launch_nt<128>(st, mw, mxh, mxl, p, grid, pdl);
  }
}

// 1 if any tensor-core GEMM on this device gave up on a barrier since the last check (clears the flag).
int gemm_tc_check_error() {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return 1;
  volatile int* f = nullptr;
  {
    std::lock_guard<std::mutex> lk(g_tc_mu);
    auto it = g_err_flags.find(dev);
    if (it == g_err_flags.end()) return 0;
    f = it->second.host;
  }
  const int h = *f;   // written by a kernel that gave up (a plain store to the mapped word); callers sync their stream first
  if (h) *f = 0;
  return h;
}

// =====================================================================================================
// Q8_0 x Q8_0 on the INT8 tensor cores: runq.c:317-342 evaluated exactly as written.
//
//   out[n, m] = sum_g  (float)( sum_{k in g} Wq[m,k] * Xq[n,k] )  *  Ws[m,g]  *  Xs[n,g]        (groups ascending)
//
// The inner integer sum of a group is ONE accumulator pass of tcgen05.mma.kind::i8 (int8 x int8 -> int32, exact;
// K = 32 per instruction, gs/32 instructions per group) with BOTH operands read straight from the TMA-written
// shared-memory tiles: no conversion warps, weights cross HBM and shared memory as the int8 bytes of the file.
// Each finished group accumulator [128 rows x NT columns] is pulled out of TMEM by the epilogue warps
// (thread = weight row), converted exactly, scaled by the row's weight scale and the column's activation scale in
// the reference's order and added in fp32 -- so without split-K the result is bit-identical to runq.c's loop
// (compiled without FMA contraction).  TMEM holds a ring of group accumulators so the MMAs of the next groups
// overlap the scaling of the previous one; shared memory holds a ring of [W 128x128 | X NTx128] int8 slots.
//   warp 0  TMA producer (weights first: they do not depend on the previous kernel)   warps 1, 3  MMA issuers
//   warp 2  TMEM allocation     warps 4-7  epilogue: one warp per 32-lane TMEM quarter, thread = weight row
// Every tcgen05 instruction costs its issuing thread a few hundred cycles whatever its size (measured: ~300 per
// kind::i8 MMA, ~200 per tcgen05.ld, ~450 per wait::ld), so the epilogue reads 128 accumulator columns -- 128/NT
// consecutive groups, laid out back to back in TMEM -- per round trip.
// Epilogue warps: four (one per 32-lane TMEM quarter) for narrow column tiles; EIGHT for NT >= 32 -- two per quarter, each with
// half of the tile's columns.  At NT = 64 a slot of 16 KB of weights leaves 64 KB of int32 group sums in TMEM (read at 64 B per
// cycle: ~1000 cycles) and costs 640 scaling instructions per row; with one warp per quarter the two are strictly serial
// (~1650 cycles per slot, 2.0 TB/s of weights); with two, one warp's TMEM read runs under the other's arithmetic.
__host__ __device__ constexpr int q8i_epi_warps(int nt) { return nt >= 32 ? 8 : 4; }
__host__ __device__ constexpr int q8i_threads(int nt) { return 128 + 32 * q8i_epi_warps(nt); }

struct Q8iP {
  int M[3]; int tiles[3]; int out_off[3];
  const float* S[3];                 // weight group scales [M, ngrp]
  const float* xsT; int ldxs;        // activation group scales, transposed: xsT[g * ldxs + n]
  int K, N, ny, gs, ngrp;
  int slots_total, slots_per_split;  // ring slots of 128 values
  float* out; long long ldo; long long zstride;
  int* error_flag;
  int nring;                         // shared-memory ring slots in use
  int splits, n_items;               // work items = (tiles of all matrices) x column tiles x splits
};

template <int NT>
struct Q8iSmem {
  static constexpr int W_BYTES = TC_BM * 128, X_BYTES = NT * 128, SLOT = W_BYTES + X_BYTES;
  static constexpr int NSLOT_RAW = (TC_SMEM_LIMIT - 3072) / SLOT;
  static constexpr int NSLOT = NSLOT_RAW > 10 ? 10 : NSLOT_RAW;
  // One tcgen05.commit per STAGE of SPC slots (a commit costs the issuing thread a few hundred cycles; the same
  // barrier releases the stage's shared-memory slots to the producer and hands its accumulators to the epilogue).
  static constexpr int SPC = 1;
  // MMA-issuing warps: a tcgen05.mma.kind::i8 costs its issuing thread ~300 cycles whatever its size (measured), so
  // the slots are dealt round-robin to two issuers (each owns whole groups: accumulation order is untouched).  The number
  // of accumulator stages is a multiple of the issuers, so a stage is always refilled by the same thread (a waiter
  // that skipped a phase of the release barrier would read its parity wrongly).
  static constexpr int NISS = 2;
  static constexpr int TMEM_COLS = 512;
  static constexpr int NMD = 64;                  // commit barriers (more than the slots that can be in flight)
  static constexpr int NACC_MAX = 32;
  static constexpr int TOTAL = NSLOT * SLOT + 1024 + (NSLOT + NMD + NACC_MAX) * 8 + 64;
  static_assert(TOTAL <= TC_SMEM_LIMIT, "shared memory overflow");
  static_assert(NSLOT + NACC_MAX + 1 <= NMD, "commit barrier ring too short");
};

__host__ __device__ constexpr uint32_t q8i_idesc(int nt) {   // kind::i8: D = S32 (2), A = B = signed 8-bit (1), K-major, M = 128
  return (2u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(nt >> 3) << 17) | ((uint32_t)(TC_BM >> 4) << 24);
}
__device__ __forceinline__ void q8i_mma(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void tc_tmem_ld8(uint32_t taddr, uint32_t (&v)[8]) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
               : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]) : "r"(taddr));
}

__device__ __forceinline__ void tc_tmem_ld64(uint32_t taddr, uint32_t* v) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x64.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32, %33, %34, %35, %36, %37, %38, %39, %40, %41, %42, %43, %44, %45, %46, %47, %48, %49, %50, %51, %52, %53, %54, %55, %56, %57, %58, %59, %60, %61, %62, %63}, [%64];"
               : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31]), "=r"(v[32]), "=r"(v[33]), "=r"(v[34]), "=r"(v[35]), "=r"(v[36]), "=r"(v[37]), "=r"(v[38]), "=r"(v[39]), "=r"(v[40]), "=r"(v[41]), "=r"(v[42]), "=r"(v[43]), "=r"(v[44]), "=r"(v[45]), "=r"(v[46]), "=r"(v[47]), "=r"(v[48]), "=r"(v[49]), "=r"(v[50]), "=r"(v[51]), "=r"(v[52]), "=r"(v[53]), "=r"(v[54]), "=r"(v[55]), "=r"(v[56]), "=r"(v[57]), "=r"(v[58]), "=r"(v[59]), "=r"(v[60]), "=r"(v[61]), "=r"(v[62]), "=r"(v[63]) : "r"(taddr));
}

// Persistent: a CTA walks the work items (weight tile, column tile, K split) blockIdx.x, blockIdx.x + gridDim.x, ...
// with ONE running slot counter, so the TMA ring, the MMA issuers and the epilogue flow from one item into the next
// without draining (a fresh CTA per item paid ~3 us of ramp for ~6 us of streaming).  Every item is padded to a whole
// number of TMEM read batches with slots past the end of K: TMA zero-fills them, their integer sums are 0 and the
// epilogue skips their groups.
template <int NT>
__global__ void __launch_bounds__(q8i_threads(NT), 1)
gemm_q8i_kernel(const __grid_constant__ CUtensorMap mw0, const __grid_constant__ CUtensorMap mw1,
                const __grid_constant__ CUtensorMap mw2, const __grid_constant__ CUtensorMap mx, const Q8iP p) {
  using S = Q8iSmem<NT>;
  extern __shared__ unsigned char tc_smem_raw[];
  unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(tc_smem_raw) + 1023) & ~(uintptr_t)1023);
  const int nring = p.nring;             // ring slots actually used (the activation scales take the rest)
  float* xs_sm = reinterpret_cast<float*>(smem + nring * S::SLOT);   // [groups of one item][NT]
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + S::NSLOT * S::SLOT);
  uint64_t* full = bars;                 // slot landed (W + X bytes)
  uint64_t* mdone = full + S::NSLOT;     // the MMAs of a slot have completed: slot reusable, accumulators readable
  uint64_t* aempty = mdone + S::NMD;     // a slot's accumulators have been read out by the 4 epilogue warps
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(aempty + S::NACC_MAX);
  int* abort_flag = reinterpret_cast<int*>(tmem_slot + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int mpg = p.gs >> 5;                     // MMAs per group
  const int gps = 128 / p.gs;                    // groups per slot
  const int nacc = S::TMEM_COLS / (gps * NT);    // accumulator stages (slots) in TMEM: 4 .. 32, a power of two
  const int nacc_sh = 31 - __clz(nacc);
  constexpr int LG = 128 / NT;                   // groups per TMEM read (128 columns)
  const int sb = LG / gps;                       // slots per TMEM read batch (NT <= gs: at least 1)
  const int tiles01 = p.tiles[0] + p.tiles[1];
  __shared__ int kt_q8_s;

  // work item -> (matrix, row0, n0, first slot, real slots, padded slots, real groups)
  struct Item { int mat, row0, n0, slot0, nslots, npad, ng; };
  auto item_of = [&](int it) -> Item {
    Item w;
    const int z = it % p.splits;
    const int t = it / p.splits;
    int tile = t / p.ny;
    w.n0 = (t - tile * p.ny) * NT;
    w.mat = 0;
    if (tile >= p.tiles[0]) { w.mat = 1; if (tile >= tiles01) { w.mat = 2; tile -= tiles01; } else tile -= p.tiles[0]; }
    w.row0 = tile * TC_BM;
    w.slot0 = z * p.slots_per_split;
    w.nslots = p.slots_total - w.slot0;
    if (w.nslots > p.slots_per_split) w.nslots = p.slots_per_split;
    w.npad = ((w.nslots + sb - 1) / sb) * sb;
    const int g0 = w.slot0 * gps;
    w.ng = p.ngrp - g0;
    if (w.ng > w.nslots * gps) w.ng = w.nslots * gps;
    return w;
  };

  if (threadIdx.x == 0) {
    for (int s = 0; s < S::NSLOT; s++) tc_mbar_init(&full[s], 1);
    for (int b = 0; b < S::NMD; b++) tc_mbar_init(&mdone[b], 1);
    for (int b = 0; b < S::NACC_MAX; b++) tc_mbar_init(&aempty[b], q8i_epi_warps(NT));
    *abort_flag = 0;
    kt_q8_s = kt_begin((unsigned long long)KT_GEMM_Q8 | ((unsigned long long)p.splits << 16) | ((unsigned long long)p.n_items << 32));
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
  const int kt = kt_q8_s;

  if (warp == 0) {
    // ===== TMA producer
    if (lane == 0) {
      uint64_t pol_w;
      asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol_w));
      int jb = 0;                                   // running slot counter at the start of the item
      bool ok = true;
      for (int it = blockIdx.x; it < p.n_items && ok; it += gridDim.x) {
        const Item w = item_of(it);
        const CUtensorMap* mw = (w.mat == 0) ? &mw0 : (w.mat == 1 ? &mw1 : &mw2);
        int j = 0;
        if (jb == 0) {
          // the first ring of weight tiles goes out before the dependency wait (transaction bytes may land before
          // the matching expect_tx: the phase cannot complete until the arrival below)
          const int pre = w.npad < nring ? w.npad : nring;
          for (; j < pre; j++) tc_tma_2d_hint(smem + j * S::SLOT, mw, (w.slot0 + j) * 32, w.row0, &full[j], pol_w);
          pdl_wait();
          kt_stamp(kt, 1);
          for (j = 0; j < pre; j++) {
            tc_mbar_expect_tx(&full[j], S::SLOT);
            tc_tma_2d(smem + j * S::SLOT + S::W_BYTES, &mx, (w.slot0 + j) * 32, w.n0, &full[j]);
          }
        }
        for (; j < w.npad; j++) {
          const int jg = jb + j, s = jg % nring;
          if (jg >= nring) {
            const int st = jg - nring;              // the slot use whose MMAs must have completed
            if (!tc_mbar_wait(&mdone[st % S::NMD], (st / S::NMD) & 1u, abort_flag)) { ok = false; break; }
          }
          tc_mbar_expect_tx(&full[s], S::SLOT);
          tc_tma_2d_hint(smem + s * S::SLOT, mw, (w.slot0 + j) * 32, w.row0, &full[s], pol_w);
          tc_tma_2d(smem + s * S::SLOT + S::W_BYTES, &mx, (w.slot0 + j) * 32, w.n0, &full[s]);
        }
        jb += w.npad;
      }
    }
  } else if (warp == 1 || warp == 3) {
    // ===== MMA issuers (warps 1 and 3): issuer i takes the slots with running index = i (mod 2)
    if (lane == 0) {
      constexpr uint32_t idesc = q8i_idesc(NT);
      const int mpg_sh = 31 - __clz(mpg), gps_sh = 31 - __clz(gps), nacc_mask = nacc - 1;   // all powers of two
      const int me = warp >> 1;
      int jb = 0;
      bool ok = true;
      for (int it = blockIdx.x; it < p.n_items && ok; it += gridDim.x) {
        const Item w = item_of(it);
        for (int j = (jb + me) & 1; j < w.npad; j += 2) {
          const int jg = jb + j, s = jg % nring;
          if (jg >= nacc) {                         // the accumulator stage must have been drained
            if (!tc_mbar_wait(&aempty[jg & nacc_mask], ((jg >> nacc_sh) & 1u) ^ 1u, abort_flag)) { ok = false; break; }
          }
          if (!tc_mbar_wait(&full[s], (jg / nring) & 1u, abort_flag)) { ok = false; break; }
          if (jg == 0) kt_stamp(kt, 2);
          asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
          const uint64_t adesc = tc_smem_desc(tc_smem_u32(smem + s * S::SLOT));
          const uint64_t bdesc = tc_smem_desc(tc_smem_u32(smem + s * S::SLOT + S::W_BYTES));
          const uint32_t acc0 = tmem_base + (uint32_t)(((jg & nacc_mask) << gps_sh) * NT);
#pragma unroll
          for (int sub = 0; sub < 4; sub++)       // past the end of K the tiles are zero-filled: always four MMAs
            q8i_mma(acc0 + (uint32_t)((sub >> mpg_sh) * NT), adesc + (uint64_t)(sub * 2), bdesc + (uint64_t)(sub * 2), idesc,
                    (uint32_t)((sub & (mpg - 1)) != 0));
          tc_commit(&mdone[jg % S::NMD]);
        }
        if (jb == 0 && warp == 1) kt_stamp(kt, 3);
        jb += w.npad;
      }
    }
  } else if (warp >= 4) {
    // ===== epilogue: thread = weight row (TMEM lane) x CH columns (all NT, or one half of them: q8i_epi_warps)
    constexpr int EW = q8i_epi_warps(NT), ET = 32 * EW, CH = NT * 4 / EW;
    const int q = warp & 3;
    const int c0 = ((warp - 4) >> 2) * CH;       // first column of this warp's share
    const int r = q * 32 + lane;
    const uint32_t tq = tmem_base + ((uint32_t)(q * 32) << 16);
    const int gps_sh = 31 - __clz(gps);
    pdl_trigger();
    pdl_wait();                                     // the activation scales come from the previous kernel
    int jb = 0;
    bool ok = true;
    uint32_t v[LG * CH];
    for (int it = blockIdx.x; it < p.n_items && ok; it += gridDim.x) {
      const Item w = item_of(it);
      const int ng = w.ng;
      const int row = w.row0 + r;
      const bool row_ok = row < p.M[w.mat];
      const int g0 = w.slot0 * gps;
      const float* srow = p.S[w.mat] + (size_t)(row_ok ? row : p.M[w.mat] - 1) * p.ngrp + g0;
      float acc[CH];
#pragma unroll
      for (int c = 0; c < CH; c++) acc[c] = 0.0f;
      // Weight scales: this thread's row, four groups per 128-bit load, requested two to three chunks (8-12 groups,
      // well over a thousand cycles) before their use -- with ~200 KB of the SM's memory carved out as shared, L1
      // keeps next to nothing and every one of these loads is an L2 round trip.
      const bool ws_vec = ((reinterpret_cast<uintptr_t>(srow) & 15u) == 0) && (p.ngrp % 4 == 0);
      auto load_ws4 = [&](int chunk) -> float4 {
        const int g = 4 * chunk;
        float4 r4 = make_float4(0.f, 0.f, 0.f, 0.f);
        if (g + 3 < ng && ws_vec) return __ldg(reinterpret_cast<const float4*>(srow + g));
        if (g < ng) r4.x = __ldg(srow + g);
        if (g + 1 < ng) r4.y = __ldg(srow + g + 1);
        if (g + 2 < ng) r4.z = __ldg(srow + g + 2);
        if (g + 3 < ng) r4.w = __ldg(srow + g + 3);
        return r4;
      };
      float4 ws_cur = load_ws4(0), ws_n1 = load_ws4(1), ws_n2 = load_ws4(2);
      {  // the activation scales of this item's K range -> shared memory (NT floats per group)
        asm volatile("bar.sync 1, %0;" ::"n"(ET) : "memory");   // the epilogue warps: everybody is done with the previous item's
        const float* src = p.xsT + (size_t)g0 * p.ldxs + w.n0;
        const int t = threadIdx.x - 128;            // 0..ET-1
        for (int i = t; i < ng * (NT / 4); i += ET) {
          const int g = i / (NT / 4), c4 = i - g * (NT / 4);
          reinterpret_cast<float4*>(xs_sm)[i] = __ldg(reinterpret_cast<const float4*>(src + (size_t)g * p.ldxs) + c4);
        }
        asm volatile("bar.sync 1, %0;" ::"n"(ET) : "memory");
      }
      for (int bi = 0; bi * sb < w.npad && ok; bi++) {
        // the batch's slots: the last slot of each issuer covers the earlier ones (an issuer's commits complete in order)
        const int s_first = jb + bi * sb, s_last = s_first + sb - 1;
        int wok = 1;
        if (lane == 0) {
          wok = tc_mbar_wait(&mdone[s_last % S::NMD], (s_last / S::NMD) & 1u, abort_flag) ? 1 : 0;
          if (wok && sb > 1) wok = tc_mbar_wait(&mdone[(s_last - 1) % S::NMD], ((s_last - 1) / S::NMD) & 1u, abort_flag) ? 1 : 0;
        }
        ok = __shfl_sync(THA_FULL_MASK, wok, 0) != 0;
        if (!ok) break;
        if (jb == 0 && bi == 0 && threadIdx.x == 128) kt_stamp(kt, 4);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        const uint32_t ta = tq + (uint32_t)(((s_first & (nacc - 1)) << gps_sh) * NT);
        if (EW == 4) {
          tc_tmem_ld64(ta, v);
          tc_tmem_ld64(ta + 64, v + 64);
        } else {   // this warp's CH columns of each of the LG groups, 16 at a time
#pragma unroll
          for (int b = 0; b < LG; b++)
#pragma unroll
            for (int c = 0; c < CH; c += 16)
              tc_tmem_ld16(ta + (uint32_t)(b * NT + c0 + c), *reinterpret_cast<uint32_t(*)[16]>(v + b * CH + c));
        }
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
        __syncwarp();
        if (lane == 0)
          for (int sl = s_first; sl <= s_last; sl++) tc_mbar_arrive(&aempty[sl & (nacc - 1)]);
#pragma unroll
        for (int b = 0; b < LG; b++) {
          const int gi = bi * LG + b;
          if (gi < ng) {
            if ((gi & 3) == 0 && gi > 0) { ws_cur = ws_n1; ws_n1 = ws_n2; ws_n2 = load_ws4((gi >> 2) + 2); }
            const int wi = gi & 3;
            const float ws = wi == 0 ? ws_cur.x : (wi == 1 ? ws_cur.y : (wi == 2 ? ws_cur.z : ws_cur.w));
#pragma unroll
            for (int c = 0; c < CH; c += 4) {
              const float4 t = *reinterpret_cast<const float4*>(xs_sm + gi * NT + c0 + c);   // same address in every lane: a broadcast
              const float xs[4] = {t.x, t.y, t.z, t.w};
#pragma unroll
              for (int e = 0; e < 4; e++) {
                // |group sum| <= 128 * 127 * 127 < 2^22: 1.5 * 2^23 + i is exact in fp32, so this IS (float)i -- two
                // full-rate instructions instead of a quarter-rate I2F
                const float f = __int_as_float(0x4B400000 + (int)v[b * CH + c + e]) - 12582912.0f;
                acc[c + e] = __fadd_rn(acc[c + e], __fmul_rn(__fmul_rn(f, ws), xs[e]));   // runq.c:334, no contraction
              }
            }
          }
        }
      }
      float* outz = p.out + (size_t)(it % p.splits) * p.zstride + p.out_off[w.mat] + row;
      if (ok && row_ok) {
#pragma unroll
        for (int c = 0; c < CH; c++) {
          const int n = w.n0 + c0 + c;
          if (n < p.N) outz[(size_t)n * p.ldo] = acc[c];
        }
      }
      if (jb == 0 && threadIdx.x == 128) kt_stamp(kt, 5);
      jb += w.npad;
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  }
  __syncthreads();
  if (threadIdx.x == 0) kt_stamp(kt, 6);
  if (warp == 2) {
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(S::TMEM_COLS));
  }
  if (threadIdx.x == 0 && *abort_flag && p.error_flag) *((volatile int*)p.error_flag) = 1;
}

bool gemm_q8i_ok(const Q8iArgs& a) {
  if (a.gs != 32 && a.gs != 64 && a.gs != 128) return false;
  if (a.K % a.gs || a.K < 128 || a.K % 16 || a.N < 1 || a.n_mats < 1 || a.n_mats > 3 || !a.Xq || !a.xsT || !a.out) return false;
  if ((reinterpret_cast<uintptr_t>(a.Xq) & 15u) || (reinterpret_cast<uintptr_t>(a.xsT) & 15u) || a.ldxs % 4) return false;
  for (int i = 0; i < a.n_mats; i++)
    if (!a.Wq[i] || !a.Ws[i] || (reinterpret_cast<uintptr_t>(a.Wq[i]) & 15u) || a.M[i] < 1) return false;
  return true;
}

int gemm_q8i_pick_nt(int N, int gs) {   // a slot's groups must fit one 128-column TMEM read: NT <= gs
  const int nt = N <= 16 ? 16 : (N <= 32 ? 32 : 64);
  return nt > gs ? gs : nt;
}

template <int NT>
static int q8i_launch_nt(cudaStream_t st, const CUtensorMap* mw, const CUtensorMap& mx, const Q8iP& p, dim3 grid, bool pdl) {
  using S = Q8iSmem<NT>;
  auto k = gemm_q8i_kernel<NT>;
  THA_CUDA_OK(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, S::TOTAL));
  THA_CUDA_OK(launch(k, grid, dim3(q8i_threads(NT)), (size_t)S::TOTAL, st, pdl, mw[0], mw[1], mw[2], mx, p));
  return THA_STATUS_SUCCESS;
}

int gemm_q8i_launch(cudaStream_t st, const Q8iArgs& a, bool pdl) {
  if (!gemm_q8i_ok(a)) return THA_STATUS_INVALID_ARG;
  const int NT = gemm_q8i_pick_nt(a.N, a.gs);
  if (a.ldxs < ceil_div(a.N, NT) * NT) return THA_STATUS_INVALID_ARG;   // a column tile reads NT scales per group
  CUtensorMap mw[3], mx;
  Q8iP p = {};
  int tiles = 0;
  for (int i = 0; i < 3; i++) {
    if (i < a.n_mats) {
      THA_TRY_RC(tc_map(a.Wq[i], a.M[i], a.K, a.K, TC_BM, &mw[i], 1));
      p.M[i] = a.M[i]; p.tiles[i] = ceil_div(a.M[i], TC_BM); p.out_off[i] = a.out_off[i]; p.S[i] = a.Ws[i];
    } else {
      mw[i] = mw[0];
    }
    tiles += p.tiles[i];
  }
  THA_TRY_RC(tc_map(a.Xq, a.N, a.K, a.K, NT, &mx, 1));
  p.xsT = a.xsT; p.ldxs = a.ldxs; p.K = a.K; p.N = a.N; p.gs = a.gs; p.ngrp = a.K / a.gs;
  p.slots_total = ceil_div(a.K, 128);
  const int splits = a.splits < 1 ? 1 : a.splits;
  p.slots_per_split = ceil_div(p.slots_total, splits);
  if (ceil_div(p.slots_total, p.slots_per_split) != splits) return THA_STATUS_INVALID_ARG;
  p.out = a.out; p.ldo = a.ldo; p.zstride = a.zstride;
  THA_TRY_RC(gemm_tc_init(&p.error_flag));
  {  // the activation scales of one item sit behind the ring: [groups per (padded) split][NT] floats
    const int slot_bytes = TC_BM * 128 + NT * 128;
    const int nslot_max = (TC_SMEM_LIMIT - 3072) / slot_bytes > 10 ? 10 : (TC_SMEM_LIMIT - 3072) / slot_bytes;
    const int sb = a.gs / NT;
    const int xs_bytes = ceil_div(p.slots_per_split, sb) * sb * (128 / a.gs) * NT * 4;
    p.nring = nslot_max - ceil_div(xs_bytes, slot_bytes);
    if (p.nring < 3) return THA_STATUS_NOT_SUPPORTED;
  }
  p.ny = ceil_div(a.N, NT);
  p.splits = splits;
  p.n_items = tiles * p.ny * splits;
  const int sms = sm_budget();
  dim3 grid(p.n_items < sms ? p.n_items : sms, 1, 1);
  switch (NT) {
    case 16: return q8i_launch_nt<16>(st, mw, mx, p, grid, pdl);
    case 32: return q8i_launch_nt<32>(st, mw, mx, p, grid, pdl);
    default: return q8i_launch_nt<64>(st, mw, mx, p, grid, pdl);
  }
}

static std::map<int, std::pair<float*, size_t>> g_split_scratch;   // per device: scratch of the C-ABI matmuls ([hi | lo], [Bt | hi | lo])

// Forget tensor maps (call when buffers they describe are freed and their addresses may be reused
// with another shape -- the key includes the shape, so reuse with the same shape stays valid).
void gemm_tc_forget() {
  std::lock_guard<std::mutex> lk(g_tc_mu);
  g_maps.clear();
}

}  // namespace tha

using namespace tha;

// C-ABI: out[n*ldo + row] = sum_k W[row,k] X[n,k]   (N rows of activations against one weight matrix)

extern "C" int tha_gemm_tf32x3(void* stream, float* out, long long ldo, const float* W, int M, int K, const float* X, int N,
                               long long ldx) {
  if (!out || !W || !X || M < 1 || N < 1 || K < 1) return THA_STATUS_INVALID_ARG;
  if (K % 4 || K < 32 || (reinterpret_cast<uintptr_t>(W) & 15u)) return THA_STATUS_NOT_SUPPORTED;
  cudaStream_t st = (cudaStream_t)stream;
  int dev = 0;
  THA_CUDA_OK(cudaGetDevice(&dev));
  float* scratch = nullptr;
  {
    std::lock_guard<std::mutex> lk(g_tc_mu);
    auto& e = g_split_scratch[dev];
    const size_t need = (size_t)2 * N * K;
    if (need > e.second) {
      if (e.first) THA_CUDA_OK(cudaFree(e.first));
      THA_CUDA_OK(cudaMalloc(&e.first, need * sizeof(float)));
      e.second = need;
    }
    scratch = e.first;
  }
  float* hi = scratch;
  float* lo = scratch + (size_t)N * K;
  split_hilo_kernel<<<dim3(ceil_div(K, 256) < 8 ? ceil_div(K, 256) : 8, N), 256, 0, st>>>(X, ldx, hi, lo, K);
  count_launch();
  THA_LAUNCH_OK();
  TcArgs a = {};
  a.n_mats = 1; a.W[0] = W; a.M[0] = M; a.out_off[0] = 0; a.K = K; a.Xhi = hi; a.Xlo = lo; a.N = N; a.ldx = K;
  a.out = out; a.ldo = ldo; a.zstride = 0; a.splits = 1; a.colscale = nullptr;
  if (!gemm_tc_ok(a)) return THA_STATUS_NOT_SUPPORTED;
  return gemm_tc_launch(st, a, false);
}
// thaBLAS_s_matmul (src/thaBLAS.cpp:134-170: C[m, n] = A[m, k] . B[k, n], all row-major fp32) on the tensor cores.
// The tcgen05 kernel wants both operands K-major: A already is (it plays the activations: one output row per A row), B is
// transposed once into scratch and plays the weights, so out[i * n + j] = sum_k Bt[j, k] A[i, k] = C[i, j] lands in place.
// Returns NOT_SUPPORTED for shapes the caller should leave to the CUDA-core kernel (tiny, or K not a multiple of 4).
namespace tha {
__global__ void __launch_bounds__(256) transpose_kernel(const float* __restrict__ in, float* __restrict__ out, int rows, int cols) {
  __shared__ float t[32][33];
  const int c0 = blockIdx.x * 32, r0 = blockIdx.y * 32;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;   // 32 x 8
#pragma unroll
  for (int i = 0; i < 32; i += 8) {
    const int r = r0 + ty + i, c = c0 + tx;
    t[ty + i][tx] = (r < rows && c < cols) ? in[(size_t)r * cols + c] : 0.0f;
  }
  __syncthreads();
#pragma unroll
  for (int i = 0; i < 32; i += 8) {
    const int c = c0 + ty + i, r = r0 + tx;                 // out[c][r]
    if (c < cols && r < rows) out[(size_t)c * rows + r] = t[tx][ty + i];
  }
}

int matmul_tc_route(cudaStream_t st, int m, int n, int k, const float* A, const float* B, float* C) {
  if (m < 32 || n < 64 || k < 64 || (k & 3) || (long long)m * n * k < (1ll << 21)) return THA_STATUS_NOT_SUPPORTED;
  if ((reinterpret_cast<uintptr_t>(A) & 15u) || (reinterpret_cast<uintptr_t>(C) & 3u)) return THA_STATUS_NOT_SUPPORTED;
  int dev = 0;
  THA_CUDA_OK(cudaGetDevice(&dev));
  float* scratch = nullptr;
  {
    std::lock_guard<std::mutex> lk(g_tc_mu);
    auto& e = g_split_scratch[dev];
    const size_t need = (size_t)n * k + (size_t)2 * m * k;
    if (need > e.second) {
      if (e.first) THA_CUDA_OK(cudaFree(e.first));
      THA_CUDA_OK(cudaMalloc(&e.first, need * sizeof(float)));
      e.second = need;
    }
    scratch = e.first;
  }
  float* bt = scratch;
  float* hi = scratch + (size_t)n * k;
  float* lo = hi + (size_t)m * k;
  transpose_kernel<<<dim3(ceil_div(n, 32), ceil_div(k, 32)), 256, 0, st>>>(B, bt, k, n);
  count_launch();
  split_hilo_kernel<<<dim3(ceil_div(k, 256) < 8 ? ceil_div(k, 256) : 8, m), 256, 0, st>>>(A, k, hi, lo, k);
  count_launch();
  THA_LAUNCH_OK();
  TcArgs a = {};
  a.n_mats = 1; a.W[0] = bt; a.M[0] = n; a.out_off[0] = 0; a.K = k; a.Xhi = hi; a.Xlo = lo; a.N = m; a.ldx = k;
  a.out = C; a.ldo = n; a.zstride = 0; a.splits = 1; a.colscale = nullptr;
  if (!gemm_tc_ok(a)) return THA_STATUS_NOT_SUPPORTED;
  return gemm_tc_launch(st, a, false);
}
}  // namespace tha

extern "C" int tha_matmul_q8_tc(void* stream, float* out, long long ldo, const int8_t* Wq, const float* Ws, int gs, int M, int K,
                                const int8_t* Xq, const float* XsT, int ldxs, int N) {
  if (!out || !Wq || !Ws || !Xq || !XsT || M < 1 || N < 1 || K < 1 || gs < 1) return THA_STATUS_INVALID_ARG;
  Q8iArgs a = {};
  a.n_mats = 1; a.Wq[0] = Wq; a.Ws[0] = Ws; a.M[0] = M; a.out_off[0] = 0; a.K = K; a.gs = gs; a.Xq = Xq; a.xsT = XsT;
  a.ldxs = ldxs; a.N = N; a.out = out; a.ldo = ldo; a.zstride = 0; a.splits = 1;
  if (!gemm_q8i_ok(a)) return THA_STATUS_NOT_SUPPORTED;
  return gemm_q8i_launch((cudaStream_t)stream, a, false);
}
extern "C" int tha_gemm_check_error(void) { return gemm_tc_check_error(); }

// Development aid: per-k-block clock64 stamps of CTA (0,0,0) of the next tensor-core matmuls.
// out_host[8][64]: 0 W load issued, 1 X load issued, 2 converter saw W, 3 converter got the TMEM slot,
// 4 converter stores complete, 5 MMA thread saw the slot full, 6 MMAs issued, 7 commit issued.
extern "C" int tha_gemm_trace(long long* out_host, int enable) {
  if (enable && !g_trace) {
    THA_CUDA_OK(cudaMalloc(&g_trace, 10 * 64 * sizeof(long long)));
    THA_CUDA_OK(cudaMemset(g_trace, 0, 10 * 64 * sizeof(long long)));
  }
  if (out_host && g_trace) THA_CUDA_OK(cudaMemcpy(out_host, g_trace, 10 * 64 * sizeof(long long), cudaMemcpyDeviceToHost));
  if (!enable && g_trace) { cudaFree(g_trace); g_trace = nullptr; }
  return THA_STATUS_SUCCESS;
}
