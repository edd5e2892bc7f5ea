// A chain of dependent tensor-core weight matmuls in ONE persistent launch (fp32 weights, 3xTF32 as gemm_tc.cu).
//
// The decode step of the reference is, per layer, rmsnorm -> wq|wk|wv -> RoPE -> attention -> wo -> residual ->
// rmsnorm -> w1|w3 -> SwiGLU -> w2 -> residual (src/thaDNN.cpp:39-73).  Everything between two attention calls is a
// chain of four matmuls whose elementwise neighbours only touch the matmul's own output tile (or need one number per
// row: 1/rms).  Issued as four launches plus four glue launches that chain costs ~9 x 9 us of dependent ramp for
// ~15 us of work (stories110M, 256 slots).  Here it is one launch:
//
//   * 148 persistent CTAs walk ONE list of work items (phase, weight tile, column tile, K split) in order, with the
//     roles and rings of gemm_tf32x3_kernel (TMA producers, converter warps, one MMA thread, TMEM accumulator);
//     the weight ring never drains between phases -- weights do not depend on anything, so while a phase waits for its
//     input the next tiles are already in shared memory.
//   * Split-K partials go to global memory (L2); the LAST split to arrive for a tile (arrival counter) sums them in
//     split order -- bit-reproducible -- and runs the tile's fused epilogue:
//       CH_RESID   x += wo/w2 output; the next matmul's input x*gain as tf32 hi/lo; the tile's sums of squares (the
//                  last tile of a column tile turns them into 1/rms, which the next matmul applies as a column scale)
//       CH_SWIGLU  w1 and w3 rows interleaved 16 by 16 inside every tile (8 TMA boxes), so gate and up of a hidden
//                  unit sit 16 lanes apart in one warp: silu(gate) * up as tf32 hi/lo
//       CH_ROPE    RoPE on the q / k pairs (adjacent rows = adjacent lanes), q -> state, k / v -> cache row of pos[n]
//       CH_PLAIN   logits
//   * A phase's activation loads wait on a done-counter of the phase before (per column tile: slots 0..127 and
//     128..255 advance independently), release/acquire at gpu scope plus a generic->async proxy fence because the
//     consumer reads through TMA.
//   * Counters are zero on entry; the last CTA to leave zeroes them again (every CTA has passed its last use by then).
// All CTAs must be co-resident: grid <= SM count, one CTA per SM (227 KB of shared memory each).
#include <stdlib.h>
#include <string.h>

#include <mutex>

#include "chain_tc.h"
#include "gemm_tc.h"
#include "tc_common.cuh"

namespace tha {

constexpr int CH_MAX_MAPS = 16;
constexpr long long CH_DEP_TIMEOUT = 6000000000ll;   // ~3 s of spinning on a done-counter: error flag, not a hang

struct ChainPhase {
  int mode, n_mats;
  int M[3], tiles[3], wmap[3];
  int xmap;                               // tensor maps xmap (hi) and xmap + 1 (lo)
  int kblocks_total, kblocks_per_split, splits;
  int item0, tiles_total;                 // first global work item; items = tiles_total * ny * splits
  int dep, dep_target;                    // done-counter (dep + column tile) the activation loads wait for; -1: none
  int sig, ctr0;                          // done-counter this phase raises (sig + column tile); first arrival counter
  int kvd, hs;
  const float* colscale;
  float* out; long long ldo;
  float* oh; float* ol; long long ldh;
  const float* gain; float* scale_out;
  float* kc; float* vc; long long kv_bs; long long loff;
  const float* rc; const float* rs;
};

struct ChainP {
  CUtensorMap maps[CH_MAX_MAPS];
  ChainPhase ph[CH_MAX_PHASES];
  int n_phases, n_items, N, ny, npad;
  float* part; long long part_ct_stride; float* ss_part;
  const int* pos; const int* slot;
  int* ctr; int n_ctr;
  int* error_flag;
  unsigned long long* trace;   // development aid (tha_chain_trace): [2 CTAs][64 items][16 stamps] of %globaltimer
};

__device__ __forceinline__ unsigned ch_ld_acquire(const int* p) {
  unsigned v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void ch_fence_async_global() { asm volatile("fence.proxy.async.global;" ::: "memory"); }
__device__ __forceinline__ unsigned long long ch_gtime() { unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); return t; }
// stamps: 0 X dependency satisfied, 1 X loads issued, 2 first operand stage ready (MMA thread), 3 MMAs issued,
// 4 accumulator complete (epilogue), 5 TMEM read out / partials stored, 6 all splits of the tile arrived, 7 share finished,
// 8 done-counter raised, 9 W loads of the item issued, 10 item header (phase * 1000 + splits)
#define CH_TRACE(slot, n_it) do { if (P.trace && (blockIdx.x == 0 || blockIdx.x == 97) && (n_it) < 64) \
    P.trace[((blockIdx.x ? 1 : 0) * 64 + (n_it)) * 16 + (slot)] = ch_gtime(); } while (0)
template <int ID>
__device__ __forceinline__ void ch_bar(int n) { asm volatile("bar.sync %0, %1;" ::"n"(ID), "r"(n) : "memory"); }
__device__ __forceinline__ void ch_bar_id(int id, int n) { asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(n) : "memory"); }

// Sum of 16 per-lane values over the 32 lanes of a warp, result for value index (lane >> 1) in s[0] of every lane
// (recursive halving: 16 shuffles, fixed order).
__device__ __forceinline__ float ch_transpose_sum16(float (&s)[16], int lane) {
  const bool b4 = lane & 16, b3 = lane & 8, b2 = lane & 4, b1 = lane & 2;
#pragma unroll
  for (int i = 0; i < 8; i++) {
    const float keep = b4 ? s[i + 8] : s[i], send = b4 ? s[i] : s[i + 8];
    s[i] = keep + __shfl_xor_sync(THA_FULL_MASK, send, 16);
  }
#pragma unroll
  for (int i = 0; i < 4; i++) {
    const float keep = b3 ? s[i + 4] : s[i], send = b3 ? s[i] : s[i + 4];
    s[i] = keep + __shfl_xor_sync(THA_FULL_MASK, send, 8);
  }
#pragma unroll
  for (int i = 0; i < 2; i++) {
    const float keep = b2 ? s[i + 2] : s[i], send = b2 ? s[i] : s[i + 2];
    s[i] = keep + __shfl_xor_sync(THA_FULL_MASK, send, 4);
  }
  {
    const float keep = b1 ? s[1] : s[0], send = b1 ? s[0] : s[1];
    s[0] = keep + __shfl_xor_sync(THA_FULL_MASK, send, 2);
  }
  s[0] += __shfl_xor_sync(THA_FULL_MASK, s[0], 1);
  return s[0];
}

// Columns of a split-K share one epilogue group handles per chunk: the share has NT / Z columns, dealt to the 3 groups.
__host__ __device__ constexpr int ch_cm(int nt, int z) {
  return ((nt / z + TC_CGROUPS - 1) / TC_CGROUPS) > 16 ? 16 : ((nt / z + TC_CGROUPS - 1) / TC_CGROUPS < 1 ? 1 : (nt / z + TC_CGROUPS - 1) / TC_CGROUPS);
}
// v[j] = sum over the Z splits (in split order) of the partial at column j of the chunk, j < cok; all CM x Z loads first.
template <int Z, int CM>
__device__ __forceinline__ void ch_sum_partials(const float* p0, long long zstride, long long ldp, int cok, float (&v)[16]) {
  float t[CM][Z];
#pragma unroll
  for (int j = 0; j < CM; j++) {
#pragma unroll
    for (int z = 0; z < Z; z++) t[j][z] = j < cok ? __ldcg(p0 + (long long)z * zstride + (long long)j * ldp) : 0.0f;
  }
#pragma unroll
  for (int j = 0; j < 16; j++) v[j] = 0.0f;
#pragma unroll
  for (int j = 0; j < CM; j++) {
    float a = t[j][0];
#pragma unroll
    for (int z = 1; z < Z; z++) a += t[j][z];
    v[j] = a;
  }
}

template <int NT>
__global__ void __launch_bounds__(TC_THREADS, 1) chain_tf32x3_kernel(const __grid_constant__ ChainP P) {
  using S = TcSmem<NT>;
  extern __shared__ unsigned char tc_smem_raw[];
  __shared__ float s_red[TC_CGROUPS][3][4][16];     // sums of squares: [group][chunk of the group][warp quarter][column]
  __shared__ float s_cscale[NT];                    // per column of the current item: 1/rms, position, cache slab
  __shared__ int s_cpos[NT], s_cslot[NT];
  __shared__ int s_flag[2];                         // [0] last CTA to leave the kernel, [1] last share of a column tile (CH_RESID)
  unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(tc_smem_raw) + 1023) & ~(uintptr_t)1023);
  unsigned char* wring = smem;
  unsigned char* xring = smem + S::SW * S::W_STAGE;
  uint64_t* bars = reinterpret_cast<uint64_t*>(xring + S::SX * S::XS_STAGE);
  uint64_t* wfull = bars;
  uint64_t* wfree = wfull + S::SW;
  uint64_t* ready = wfree + S::SW;
  uint64_t* aempty = ready + S::SA;
  uint64_t* accum = aempty + S::SA;
  uint64_t* accfree = accum + 1;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(accfree + 1);
  int* abort_flag = reinterpret_cast<int*>(tmem_slot + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  struct Item { int ph, mat, tile, row0, n0, ct, z, kb0, nkb, nst; };
  auto item_of = [&](int it) -> Item {
    Item w;
    w.ph = 0;
#pragma unroll
    for (int k = 1; k < CH_MAX_PHASES; k++) if (k < P.n_phases && it >= P.ph[k].item0) w.ph = k;
    const ChainPhase& f = P.ph[w.ph];
    const int li = it - f.item0;
    w.z = li % f.splits;
    const int t = li / f.splits;
    w.tile = t / P.ny;
    w.ct = t - w.tile * P.ny;
    w.n0 = w.ct * NT;
    int tile = w.tile;
    w.mat = 0;
    if (f.mode != CH_SWIGLU) {
      if (tile >= f.tiles[0]) { w.mat = 1; tile -= f.tiles[0]; if (tile >= f.tiles[1]) { w.mat = 2; tile -= f.tiles[1]; } }
      w.row0 = tile * TC_BM;
    } else {
      w.row0 = tile * (TC_BM / 2);           // hidden units of this tile: 64 (w1 rows and the same w3 rows)
    }
    w.kb0 = w.z * f.kblocks_per_split;
    w.nkb = f.kblocks_total - w.kb0;
    if (w.nkb > f.kblocks_per_split) w.nkb = f.kblocks_per_split;
    w.nst = (w.nkb + S::KG - 1) / S::KG;
    return w;
  };

  if (threadIdx.x == 0) {
    for (int s = 0; s < S::SW; s++) { tc_mbar_init(&wfull[s], 1); tc_mbar_init(&wfree[s], 4); }
    for (int s = 0; s < S::SA; s++) { tc_mbar_init(&ready[s], 5); tc_mbar_init(&aempty[s], 1); }
    tc_mbar_init(accum, 1);
    tc_mbar_init(accfree, 4 * TC_CGROUPS);
    *abort_flag = 0;
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  if (warp == 2) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tc_smem_u32(tmem_slot)), "r"(S::TMEM_COLS));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::);
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem_base = *tmem_slot;
  const uint32_t tmem_d = tmem_base;
  const uint32_t tmem_dl = tmem_base + NT;
  const uint32_t tmem_a = tmem_base + S::ACC_COLS;

  if (warp == 0) {
    // ===== W producer: never waits for anything but ring space -- runs ahead through the phases
    if (lane == 0) {
      int gb = 0, n_it = 0;
      bool ok = true;
      for (int it = blockIdx.x; it < P.n_items && ok; it += gridDim.x, n_it++) {
        const Item w = item_of(it);
        const ChainPhase& f = P.ph[w.ph];
        for (int i = 0; i < w.nst; i++) {
          const int g = gb + i, s = g % S::SW;
          if (i == w.nst - 1) CH_TRACE(9, n_it);
          if (g >= S::SW && !tc_mbar_wait(&wfree[s], ((g / S::SW) & 1u) ^ 1u, abort_flag)) { ok = false; break; }
          const int nk = (w.nkb - i * S::KG) < S::KG ? (w.nkb - i * S::KG) : S::KG;
          tc_mbar_expect_tx(&wfull[s], nk * S::W_BYTES);
          for (int kk = 0; kk < nk; kk++) {
            unsigned char* dst = wring + s * S::W_STAGE + kk * S::W_BYTES;
            const int kc = (w.kb0 + i * S::KG + kk) * TC_BK;
            if (f.mode != CH_SWIGLU) {
              tc_tma_2d(dst, &P.maps[f.wmap[w.mat]], kc, w.row0, &wfull[s]);
            } else {
#pragma unroll
              for (int q = 0; q < 4; q++) {   // lanes 32q..32q+15: w1 rows, 32q+16..32q+31: the same rows of w3
                tc_tma_2d(dst + (q * 32) * 128, &P.maps[f.wmap[0]], kc, w.row0 + 16 * q, &wfull[s]);
                tc_tma_2d(dst + (q * 32 + 16) * 128, &P.maps[f.wmap[1]], kc, w.row0 + 16 * q, &wfull[s]);
              }
            }
          }
        }
        gb += w.nst;
      }
    }
  } else if (warp == 3) {
    // ===== X producer: the activations of a phase exist once the phase before has finished this column tile
    if (lane == 0) {
      pdl_wait();
      int gb = 0, n_it = 0;
      bool ok = true;
      for (int it = blockIdx.x; it < P.n_items && ok; it += gridDim.x, n_it++) {
        const Item w = item_of(it);
        const ChainPhase& f = P.ph[w.ph];
        if (P.trace && (blockIdx.x == 0 || blockIdx.x == 97) && n_it < 64)
          P.trace[((blockIdx.x ? 1 : 0) * 64 + n_it) * 16 + 10] = (unsigned long long)(w.ph * 1000 + f.splits);
        if (f.dep >= 0) {
          const int* c = P.ctr + f.dep + w.ct;
          const long long t0 = tc_clock();
          while ((int)(ch_ld_acquire(c) - (unsigned)f.dep_target) < 0) {
            if (*((volatile int*)abort_flag)) { ok = false; break; }
            if (tc_clock() - t0 > CH_DEP_TIMEOUT) { *abort_flag = 1; ok = false; break; }
          }
          if (!ok) break;
          ch_fence_async_global();   // the rows were written with generic stores; TMA reads them through the async proxy
        }
        CH_TRACE(0, n_it);
        const CUtensorMap* mxh = &P.maps[f.xmap];
        const CUtensorMap* mxl = &P.maps[f.xmap + 1];
        for (int i = 0; i < w.nst; i++) {
          const int g = gb + i, s = g % S::SX;
          if (g >= S::SX && !tc_mbar_wait(&aempty[s], ((g / S::SX) & 1u) ^ 1u, abort_flag)) { ok = false; break; }
          const int nk = (w.nkb - i * S::KG) < S::KG ? (w.nkb - i * S::KG) : S::KG;
          tc_mbar_expect_tx(&ready[s], nk * S::XS_BYTES);
          for (int kk = 0; kk < nk; kk++) {
            unsigned char* st = xring + s * S::XS_STAGE + kk * S::XS_BYTES;
            tc_tma_2d(st, mxh, (w.kb0 + i * S::KG + kk) * TC_BK, w.n0, &ready[s]);
            tc_tma_2d(st + S::X_BYTES, mxl, (w.kb0 + i * S::KG + kk) * TC_BK, w.n0, &ready[s]);
          }
        }
        CH_TRACE(1, n_it);
        gb += w.nst;
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer (as gemm_tf32x3_kernel)
    if (lane == 0) {
      constexpr uint32_t idesc = tc_idesc(NT), idesc2 = tc_idesc(2 * NT);
      int gb = 0, n_it = 0;
      bool ok = true;
      for (int it = blockIdx.x; it < P.n_items && ok; it += gridDim.x, n_it++) {
        const Item w = item_of(it);
        if (n_it > 0) {
          if (!tc_mbar_wait(accfree, (n_it - 1) & 1u, abort_flag)) break;
          asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        }
        for (int i = 0; i < w.nst; i++) {
          const int g = gb + i, sa = g % S::SA;
          if (!tc_mbar_wait(&ready[sa], (g / S::SA) & 1u, abort_flag)) { ok = false; break; }
          asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
          if (i == 0) CH_TRACE(2, n_it);
          const int nk = (w.nkb - i * S::KG) < S::KG ? (w.nkb - i * S::KG) : S::KG;
#pragma unroll
          for (int kk = 0; kk < S::KG; kk++) {
            if (kk < nk) {
              const uint64_t xhi = tc_smem_desc(tc_smem_u32(xring + sa * S::XS_STAGE + kk * S::XS_BYTES));
              const uint32_t ahi = tmem_a + sa * S::A_STAGE + kk * S::A_COLS, alo = ahi + TC_BK;
#pragma unroll
              for (int k = 0; k < TC_BK / 8; k++) {
                const uint64_t adv = (uint64_t)(k * 2);
                tc_mma_tf32_ts(tmem_d, ahi + 8 * k, xhi + adv, idesc2, (uint32_t)((i | kk | k) != 0));
                tc_mma_tf32_ts(tmem_dl, alo + 8 * k, xhi + adv, idesc, 1u);
              }
            }
          }
          tc_commit(&aempty[sa]);
        }
        if (ok) tc_commit(accum);
        CH_TRACE(3, n_it);
        gb += w.nst;
      }
    }
  } else if (warp >= 4) {
    // ===== converters, then the item's epilogue.  Thread = row of the tile (TMEM lane).
    const int q = warp & 3;
    const int r = q * 32 + lane;
    const int grp = (warp >> 2) - 1;
    const int et = threadIdx.x - 128;                  // 0 .. 383 among the epilogue threads
    constexpr int NEPI = 128 * TC_CGROUPS;
    // After a timeout (ok == false) these warps keep walking the items with every wait short-circuited: the named
    // barriers below must see all 12 warps the same number of times, or a broken pipeline would hang instead of ending.
    bool ok = true;
    int gb = 0, n_it = 0;
    for (int it = blockIdx.x; it < P.n_items; it += gridDim.x, n_it++) {
      const Item w = item_of(it);
      const ChainPhase& f = P.ph[w.ph];
      int i0 = (grp - gb % TC_CGROUPS + TC_CGROUPS) % TC_CGROUPS;
      for (int i = i0; i < w.nst && ok; i += TC_CGROUPS) {
        const int g = gb + i, sw = g % S::SW, sa = g % S::SA;
        const int nk = (w.nkb - i * S::KG) < S::KG ? (w.nkb - i * S::KG) : S::KG;
        ok = tc_mbar_wait(&wfull[sw], (g / S::SW) & 1u, abort_flag);
#pragma unroll
        for (int kk = 0; kk < S::KG; kk++) {
          if (kk < nk && ok) {
            uint32_t hi[32], lo[32];
            const unsigned char* wrow = wring + sw * S::W_STAGE + kk * S::W_BYTES + r * 128;
#pragma unroll
            for (int j = 0; j < 8; j++) {
              const float4 v = *reinterpret_cast<const float4*>(wrow + ((j ^ (r & 7)) << 4));
              const float h0 = tc_tf32(v.x), h1 = tc_tf32(v.y), h2 = tc_tf32(v.z), h3 = tc_tf32(v.w);
              hi[4 * j + 0] = __float_as_uint(h0); hi[4 * j + 1] = __float_as_uint(h1);
              hi[4 * j + 2] = __float_as_uint(h2); hi[4 * j + 3] = __float_as_uint(h3);
              lo[4 * j + 0] = __float_as_uint(v.x - h0); lo[4 * j + 1] = __float_as_uint(v.y - h1);
              lo[4 * j + 2] = __float_as_uint(v.z - h2); lo[4 * j + 3] = __float_as_uint(v.w - h3);
            }
            if (kk == 0 && g >= S::SA && ok) ok = tc_mbar_wait(&aempty[sa], ((g / S::SA) & 1u) ^ 1u, abort_flag);
            const uint32_t ta = tmem_a + sa * S::A_STAGE + kk * S::A_COLS + ((uint32_t)(q * 32) << 16);
            tc_tmem_st32(ta, hi);
            tc_tmem_st32(ta + TC_BK, lo);
          }
        }
        __syncwarp();
        if (lane == 0) tc_mbar_arrive(&wfree[sw]);
        asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
        __syncwarp();
        if (lane == 0) tc_mbar_arrive(&ready[sa]);
      }
      // ---------------------------------------------------------------- epilogue of the item
      if (n_it == 0) {
        pdl_trigger();
        pdl_wait();
      }
      if (ok) ok = tc_mbar_wait(accum, n_it & 1u, abort_flag);
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      if (et == 0) CH_TRACE(4, n_it);
      const int mode = f.mode;
      // row of this thread in the phase's output, and whether it exists
      int orow;                                       // index along the output feature axis
      bool row_ok;
      if (mode == CH_SWIGLU) { orow = w.row0 + q * 16 + (lane & 15); row_ok = orow < f.M[0]; }
      else { orow = w.row0 + r; row_ok = orow < f.M[w.mat]; }
      // split-K partials: one region per column tile (column tiles run through the phases independently), inside it
      // [split][column of the tile][row of the phase's output, tile-major]
      const long long ldp = (long long)f.tiles_total * TC_BM;
      const long long zstride = (long long)NT * ldp;
      float* prow = P.part + (long long)w.ct * P.part_ct_stride + (long long)w.tile * TC_BM + r;   // + z * zstride + (n - n0) * ldp
      const uint32_t tq = tmem_d + ((uint32_t)(q * 32) << 16);
      // what every thread needs to know about a column (the slot's 1/rms, position, cache slab): one L2 round trip for the
      // CTA instead of one per thread and chunk.  Written here, read after the next epilogue-wide barrier.
      if (et < NT) {
        const int n = w.n0 + et;
        const bool in = n < P.N;
        s_cscale[et] = (in && f.colscale) ? __ldcg(f.colscale + n) : 1.0f;
        s_cpos[et] = in ? __ldg(P.pos + n) : -1;
        s_cslot[et] = (in && P.slot) ? __ldg(P.slot + n) : n;
      }

      // the fused epilogue of one chunk: v[j] = full-K sum of (this thread's row, column n0 + c + j), j < cmax;
      // xin[j] (CH_RESID): the residual stream's current value at that place
      auto finish_chunk = [&](int c, int ci, float (&v)[16], int cmax, const float (&xin)[16]) {
        if (mode == CH_PLAIN) {
#pragma unroll
          for (int j = 0; j < 16; j++) {
            const int n = w.n0 + c + j;
            if (j < cmax && n < P.N && row_ok) f.out[(size_t)n * f.ldo + orow] = v[j] * s_cscale[c + j];
          }
        } else if (mode == CH_RESID) {
          float ssq[16];
          const float gn = row_ok ? __ldg(f.gain + orow) : 0.0f;
#pragma unroll
          for (int j = 0; j < 16; j++) {
            const int n = w.n0 + c + j;
            ssq[j] = 0.0f;
            if (j < cmax && n < P.N && row_ok) {
              const float xo = xin[j] + v[j];
              f.out[(size_t)n * f.ldo + orow] = xo;
              ssq[j] = xo * xo;
              float gx = xo * gn;
              if (s_cpos[c + j] < 0) gx = 0.0f;              // idle slot: zeros into the next matmul
              const float h = tc_tf32(gx);
              f.oh[(size_t)n * f.ldh + orow] = h;
              f.ol[(size_t)n * f.ldh + orow] = gx - h;
            }
          }
          const float tot = ch_transpose_sum16(ssq, lane);
          if ((lane & 1) == 0) s_red[grp][ci][q][lane >> 1] = tot;
          ch_bar_id(2 + grp, 128);
          if (q == 0 && lane < 16) {
            const int n = w.n0 + c + lane;
            const float t4 = ((s_red[grp][ci][0][lane] + s_red[grp][ci][1][lane]) + s_red[grp][ci][2][lane]) + s_red[grp][ci][3][lane];
            if (lane < cmax) P.ss_part[(size_t)w.tile * P.npad + n] = t4;
          }
        } else if (mode == CH_SWIGLU) {
#pragma unroll
          for (int j = 0; j < 16; j++) {
            const int n = w.n0 + c + j;
            const bool col_ok = j < cmax && n < P.N;
            const float a = v[j] * (j < cmax ? s_cscale[c + j] : 1.0f);
            const float up = __shfl_down_sync(THA_FULL_MASK, a, 16);
            if (lane < 16 && col_ok && row_ok) {
              float rr = a;
              rr *= (1.0f / (1.0f + expf(-rr)));      // run.cc:344-347 operation order
              rr *= up;
              if (s_cpos[c + j] < 0) rr = 0.0f;
              const float h = tc_tf32(rr);
              f.oh[(size_t)n * f.ldh + orow] = h;
              f.ol[(size_t)n * f.ldh + orow] = rr - h;
            }
          }
        } else {   // CH_ROPE
          const int half = f.hs >> 1;
          const int jj = (orow % f.hs) >> 1;
          float cs[16], sn[16];
          if (w.mat < 2) {
#pragma unroll
            for (int j = 0; j < 16; j++) {               // all table loads of the chunk in flight together
              const int pb = j < cmax ? s_cpos[c + j] : -1;
              cs[j] = (pb >= 0 && row_ok) ? __ldg(f.rc + (size_t)pb * half + jj) : 1.0f;
              sn[j] = (pb >= 0 && row_ok) ? __ldg(f.rs + (size_t)pb * half + jj) : 0.0f;
            }
          }
#pragma unroll
          for (int j = 0; j < 16; j++) {
            const int n = w.n0 + c + j;
            const int pb = j < cmax ? s_cpos[c + j] : -1;
            const float a = v[j] * (j < cmax ? s_cscale[c + j] : 1.0f);
            const float other = __shfl_xor_sync(THA_FULL_MASK, a, 1);
            if (pb >= 0 && n < P.N && row_ok) {
              float o = a;
              if (w.mat < 2) o = (lane & 1) ? (other * sn[j] + a * cs[j]) : (a * cs[j] - other * sn[j]);   // run.cc:271-286
              if (w.mat == 0) {
                f.out[(size_t)n * f.ldo + orow] = o;
              } else {
                float* dst = (w.mat == 1 ? f.kc : f.vc) + (size_t)s_cslot[c + j] * f.kv_bs + f.loff + (size_t)pb * f.kvd + orow;
                *dst = o;
              }
            }
          }
        }
      };

      const bool direct = f.splits == 1;
      if (direct) ch_bar<1>(NEPI);                    // the per-column values above are in place
      {
        int ci = 0;
#pragma unroll 1
        for (int c = 16 * grp; c < NT; c += 16 * TC_CGROUPS, ci++) {
          uint32_t a0[16], a1[16];
          tc_tmem_ld16(tq + (uint32_t)c, a0);
          tc_tmem_ld16(tq + (uint32_t)(c + NT), a1);
          float xin[16];
#pragma unroll
          for (int j = 0; j < 16; j++)
            xin[j] = (direct && mode == CH_RESID && (w.n0 + c + j) < P.N && row_ok) ? __ldcg(f.out + (size_t)(w.n0 + c + j) * f.ldo + orow) : 0.0f;
          asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
          float v[16];
#pragma unroll
          for (int j = 0; j < 16; j++) v[j] = __uint_as_float(a0[j]) + __uint_as_float(a1[j]);
          if (direct) {
            finish_chunk(c, ci, v, 16, xin);
          } else {
#pragma unroll
            for (int j = 0; j < 16; j++) {
              const int n = w.n0 + c + j;
              if (n < P.N) __stcg(prow + (long long)w.z * zstride + (long long)(c + j) * ldp, v[j]);
            }
          }
        }
      }
      asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
      __syncwarp();
      if (lane == 0) tc_mbar_arrive(accfree);           // the next item's MMAs may overwrite the accumulator
      if (et == 0) CH_TRACE(5, n_it);

      if (!direct) {
        // ---- split-K: once every split of this tile has stored its partials, each of them finishes ITS share of the
        // tile's columns (NT / splits of them): the reduction and the fused epilogue of a tile are spread over the CTAs
        // that computed it instead of falling to whoever arrives last.  The splits of a tile are consecutive work items
        // (= neighbouring CTAs at the same point of their lists), so the wait is short; it cannot deadlock: every CTA
        // walks its items in order, and an unfinished item with the lowest index only ever waits for splits on CTAs
        // that have nothing older left.
        ch_bar<1>(NEPI);
        if (et == 0) {
          __threadfence();
          int* tc = P.ctr + f.ctr0 + w.tile * P.ny + w.ct;
          atomicAdd(tc, 1);
          const long long t0 = tc_clock();
          while ((int)(ch_ld_acquire(tc) - (unsigned)f.splits) < 0) {
            if (*((volatile int*)abort_flag)) break;
            if (tc_clock() - t0 > CH_DEP_TIMEOUT) { *abort_flag = 1; break; }
          }
          __threadfence();
          CH_TRACE(6, n_it);
        }
        ch_bar<1>(NEPI);
        // This group's columns of the share, at most 16 at a time; every load of a chunk (all splits, and the residual
        // stream's values) is issued before the first one is used: one L2 round trip per chunk.
        const int ncol = NT / f.splits;                 // host: splits divides NT
        const int per_g = (ncol + TC_CGROUPS - 1) / TC_CGROUPS;
        const int g0 = grp * per_g;
        int g1 = g0 + per_g; if (g1 > ncol) g1 = ncol;
        int ci = 0;
#pragma unroll 1
        for (int cc = g0; cc < g1 || ci == 0; cc += 16, ci++) {   // at least one (possibly empty) chunk: the group barriers inside must match
          const int cmax = (g1 - cc) < 0 ? 0 : ((g1 - cc) > 16 ? 16 : (g1 - cc));
          const int c = w.z * ncol + cc;
          int cok = P.N - (w.n0 + c);                   // columns of the chunk that exist
          if (cok > cmax) cok = cmax;
          const float* p0 = prow + (long long)c * ldp;
          float v[16], xin[16];
#pragma unroll
          for (int j = 0; j < 16; j++)
            xin[j] = (mode == CH_RESID && j < cok && row_ok) ? __ldcg(f.out + (size_t)(w.n0 + c + j) * f.ldo + orow) : 0.0f;
          switch (f.splits) {
            case 2: ch_sum_partials<2, ch_cm(NT, 2)>(p0, zstride, ldp, cok, v); break;
            case 4: ch_sum_partials<4, ch_cm(NT, 4)>(p0, zstride, ldp, cok, v); break;
            case 8: ch_sum_partials<8, ch_cm(NT, 8)>(p0, zstride, ldp, cok, v); break;
            default: ch_sum_partials<16, ch_cm(NT, 16)>(p0, zstride, ldp, cok, v); break;
          }
          finish_chunk(c, ci, v, cmax, xin);
        }
      }
      {
        // ---- this item's share of the tile is in place: raise the phase's done-counter of this column tile
        ch_fence_async_global();
        ch_bar<1>(NEPI);
        if (et == 0) {
          CH_TRACE(7, n_it);
          __threadfence();
          const int old = atomicAdd(P.ctr + f.sig + w.ct, 1);
          s_flag[1] = (mode == CH_RESID && old == f.tiles_total * f.splits - 1) ? 1 : 0;
          __threadfence();
        }
        if (mode == CH_RESID) {
          ch_bar<1>(NEPI);
          if (s_flag[1]) {
            // last share of the column tile: sums of squares of every row tile -> 1/rms (run.cc:196-209)
            if (et < NT) {
              const int n = w.n0 + et;
              if (n < P.N) {
                float ss = 0.0f;
                for (int t = 0; t < f.tiles_total; t += 8) {    // eight loads in flight, added in tile order
                  float p8[8];
#pragma unroll
                  for (int u = 0; u < 8; u++) p8[u] = (t + u < f.tiles_total) ? __ldcg(P.ss_part + (size_t)(t + u) * P.npad + n) : 0.0f;
#pragma unroll
                  for (int u = 0; u < 8; u++) ss += p8[u];
                }
                ss /= (float)f.M[0];
                ss += 1e-5f;
                f.scale_out[n] = 1.0f / sqrtf(ss);
              }
            }
            ch_bar<1>(NEPI);
            if (et == 0) {
              __threadfence();
              atomicAdd(P.ctr + f.sig + w.ct, 1);
            }
          }
        }
      }
      if (et == 0) CH_TRACE(8, n_it);
      ch_bar<1>(NEPI);    // s_flag / s_red / the per-column values are reused by the next item
      gb += w.nst;
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  }
  __syncthreads();
  if (warp == 2) {
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(S::TMEM_COLS));
  }
  if (threadIdx.x == 0 && *abort_flag && P.error_flag) *((volatile int*)P.error_flag) = 1;
  // ---- the last CTA to leave zeroes the counters for the next launch
  if (threadIdx.x == 0) {
    __threadfence();
    s_flag[0] = (atomicAdd(P.ctr, 1) == (int)gridDim.x - 1) ? 1 : 0;
  }
  __syncthreads();
  if (s_flag[0]) {
    for (int i = threadIdx.x; i < P.n_ctr; i += blockDim.x) P.ctr[i] = 0;
  }
}

// ------------------------------------------------------------------------------------------------ host side ----
// Split-K factor of a phase.  CTAs that have no item in a phase are not lost -- they run ahead into the next phase's
// weight loads -- and the splits of a tile share its reduction and epilogue, so a phase is fastest when its items fill
// ONE wave of the SMs: the largest power of two (it has to divide the column tile) <= 148 / (tiles x column tiles), with
// at least two pipeline stages per item; 1 for phases with more tiles than SMs (the classifier).
static int chain_splits(int kb, int tile_items, int kg, int sms, int nt) {
  int zmax = kb / (2 * kg);
  if (zmax > 16) zmax = 16;
  if (zmax > nt) zmax = nt;
  int z = 1;
  while (2 * z <= zmax && 2 * z * tile_items <= sms) z *= 2;
  return z;
}

struct ChainPlan {
  int NT, ny, npad, n_items, n_ctr;
  int tiles_total[CH_MAX_PHASES], tiles[CH_MAX_PHASES][3], splits[CH_MAX_PHASES], per[CH_MAX_PHASES], item0[CH_MAX_PHASES];
  int sig[CH_MAX_PHASES], ctr0[CH_MAX_PHASES];
  size_t part_floats, part_ct_stride, ss_floats;
};

static void chain_plan(const ChainArgs& a, ChainPlan* p) {
  p->NT = gemm_tc_pick_nt(a.N);
  p->ny = ceil_div(a.N, p->NT);
  p->npad = p->ny * p->NT;
  const int kg = p->NT <= 64 ? 2 : 1;
  int items = 0, ctr = 1;          // ctr[0]: CTAs that have left
  p->part_floats = 0; p->part_ct_stride = 0; p->ss_floats = 0;
  for (int i = 0; i < a.n; i++) { p->sig[i] = ctr; ctr += p->ny; }
  for (int i = 0; i < a.n; i++) {
    const ChainMatmul& m = a.mm[i];
    int tt = 0;
    for (int k = 0; k < 3; k++) p->tiles[i][k] = 0;
    if (m.mode == CH_SWIGLU) { p->tiles[i][0] = ceil_div(m.M[0], TC_BM / 2); tt = p->tiles[i][0]; }
    else for (int k = 0; k < m.n_mats; k++) { p->tiles[i][k] = ceil_div(m.M[k], TC_BM); tt += p->tiles[i][k]; }
    p->tiles_total[i] = tt;
    const int kb = ceil_div(m.K, TC_BK);
    int z = m.splits > 0 ? m.splits : chain_splits(kb, tt * p->ny, kg, 148, p->NT);
    int per = kb;
    for (; z >= 1; z--) {   // the splits share the tile's columns (z divides the column tile) and are whole pipeline stages, none empty
      per = ceil_div(ceil_div(kb, z), kg) * kg;
      if (p->NT % z == 0 && ceil_div(kb, per) == z) break;
    }
    if (z < 1) { z = 1; per = ceil_div(kb, kg) * kg; }
    p->splits[i] = z; p->per[i] = per;
    p->item0[i] = items;
    items += tt * p->ny * z;
    p->ctr0[i] = ctr;
    if (z > 1) {
      ctr += tt * p->ny;
      const size_t pf = (size_t)z * p->NT * tt * TC_BM;
      if (pf > p->part_ct_stride) p->part_ct_stride = pf;
    }
    if (m.mode == CH_RESID) {
      const size_t sf = (size_t)tt * p->npad;
      if (sf > p->ss_floats) p->ss_floats = sf;
    }
  }
  p->part_floats = p->part_ct_stride * p->ny;
  p->n_items = items;
  p->n_ctr = ctr;
}

bool chain_tc_ok(const ChainArgs& a) {
  if (a.n < 1 || a.n > CH_MAX_PHASES || a.N < 1 || !a.pos) return false;
  int maps = 0;
  for (int i = 0; i < a.n; i++) {
    const ChainMatmul& m = a.mm[i];
    if (m.n_mats < 1 || m.n_mats > 3 || m.K % 4 || m.K < TC_BK || !m.Xhi || !m.Xlo) return false;
    if ((reinterpret_cast<uintptr_t>(m.Xhi) & 15u) || (reinterpret_cast<uintptr_t>(m.Xlo) & 15u)) return false;
    for (int k = 0; k < m.n_mats; k++)
      if (!m.W[k] || (reinterpret_cast<uintptr_t>(m.W[k]) & 15u) || m.M[k] < 1) return false;
    maps += m.n_mats + 2;
    switch (m.mode) {
      case CH_PLAIN: if (!m.out || m.n_mats != 1) return false; break;
      case CH_RESID: if (!m.out || !m.oh || !m.ol || !m.gain || !m.scale_out || m.n_mats != 1) return false; break;
      case CH_SWIGLU: if (!m.oh || !m.ol || m.n_mats != 2 || m.M[0] != m.M[1] || m.M[0] % 16) return false; break;
      case CH_ROPE:
        if (!m.out || !m.kc || !m.vc || !m.rc || !m.rs || m.n_mats != 3 || (m.hs & 1) || m.hs < 2 || m.M[0] % m.hs || m.M[1] % m.hs ||
            m.M[1] != m.kvd || m.M[2] != m.kvd) return false;
        break;
      default: return false;
    }
  }
  return maps <= CH_MAX_MAPS;
}

void chain_tc_need(const ChainArgs& a, size_t* part_floats, size_t* ss_floats, int* ctr_ints) {
  ChainPlan p;
  chain_plan(a, &p);
  if (part_floats) *part_floats = p.part_floats;
  if (ss_floats) *ss_floats = p.ss_floats;
  if (ctr_ints) *ctr_ints = p.n_ctr;
}

static unsigned long long* g_chain_trace = nullptr;
static int g_chain_trace_countdown = 0;   // trace the n-th chain launch from now

template <int NT>
static int chain_launch_nt(cudaStream_t st, const ChainP& P, int grid, bool pdl) {
  using S = TcSmem<NT>;
  auto k = chain_tf32x3_kernel<NT>;
  THA_CUDA_OK(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, S::TOTAL));
  THA_CUDA_OK(launch(k, dim3(grid), dim3(TC_THREADS), (size_t)S::TOTAL, st, pdl, P));
  return THA_STATUS_SUCCESS;
}

int chain_tc_launch(cudaStream_t st, const ChainArgs& a, bool pdl) {
  if (!chain_tc_ok(a)) return THA_STATUS_INVALID_ARG;
  ChainPlan pl;
  chain_plan(a, &pl);
  if (pl.part_floats > a.part_floats || pl.ss_floats > a.ss_floats || pl.n_ctr > a.ctr_ints || !a.ctr) return THA_STATUS_INVALID_ARG;
  if (pl.part_floats && !a.part) return THA_STATUS_INVALID_ARG;
  if (pl.ss_floats && !a.ss_part) return THA_STATUS_INVALID_ARG;
  static ChainP P;   // 3.3 KB: kept off the stack; filled and consumed under the lock below
  static std::mutex mu;
  std::lock_guard<std::mutex> lk(mu);
  memset(&P, 0, sizeof(P));
  int nm = 0;
  for (int i = 0; i < a.n; i++) {
    const ChainMatmul& m = a.mm[i];
    ChainPhase& f = P.ph[i];
    f.mode = m.mode; f.n_mats = m.n_mats;
    for (int k = 0; k < 3; k++) { f.M[k] = k < m.n_mats ? m.M[k] : 0; f.tiles[k] = pl.tiles[i][k]; f.wmap[k] = 0; }
    for (int k = 0; k < m.n_mats; k++) {
      THA_TRY_RC(tc_map(m.W[k], m.M[k], m.K, m.K, m.mode == CH_SWIGLU ? 16 : TC_BM, &P.maps[nm]));
      f.wmap[k] = nm++;
    }
    THA_TRY_RC(tc_map(m.Xhi, a.N, m.K, m.K, pl.NT, &P.maps[nm]));
    f.xmap = nm++;
    THA_TRY_RC(tc_map(m.Xlo, a.N, m.K, m.K, pl.NT, &P.maps[nm]));
    nm++;
    f.kblocks_total = ceil_div(m.K, TC_BK); f.kblocks_per_split = pl.per[i]; f.splits = pl.splits[i];
    f.item0 = pl.item0[i]; f.tiles_total = pl.tiles_total[i];
    f.dep = i == 0 ? -1 : pl.sig[i - 1];
    f.dep_target = i == 0 ? 0 : pl.tiles_total[i - 1] * pl.splits[i - 1] + (a.mm[i - 1].mode == CH_RESID ? 1 : 0);
    f.sig = pl.sig[i]; f.ctr0 = pl.ctr0[i];
    f.kvd = m.kvd; f.hs = m.hs;
    f.colscale = m.colscale; f.out = m.out; f.ldo = m.ldo; f.oh = m.oh; f.ol = m.ol;
    f.ldh = m.mode == CH_SWIGLU ? m.M[0] : (m.mode == CH_RESID ? m.M[0] : 0);
    f.gain = m.gain; f.scale_out = m.scale_out;
    f.kc = m.kc; f.vc = m.vc; f.kv_bs = m.kv_bs; f.loff = m.loff; f.rc = m.rc; f.rs = m.rs;
  }
  P.n_phases = a.n; P.n_items = pl.n_items; P.N = a.N; P.ny = pl.ny; P.npad = pl.npad;
  P.part = a.part; P.part_ct_stride = (long long)pl.part_ct_stride; P.ss_part = a.ss_part; P.pos = a.pos; P.slot = a.slot; P.ctr = a.ctr; P.n_ctr = pl.n_ctr;
  THA_TRY_RC(gemm_tc_init(&P.error_flag));
  P.trace = (g_chain_trace && --g_chain_trace_countdown == 0) ? g_chain_trace : nullptr;
  int dev = 0, sms = 148;
  if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  const int grid = pl.n_items < sms ? pl.n_items : sms;
  switch (pl.NT) {
    case 16: return chain_launch_nt<16>(st, P, grid, pdl);
    case 32: return chain_launch_nt<32>(st, P, grid, pdl);
    case 64: return chain_launch_nt<64>(st, P, grid, pdl);
    default: return chain_launch_nt<128>(st, P, grid, pdl);
  }
}

}  // namespace tha

// Development aid: %globaltimer stamps of CTAs 0 and 97 of the LAST chain launched while tracing was on
// (out_host[2][64][16], see CH_TRACE).  enable = n > 0: allocate and trace the n-th chain launch from now; 0 frees.
extern "C" int tha_chain_trace(unsigned long long* out_host, int enable) {
  using namespace tha;
  const size_t n = 2 * 64 * 16;
  if (enable && !g_chain_trace) {
    THA_CUDA_OK(cudaMalloc(&g_chain_trace, n * sizeof(unsigned long long)));
    THA_CUDA_OK(cudaMemset(g_chain_trace, 0, n * sizeof(unsigned long long)));
  }
  if (enable) g_chain_trace_countdown = enable;
  if (out_host && g_chain_trace) THA_CUDA_OK(cudaMemcpy(out_host, g_chain_trace, n * sizeof(unsigned long long), cudaMemcpyDeviceToHost));
  if (!enable && g_chain_trace) { cudaFree(g_chain_trace); g_chain_trace = nullptr; }
  return THA_STATUS_SUCCESS;
}
