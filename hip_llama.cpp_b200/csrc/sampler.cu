// The sampler of the driver on the device (reference: src/llama.cpp:262-399 == run.cc:566-720): temperature softmax,
// multinomial or top-p (nucleus) selection, xorshift* coin per request.  One CTA per slot; the B x vocab logits never
// leave HBM (the host sampler reads 32.8 MB per step at 256 slots and qsorts up to 32000 candidates per slot).
//
// What is kept from the reference, and what cannot be:
//   * the random stream is identical (integer xorshift*, coin = (u32 >> 8) / 2^24, one coin per SAMPLED token, none while
//     a prompt token is forced);
//   * the selection rule is identical as a statement about real numbers: candidates p >= (1 - topp) / (n - 1), sorted by
//     descending probability (equal probabilities in index order, as a stable sort leaves them), nucleus = shortest prefix
//     whose mass exceeds topp, token = first of the prefix whose cumulative mass exceeds coin * nucleus mass;
//   * the reference evaluates those sums in fp32, sequentially in sorted order, on probabilities from the host's expf.
//     Neither the libm bits nor a 32000-term sequential rounding pattern can be reproduced by a parallel device
//     kernel, so masses are accumulated here as 64-bit integers in units of 2^-40 (exact, order-independent, bit-
//     reproducible run to run).  The chosen token therefore equals the reference's unless the coin falls within fp32
//     rounding (~1e-5 of the mass) of a boundary between two tokens (tests/test_gpu_sampler.py states exactly that).
// No sort: the position where a prefix mass crosses a target is found by a three-level radix descent over the
// probability's bit pattern (12 + 10 + 10 bits, mass histograms in shared memory), then by rank among equal keys.
#include <stdlib.h>

#include "common.cuh"
#include "workspace.h"
#include "ktrace.cuh"

namespace tha {

constexpr int SMP_THREADS = 512;
constexpr int SMP_BINS = 4096;
typedef unsigned long long u64;

struct SampleP {
  float* logits;              // [B, vocab]; overwritten with the probabilities (the reference's sample() does the same)
  int vocab;
  float temperature, topp;
  u64* rng;                   // [B] xorshift* state per slot
  const u64* reseed;          // nullable [B]: non-zero = the slot starts a new request with this seed
  const int* force;           // nullable [B]: >= 0 = the next token is this prompt token (no coin is drawn)
  const int* pos;             // [B]: < 0 = idle slot
  int* next;                  // [B]
};

__device__ __forceinline__ u64 smp_fixed(float p) { return __float2ull_rn(p * 1099511627776.0f); }   // units of 2^-40

// exclusive prefix sum over the block in thread order; *total = sum of all
__device__ __forceinline__ u64 smp_scan(u64 v, u64* s_warp, u64* total) {
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  u64 inc = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const u64 t = __shfl_up_sync(THA_FULL_MASK, inc, o);
    if (lane >= o) inc += t;
  }
  __syncthreads();                       // s_warp may still be read from a previous call
  if (lane == 31) s_warp[wid] = inc;
  __syncthreads();
  if (wid == 0) {
    u64 w = lane < SMP_THREADS / 32 ? s_warp[lane] : 0ull;
    u64 winc = w;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const u64 t = __shfl_up_sync(THA_FULL_MASK, winc, o);
      if (lane >= o) winc += t;
    }
    s_warp[lane] = winc - w;             // exclusive offset of each warp
    if (lane == 31) s_warp[32] = winc;
  }
  __syncthreads();
  *total = s_warp[32];
  return s_warp[wid] + inc - v;
}

struct SmpShared {
  u64 bins[SMP_BINS];
  u64 warp[33];
  u64 base, mass;
  int bin, found, idx;
  unsigned key;
  float red[33];
  u64 lv0[16];     // resident selection: the sixteen first-level sums (the same for both selections of a sample)
};

// The k-th (0-based) index i, ascending, with match(i) true.  Indices are dealt to threads round-robin (coalesced loads),
// so index order = round by round, thread order inside a round: one counting barrier per round, then a scan inside the
// round that holds the answer.  All threads call this; the result is broadcast through sh.idx (-1: fewer than k+1 matches).
template <typename Match>
__device__ int smp_kth(int n, u64 k, SmpShared& sh, Match match) {
  u64 seen = 0;
  int answer = -1;
  for (int base = 0; base < n; base += SMP_THREADS) {
    const int i = base + threadIdx.x;
    const bool f = i < n && match(i);
    const int cnt = __syncthreads_count(f);
    if (seen + (u64)cnt > k) {
      u64 total;
      const u64 before = smp_scan(f ? 1ull : 0ull, sh.warp, &total);
      if (f && seen + before == k) sh.idx = i;
      __syncthreads();
      answer = sh.idx;
      break;
    }
    seen += (u64)cnt;
  }
  __syncthreads();
  return answer;
}

// First element, in the order (probability descending, index ascending) over the candidates p >= cutoff, whose inclusive
// prefix mass exceeds `target`.  Returns false when the candidates' total mass does not (then *idx is the LAST candidate
// in that order and *incl the total).  All threads of the block call this with the same arguments.
// Keys: candidates lie in [cutoff, 1], i.e. in at most 32 binades, so the bit pattern minus the cutoff's exponent has 28
// significant bits: 5 of exponent + 23 of mantissa, split 12 / 8 / 8.  Seven mantissa bits in the first level spread
// even a near-uniform distribution (random-init models) over hundreds of bins instead of two.
__device__ bool smp_select(const float* pr, int n, float cutoff, u64 target, SmpShared& sh, int* idx, u64* incl) {
  unsigned pmask = 0, pval = 0;
  u64 base = 0;
  const unsigned e0 = __float_as_uint(cutoff) >> 23;                  // exponent of the cutoff (cutoff > 0)
  const bool narrow = e0 > 95 && e0 <= 127;                           // [cutoff, 1] within 32 binades (always, unless topp is within 1e-5 of 1)
  const unsigned rebase = narrow ? e0 << 23 : 0u;                      // otherwise: the plain bit pattern, 12 / 10 / 10
  const int shifts[3] = {narrow ? 16 : 20, narrow ? 8 : 10, 0}, nbins[3] = {4096, narrow ? 256 : 1024, narrow ? 256 : 1024};
  for (int lv = 0; lv < 3; lv++) {
    const int shift = shifts[lv], nb = nbins[lv];
    for (int i = threadIdx.x; i < nb; i += SMP_THREADS) sh.bins[i] = 0;
    if (threadIdx.x == 0) sh.found = 0;
    __syncthreads();
    {   // mass histogram of the elements that match the prefix; consecutive hits of one bin are merged before the atomic
      int cur = -1, old = -1;
      u64 acc = 0, oacc = 0;
      for (int i = threadIdx.x; i < n; i += SMP_THREADS) {
        const float p = pr[i];
        if (!(p >= cutoff)) continue;
        const unsigned key = __float_as_uint(p) - rebase;
        if ((key & pmask) != pval) continue;
        const int b = (int)((key >> shift) & (unsigned)(nb - 1));
        if (b != cur) {
          if (b == old) {           // two-entry cache: a thread alternating between two bins issues no atomics until the end
            const int tb = cur; cur = old; old = tb;
            const u64 ta = acc; acc = oacc; oacc = ta;
          } else {
            if (old >= 0) atomicAdd(&sh.bins[old], oacc);
            old = cur; oacc = acc;
            cur = b; acc = 0;
          }
        }
        acc += smp_fixed(p);
      }
      if (cur >= 0) atomicAdd(&sh.bins[cur], acc);
      if (old >= 0) atomicAdd(&sh.bins[old], oacc);
    }
    __syncthreads();
    // descending walk over the bins: thread t owns bins nb-1 - t*bp ... (bp each)
    const int bp = (nb + SMP_THREADS - 1) / SMP_THREADS;
    u64 local = 0;
    for (int k = 0; k < bp; k++) {
      const int bi = (int)threadIdx.x * bp + k;
      if (bi < nb) local += sh.bins[nb - 1 - bi];
    }
    u64 total;
    const u64 excl = smp_scan(local, sh.warp, &total);
    if (lv == 0 && base + total <= target) {
      // not enough mass: the last candidate in sorted order = lowest key, highest index among its equals
      unsigned best = 0xFFFFFFFFu; int bi = -1;
      for (int i = threadIdx.x; i < n; i += SMP_THREADS) {
        const float p = pr[i];
        if (!(p >= cutoff)) continue;
        const unsigned key = __float_as_uint(p);
        if (key < best || (key == best && i > bi)) { best = key; bi = i; }
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        const unsigned ok = __shfl_xor_sync(THA_FULL_MASK, best, o);
        const int oi = __shfl_xor_sync(THA_FULL_MASK, bi, o);
        if (ok < best || (ok == best && oi > bi)) { best = ok; bi = oi; }
      }
      __syncthreads();
      unsigned* wk = reinterpret_cast<unsigned*>(sh.bins);
      int* wi = reinterpret_cast<int*>(sh.bins) + 64;
      if ((threadIdx.x & 31) == 0) { wk[threadIdx.x >> 5] = best; wi[threadIdx.x >> 5] = bi; }
      __syncthreads();
      if (threadIdx.x == 0) {
        for (int k = 1; k < SMP_THREADS / 32; k++)
          if (wk[k] < best || (wk[k] == best && wi[k] > bi)) { best = wk[k]; bi = wi[k]; }
        sh.idx = bi;
      }
      __syncthreads();
      *idx = sh.idx;
      *incl = base + total;
      return false;
    }
    if (local > 0 && base + excl <= target && target < base + excl + local) {
      u64 run = base + excl;
      for (int k = 0; k < bp; k++) {
        const int b = nb - 1 - ((int)threadIdx.x * bp + k);
        if (b < 0) break;
        const u64 m = sh.bins[b];
        if (m > 0 && run + m > target) { sh.bin = b; sh.base = run; sh.found = 1; break; }
        run += m;
      }
    }
    __syncthreads();
    pval |= (unsigned)sh.bin << shift;
    pmask |= (unsigned)(nb - 1) << shift;
    base = sh.base;
    __syncthreads();
  }
  // every element of the class has the key `pval` (one probability, one mass q): the (k+1)-th of them in index order
  // is the first whose inclusive prefix exceeds the target
  const unsigned bits = pval + rebase;
  const u64 q = smp_fixed(__uint_as_float(bits));
  const u64 k = q ? (target - base) / q : 0;
  *idx = smp_kth(n, k, sh, [&](int i) { return __float_as_uint(pr[i]) == bits; });
  *incl = base + (k + 1) * q;
  return true;
}

// The same selection for the shared-memory-resident row, WITHOUT atomics and on integer KEYS.  The histogram form above pays
// one 64-bit shared atomic (a compare-and-swap loop) per candidate and level and re-derives the key from the probability in
// every pass: ~100 us per CTA at 32000 near-uniform candidates.  Here
//   * the normalising pass leaves the rebased KEY of each candidate in the shared row (0xFFFFFFFF for p < cutoff); a later
//     pass costs a load, a mask and a compare per element, the rare match takes a real branch (smp_hit / smp_keep are not
//     inlined, so the compiler cannot turn the rare path into sixty predicated instructions per element);
//   * a level resolves FOUR key bits: every thread adds the masses of its matches into 16 counters of its own
//     (priv[bin][thread]: no sharing, no conflicts), sixteen warps reduce one bin each, every thread walks the 16 sums;
//   * after two levels over the whole row the keys that still match (1 / 256 of the candidates for a smooth distribution) are
//     compacted into a list, and the remaining levels run over the list.
// Same keys, same masses, same tie rule as smp_select: same answer.
constexpr int SMP_LIST = SMP_BINS;   // the compacted list lives in sh.bins: (key, index) per entry
constexpr unsigned SMP_NOKEY = 0xFFFFFFFFu;
struct SmpKeys { const unsigned* key; int n; unsigned rebase; int nlev; };

// this thread's keys (i = thread, thread + 512, ...) EIGHT loads at a time, f(key, i) on those with (key & pmask) == pval
template <typename F>
__device__ __forceinline__ void smp_walk(const SmpKeys& K, unsigned pmask, unsigned pval, F f) {
  for (int i0 = threadIdx.x; i0 < K.n; i0 += 8 * SMP_THREADS) {
    unsigned k[8];
#pragma unroll
    for (int u = 0; u < 8; u++) { const int i = i0 + u * SMP_THREADS; k[u] = i < K.n ? K.key[i] : SMP_NOKEY; }
#pragma unroll
    for (int u = 0; u < 8; u++)
      if (k[u] != SMP_NOKEY && (k[u] & pmask) == pval) f(k[u], i0 + u * SMP_THREADS);
  }
}
__device__ __noinline__ void smp_hit(u64* mine, unsigned key, int shift, unsigned rebase) {
  mine[((key >> shift) & 15u) * SMP_THREADS] += smp_fixed(__uint_as_float(key + rebase));
}
__device__ __noinline__ void smp_keep(int* count, unsigned long long* list, int cap, unsigned key, int i) {
  const int at = atomicAdd(count, 1);
  if (at < cap) list[at] = ((unsigned long long)key << 32) | (unsigned)i;
}
// The k-th (0-based, index order) element whose key is `key`: the matches (usually one) are collected in shared memory and
// ranked by index; -1 if there are fewer than k + 1.  More matches than the list holds: the round-by-round scan.
__device__ int smp_kth_r(const SmpKeys& K, u64 k, unsigned key, SmpShared& sh) {
  unsigned long long* list = reinterpret_cast<unsigned long long*>(sh.bins);
  constexpr int CAP = 2048;
  if (threadIdx.x == 0) { sh.found = 0; sh.idx = -1; }
  __syncthreads();
  smp_walk(K, 0xFFFFFFFFu, key, [&](unsigned kk, int i) { smp_keep(&sh.found, list, CAP, kk, i); });
  __syncthreads();
  const int cnt = sh.found;
  __syncthreads();
  if (cnt > CAP) return smp_kth(K.n, k, sh, [&](int i) { return K.key[i] == key; });
  if ((u64)cnt <= k) return -1;
  for (int j = threadIdx.x; j < cnt; j += SMP_THREADS) {   // rank of my entry = entries with a smaller index
    const int mine_i = (int)(unsigned)list[j];
    int rank = 0;
    for (int t = 0; t < cnt; t++) rank += (int)(unsigned)list[t] < mine_i;
    if ((u64)rank == k) sh.idx = mine_i;
  }
  __syncthreads();
  const int r = sh.idx;
  __syncthreads();
  return r;
}

// reuse0: the first level's sums are already in sh.lv0 (left there by an earlier call over the same row)
__device__ bool smp_select_r(const SmpKeys& K, u64 target, SmpShared& sh, u64* priv, int* idx, u64* incl, bool reuse0, int kt = -1) {
  static_assert(SMP_THREADS == 512, "sixteen warps reduce sixteen bins");
  unsigned pmask = 0, pval = 0;
  u64 base = 0;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  u64* mine = priv + threadIdx.x;                        // my counter of bin b: mine[b * SMP_THREADS]
  unsigned long long* list = reinterpret_cast<unsigned long long*>(sh.bins);
  int n_list = -1;                                       // < 0: levels still run over the whole row
  for (int lv = 0; lv < K.nlev; lv++) {
    const int shift = 4 * (K.nlev - 1 - lv);
    if (lv == 2) {   // ---- compact the keys that match the first eight bits
      if (threadIdx.x == 0) sh.found = 0;
      __syncthreads();
      smp_walk(K, pmask, pval, [&](unsigned key, int i) { smp_keep(&sh.found, list, SMP_LIST, key, i); });
      __syncthreads();
      n_list = sh.found <= SMP_LIST ? sh.found : -1;     // too many equal-prefix elements: keep walking the row
      __syncthreads();
    }
    const bool skip = lv == 0 && reuse0;
#pragma unroll
    for (int b = 0; b < 16; b++) mine[b * SMP_THREADS] = 0;
    if (skip) {
    } else if (n_list < 0) {
      if (lv == 0) {   // every candidate matches: the update inline, the loads of eight elements in flight
        for (int i0 = threadIdx.x; i0 < K.n; i0 += 8 * SMP_THREADS) {
          unsigned k[8];
#pragma unroll
          for (int u = 0; u < 8; u++) { const int i = i0 + u * SMP_THREADS; k[u] = i < K.n ? K.key[i] : SMP_NOKEY; }
#pragma unroll
          for (int u = 0; u < 8; u++)
            if (k[u] != SMP_NOKEY) mine[((k[u] >> shift) & 15u) * SMP_THREADS] += smp_fixed(__uint_as_float(k[u] + K.rebase));
        }
      } else {
        smp_walk(K, pmask, pval, [&](unsigned key, int) { smp_hit(mine, key, shift, K.rebase); });
      }
    } else {
      for (int j = threadIdx.x; j < n_list; j += SMP_THREADS) {
        const unsigned key = (unsigned)(list[j] >> 32);
        if ((key & pmask) != pval) continue;
        mine[((key >> shift) & 15u) * SMP_THREADS] += smp_fixed(__uint_as_float(key + K.rebase));
      }
    }
    __syncthreads();
    if (skip) {
      if (threadIdx.x < 16) sh.warp[threadIdx.x] = sh.lv0[threadIdx.x];
    } else {   // warp w: the sum of bin w over the 512 threads
      u64 t = 0;
#pragma unroll
      for (int j = 0; j < SMP_THREADS / 32; j++) t += priv[warp * SMP_THREADS + lane + 32 * j];
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(THA_FULL_MASK, t, o);
      if (lane == 0) { sh.warp[warp] = t; if (lv == 0) sh.lv0[warp] = t; }
    }
    __syncthreads();
    u64 total = 0;
#pragma unroll
    for (int b = 0; b < 16; b++) total += sh.warp[b];
    if (lv == 0 && base + total <= target) {
      // not enough mass: the last candidate in sorted order = lowest key, highest index among its equals
      unsigned best = SMP_NOKEY; int bi = -1;
      for (int i = threadIdx.x; i < K.n; i += SMP_THREADS) {
        const unsigned key = K.key[i];
        if (key == SMP_NOKEY) continue;
        if (key < best || (key == best && i > bi)) { best = key; bi = i; }
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        const unsigned ok = __shfl_xor_sync(THA_FULL_MASK, best, o);
        const int oi = __shfl_xor_sync(THA_FULL_MASK, bi, o);
        if (ok < best || (ok == best && oi > bi)) { best = ok; bi = oi; }
      }
      __syncthreads();
      unsigned* wk = reinterpret_cast<unsigned*>(sh.bins);
      int* wi = reinterpret_cast<int*>(sh.bins) + 64;
      if (lane == 0) { wk[warp] = best; wi[warp] = bi; }
      __syncthreads();
      if (threadIdx.x == 0) {
        for (int k = 1; k < SMP_THREADS / 32; k++)
          if (wk[k] < best || (wk[k] == best && wi[k] > bi)) { best = wk[k]; bi = wi[k]; }
        sh.idx = bi;
      }
      __syncthreads();
      *idx = sh.idx;
      *incl = base + total;
      return false;
    }
    // descending walk over the sixteen sums (every thread, same result): the bin in which the prefix mass crosses the target
    int bin = 0;
    u64 run = base;
#pragma unroll
    for (int b = 15; b >= 0; b--) {
      const u64 m = sh.warp[b];
      if (m > 0 && run + m > target) { bin = b; break; }
      run += m;
    }
    base = run;
    pval |= (unsigned)bin << shift;
    pmask |= 15u << shift;
    __syncthreads();                                      // sh.warp and priv are rewritten by the next level
    kt_stamp(kt, 7 + lv);
  }
  const u64 q = smp_fixed(__uint_as_float(pval + K.rebase));
  const u64 k = q ? (target - base) / q : 0;
  *idx = smp_kth_r(K, k, pval, sh);
  *incl = base + (k + 1) * q;
  return true;
}

// RESIDENT: the slot's row is copied into (dynamic) shared memory once and every later pass -- scaling, exponentials,
// normalisation, two selections of three histogram levels each -- runs on that copy; the probabilities go back to HBM at the
// end (the reference's sample() leaves them in the logits buffer too).  The fourteen passes over global memory of the other form
// cost ~0.2 ms per step at 256 slots x 32000 entries (latency-bound: L2 round trips with one CTA's loads in flight).
template <bool RESIDENT>
__global__ void __launch_bounds__(SMP_THREADS, RESIDENT ? 1 : 2) sample_kernel(const SampleP P) {
  __shared__ SmpShared sh;
  extern __shared__ float smp_row[];
  const int kt = threadIdx.x == 0 ? kt_begin(KT_OTHER) : -1;
  pdl_trigger();   // the next kernel may start its launch (it waits for this grid to complete before touching memory)
  pdl_wait();
  const int b = blockIdx.x, n = P.vocab;
  if (P.pos[b] < 0) {
    if (threadIdx.x == 0) { P.next[b] = 0; if (P.reseed && P.reseed[b]) P.rng[b] = P.reseed[b]; }
    return;
  }
  u64 state = 0;
  if (threadIdx.x == 0) {
    state = (P.reseed && P.reseed[b]) ? P.reseed[b] : P.rng[b];
  }
  if (P.force && P.force[b] >= 0) {           // inside the prompt: the reference does not call sample() (src/llama.cpp:1029-1030)
    if (threadIdx.x == 0) { P.next[b] = P.force[b]; P.rng[b] = state; }
    return;
  }
  float* gpr = P.logits + (size_t)b * n;
  float* pr = gpr;
  // ---- greedy (temperature 0): first maximum (src/llama.cpp:275-286)
  if (P.temperature == 0.0f) {
    float bv = -INFINITY; int bi = 0x7fffffff;
    for (int i = threadIdx.x; i < n; i += SMP_THREADS) { const float v = pr[i]; if (v > bv) { bv = v; bi = i; } }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float ov = __shfl_xor_sync(THA_FULL_MASK, bv, o); const int oi = __shfl_xor_sync(THA_FULL_MASK, bi, o);
      if (ov > bv || (ov == bv && oi < bi)) { bv = ov; bi = oi; }
    }
    float* wv = sh.red; int* wi = reinterpret_cast<int*>(sh.bins);
    if ((threadIdx.x & 31) == 0) { wv[threadIdx.x >> 5] = bv; wi[threadIdx.x >> 5] = bi; }
    __syncthreads();
    if (threadIdx.x == 0) {
      for (int k = 1; k < SMP_THREADS / 32; k++) if (wv[k] > bv || (wv[k] == bv && wi[k] < bi)) { bv = wv[k]; bi = wi[k]; }
      P.next[b] = bi == 0x7fffffff ? 0 : bi;
      P.rng[b] = state;
    }
    return;
  }
  // ---- softmax(logits / temperature)  (src/llama.cpp:376-384, :262-274)
  float m = -INFINITY;
  if (RESIDENT) {
    pr = smp_row;
    if ((n & 3) == 0 && (reinterpret_cast<uintptr_t>(gpr) & 15) == 0) {
      for (int i0 = threadIdx.x; i0 < (n >> 2); i0 += 4 * SMP_THREADS) {   // four 128-bit loads in flight per thread
        float4 v[4];
#pragma unroll
        for (int u = 0; u < 4; u++) { const int i = i0 + u * SMP_THREADS; if (i < (n >> 2)) v[u] = reinterpret_cast<const float4*>(gpr)[i]; }
#pragma unroll
        for (int u = 0; u < 4; u++) {
          const int i = i0 + u * SMP_THREADS;
          if (i < (n >> 2)) {
            v[u].x = __fdiv_rn(v[u].x, P.temperature); v[u].y = __fdiv_rn(v[u].y, P.temperature);
            v[u].z = __fdiv_rn(v[u].z, P.temperature); v[u].w = __fdiv_rn(v[u].w, P.temperature);
            reinterpret_cast<float4*>(smp_row)[i] = v[u];
            m = fmaxf(fmaxf(m, fmaxf(v[u].x, v[u].y)), fmaxf(v[u].z, v[u].w));
          }
        }
      }
    } else {
      for (int i = threadIdx.x; i < n; i += SMP_THREADS) { const float v = __fdiv_rn(gpr[i], P.temperature); smp_row[i] = v; m = fmaxf(m, v); }
    }
  } else {
    for (int i = threadIdx.x; i < n; i += SMP_THREADS) { const float v = __fdiv_rn(pr[i], P.temperature); pr[i] = v; m = fmaxf(m, v); }
  }
  m = block_max(m, sh.red);
  kt_stamp(kt, 2);
  float sum = 0.0f;
  for (int i = threadIdx.x; i < n; i += SMP_THREADS) { const float e = expf(pr[i] - m); pr[i] = e; sum += e; }
  sum = block_sum(sum, sh.red);
  kt_stamp(kt, 3);
  // top-p on the resident row: the probabilities go back to HBM, the row keeps each candidate's rebased key
  const float cutoff_r = (1.0f - P.topp) / (float)(n - 1);
  const bool keyed = RESIDENT && P.topp > 0.0f && P.topp < 1.0f;
  const unsigned e0_r = __float_as_uint(cutoff_r) >> 23;
  const bool narrow_r = e0_r > 95 && e0_r <= 127;
  const unsigned rebase_r = narrow_r ? e0_r << 23 : 0u;
  for (int i = threadIdx.x; i < n; i += SMP_THREADS) {
    const float q = __fdiv_rn(pr[i], sum);
    if (RESIDENT) gpr[i] = q;
    if (keyed) reinterpret_cast<unsigned*>(pr)[i] = (q >= cutoff_r) ? __float_as_uint(q) - rebase_r : SMP_NOKEY;
    else pr[i] = q;
  }
  __syncthreads();
  kt_stamp(kt, 4);
  // ---- the coin (xorshift*, src/llama.cpp:401-411)
  if (threadIdx.x == 0) {
    state ^= state >> 12; state ^= state << 25; state ^= state >> 27;
    const unsigned u = (unsigned)((state * 0x2545F4914F6CDD1Dull) >> 32);
    sh.key = u >> 8;
    P.rng[b] = state;
  }
  __syncthreads();
  const u64 coin24 = sh.key;                  // coin = coin24 / 2^24
  int tok = n - 1;
  if (P.topp <= 0.0f || P.topp >= 1.0f) {
    // ---- multinomial in index order (src/llama.cpp:288-307): first i with coin < cdf(i); indices are dealt round-robin,
    // so the cdf is scanned round by round (thread order inside a round = index order)
    const u64 target = coin24 << 16;
    u64 run = 0;
    if (threadIdx.x == 0) sh.idx = n - 1;
    __syncthreads();
    for (int base = 0; base < n; base += SMP_THREADS) {
      const int i = base + threadIdx.x;
      const u64 mq = i < n ? smp_fixed(pr[i]) : 0ull;
      u64 total;
      const u64 excl = smp_scan(mq, sh.warp, &total);
      if (run + total > target) {
        if (mq > 0 && run + excl <= target && target < run + excl + mq) sh.idx = i;
        break;                                   // block-uniform: run and total are the same in every thread
      }
      run += total;
    }
    __syncthreads();
    tok = sh.idx;
  } else {
    // ---- top-p (src/llama.cpp:309-374)
    const float cutoff = (1.0f - P.topp) / (float)(n - 1);
    const u64 t1 = (u64)((double)P.topp * 1099511627776.0);
    int last; u64 cumulative;
    u64* priv = reinterpret_cast<u64*>(smp_row + ((n + 3) & ~3));   // RESIDENT: 16 x 512 counters behind the row
    const SmpKeys K = {reinterpret_cast<const unsigned*>(pr), n, rebase_r, narrow_r ? 7 : 8};
    if (RESIDENT) smp_select_r(K, t1, sh, priv, &last, &cumulative, false, kt);
    else smp_select(pr, n, cutoff, t1, sh, &last, &cumulative);
    __syncthreads();
    kt_stamp(kt, 5);
    const u64 r = (coin24 * cumulative) >> 24;        // coin * cumulative; < 2^64: cumulative <= ~2^40, coin24 < 2^24
    int pick; u64 incl;
    const bool ok = RESIDENT ? smp_select_r(K, r, sh, priv, &pick, &incl, true) : smp_select(pr, n, cutoff, r, sh, &pick, &incl);
    tok = ok ? pick : last;
    if (last < 0) tok = n - 1;                        // no candidate at all (cannot happen for a normalised distribution)
  }
  if (threadIdx.x == 0) P.next[b] = tok;
  kt_stamp(kt, 6);
}

int sample_launch(cudaStream_t st, int B, float* logits, int vocab, float temperature, float topp, unsigned long long* rng,
                  const unsigned long long* reseed, const int* force, const int* pos, int* next, bool pdl) {
  SampleP p;
  p.logits = logits; p.vocab = vocab; p.temperature = temperature; p.topp = topp; p.rng = rng; p.reseed = reseed;
  p.force = force; p.pos = pos; p.next = next;
  // the row in shared memory when it fits beside the histograms (vocab <= ~48 k); THA_SAMPLE_RESIDENT=0 keeps it in HBM / L2
  const int resident = sample_resident_mode();
  const size_t row = (size_t)((vocab + 3) & ~3) * sizeof(float) + (size_t)16 * SMP_THREADS * sizeof(u64);   // the row + the private counters
  if (resident && temperature != 0.0f && row + sizeof(SmpShared) + 1024 <= (size_t)227 * 1024) {
    THA_CUDA_OK(cudaFuncSetAttribute(sample_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)row));
    THA_CUDA_OK(launch(sample_kernel<true>, dim3(B), dim3(SMP_THREADS), row, st, pdl, p));
  } else {
    THA_CUDA_OK(launch(sample_kernel<false>, dim3(B), dim3(SMP_THREADS), 0, st, pdl, p));
  }
  return THA_STATUS_SUCCESS;
}

}  // namespace tha

using namespace tha;

// The driver's sampler for B slots at once.  logits_d [B, vocab] is overwritten with probabilities.  rng_d [B]: the
// xorshift* state of the request in each slot (advanced once per sampled token); reseed_d (nullable) [B]: non-zero =
// (re)start the slot's stream from this seed before sampling; force_d (nullable) [B]: >= 0 = take this token (prompt
// still being fed: no coin); pos_d [B]: < 0 = idle slot.  temperature 0 = greedy.
extern "C" int tha_sample(void* stream, int n_batches, float* logits_d, int vocab, float temperature, float topp,
                          unsigned long long* rng_d, const unsigned long long* reseed_d, const int* force_d, const int* pos_d,
                          int* next_d) {
  if (n_batches < 1 || !logits_d || vocab < 2 || !rng_d || !pos_d || !next_d || temperature < 0.0f) return THA_STATUS_INVALID_ARG;
  return sample_launch((cudaStream_t)stream, n_batches, logits_d, vocab, temperature, topp, rng_d, reseed_d, force_d, pos_d, next_d,
                       false);
}
