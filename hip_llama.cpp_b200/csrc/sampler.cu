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
#include "common.cuh"
#include "workspace.h"

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

__global__ void __launch_bounds__(SMP_THREADS, 2) sample_kernel(const SampleP P) {
  __shared__ SmpShared sh;
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
  float* pr = P.logits + (size_t)b * n;
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
  for (int i = threadIdx.x; i < n; i += SMP_THREADS) { const float v = __fdiv_rn(pr[i], P.temperature); pr[i] = v; m = fmaxf(m, v); }
  m = block_max(m, sh.red);
  float sum = 0.0f;
  for (int i = threadIdx.x; i < n; i += SMP_THREADS) { const float e = expf(pr[i] - m); pr[i] = e; sum += e; }
  sum = block_sum(sum, sh.red);
  for (int i = threadIdx.x; i < n; i += SMP_THREADS) pr[i] = __fdiv_rn(pr[i], sum);
  __syncthreads();
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
    smp_select(pr, n, cutoff, t1, sh, &last, &cumulative);
    __syncthreads();
    const u64 r = (coin24 * cumulative) >> 24;        // coin * cumulative; < 2^64: cumulative <= ~2^40, coin24 < 2^24
    int pick; u64 incl;
    const bool ok = smp_select(pr, n, cutoff, r, sh, &pick, &incl);
    tok = ok ? pick : last;
    if (last < 0) tok = n - 1;                        // no candidate at all (cannot happen for a normalised distribution)
  }
  if (threadIdx.x == 0) P.next[b] = tok;
}

int sample_launch(cudaStream_t st, int B, float* logits, int vocab, float temperature, float topp, unsigned long long* rng,
                  const unsigned long long* reseed, const int* force, const int* pos, int* next, bool pdl) {
  SampleP p;
  p.logits = logits; p.vocab = vocab; p.temperature = temperature; p.topp = topp; p.rng = rng; p.reseed = reseed;
  p.force = force; p.pos = pos; p.next = next;
  THA_CUDA_OK(launch(sample_kernel, dim3(B), dim3(SMP_THREADS), 0, st, pdl, p));
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
