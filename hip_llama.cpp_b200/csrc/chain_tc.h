// Host-side interface of the chained tensor-core matmuls (chain_tc.cu): up to four dependent weight matmuls of the
// decode step -- wo, w1|w3, w2 and the next layer's wq|wk|wv (or the classifier) -- in ONE persistent launch, the
// elementwise neighbours of each fused into its epilogue (src/thaDNN.cpp:39-73 is the loop this replaces).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace tha {

enum ChainEpilogue {
  CH_PLAIN = 0,    // out[n, off + row] = acc * colscale[n]                                   (classifier -> logits)
  CH_RESID = 1,    // x[n, row] += acc; next matmul input (x * gain) as tf32 hi/lo; sums of squares -> 1/rms of the row
  CH_SWIGLU = 2,   // w1 | w3 interleaved inside every tile: hb = silu(w1.x) * (w3.x) as tf32 hi/lo
  CH_ROPE = 3,     // wq | wk | wv: RoPE on q and k pairs, q -> state, k / v -> the cache row of pos[n]
};

constexpr int CH_MAX_PHASES = 4;

struct ChainMatmul {
  int mode;
  int n_mats;
  const float* W[3]; int M[3];                 // row-major [M, K]; CH_SWIGLU: W[0] = w1, W[1] = w3, M[0] = M[1] = hidden
  int K;
  const float* Xhi; const float* Xlo;          // [N, K] tf32 hi / lo of the input rows (written by the previous phase / kernel)
  const float* colscale;                       // nullable [N]
  float* out; long long ldo;                   // CH_PLAIN logits / CH_RESID x / CH_ROPE q
  float* oh; float* ol;                        // CH_RESID [N, M], CH_SWIGLU [N, hidden]
  const float* gain; float* scale_out;         // CH_RESID
  float* kc; float* vc; long long kv_bs; long long loff; int kvd, hs; const float* rc; const float* rs;   // CH_ROPE
  int splits;                                  // 0: let the planner choose
};

struct ChainArgs {
  int n; ChainMatmul mm[CH_MAX_PHASES];
  int N;                                       // rows (slots)
  const int* pos; const int* slot;             // [N]; slot == nullptr: row n lives in cache slab n
  float* part; size_t part_floats;             // split-K partials (shared by the phases; they run one after another)
  float* ss_part; size_t ss_floats;            // CH_RESID: sums of squares per (row tile, column)
  int* ctr; int ctr_ints;                      // counter block of THIS launch (zero on entry; the kernel leaves it zero)
};

bool chain_tc_ok(const ChainArgs& a);
// scratch a chain needs (floats of partials, floats of ss, ints of counters)
void chain_tc_need(const ChainArgs& a, size_t* part_floats, size_t* ss_floats, int* ctr_ints);
int chain_tc_launch(cudaStream_t st, const ChainArgs& a, bool pdl);

}  // namespace tha
