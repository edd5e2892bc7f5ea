// Tensor-parallel context of the row pipeline (step_tc.cu): Llama-2-7B split 2/4/8 ways over the GPUs of one NVLink /
// NVSwitch box (SURVEY.md 8e "Optional TP").  wq/wk/wv/w1/w3 are split by output rows (heads, hidden units), wo/w2 by
// input columns; each rank's wo / w2 matmul therefore yields PARTIAL sums of the residual update.  Those partials are
// never all-reduced by a separate collective: every rank's matmul writes its split-K partials into an exchange buffer
// in its own HBM that its peers have mapped (cudaIpc / peer access), and the kernel that consumes them anyway
// (tc_prep*: residual add + RMSNorm + split / quantise) reads all ranks' partials over NVLink in a fixed order --
// the all-reduce is fused into its consumer, and every rank computes a bit-identical residual stream.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "workspace.h"

namespace tha {

constexpr int TP_MAX = 8;
constexpr int TP_HDR_BYTES = 1024;   // [flags: TP_MAX x u32][epoch u32][err i32] at the head of the exchange block

// What a consumer kernel needs (passed by value).  n == 0: no tensor parallelism.
struct TpPeers {
  int n, rank;
  int k, nx;                              // index of this exchange within a step, exchanges per step
  const float* part[TP_MAX];              // rank r's exchange buffer for this exchange (own one included)
  int nz[TP_MAX];                         // split-K partials rank r wrote
  unsigned* peer_flags[TP_MAX];           // rank r's flag array (own one included)
  const unsigned* epoch;                  // local: steps completed
  int* err;                               // local: set when a wait timed out
};

struct TpCtx {
  int tp = 1, rank = 0, head_size = 0;
  int hidden_of[TP_MAX] = {0};            // local hidden_dim of every rank
  size_t xchg_floats = 0;                 // capacity of EACH of the two exchange buffers
  unsigned char* base = nullptr;          // local block: header, xchg0, xchg1
  unsigned char* peer[TP_MAX] = {nullptr};   // every rank's block as mapped here (peer[rank] == base)
  bool ipc_opened[TP_MAX] = {false};
  bool connected = false;
  int mode = 0;                           // 1: prepare only (tha_tp_set_mode)
  int device = -1;
  Workspace ws;                           // this rank's own scratch (two ranks may share a device in tests)
  size_t block_bytes() const { return TP_HDR_BYTES + 2 * xchg_floats * sizeof(float); }
  float* xchg(int r, int which) const { return reinterpret_cast<float*>(peer[r] + TP_HDR_BYTES) + (size_t)which * xchg_floats; }
  unsigned* flags(int r) const { return reinterpret_cast<unsigned*>(peer[r]); }
  unsigned* epoch() const { return reinterpret_cast<unsigned*>(base) + TP_MAX; }
  int* err() const { return reinterpret_cast<int*>(base) + TP_MAX + 1; }
};

}  // namespace tha
