// Per-device scratch owned by the library (grown lazily, never inside a stream
// capture): split-attention partials, device mirrors of token[]/pos[], RoPE
// tables, Q8_0 activation scratch, cached CUDA graphs.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "common.cuh"

#include <map>
#include <mutex>
#include <tuple>
#include <vector>

namespace tha {

struct RopeTable { float* cos_d = nullptr; float* sin_d = nullptr; };

struct StepGraph {
  cudaGraphExec_t exec = nullptr;
  int launches = 0;
  std::vector<unsigned char> key;        // the argument bytes the graph was captured for (the map key is only their hash)
  unsigned long long last_use = 0;       // for eviction: ws->graph_clock at the last replay
};
constexpr size_t WS_MAX_GRAPHS = 64;     // cached executable graphs per workspace; the least recently used one goes first

struct Workspace {
  int device = -1;
  // split-attention partials
  float4* part_acc = nullptr; float2* part_ml = nullptr; size_t part_cap = 0;  // in slots
  int* counters = nullptr; int counters_cap = 0;
  // device mirrors of the host token[] / pos[] arrays of thaDNN_s_forward_batch
  int* tok_d = nullptr; int* pos_d = nullptr; int* tokpos_h = nullptr; int tokpos_cap = 0;
  int* next_d = nullptr; int* next_h = nullptr;   // next token per slot (device, pinned host)
  int* force_d = nullptr;                         // sampled steps: prompt token to take instead of sampling (-1: sample)
  unsigned long long* reseed_d = nullptr;         // ... non-zero: the slot starts a request with this sampler seed
  unsigned long long* rng_d = nullptr;            // xorshift* state of the request in each slot (sampler.cu)
  // Q8_0: quantised activations of the current matmul input ([B, K] int8 + [B, K/gs] scales)
  int8_t* q8_xq = nullptr; float* q8_xs = nullptr; size_t q8_cap = 0;
  // RoPE cos/sin tables keyed by (head_size, seq_len); built on the host with libm so
  // they are bit-identical to what run.cc:271-277 computes
  std::map<std::pair<int, int>, RopeTable> rope;
  // cached graphs of the decode step keyed by a hash of every launch argument
  std::map<unsigned long long, StepGraph> graphs;
  unsigned long long graph_clock = 0;
  // One step at a time: the scratch buffers, the attention split counters and the graph cache are shared by every stream and
  // host thread that uses this device.  run_step() serialises host threads on `mu`, records `step_done` behind every step and,
  // when the stream differs from the previous step's, makes the new stream wait for that event first -- two streams never have
  // steps in flight over the same scratch.  (last_stream is only compared, never dereferenced: it may have been destroyed.)
  std::mutex mu;
  cudaStream_t last_stream = nullptr;
  cudaEvent_t step_done = nullptr;
  // two half-batches of a step as two concurrent branches of the graph (forward.cu, "dual"): the second branch's
  // scratch owner, its stream and the fork / join events
  Workspace* sibling = nullptr;
  Workspace* parent = nullptr;            // set on a sibling: graphs that captured its pointers live in the parent
  cudaStream_t side_stream = nullptr;
  cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
};

// Registered block pools (tha_kv_pool_register): a tha_state_t whose key_cache is a registered pool is PAGED
struct KvPool { const float* k; const float* v; const int* table; int stride; int block_size; int n_blocks; };
bool kv_pool_find(const float* key_cache, KvPool* out);
KvPaged kv_paged_of(const KvPool& pool, int n_layers, int kv_dim, int layer);

Workspace* workspace();   // for the current device (thread-safe)
int ws_ensure_attn(Workspace* ws, size_t part_slots, int n_counters);
int ws_ensure_tokpos(Workspace* ws, int n_batches);
int ws_ensure_q8(Workspace* ws, int n_batches, int kmax, int gs);
int ws_ensure_dual(Workspace* ws);
// cached graphs captured pointers that are about to change
inline void ws_drop_graphs(Workspace* ws) {
  for (Workspace* w = ws; w; w = w->parent) {
    for (auto& g : w->graphs) if (g.second.exec) cudaGraphExecDestroy(g.second.exec);
    w->graphs.clear();
  }
}
void ws_release(Workspace* ws);        // free everything a (tensor-parallel rank's) workspace owns
void step_tc_release(Workspace* ws);   // ... and the row pipeline's scratch keyed by it
int skip_mode();       // bit mask of kernel classes left out of the row pipeline (live-share measurements): 1 layer matmuls, 2 attention,
                       // 4 prep, 8 RoPE / scatter, 16 SwiGLU / quantise, 32 classifier, 64 argmax
int sample_resident_mode();   // 1: the device sampler keeps a slot's row in shared memory
int rope_fuse_mode();  // 1: a decode step's RoPE + KV scatter run inside the attention kernel
int chain_mode();      // 1: matmul chains with fused epilogues (chain_tc.cu), 0: one launch per matmul (the round-1 pipeline)
int sm_budget();            // SMs one matmul launch may occupy: all of them, or a branch's share while a dual step is planned / issued
void set_sm_budget(int n);  // 0 = all
int dual_min_rows();   // rows from which a step is issued as two concurrent half-batches (0 = never)
int ws_rope(Workspace* ws, int head_size, int seq_len, const float** cos_d, const float** sin_d);

}  // namespace tha
