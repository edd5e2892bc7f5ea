// Library state: options, last error, per-device workspace.
#include <math.h>
#include <stdlib.h>
#include <string.h>

#include <map>
#include <mutex>
#include <string>
#include <vector>

#include "common.cuh"
#include "workspace.h"
#include "ktrace.cuh"

namespace tha {

void mega_set_enabled(int v);

static thread_local char g_err[512] = "";
static thread_local int g_launches = 0;
static int g_pdl = -1, g_graph = -1, g_tc = -2;
static std::mutex g_mu;
static std::map<int, Workspace*> g_ws;

void set_last_error(const char* where, cudaError_t e) {
  snprintf(g_err, sizeof(g_err), "%s: %s (%s)", where, cudaGetErrorString(e), cudaGetErrorName(e));
  cudaGetLastError();  // clear the sticky-free error so later calls are not poisoned
}

static int env_flag(const char* name, int dflt);
void prefer_max_smem(const void* fn) {
  static int on = -1;
  if (on < 0) on = env_flag("THA_CARVEOUT", 0);   // measured: 0.883 -> 0.928 ms/step with it on (the glue kernels lose their L1)
  if (!on) return;
  static std::mutex mu;
  static std::map<std::pair<int, const void*>, bool> done;   // per (device, kernel)
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return;
  std::lock_guard<std::mutex> lk(mu);
  bool& d = done[std::make_pair(dev, fn)];
  if (d) return;
  d = true;
  if (cudaFuncSetAttribute(fn, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared) != cudaSuccess) cudaGetLastError();
}
static int env_flag(const char* name, int dflt) {
  const char* v = getenv(name);
  if (!v || !*v) return dflt;
  return atoi(v) != 0;
}
int pdl_enabled() {
  if (g_pdl < 0) g_pdl = env_flag("THA_PDL", 1);
  return g_pdl;
}
int graph_enabled() {
  if (g_graph < 0) g_graph = env_flag("THA_GRAPH", 1);
  return g_graph;
}
int tc_mode() {
  if (g_tc == -2) { const char* v = getenv("THA_TC"); g_tc = (v && *v) ? atoi(v) : -1; }
  return g_tc;
}
int* opt_tc() { return &g_tc; }
static int g_dual = -2;
int dual_min_rows() {
  // From 256 rows on a step is issued as two concurrent half-batch branches, each branch's matmuls planned for and limited to
  // half of the SMs (forward.cu dual_plan): the two dependency chains overlap each other's latency.  Measured after the
  // round-2 fixes: stories110M fp32 at 256 slots 0.811 -> 0.793 ms per step; at 128 / 64 slots it loses (0.645 -> 0.670,
  // 0.541 -> 0.578: a chain of half the slots is barely shorter), hence the threshold.  THA_DUAL=0 / option "dual" 0: never.
  if (g_dual == -2) { const char* v = getenv("THA_DUAL"); g_dual = (v && *v) ? atoi(v) : 256; }
  return g_dual;
}
int* opt_dual() { return &g_dual; }
static thread_local int g_sm_budget = 0;
int sm_budget() {
  static int sms = 0;
  if (!sms) { int dev = 0, n = 148; if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev); sms = n; }
  return (g_sm_budget > 0 && g_sm_budget < sms) ? g_sm_budget : sms;
}
void set_sm_budget(int n) { g_sm_budget = n; }
static int g_skip = -1;
int skip_mode() {   // development / measurement aid: leave a class of kernels out of the step (step_tc.cu); results are then garbage
  if (g_skip < 0) { const char* v = getenv("THA_SKIP"); g_skip = (v && *v) ? atoi(v) : 0; }
  return g_skip;
}
static int g_chain = -2;
int chain_mode() {   // 1: matmul chains with fused epilogues (chain_tc.cu) where the shape allows; 0 (default): one launch per matmul
  // Off by default: measured on stories110M fp32 at 256 slots the chained step takes 1.29 ms against 0.99 ms -- a phase
  // costs ~20 us whether its synchronisation points are kernel boundaries or counters in L2 (profiles/r02_chain_trace.txt).
  if (g_chain == -2) { const char* v = getenv("THA_CHAIN"); g_chain = (v && *v) ? atoi(v) : 0; }
  return g_chain;
}
int* opt_chain() { chain_mode(); return &g_chain; }
static int g_rope_fuse = -2;
int rope_fuse_mode() {   // 1 (default): RoPE + KV scatter of a decode step inside the attention kernel; 0: tc_rope_scatter_kernel
  if (g_rope_fuse == -2) { const char* v = getenv("THA_ROPE_FUSE"); g_rope_fuse = (v && *v) ? atoi(v) : 1; }
  return g_rope_fuse;
}
int* opt_rope_fuse() { rope_fuse_mode(); return &g_rope_fuse; }
static int g_sample_resident = -2;
int sample_resident_mode() {   // 1 (default): the sampler keeps a slot's row in shared memory; 0: every pass over HBM / L2 (the first form)
  if (g_sample_resident == -2) { const char* v = getenv("THA_SAMPLE_RESIDENT"); g_sample_resident = (v && *v) ? atoi(v) : 1; }
  return g_sample_resident;
}
int* opt_sample_resident() { sample_resident_mode(); return &g_sample_resident; }
void count_launch(int n) { g_launches += n; }
int launches_reset() { int v = g_launches; g_launches = 0; return v; }
int launches_get() { return g_launches; }

static int g_trace = -1;
static std::vector<std::pair<const void*, cudaEvent_t>> g_trace_ev;
void trace_launch(const void* fn, cudaStream_t st) {
  if (g_trace < 0) g_trace = env_flag("THA_TRACE", 0);
  if (!g_trace) return;
  cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
  if (cudaStreamIsCapturing(st, &cs) != cudaSuccess || cs != cudaStreamCaptureStatusNone) { cudaGetLastError(); return; }
  cudaEvent_t ev;
  if (cudaEventCreate(&ev) != cudaSuccess) return;
  cudaEventRecord(ev, st);
  std::lock_guard<std::mutex> lk(g_mu);
  g_trace_ev.emplace_back(fn, ev);
}

static std::map<const float*, KvPool> g_kv_pools;
bool kv_pool_find(const float* key_cache, KvPool* out) {
  std::lock_guard<std::mutex> lk(g_mu);
  auto it = g_kv_pools.find(key_cache);
  if (it == g_kv_pools.end()) return false;
  if (out) *out = it->second;
  return true;
}
KvPaged kv_paged_of(const KvPool& pool, int n_layers, int kv_dim, int layer) {
  KvPaged g;
  g.table = pool.table; g.stride = pool.stride;
  int sh = 0;
  while ((1 << sh) < pool.block_size) sh++;
  g.bs_shift = sh; g.bs_mask = pool.block_size - 1;
  g.blk_stride = (long long)n_layers * pool.block_size * kv_dim;
  g.layer_off = (long long)layer * pool.block_size * kv_dim;
  return g;
}

Workspace* workspace() {
  int dev = -1;
  if (cudaGetDevice(&dev) != cudaSuccess) { set_last_error("cudaGetDevice", cudaGetLastError()); return nullptr; }
  std::lock_guard<std::mutex> lk(g_mu);
  auto it = g_ws.find(dev);
  if (it != g_ws.end()) return it->second;
  Workspace* ws = new Workspace();
  ws->device = dev;
  g_ws[dev] = ws;
  return ws;
}

void ws_release(Workspace* ws) {
  ws_drop_graphs(ws);
  step_tc_release(ws);
  if (ws->part_acc) { cudaFree(ws->part_acc); cudaFree(ws->part_ml); ws->part_acc = nullptr; ws->part_ml = nullptr; ws->part_cap = 0; }
  if (ws->counters) { cudaFree(ws->counters); ws->counters = nullptr; ws->counters_cap = 0; }
  if (ws->tok_d) { cudaFree(ws->tok_d); cudaFreeHost(ws->tokpos_h); ws->tok_d = nullptr; ws->tokpos_cap = 0; }
  if (ws->q8_xq) { cudaFree(ws->q8_xq); cudaFree(ws->q8_xs); ws->q8_xq = nullptr; ws->q8_cap = 0; }
  for (auto& r : ws->rope) { cudaFree(r.second.cos_d); cudaFree(r.second.sin_d); }
  ws->rope.clear();
}

int ws_ensure_dual(Workspace* ws) {
  if (ws->sibling) return THA_STATUS_SUCCESS;
  Workspace* sib = new Workspace();
  sib->device = ws->device;
  sib->parent = ws;
  THA_CUDA_OK(cudaStreamCreateWithFlags(&ws->side_stream, cudaStreamNonBlocking));
  THA_CUDA_OK(cudaEventCreateWithFlags(&ws->ev_fork, cudaEventDisableTiming));
  THA_CUDA_OK(cudaEventCreateWithFlags(&ws->ev_join, cudaEventDisableTiming));
  ws->sibling = sib;
  return THA_STATUS_SUCCESS;
}

int ws_ensure_attn(Workspace* ws, size_t part_slots, int n_counters) {
  if (part_slots > ws->part_cap) {
    if (ws->part_acc) { cudaFree(ws->part_acc); cudaFree(ws->part_ml); }
    size_t cap = part_slots + part_slots / 4;
    THA_CUDA_OK(cudaMalloc(&ws->part_acc, cap * sizeof(float4)));
    THA_CUDA_OK(cudaMalloc(&ws->part_ml, cap * sizeof(float2)));
    ws->part_cap = cap;
    ws_drop_graphs(ws);  // cached graphs captured the old pointers
  }
  if (n_counters > ws->counters_cap) {
    if (ws->counters) cudaFree(ws->counters);
    int cap = n_counters * 2;
    THA_CUDA_OK(cudaMalloc(&ws->counters, cap * sizeof(int)));
    THA_CUDA_OK(cudaMemset(ws->counters, 0, cap * sizeof(int)));
    ws->counters_cap = cap;
    ws_drop_graphs(ws);
  }
  return THA_STATUS_SUCCESS;
}

int ws_ensure_tokpos(Workspace* ws, int n_batches) {
  if (n_batches <= ws->tokpos_cap) return THA_STATUS_SUCCESS;
  if (ws->tok_d) { cudaFree(ws->tok_d); cudaFreeHost(ws->tokpos_h); }
  int cap = n_batches < 64 ? 64 : n_batches * 2;
  // device: [tok | pos | force | reseed (u64) | next | rng (u64)]; the pinned mirror has the same first five blocks, so ONE
  // copy of 5 * cap ints stages everything a sampled step takes from the host (greedy steps copy the first two)
  THA_CUDA_OK(cudaMalloc(&ws->tok_d, 8 * cap * sizeof(int)));
  THA_CUDA_OK(cudaMemset(ws->tok_d, 0, 8 * cap * sizeof(int)));
  ws->pos_d = ws->tok_d + cap;
  ws->force_d = ws->tok_d + 2 * cap;
  ws->reseed_d = reinterpret_cast<unsigned long long*>(ws->tok_d + 3 * cap);
  ws->next_d = ws->tok_d + 5 * cap;
  ws->rng_d = reinterpret_cast<unsigned long long*>(ws->tok_d + 6 * cap);
  THA_CUDA_OK(cudaMallocHost(&ws->tokpos_h, 6 * cap * sizeof(int)));
  ws->next_h = ws->tokpos_h + 5 * cap;
  ws->tokpos_cap = cap;
  ws_drop_graphs(ws);
  return THA_STATUS_SUCCESS;
}

int ws_ensure_q8(Workspace* ws, int n_batches, int kmax, int gs) {
  const size_t need = (size_t)n_batches * kmax;
  if (need <= ws->q8_cap) return THA_STATUS_SUCCESS;
  if (ws->q8_xq) { cudaFree(ws->q8_xq); cudaFree(ws->q8_xs); }
  THA_CUDA_OK(cudaMalloc(&ws->q8_xq, need));
  (void)gs;   // groups are at least 16 values (the kernel's requirement): need / 16 scales always suffice
  THA_CUDA_OK(cudaMalloc(&ws->q8_xs, (need / 16 + 1) * sizeof(float)));
  ws->q8_cap = need;
  ws_drop_graphs(ws);
  return THA_STATUS_SUCCESS;
}

// cos/sin of pos * 10000^(-(2j)/hs), exactly as run.cc:271-277 evaluates them (float
// powf / cosf / sinf of the host libm), so the fused RoPE is bit-identical to the CPU path.
int ws_rope(Workspace* ws, int head_size, int seq_len, const float** cos_d, const float** sin_d) {
  std::lock_guard<std::mutex> lk(g_mu);
  auto key = std::make_pair(head_size, seq_len);
  auto it = ws->rope.find(key);
  if (it == ws->rope.end()) {
    const int half = head_size / 2;
    std::vector<float> c((size_t)seq_len * half), s((size_t)seq_len * half);
    for (int pos = 0; pos < seq_len; pos++)
      for (int j = 0; j < half; j++) {
        int head_dim = 2 * j;
        float freq = 1.0f / powf(10000.0f, head_dim / (float)head_size);
        float val = pos * freq;
        c[(size_t)pos * half + j] = cosf(val);
        s[(size_t)pos * half + j] = sinf(val);
      }
    RopeTable t;
    THA_CUDA_OK(cudaMalloc(&t.cos_d, c.size() * sizeof(float)));
    THA_CUDA_OK(cudaMalloc(&t.sin_d, s.size() * sizeof(float)));
    THA_CUDA_OK(cudaMemcpy(t.cos_d, c.data(), c.size() * sizeof(float), cudaMemcpyHostToDevice));
    THA_CUDA_OK(cudaMemcpy(t.sin_d, s.data(), s.size() * sizeof(float), cudaMemcpyHostToDevice));
    it = ws->rope.emplace(key, t).first;
  }
  *cos_d = it->second.cos_d;
  *sin_d = it->second.sin_d;
  return THA_STATUS_SUCCESS;
}

const char* last_error_text() { return g_err; }
void set_opt(int* which, int v) { *which = v; }
int* opt_pdl() { return &g_pdl; }
int* opt_graph() { return &g_graph; }

}  // namespace tha

extern "C" int tha_version(void) { return 100; }

// ---- ktrace.cuh: device-side step timeline (development aid) ----
namespace tha {
static std::vector<int (*)(const KtSlots&)>& kt_setters() { static std::vector<int (*)(const KtSlots&)> v; return v; }
void kt_register(int (*setter)(const KtSlots&)) { kt_setters().push_back(setter); }
static KtSlots g_kt = {nullptr, nullptr};
}  // namespace tha
// enable = 1: arm (records start at 0).  enable = 0: copy up to max_records records of KT_STAMPS u64 into out_host, disarm;
// returns the record count through *n_out.
extern "C" int tha_ktrace(unsigned long long* out_host, int max_records, int enable, int* n_out) {
  using namespace tha;
  if (enable) {
    if (!g_kt.buf) {
      THA_CUDA_OK(cudaMalloc(&g_kt.buf, (size_t)KT_MAX * KT_STAMPS * sizeof(unsigned long long)));
      THA_CUDA_OK(cudaMalloc(&g_kt.ctr, sizeof(unsigned int)));
    }
    THA_CUDA_OK(cudaMemset(g_kt.buf, 0, (size_t)KT_MAX * KT_STAMPS * sizeof(unsigned long long)));
    THA_CUDA_OK(cudaMemset(g_kt.ctr, 0, sizeof(unsigned int)));
    for (auto f : kt_setters()) if (f(g_kt)) return THA_STATUS_EXECUTION_FAILED;
    return THA_STATUS_SUCCESS;
  }
  if (!g_kt.buf) return THA_STATUS_INVALID_ARG;
  THA_CUDA_OK(cudaDeviceSynchronize());
  unsigned int n = 0;
  THA_CUDA_OK(cudaMemcpy(&n, g_kt.ctr, sizeof(n), cudaMemcpyDeviceToHost));
  if (n > (unsigned)KT_MAX) n = KT_MAX;
  if ((int)n > max_records) n = max_records > 0 ? max_records : 0;
  if (out_host && n) THA_CUDA_OK(cudaMemcpy(out_host, g_kt.buf, (size_t)n * KT_STAMPS * sizeof(unsigned long long), cudaMemcpyDeviceToHost));
  if (n_out) *n_out = (int)n;
  const KtSlots off = {nullptr, nullptr};
  for (auto f : kt_setters()) f(off);
  cudaFree(g_kt.buf); cudaFree(g_kt.ctr); g_kt = off;
  return THA_STATUS_SUCCESS;
}

// Paged KV cache.  key_pool / value_pool: [n_blocks][n_layers][block_size][kv_dim] fp32 on the device (the caller allocates
// them; a tha_state_t is paged when its key_cache == key_pool and value_cache == value_pool); block_table_d: [slots][stride]
// block ids on the device, row s listing the blocks that hold positions 0.., block_size positions each (a power of two).
// The kernels read the table at run time: update it (stream-ordered) before a step that needs a new block.
extern "C" int tha_kv_pool_register(float* key_pool, float* value_pool, const int* block_table_d, int table_stride, int block_size,
                                    int n_blocks) {
  if (!key_pool || !value_pool || !block_table_d || table_stride < 1 || block_size < 1 || (block_size & (block_size - 1)) || n_blocks < 1)
    return THA_STATUS_INVALID_ARG;
  std::lock_guard<std::mutex> lk(tha::g_mu);
  tha::KvPool p = {key_pool, value_pool, block_table_d, table_stride, block_size, n_blocks};
  tha::g_kv_pools[key_pool] = p;
  for (auto& kv : tha::g_ws) tha::ws_drop_graphs(kv.second);   // a graph may have captured this address as a dense slab
  return THA_STATUS_SUCCESS;
}
extern "C" int tha_kv_pool_unregister(float* key_pool) {
  std::lock_guard<std::mutex> lk(tha::g_mu);
  tha::g_kv_pools.erase(key_pool);
  for (auto& kv : tha::g_ws) tha::ws_drop_graphs(kv.second);
  return THA_STATUS_SUCCESS;
}
extern "C" const char* tha_last_error(void) { return tha::last_error_text(); }
extern "C" int tha_set_option(const char* name, int value) {
  if (!name) return THA_STATUS_INVALID_ARG;
  if (!strcmp(name, "pdl")) { *tha::opt_pdl() = value != 0; return THA_STATUS_SUCCESS; }
  if (!strcmp(name, "graph")) { *tha::opt_graph() = value != 0; return THA_STATUS_SUCCESS; }
  if (!strcmp(name, "tc")) { *tha::opt_tc() = value; return THA_STATUS_SUCCESS; }
  if (!strcmp(name, "dual")) { tha::dual_min_rows(); *tha::opt_dual() = value; return THA_STATUS_SUCCESS; }
  if (!strcmp(name, "skip")) {   // cached graphs were captured with the old mask
    tha::skip_mode();
    tha::g_skip = value < 0 ? 0 : value;
    std::lock_guard<std::mutex> lk(tha::g_mu);
    for (auto& kv : tha::g_ws) tha::ws_drop_graphs(kv.second);
    return THA_STATUS_SUCCESS;
  }
  if (!strcmp(name, "chain")) { *tha::opt_chain() = value; return THA_STATUS_SUCCESS; }
  if (!strcmp(name, "sample_resident")) {   // graphs that contain the sampler were captured with the other kernel
    *tha::opt_sample_resident() = value != 0;
    std::lock_guard<std::mutex> lk(tha::g_mu);
    for (auto& kv : tha::g_ws) tha::ws_drop_graphs(kv.second);
    return THA_STATUS_SUCCESS;
  }
  if (!strcmp(name, "ropefuse")) {   // cached graphs were captured with the other kernel sequence
    *tha::opt_rope_fuse() = value != 0;
    std::lock_guard<std::mutex> lk(tha::g_mu);
    for (auto& kv : tha::g_ws) tha::ws_drop_graphs(kv.second);
    return THA_STATUS_SUCCESS;
  }
  if (!strcmp(name, "mega")) { tha::mega_set_enabled(value); return THA_STATUS_SUCCESS; }
  return THA_STATUS_INVALID_ARG;
}

// Development aid: per-kernel live time (event to event) of everything launched since the last dump; skips the first
// `skip` launches.  Prints "name  count  avg_us  total_us" lines to stdout.
extern "C" int tha_trace_dump(int skip) {
  cudaDeviceSynchronize();
  std::map<std::string, std::pair<int, double>> agg;
  double total = 0.0;
  {
    std::lock_guard<std::mutex> lk(tha::g_mu);
    auto& v = tha::g_trace_ev;
    for (size_t i = 1; i < v.size(); i++) {
      if ((int)i < skip) continue;
      float ms = 0.f;
      if (cudaEventElapsedTime(&ms, v[i - 1].second, v[i].second) != cudaSuccess) { cudaGetLastError(); continue; }
      const char* name = nullptr;
      if (cudaFuncGetName(&name, v[i].first) != cudaSuccess || !name) { cudaGetLastError(); name = "?"; }
      auto& a = agg[std::string(name).substr(0, 60)];
      a.first++; a.second += ms * 1e3; total += ms * 1e3;
    }
    for (auto& e : v) cudaEventDestroy(e.second);
    v.clear();
  }
  for (auto& kv : agg) printf("%-62s n=%5d avg %8.2f us total %10.1f us (%.1f%%)\n", kv.first.c_str(), kv.second.first, kv.second.second / kv.second.first, kv.second.second, 100.0 * kv.second.second / total);
  printf("total %.1f us\n", total);
  fflush(stdout);
  return 0;
}
