// The batched decode step (thaDNN_s_forward_batch, src/thaDNN.cpp:13-81) as
// 5 fused kernels per layer + embedding + classifier, replayed from a CUDA graph:
//
//   embed                       x[b] = tok_emb[token[b]]
//   per layer l:
//     gemv<QKV>     rmsnorm(x, rms_att[l]) -> wq|wk|wv -> RoPE -> q, K-cache[pos], V-cache[pos]
//     attn_fused    QK^T / online softmax / PV over the cache rows 0..pos           -> xb
//     gemv<RESID>   x += wo . xb
//     gemv<SWIGLU>  rmsnorm(x, rms_ffn[l]) -> silu(w1 . xb) * (w3 . xb)             -> hb
//     gemv<RESID>   x += w2 . hb
//   gemv<PLAIN>    rmsnorm(x, rms_final) -> wcls                                   -> logits
//   [argmax]       greedy next token on the device
//
// versus the reference's L*(12+4B)+2 launches, B blocking D2D copies, a per-step
// hipMalloc/hipFree and a device-wide sync.  Every kernel carries the programmatic
// dependent-launch attribute: its weight prefetch overlaps its predecessor's tail.
#include "common.cuh"
#include "gemv.cuh"
#include "tp.h"
#include "workspace.h"

namespace tha {

int embed_launch(cudaStream_t st, int B, float* x, const float* table, const int* token, const int* pos, int dim, bool pdl);
int argmax_launch(cudaStream_t st, int B, const float* logits, int vocab, int* next, bool pdl);
bool attn_fused_supported(int n_heads, int head_size, int dim, int kv_dim);
int attn_fused_prepare(Workspace* ws, int B, int n_heads, int n_kv_heads, int head_size, int seq_len);
int attn_fused_launch(cudaStream_t stream, Workspace* ws, int B, const int* pos_d, int n_heads, int n_kv_heads,
                      const float* q, float* out, const float* kc, const float* vc, long long kv_batch_stride,
                      int head_size, int seq_len, bool pdl, float* out_hi = nullptr, float* out_lo = nullptr,
                      int8_t* out_q = nullptr, float* out_sT = nullptr, int ldxs = 0, int gs = 0, const KvPaged* pg = nullptr);
int launches_reset();
int launches_get();
// megastep.cu: the whole step as one persistent kernel (small models, B <= 8)
int mega_enabled();
void mega_set_enabled(int v);
bool mega_supported(const tha_config_t* c, int B);
int mega_step_launch(cudaStream_t st, Workspace* ws, int B, const tha_config_t* cfg, const tha_weights_f32_t* w,
                     const tha_state_t* s, int* token_d, int* pos_d, float* logits_d, int* next_d, int advance,
                     const float* rope_cos, const float* rope_sin, bool prepare_only);
int mega_check_abort(int dev);
// step_tc.cu: the step with tensor-core weight matmuls (more than 8 slots)
bool step_tc_supported(const tha_config_t* c, int B);
int step_tc_prepare(Workspace* ws, const tha_config_t* c, int B, const TpCtx* tp = nullptr);
int step_tc_issue(cudaStream_t st, Workspace* ws, int B, const tha_config_t& c, const tha_weights_f32_t& w,
                  const tha_state_t& s, const int* token_d, const int* pos_d, float* logits, int* next_d,
                  const float* rc, const float* rs, bool pdl, const TpCtx* tp = nullptr);
bool prefill_tc_supported(const tha_config_t* c, const TpCtx* tp = nullptr);
int step_tc_prefill(cudaStream_t st, Workspace* ws, const tha_config_t& c, const tha_weights_f32_t* w32,
                    const tha_weights_q8_t* wq, const tha_state_t& s, int n_seqs, const int* slot, const int* pos0,
                    const int* len, const int* tokens, float* logits, int* next_d, const float* rc, const float* rs,
                    const TpCtx* tp = nullptr);
bool step_tc_q8_supported(const tha_config_t* c, int B, int gs, const TpCtx* tp = nullptr);
bool step_tc_tp_supported(const tha_config_t* c, int B, const TpCtx* tp);
int step_tc_q8_prepare(Workspace* ws, const tha_config_t* c, int B, int gs, const TpCtx* tp = nullptr);
int step_tc_q8_issue(cudaStream_t st, Workspace* ws, int B, const tha_config_t& c, const tha_weights_q8_t& wq,
                     const tha_state_t& s, const int* token_d, const int* pos_d, float* logits, int* next_d,
                     const float* rc, const float* rs, bool pdl, const TpCtx* tp = nullptr);
TpCtx* tp_ctx(tha_tp_t* t);
int gemm_tc_check_error();
int tc_mode();   // -1 auto (more than TC_MIN_SLOTS slots), 0 never, 1 whenever the shape allows
// From five slots on the tensor-core row pipeline wins over both small-batch paths (measured after the round-2 read-out fixes):
// stories110M fp32 at 5 / 8 slots 0.476 / 0.485 ms per step against the persistent step kernel's 0.537 / 0.559; Llama-2-7B fp32 at
// 8 slots 6.83 ms against the fused GEMV step's 10.71 (0.60 vs 0.38 of HBM); at 4 slots the two are level, below the GEMV / mega paths win.
constexpr int TC_MIN_SLOTS = 4;
int advance_launch(cudaStream_t st, int B, int* token, int* pos, const int* next, int seq_len, bool pdl);
int sample_launch(cudaStream_t st, int B, float* logits, int vocab, float temperature, float topp, unsigned long long* rng,
                  const unsigned long long* reseed, const int* force, const int* pos, int* next, bool pdl);

static thread_local int g_last_step_launches = 0;
#define THA_TRY_RC2(expr) do { int _st = (expr); if (_st != THA_STATUS_SUCCESS) return _st; } while (0)

// A step of R >= dual_min_rows() rows is issued as TWO concurrent branches of half the rows each (fork / join inside the
// captured graph).  Every kernel of the row pipeline costs ~5-10 us of dependent latency whatever its size (launch ->
// griddepcontrol.wait -> load -> compute -> flush), 9-10 kernels per layer; two independent chains overlap each other's
// latency and the attention kernel of one (HBM-bound) runs under the matmuls of the other.  The weights are read by
// both branches; they run a layer apart at most, so the second read is an L2 hit (28 MB per layer of stories110M).
struct DualPlan { bool on; int B0, B1; Workspace* ws2; cudaStream_t st2; int sms; };
// The two branches must RUN side by side, not one after the other: every matmul launch of a branch is planned for, and
// limited to, half of the SMs (THA_DUAL_SMS overrides) -- a persistent matmul that takes all 148 SMs leaves nothing for its twin.
static int dual_sms() {
  static int v = -1;
  if (v < 0) { const char* e = getenv("THA_DUAL_SMS"); v = (e && *e) ? atoi(e) : 74; }
  return v;
}
struct SmBudgetScope { explicit SmBudgetScope(int n) { set_sm_budget(n); } ~SmBudgetScope() { set_sm_budget(0); } };
// Not for paged states (the second branch addresses its slots through the same block table rows) and not with chained matmuls.
static int dual_plan(Workspace* ws, int B, DualPlan* d, bool paged) {
  d->on = false; d->B0 = B; d->B1 = 0; d->ws2 = nullptr; d->st2 = nullptr; d->sms = 0;
  const int m = dual_min_rows();
  if (m <= 0 || B < m || B < 2 || paged || chain_mode()) return THA_STATUS_SUCCESS;
  THA_TRY_RC2(ws_ensure_dual(ws));
  d->on = true; d->sms = dual_sms(); d->B0 = (B + 1) / 2; d->B1 = B - d->B0; d->ws2 = ws->sibling; d->st2 = ws->side_stream;
  return THA_STATUS_SUCCESS;
}
static tha_state_t state_rows(const tha_state_t& s, const tha_config_t& c, int row0) {
  const int hs = c.dim / c.n_heads, kvd = hs * c.n_kv_heads;
  const size_t kv_bs = (size_t)c.n_layers * c.seq_len * kvd;
  tha_state_t t = s;
  t.x = s.x + (size_t)row0 * c.dim; t.q = s.q + (size_t)row0 * c.dim; t.xb = s.xb + (size_t)row0 * c.dim;
  t.key_cache = s.key_cache + row0 * kv_bs; t.value_cache = s.value_cache + row0 * kv_bs;
  return t;
}

struct StepArgs {
  int B; tha_config_t cfg; tha_weights_f32_t w; tha_state_t s;
  const int* token_d; const int* pos_d; float* logits_d; int* next_d; int pdl; int mega; int advance;
};

static unsigned long long fnv1a(const void* data, size_t n, unsigned long long h = 1469598103934665603ull) {
  const unsigned char* p = (const unsigned char*)data;
  for (size_t i = 0; i < n; i++) { h ^= p[i]; h *= 1099511628211ull; }
  return h;
}

#define THA_TRY(expr) do { int _st = (expr); if (_st != THA_STATUS_SUCCESS) return _st; } while (0)


// Issue the step directly, or (default) capture it once per distinct argument set into a
// CUDA graph and replay it.  `key`/`key_len`: every value the launches depend on.
template <typename IssueFn>
static int run_step(cudaStream_t st, Workspace* ws, const void* key_data, size_t key_len, IssueFn issue, bool do_launch = true) {
  std::lock_guard<std::mutex> guard(ws->mu);
  if (!ws->step_done) THA_CUDA_OK(cudaEventCreateWithFlags(&ws->step_done, cudaEventDisableTiming));
  else if (ws->last_stream != st) THA_CUDA_OK(cudaStreamWaitEvent(st, ws->step_done, 0));   // the previous step (another stream) still owns the scratch
  ws->last_stream = st;
  struct StepDone {   // recorded behind whatever this call put on the stream
    Workspace* ws; cudaStream_t st; bool armed;
    ~StepDone() { if (armed) cudaEventRecord(ws->step_done, st); }
  } step_done_guard{ws, st, do_launch};
  launches_reset();
  const bool use_graph = graph_enabled() && st != nullptr && st != cudaStreamLegacy;
  if (!use_graph) {
    if (!do_launch) return THA_STATUS_SUCCESS;
    int rc2 = issue();
    g_last_step_launches = launches_get();
    return rc2;
  }
  // The cache is keyed by a hash of the argument bytes; a hit is only trusted after comparing the bytes themselves (the
  // argument structs of the step variants differ in size, and the size is compared first): a colliding entry is recaptured.
  const unsigned long long key = fnv1a(key_data, key_len);
  auto it = ws->graphs.find(key);
  if (it != ws->graphs.end() && (it->second.key.size() != key_len || memcmp(it->second.key.data(), key_data, key_len) != 0)) {
    if (it->second.exec) cudaGraphExecDestroy(it->second.exec);
    ws->graphs.erase(it);
    it = ws->graphs.end();
  }
  const bool fresh = it == ws->graphs.end();
  if (fresh) {
    if (ws->graphs.size() >= WS_MAX_GRAPHS) {   // bounded: drop the least recently replayed graph
      auto victim = ws->graphs.begin();
      for (auto g = ws->graphs.begin(); g != ws->graphs.end(); ++g)
        if (g->second.last_use < victim->second.last_use) victim = g;
      if (victim->second.exec) cudaGraphExecDestroy(victim->second.exec);
      ws->graphs.erase(victim);
    }
    cudaGraph_t graph = nullptr;
    THA_CUDA_OK(cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal));
    int rc2 = issue();
    cudaError_t e = cudaStreamEndCapture(st, &graph);
    if (rc2 != THA_STATUS_SUCCESS) { if (graph) cudaGraphDestroy(graph); return rc2; }
    if (e != cudaSuccess) { set_last_error("cudaStreamEndCapture", e); return THA_STATUS_EXECUTION_FAILED; }
    StepGraph sg;
    sg.launches = launches_get();
    sg.key.assign(static_cast<const unsigned char*>(key_data), static_cast<const unsigned char*>(key_data) + key_len);
    e = cudaGraphInstantiate(&sg.exec, graph, 0);
    cudaGraphDestroy(graph);
    if (e != cudaSuccess) { set_last_error("cudaGraphInstantiate", e); return THA_STATUS_EXECUTION_FAILED; }
    it = ws->graphs.emplace(key, sg).first;
  }
  it->second.last_use = ++ws->graph_clock;
  if (do_launch) THA_CUDA_OK(cudaGraphLaunch(it->second.exec, st));
  else if (fresh) THA_CUDA_OK(cudaGraphUpload(it->second.exec, st));   // prepare only: the executable graph is resident, nothing runs
  g_last_step_launches = it->second.launches;
  return THA_STATUS_SUCCESS;
}

static int issue_step_f32(cudaStream_t st, Workspace* ws, const StepArgs& a, const float* rc, const float* rs) {
  const tha_config_t& c = a.cfg;
  const int d = c.dim, hd = c.hidden_dim, hs = d / c.n_heads, kvd = hs * c.n_kv_heads;
  const int kv_mul = c.n_heads / c.n_kv_heads;
  const long long kv_bs = (long long)c.n_layers * c.seq_len * kvd;
  const bool pdl = a.pdl != 0;
  const bool fused = attn_fused_supported(c.n_heads, hs, d, kvd);
  THA_TRY(embed_launch(st, a.B, a.s.x, a.w.token_embedding_table, a.token_d, a.pos_d, d, pdl));
  for (int l = 0; l < c.n_layers; l++) {
    const int loff = l * c.seq_len * kvd;
    GemvP p = {};
    p.W0 = a.w.wq + (size_t)l * d * d; p.M0 = d;
    p.W1 = a.w.wk + (size_t)l * d * kvd; p.M1 = kvd;
    p.W2 = a.w.wv + (size_t)l * d * kvd; p.M2 = kvd;
    p.K = d; p.B = a.B; p.X = a.s.x; p.x_stride = d;
    p.rms_w = a.w.rms_att_weight + (size_t)l * d; p.pos = a.pos_d;
    p.out = a.s.q; p.out_batch_stride = d;
    p.kc = a.s.key_cache; p.vc = a.s.value_cache; p.kv_batch_stride = kv_bs; p.loff = loff;
    p.kv_dim = kvd; p.head_size = hs; p.rope_cos = rc; p.rope_sin = rs;
    THA_TRY(gemv_launch(st, GEMV_QKV, p, pdl));

    if (fused) {
      THA_TRY(attn_fused_launch(st, ws, a.B, a.pos_d, c.n_heads, c.n_kv_heads, a.s.q, a.s.xb, a.s.key_cache + loff,
                                a.s.value_cache + loff, kv_bs, hs, c.seq_len, pdl));
    } else {
      THA_TRY(tha_mha_scores_v1_f32(st, a.B, a.pos_d, c.n_heads, c.n_layers, a.s.q, a.s.att, a.s.key_cache, hs,
                                    c.seq_len, loff, kvd, d, kv_mul));
      THA_TRY(tha_mha_softmax_f32(st, a.B, a.s.att, a.pos_d, c.seq_len, c.n_heads));
      THA_TRY(tha_mha_weighted_v1_f32(st, a.B, a.pos_d, c.n_heads, a.s.xb, a.s.att, a.s.value_cache, hs, c.seq_len,
                                      loff, kvd, kv_mul, d, c.n_layers));
    }

    GemvP o = {};
    o.W0 = a.w.wo + (size_t)l * d * d; o.M0 = d; o.K = d; o.B = a.B; o.X = a.s.xb; o.x_stride = d;
    o.pos = a.pos_d; o.out = a.s.x; o.out_batch_stride = d;
    THA_TRY(gemv_launch(st, GEMV_RESID, o, pdl));

    GemvP f = {};
    f.W0 = a.w.w1 + (size_t)l * d * hd; f.W1 = a.w.w3 + (size_t)l * d * hd; f.M0 = hd; f.M1 = hd;
    f.K = d; f.B = a.B; f.X = a.s.x; f.x_stride = d; f.rms_w = a.w.rms_ffn_weight + (size_t)l * d;
    f.pos = a.pos_d; f.out = a.s.hb; f.out_batch_stride = hd;
    THA_TRY(gemv_launch(st, GEMV_SWIGLU, f, pdl));

    GemvP g = {};
    g.W0 = a.w.w2 + (size_t)l * d * hd; g.M0 = d; g.K = hd; g.B = a.B; g.X = a.s.hb; g.x_stride = hd;
    g.pos = a.pos_d; g.out = a.s.x; g.out_batch_stride = d;
    THA_TRY(gemv_launch(st, GEMV_RESID, g, pdl));
  }
  float* logits = a.logits_d ? a.logits_d : a.s.logits;
  GemvP cl = {};
  cl.W0 = a.w.wcls; cl.M0 = c.vocab_size; cl.K = d; cl.B = a.B; cl.X = a.s.x; cl.x_stride = d;
  cl.rms_w = a.w.rms_final_weight; cl.pos = a.pos_d; cl.out = logits; cl.out_batch_stride = c.vocab_size;
  THA_TRY(gemv_launch(st, GEMV_PLAIN, cl, pdl));
  if (a.next_d) {
    THA_TRY(argmax_launch(st, a.B, logits, c.vocab_size, a.next_d, pdl));
  }
  return THA_STATUS_SUCCESS;
}

static int validate_f32(int B, const tha_config_t* c, const tha_weights_f32_t* w, const tha_state_t* s) {
  if (B <= 0 || !c || !w || !s) return THA_STATUS_INVALID_ARG;
  if (c->dim <= 0 || c->n_heads <= 0 || c->n_kv_heads <= 0 || c->dim % c->n_heads || c->n_heads % c->n_kv_heads)
    return THA_STATUS_INVALID_ARG;
  const int hs = c->dim / c->n_heads;
  if (c->dim % 4 || c->hidden_dim % 4 || (hs & 1) || ((hs * c->n_kv_heads) % 4)) return THA_STATUS_NOT_SUPPORTED;
  if (!w->token_embedding_table || !w->wq || !w->wk || !w->wv || !w->wo || !w->w1 || !w->w2 || !w->w3 || !w->wcls ||
      !w->rms_att_weight || !w->rms_ffn_weight || !w->rms_final_weight)
    return THA_STATUS_INVALID_ARG;
  if (!s->x || !s->xb || !s->hb || !s->q || !s->key_cache || !s->value_cache) return THA_STATUS_INVALID_ARG;
  return THA_STATUS_SUCCESS;
}

int decode_step_f32(cudaStream_t st, int B, const tha_config_t* cfg, const tha_weights_f32_t* w, const tha_state_t* s,
                    const int* token_d, const int* pos_d, float* logits_d, int* next_d, int advance) {
  THA_TRY(validate_f32(B, cfg, w, s));
  if (!token_d || !pos_d || (!logits_d && !s->logits)) return THA_STATUS_INVALID_ARG;
  if (advance && !next_d) return THA_STATUS_INVALID_ARG;
  Workspace* ws = workspace();
  if (!ws) return THA_STATUS_EXECUTION_FAILED;
  const int hs = cfg->dim / cfg->n_heads, kvd = hs * cfg->n_kv_heads;
  const float *rc, *rs;
  THA_TRY(ws_rope(ws, hs, cfg->seq_len, &rc, &rs));
  StepArgs a;
  memset(&a, 0, sizeof(a));
  a.B = B; a.cfg = *cfg; a.w = *w; a.s = *s; a.token_d = token_d; a.pos_d = pos_d; a.logits_d = logits_d;
  a.next_d = next_d; a.pdl = pdl_enabled(); a.advance = advance;
  const int tcm = tc_mode();
  // a paged KV cache (tha_kv_pool_register) is read and written by the row pipeline's kernels only
  const bool paged = kv_pool_find(s->key_cache, nullptr);
  if (paged && !step_tc_supported(cfg, B)) return THA_STATUS_NOT_SUPPORTED;
  const bool use_tc = paged || (tcm != 0 && (tcm == 1 || B > TC_MIN_SLOTS) && step_tc_supported(cfg, B));
  a.mega = !use_tc && mega_enabled() && mega_supported(cfg, B);
  if (a.mega) {
    // scratch is grown outside any capture; the kernel writes logits only where the caller wants them
    THA_TRY(mega_step_launch(st, ws, B, cfg, w, s, nullptr, nullptr, nullptr, nullptr, 0, rc, rs, true));
    float* lg = logits_d ? logits_d : (next_d ? nullptr : s->logits);
    return run_step(st, ws, &a, sizeof(a), [&]() {
      return mega_step_launch(st, ws, B, cfg, w, s, const_cast<int*>(token_d), const_cast<int*>(pos_d), lg, next_d,
                              advance, rc, rs, false);
    });
  }
  if (use_tc) {
    DualPlan dp;
    THA_TRY(dual_plan(ws, B, &dp, paged));
    SmBudgetScope budget(dp.sms);
    THA_TRY(step_tc_prepare(ws, cfg, dp.B0));
    if (dp.on) THA_TRY(step_tc_prepare(dp.ws2, cfg, dp.B1));
    a.mega = dp.on ? 3 : 2;   // part of the graph key
    float* lg = logits_d ? logits_d : s->logits;
    return run_step(st, ws, &a, sizeof(a), [&]() {
      const bool pdl = a.pdl != 0;
      if (dp.on) {   // the second half of the slots: a concurrent branch
        THA_CUDA_OK(cudaEventRecord(ws->ev_fork, st));
        THA_CUDA_OK(cudaStreamWaitEvent(dp.st2, ws->ev_fork, 0));
        const tha_state_t s2 = state_rows(*s, *cfg, dp.B0);
        int* nx2 = next_d ? next_d + dp.B0 : nullptr;
        THA_TRY(step_tc_issue(dp.st2, dp.ws2, dp.B1, *cfg, *w, s2, token_d + dp.B0, pos_d + dp.B0,
                              lg + (size_t)dp.B0 * cfg->vocab_size, nx2, rc, rs, pdl));
        if (advance) THA_TRY(advance_launch(dp.st2, dp.B1, const_cast<int*>(token_d) + dp.B0, const_cast<int*>(pos_d) + dp.B0, nx2, cfg->seq_len, pdl));
        THA_CUDA_OK(cudaEventRecord(ws->ev_join, dp.st2));
      }
      THA_TRY(step_tc_issue(st, ws, dp.B0, *cfg, *w, *s, token_d, pos_d, lg, next_d, rc, rs, pdl));
      if (advance) THA_TRY(advance_launch(st, dp.B0, const_cast<int*>(token_d), const_cast<int*>(pos_d), next_d, cfg->seq_len, pdl));
      if (dp.on) THA_CUDA_OK(cudaStreamWaitEvent(st, ws->ev_join, 0));
      return (int)THA_STATUS_SUCCESS;
    });
  }
  const bool fused = attn_fused_supported(cfg->n_heads, hs, cfg->dim, kvd);
  if (!fused && !s->att) return THA_STATUS_INVALID_ARG;
  if (fused) THA_TRY(attn_fused_prepare(ws, B, cfg->n_heads, cfg->n_kv_heads, hs, cfg->seq_len));   // grow the split scratch now: no allocation may happen inside a capture
  return run_step(st, ws, &a, sizeof(a), [&]() {
    THA_TRY(issue_step_f32(st, ws, a, rc, rs));
    if (advance) THA_TRY(advance_launch(st, B, const_cast<int*>(token_d), const_cast<int*>(pos_d), next_d, cfg->seq_len, a.pdl != 0));
    return (int)THA_STATUS_SUCCESS;
  });
}

// ---------------------------------------------------------------------------- Q8_0 ----
int gemv_q8_launch(cudaStream_t stream, int mode, const GemvP& p, bool pdl);
int q8_prep_launch(cudaStream_t st, int B, const float* X, int x_stride, const float* rms_w, const int* pos, int K, int gs,
                   int8_t* xq, float* xs, bool pdl);

struct StepArgsQ8 {
  int B; tha_config_t cfg; tha_weights_q8_t w; tha_state_t s;
  const int* token_d; const int* pos_d; float* logits_d; int* next_d; int pdl; int tc;
};

// runq.c:344-481 with the same 5-kernels-per-layer structure; every matmul kernel
// quantises its own input rows in its prologue (runq.c:367,437,450,465,478).
static int issue_step_q8(cudaStream_t st, Workspace* ws, const StepArgsQ8& a, const float* rc, const float* rs) {
  const tha_config_t& c = a.cfg;
  const int d = c.dim, hd = c.hidden_dim, hs = d / c.n_heads, kvd = hs * c.n_kv_heads, gs = a.w.group_size;
  const int kv_mul = c.n_heads / c.n_kv_heads;
  const long long kv_bs = (long long)c.n_layers * c.seq_len * kvd;
  const bool pdl = a.pdl != 0;
  const bool fused = attn_fused_supported(c.n_heads, hs, d, kvd);
  THA_TRY(embed_launch(st, a.B, a.s.x, a.w.token_embedding_table, a.token_d, a.pos_d, d, pdl));
  for (int l = 0; l < c.n_layers; l++) {
    const int loff = l * c.seq_len * kvd;
    GemvP p = {};
    p.gs = gs;
    p.Q0 = a.w.wq.q + (size_t)l * d * d;   p.S0 = a.w.wq.s + (size_t)l * d * d / gs;   p.M0 = d;
    p.Q1 = a.w.wk.q + (size_t)l * d * kvd; p.S1 = a.w.wk.s + (size_t)l * d * kvd / gs; p.M1 = kvd;
    p.Q2 = a.w.wv.q + (size_t)l * d * kvd; p.S2 = a.w.wv.s + (size_t)l * d * kvd / gs; p.M2 = kvd;
    p.K = d; p.B = a.B; p.X = a.s.x; p.x_stride = d;
    p.rms_w = a.w.rms_att_weight + (size_t)l * d; p.pos = a.pos_d;
    p.out = a.s.q; p.out_batch_stride = d;
    p.kc = a.s.key_cache; p.vc = a.s.value_cache; p.kv_batch_stride = kv_bs; p.loff = loff;
    p.kv_dim = kvd; p.head_size = hs; p.rope_cos = rc; p.rope_sin = rs;
    THA_TRY(q8_prep_launch(st, a.B, a.s.x, d, p.rms_w, a.pos_d, d, gs, ws->q8_xq, ws->q8_xs, pdl));
    p.XQ = ws->q8_xq; p.XS = ws->q8_xs;
    THA_TRY(gemv_q8_launch(st, GEMV_QKV, p, pdl));

    if (fused) {
      THA_TRY(attn_fused_launch(st, ws, a.B, a.pos_d, c.n_heads, c.n_kv_heads, a.s.q, a.s.xb, a.s.key_cache + loff,
                                a.s.value_cache + loff, kv_bs, hs, c.seq_len, pdl));
    } else {
      THA_TRY(tha_mha_scores_v1_f32(st, a.B, a.pos_d, c.n_heads, c.n_layers, a.s.q, a.s.att, a.s.key_cache, hs,
                                    c.seq_len, loff, kvd, d, kv_mul));
      THA_TRY(tha_mha_softmax_f32(st, a.B, a.s.att, a.pos_d, c.seq_len, c.n_heads));
      THA_TRY(tha_mha_weighted_v1_f32(st, a.B, a.pos_d, c.n_heads, a.s.xb, a.s.att, a.s.value_cache, hs, c.seq_len,
                                      loff, kvd, kv_mul, d, c.n_layers));
    }

    GemvP o = {};
    o.gs = gs; o.Q0 = a.w.wo.q + (size_t)l * d * d; o.S0 = a.w.wo.s + (size_t)l * d * d / gs; o.M0 = d; o.K = d;
    o.B = a.B; o.X = a.s.xb; o.x_stride = d; o.pos = a.pos_d; o.out = a.s.x; o.out_batch_stride = d;
    THA_TRY(q8_prep_launch(st, a.B, a.s.xb, d, nullptr, a.pos_d, d, gs, ws->q8_xq, ws->q8_xs, pdl));
    o.XQ = ws->q8_xq; o.XS = ws->q8_xs;
    THA_TRY(gemv_q8_launch(st, GEMV_RESID, o, pdl));

    GemvP f = {};
    f.gs = gs;
    f.Q0 = a.w.w1.q + (size_t)l * d * hd; f.S0 = a.w.w1.s + (size_t)l * d * hd / gs; f.M0 = hd;
    f.Q1 = a.w.w3.q + (size_t)l * d * hd; f.S1 = a.w.w3.s + (size_t)l * d * hd / gs; f.M1 = hd;
    f.K = d; f.B = a.B; f.X = a.s.x; f.x_stride = d; f.rms_w = a.w.rms_ffn_weight + (size_t)l * d;
    f.pos = a.pos_d; f.out = a.s.hb; f.out_batch_stride = hd;
    THA_TRY(q8_prep_launch(st, a.B, a.s.x, d, f.rms_w, a.pos_d, d, gs, ws->q8_xq, ws->q8_xs, pdl));
    f.XQ = ws->q8_xq; f.XS = ws->q8_xs;
    THA_TRY(gemv_q8_launch(st, GEMV_SWIGLU, f, pdl));

    GemvP g = {};
    g.gs = gs; g.Q0 = a.w.w2.q + (size_t)l * d * hd; g.S0 = a.w.w2.s + (size_t)l * d * hd / gs; g.M0 = d; g.K = hd;
    g.B = a.B; g.X = a.s.hb; g.x_stride = hd; g.pos = a.pos_d; g.out = a.s.x; g.out_batch_stride = d;
    THA_TRY(q8_prep_launch(st, a.B, a.s.hb, hd, nullptr, a.pos_d, hd, gs, ws->q8_xq, ws->q8_xs, pdl));
    g.XQ = ws->q8_xq; g.XS = ws->q8_xs;
    THA_TRY(gemv_q8_launch(st, GEMV_RESID, g, pdl));
  }
  float* logits = a.logits_d ? a.logits_d : a.s.logits;
  GemvP cl = {};
  cl.gs = gs; cl.Q0 = a.w.wcls.q; cl.S0 = a.w.wcls.s; cl.M0 = c.vocab_size; cl.K = d; cl.B = a.B; cl.X = a.s.x;
  cl.x_stride = d; cl.rms_w = a.w.rms_final_weight; cl.pos = a.pos_d; cl.out = logits; cl.out_batch_stride = c.vocab_size;
  THA_TRY(q8_prep_launch(st, a.B, a.s.x, d, cl.rms_w, a.pos_d, d, gs, ws->q8_xq, ws->q8_xs, pdl));
  cl.XQ = ws->q8_xq; cl.XS = ws->q8_xs;
  THA_TRY(gemv_q8_launch(st, GEMV_PLAIN, cl, pdl));
  if (a.next_d) THA_TRY(argmax_launch(st, a.B, logits, c.vocab_size, a.next_d, pdl));
  return THA_STATUS_SUCCESS;
}

int decode_step_q8(cudaStream_t st, int B, const tha_config_t* cfg, const tha_weights_q8_t* w, const tha_state_t* s,
                   const int* token_d, const int* pos_d, float* logits_d, int* next_d) {
  if (B <= 0 || !cfg || !w || !s || !token_d || !pos_d || (!logits_d && !s->logits)) return THA_STATUS_INVALID_ARG;
  if (cfg->dim <= 0 || cfg->n_heads <= 0 || cfg->n_kv_heads <= 0 || cfg->dim % cfg->n_heads ||
      cfg->n_heads % cfg->n_kv_heads || w->group_size <= 0)
    return THA_STATUS_INVALID_ARG;
  const int hs = cfg->dim / cfg->n_heads, kvd = hs * cfg->n_kv_heads, gs = w->group_size;
  if (cfg->dim % gs || cfg->hidden_dim % gs || cfg->dim % 16 || cfg->hidden_dim % 16 || (hs & 1) || kvd % 4)
    return THA_STATUS_NOT_SUPPORTED;
  if (!w->token_embedding_table || !w->wq.q || !w->wk.q || !w->wv.q || !w->wo.q || !w->w1.q || !w->w2.q || !w->w3.q ||
      !w->wcls.q || !w->rms_att_weight || !w->rms_ffn_weight || !w->rms_final_weight)
    return THA_STATUS_INVALID_ARG;
  if (!s->x || !s->xb || !s->hb || !s->q || !s->key_cache || !s->value_cache) return THA_STATUS_INVALID_ARG;
  Workspace* ws = workspace();
  if (!ws) return THA_STATUS_EXECUTION_FAILED;
  const float *rc, *rs;
  THA_TRY(ws_rope(ws, hs, cfg->seq_len, &rc, &rs));
  StepArgsQ8 a;
  memset(&a, 0, sizeof(a));
  a.B = B; a.cfg = *cfg; a.w = *w; a.s = *s; a.token_d = token_d; a.pos_d = pos_d; a.logits_d = logits_d;
  a.next_d = next_d; a.pdl = pdl_enabled();
  // Default for every batch size the shapes allow: the int8 tensor-core matmul (gemm_q8i_kernel) streams the weights
  // faster than the dp4a GEMV at 1 slot already (7B: 3.6 vs 3.9 ms) and does not re-read them per batch tile of 8.
  // Option "tc" = 0 keeps the GEMV step.
  const int tcm = tc_mode();
  if (tcm != 0 && step_tc_q8_supported(cfg, B, gs)) {
    DualPlan dp;
    THA_TRY(dual_plan(ws, B, &dp, kv_pool_find(s->key_cache, nullptr)));
    SmBudgetScope budget(dp.sms);
    THA_TRY(step_tc_q8_prepare(ws, cfg, dp.B0, gs));
    if (dp.on) THA_TRY(step_tc_q8_prepare(dp.ws2, cfg, dp.B1, gs));
    a.tc = dp.on ? 2 : 1;   // part of the graph key
    float* lg = logits_d ? logits_d : s->logits;
    if (!lg) return THA_STATUS_INVALID_ARG;
    return run_step(st, ws, &a, sizeof(a), [&]() {
      const bool pdl = a.pdl != 0;
      if (dp.on) {
        THA_CUDA_OK(cudaEventRecord(ws->ev_fork, st));
        THA_CUDA_OK(cudaStreamWaitEvent(dp.st2, ws->ev_fork, 0));
        const tha_state_t s2 = state_rows(*s, *cfg, dp.B0);
        THA_TRY(step_tc_q8_issue(dp.st2, dp.ws2, dp.B1, *cfg, *w, s2, token_d + dp.B0, pos_d + dp.B0,
                                 lg + (size_t)dp.B0 * cfg->vocab_size, next_d ? next_d + dp.B0 : nullptr, rc, rs, pdl));
        THA_CUDA_OK(cudaEventRecord(ws->ev_join, dp.st2));
      }
      THA_TRY(step_tc_q8_issue(st, ws, dp.B0, *cfg, *w, *s, token_d, pos_d, lg, next_d, rc, rs, pdl));
      if (dp.on) THA_CUDA_OK(cudaStreamWaitEvent(st, ws->ev_join, 0));
      return (int)THA_STATUS_SUCCESS;
    });
  }
  if (kv_pool_find(s->key_cache, nullptr)) return THA_STATUS_NOT_SUPPORTED;   // paged KV: tensor-core row pipeline only
  const bool fused = attn_fused_supported(cfg->n_heads, hs, cfg->dim, kvd);
  if (!fused && !s->att) return THA_STATUS_INVALID_ARG;
  if (fused) THA_TRY(attn_fused_prepare(ws, B, cfg->n_heads, cfg->n_kv_heads, hs, cfg->seq_len));
  THA_TRY(ws_ensure_q8(ws, B, cfg->dim > cfg->hidden_dim ? cfg->dim : cfg->hidden_dim, gs));
  return run_step(st, ws, &a, sizeof(a), [&]() { return issue_step_q8(st, ws, a, rc, rs); });
}

// ---------------------------------------------------------------------------- bf16 ----
bool step_bf16_supported(const tha_config_t* c);
int step_bf16_prepare(Workspace* ws, const tha_config_t* c, int B);
int step_bf16_issue(cudaStream_t st, Workspace* ws, int B, const tha_config_t& c, const tha_weights_bf16_t& w, const tha_state_t& s,
                    const int* token_d, const int* pos_d, float* logits, int* next_d, const float* rc, const float* rs, bool pdl);
int step_bf16_prefill(cudaStream_t st, Workspace* ws, const tha_config_t& c, const tha_weights_bf16_t& w, const tha_state_t& s,
                      int n_seqs, const int* slot, const int* pos0, const int* len, const int* tokens, float* logits, int* next_d,
                      const float* rc, const float* rs);
int bf16_convert_launch(cudaStream_t st, const float* src, uint16_t* dst, size_t n);

int embed_bf16_launch(cudaStream_t st, int B, float* x, const uint16_t* table, const int* token, const int* pos, int d, bool pdl);
int gemv_bf16_launch(cudaStream_t stream, int mode, const GemvP& p, bool pdl);

struct StepArgsBf16 {
  int B; tha_config_t cfg; tha_weights_bf16_t w; tha_state_t s;
  const int* token_d; const int* pos_d; float* logits_d; int* next_d; int pdl; int gemv;
};

// 1-8 slots: the five fused kernels per layer of issue_step_f32 over bf16 weights (gemv_bf16.cu)
static int issue_step_bf16_gemv(cudaStream_t st, Workspace* ws, const StepArgsBf16& a, float* logits, const float* rc, const float* rs) {
  const tha_config_t& c = a.cfg;
  const int d = c.dim, hd = c.hidden_dim, hs = d / c.n_heads, kvd = hs * c.n_kv_heads;
  const long long kv_bs = (long long)c.n_layers * c.seq_len * kvd;
  const bool pdl = a.pdl != 0;
  auto W = [](const uint16_t* p) { return reinterpret_cast<const float*>(p); };   // GemvP carries bf16 rows behind float pointers
  THA_TRY(embed_bf16_launch(st, a.B, a.s.x, a.w.token_embedding_table, a.token_d, a.pos_d, d, pdl));
  for (int l = 0; l < c.n_layers; l++) {
    const int loff = l * c.seq_len * kvd;
    GemvP p = {};
    p.W0 = W(a.w.wq + (size_t)l * d * d); p.M0 = d;
    p.W1 = W(a.w.wk + (size_t)l * d * kvd); p.M1 = kvd;
    p.W2 = W(a.w.wv + (size_t)l * d * kvd); p.M2 = kvd;
    p.K = d; p.B = a.B; p.X = a.s.x; p.x_stride = d;
    p.rms_w = a.w.rms_att_weight + (size_t)l * d; p.pos = a.pos_d;
    p.out = a.s.q; p.out_batch_stride = d;
    p.kc = a.s.key_cache; p.vc = a.s.value_cache; p.kv_batch_stride = kv_bs; p.loff = loff;
    p.kv_dim = kvd; p.head_size = hs; p.rope_cos = rc; p.rope_sin = rs;
    THA_TRY(gemv_bf16_launch(st, GEMV_QKV, p, pdl));
    THA_TRY(attn_fused_launch(st, ws, a.B, a.pos_d, c.n_heads, c.n_kv_heads, a.s.q, a.s.xb, a.s.key_cache + loff,
                              a.s.value_cache + loff, kv_bs, hs, c.seq_len, pdl));
    GemvP o = {};
    o.W0 = W(a.w.wo + (size_t)l * d * d); o.M0 = d; o.K = d; o.B = a.B; o.X = a.s.xb; o.x_stride = d;
    o.pos = a.pos_d; o.out = a.s.x; o.out_batch_stride = d;
    THA_TRY(gemv_bf16_launch(st, GEMV_RESID, o, pdl));
    GemvP f = {};
    f.W0 = W(a.w.w1 + (size_t)l * d * hd); f.W1 = W(a.w.w3 + (size_t)l * d * hd); f.M0 = hd; f.M1 = hd;
    f.K = d; f.B = a.B; f.X = a.s.x; f.x_stride = d; f.rms_w = a.w.rms_ffn_weight + (size_t)l * d;
    f.pos = a.pos_d; f.out = a.s.hb; f.out_batch_stride = hd;
    THA_TRY(gemv_bf16_launch(st, GEMV_SWIGLU, f, pdl));
    GemvP g = {};
    g.W0 = W(a.w.w2 + (size_t)l * d * hd); g.M0 = d; g.K = hd; g.B = a.B; g.X = a.s.hb; g.x_stride = hd;
    g.pos = a.pos_d; g.out = a.s.x; g.out_batch_stride = d;
    THA_TRY(gemv_bf16_launch(st, GEMV_RESID, g, pdl));
  }
  GemvP cl = {};
  cl.W0 = W(a.w.wcls); cl.M0 = c.vocab_size; cl.K = d; cl.B = a.B; cl.X = a.s.x; cl.x_stride = d;
  cl.rms_w = a.w.rms_final_weight; cl.pos = a.pos_d; cl.out = logits; cl.out_batch_stride = c.vocab_size;
  THA_TRY(gemv_bf16_launch(st, GEMV_PLAIN, cl, pdl));
  if (a.next_d) THA_TRY(argmax_launch(st, a.B, logits, c.vocab_size, a.next_d, pdl));
  return THA_STATUS_SUCCESS;
}

static int check_bf16(const tha_config_t* cfg, const tha_weights_bf16_t* w, const tha_state_t* s) {
  if (!cfg || !w || !s) return THA_STATUS_INVALID_ARG;
  if (!w->token_embedding_table || !w->wq || !w->wk || !w->wv || !w->wo || !w->w1 || !w->w2 || !w->w3 || !w->wcls ||
      !w->rms_att_weight || !w->rms_ffn_weight || !w->rms_final_weight)
    return THA_STATUS_INVALID_ARG;
  if (!s->x || !s->xb || !s->q || !s->key_cache || !s->value_cache) return THA_STATUS_INVALID_ARG;
  if (!step_bf16_supported(cfg)) return THA_STATUS_NOT_SUPPORTED;
  return THA_STATUS_SUCCESS;
}

int decode_step_bf16(cudaStream_t st, int B, const tha_config_t* cfg, const tha_weights_bf16_t* w, const tha_state_t* s,
                     const int* token_d, const int* pos_d, float* logits_d, int* next_d) {
  if (B <= 0 || !token_d || !pos_d) return THA_STATUS_INVALID_ARG;
  THA_TRY(check_bf16(cfg, w, s));
  float* lg = logits_d ? logits_d : s->logits;
  if (!lg) return THA_STATUS_INVALID_ARG;
  Workspace* ws = workspace();
  if (!ws) return THA_STATUS_EXECUTION_FAILED;
  const float *rc, *rs;
  THA_TRY(ws_rope(ws, cfg->dim / cfg->n_heads, cfg->seq_len, &rc, &rs));
  StepArgsBf16 a;
  memset(&a, 0, sizeof(a));
  a.B = B; a.cfg = *cfg; a.w = *w; a.s = *s; a.token_d = token_d; a.pos_d = pos_d; a.logits_d = logits_d; a.next_d = next_d;
  a.pdl = pdl_enabled();
  // 1-2 slots: the weights stream through the fused GEMV kernels (5 launches per layer); more: tensor cores (Llama-2-7B:
  // 8 slots 12.4 ms per step on the GEMV path -- shared-memory-bound at 8 FMAs per weight -- against 4.0 ms at 64 slots on
  // the tensor cores).  Option "tc" = 1 forces the tensor cores, 0 the GEMV path up to 8 slots.
  // The tensor-core row pipeline at every slot count (7 launches per layer since RoPE and the output rounding moved into the
  // attention kernel): Llama-2-7B at 1 / 2 slots 2.96 / 2.98 ms per step against the fused GEMV step's 3.53 / 4.08
  // (tools/bf16_small.py); the GEMV step remains behind tha_set_option("tc", 0).
  a.gemv = (B <= 8 && tc_mode() == 0 && s->hb && !kv_pool_find(s->key_cache, nullptr)) ? 1 : 0;
  if (a.gemv) {
    const int hs = cfg->dim / cfg->n_heads;
    THA_TRY(attn_fused_prepare(ws, B, cfg->n_heads, cfg->n_kv_heads, hs, cfg->seq_len));
    return run_step(st, ws, &a, sizeof(a), [&]() { return issue_step_bf16_gemv(st, ws, a, lg, rc, rs); });
  }
  THA_TRY(step_bf16_prepare(ws, cfg, B));
  return run_step(st, ws, &a, sizeof(a), [&]() {
    return step_bf16_issue(st, ws, B, *cfg, *w, *s, token_d, pos_d, lg, next_d, rc, rs, a.pdl != 0);
  });
}

int last_step_launches_get() { return g_last_step_launches; }
void last_step_launches_set(int v) { g_last_step_launches = v; }

// Host-side range check of a step's inputs: an idle slot carries pos < 0; anything else must index the RoPE tables,
// the slot's KV slab and the embedding table.
static int check_tokpos(int B, int vocab, int seq_len, const int* token_host, const int* pos_host) {
  for (int b = 0; b < B; b++) {
    if (pos_host[b] < 0) continue;
    if (pos_host[b] >= seq_len || token_host[b] < 0 || token_host[b] >= vocab) {
      set_last_error("token / pos out of range", cudaErrorInvalidValue);
      return THA_STATUS_INVALID_ARG;
    }
  }
  return THA_STATUS_SUCCESS;
}

// H2D of token/pos through the pinned staging block, step, D2H of the logits, sync.
template <typename StepFn>
static int forward_host(cudaStream_t st, int B, int vocab, int seq_len, float* state_logits, const int* token_host,
                        const int* pos_host, float* logits_out, StepFn step) {
  if (B <= 0 || !token_host || !pos_host || !logits_out) return THA_STATUS_INVALID_ARG;
  THA_TRY(check_tokpos(B, vocab, seq_len, token_host, pos_host));
  Workspace* ws = workspace();
  if (!ws) return THA_STATUS_EXECUTION_FAILED;
  THA_TRY(ws_ensure_tokpos(ws, B));
  const int cap = ws->tokpos_cap;
  memcpy(ws->tokpos_h, token_host, B * sizeof(int));
  memcpy(ws->tokpos_h + cap, pos_host, B * sizeof(int));
  THA_CUDA_OK(cudaMemcpyAsync(ws->tok_d, ws->tokpos_h, 2 * cap * sizeof(int), cudaMemcpyHostToDevice, st));
  float* target = state_logits ? state_logits : logits_out;
  THA_TRY(step(ws->tok_d, ws->pos_d, target));
  if (target != logits_out)
    THA_CUDA_OK(cudaMemcpyAsync(logits_out, target, (size_t)B * vocab * sizeof(float), cudaMemcpyDefault, st));
  THA_CUDA_OK(cudaStreamSynchronize(st));
  if (mega_check_abort(ws->device)) {
    set_last_error("persistent step kernel: a barrier wait timed out", cudaErrorLaunchTimeout);
    return THA_STATUS_EXECUTION_FAILED;
  }
  if (gemm_tc_check_error()) {
    set_last_error("tensor-core matmul: a pipeline barrier wait timed out", cudaErrorLaunchTimeout);
    return THA_STATUS_EXECUTION_FAILED;
  }
  return THA_STATUS_SUCCESS;
}

}  // namespace tha

using namespace tha;

extern "C" int tha_decode_step_f32(void* stream, int n_batches, const tha_config_t* cfg, const tha_weights_f32_t* w,
                                   const tha_state_t* s, const int* token_d, const int* pos_d, float* logits_d,
                                   int* next_d) {
  return decode_step_f32((cudaStream_t)stream, n_batches, cfg, w, s, token_d, pos_d, logits_d, next_d, 0);
}

extern "C" int tha_decode_greedy_f32(void* stream, int n_batches, const tha_config_t* cfg, const tha_weights_f32_t* w,
                                     const tha_state_t* s, int* token_d, int* pos_d, float* logits_d, int* next_d) {
  return decode_step_f32((cudaStream_t)stream, n_batches, cfg, w, s, token_d, pos_d, logits_d, next_d, 1);
}

extern "C" int tha_forward_batch_f32(void* stream, int n_batches, const tha_config_t* cfg, const tha_weights_f32_t* w,
                                     const tha_state_t* s, const int* token_host, const int* pos_host,
                                     float* logits_out) {
  if (!cfg || !s) return THA_STATUS_INVALID_ARG;
  cudaStream_t st = (cudaStream_t)stream;
  return forward_host(st, n_batches, cfg->vocab_size, cfg->seq_len, s->logits, token_host, pos_host, logits_out,
                      [&](const int* tok_d, const int* pos_d, float* target) {
                        return decode_step_f32(st, n_batches, cfg, w, s, tok_d, pos_d, target, nullptr, 0);
                      });
}

// Greedy variant of the host-facing step: argmax on the device, B ints back instead of B*vocab floats.
template <typename StepFn>
static int forward_host_greedy(cudaStream_t st, int B, const tha_config_t* cfg, const int* token_host, const int* pos_host,
                               int* next_host, StepFn step) {
  if (B <= 0 || !cfg || !token_host || !pos_host || !next_host) return THA_STATUS_INVALID_ARG;
  THA_TRY(check_tokpos(B, cfg->vocab_size, cfg->seq_len, token_host, pos_host));
  Workspace* ws = workspace();
  if (!ws) return THA_STATUS_EXECUTION_FAILED;
  THA_TRY(ws_ensure_tokpos(ws, B));
  const int cap = ws->tokpos_cap;
  memcpy(ws->tokpos_h, token_host, B * sizeof(int));
  memcpy(ws->tokpos_h + cap, pos_host, B * sizeof(int));
  THA_CUDA_OK(cudaMemcpyAsync(ws->tok_d, ws->tokpos_h, 2 * cap * sizeof(int), cudaMemcpyHostToDevice, st));
  THA_TRY(step(ws->tok_d, ws->pos_d, ws->next_d));
  THA_CUDA_OK(cudaMemcpyAsync(ws->next_h, ws->next_d, B * sizeof(int), cudaMemcpyDeviceToHost, st));
  THA_CUDA_OK(cudaStreamSynchronize(st));
  if (mega_check_abort(ws->device)) {
    set_last_error("persistent step kernel: a barrier wait timed out", cudaErrorLaunchTimeout);
    return THA_STATUS_EXECUTION_FAILED;
  }
  if (gemm_tc_check_error()) {
    set_last_error("tensor-core matmul: a pipeline barrier wait timed out", cudaErrorLaunchTimeout);
    return THA_STATUS_EXECUTION_FAILED;
  }
  memcpy(next_host, ws->next_h, B * sizeof(int));
  return THA_STATUS_SUCCESS;
}

// The driver's step in its default (sampled) mode: HOST token / pos / forced prompt token / sampler reseed in, HOST next
// token per slot out; softmax, top-p and the coin run on the device (sampler.cu), B ints cross PCIe instead of B * vocab
// floats (src/llama.cpp:1025-1050 around thaDNN_s_forward_batch + sample()).
template <typename StepFn>
static int forward_host_sample(cudaStream_t st, int B, const tha_config_t* cfg, float* logits_d, const int* token_host,
                               const int* pos_host, const int* force_host, const unsigned long long* reseed_host,
                               float temperature, float topp, int* next_host, StepFn step) {
  if (B <= 0 || !cfg || !logits_d || !token_host || !pos_host || !next_host || temperature < 0.0f) return THA_STATUS_INVALID_ARG;
  THA_TRY(check_tokpos(B, cfg->vocab_size, cfg->seq_len, token_host, pos_host));
  Workspace* ws = workspace();
  if (!ws) return THA_STATUS_EXECUTION_FAILED;
  THA_TRY(ws_ensure_tokpos(ws, B));
  const int cap = ws->tokpos_cap;
  int* h = ws->tokpos_h;
  memcpy(h, token_host, B * sizeof(int));
  memcpy(h + cap, pos_host, B * sizeof(int));
  for (int b = 0; b < B; b++) {
    const int f = force_host ? force_host[b] : -1;
    if (f >= cfg->vocab_size) return THA_STATUS_INVALID_ARG;
    h[2 * cap + b] = f;
  }
  unsigned long long* rs = reinterpret_cast<unsigned long long*>(h + 3 * cap);
  for (int b = 0; b < B; b++) rs[b] = reseed_host ? reseed_host[b] : 0ull;
  THA_CUDA_OK(cudaMemcpyAsync(ws->tok_d, h, 5 * cap * sizeof(int), cudaMemcpyHostToDevice, st));
  THA_TRY(step(ws->tok_d, ws->pos_d, logits_d));
  THA_TRY(sample_launch(st, B, logits_d, cfg->vocab_size, temperature, topp, ws->rng_d, ws->reseed_d, ws->force_d, ws->pos_d,
                        ws->next_d, false));
  THA_CUDA_OK(cudaMemcpyAsync(ws->next_h, ws->next_d, B * sizeof(int), cudaMemcpyDeviceToHost, st));
  THA_CUDA_OK(cudaStreamSynchronize(st));
  if (mega_check_abort(ws->device)) {
    set_last_error("persistent step kernel: a barrier wait timed out", cudaErrorLaunchTimeout);
    return THA_STATUS_EXECUTION_FAILED;
  }
  if (gemm_tc_check_error()) {
    set_last_error("tensor-core matmul: a pipeline barrier wait timed out", cudaErrorLaunchTimeout);
    return THA_STATUS_EXECUTION_FAILED;
  }
  memcpy(next_host, ws->next_h, B * sizeof(int));
  return THA_STATUS_SUCCESS;
}

extern "C" int tha_forward_sample_f32(void* stream, int n_batches, const tha_config_t* cfg, const tha_weights_f32_t* w,
                                      const tha_state_t* s, const int* token_host, const int* pos_host, const int* force_host,
                                      const unsigned long long* reseed_host, float temperature, float topp, int* next_host) {
  if (!cfg || !s || !s->logits) return THA_STATUS_INVALID_ARG;
  cudaStream_t st = (cudaStream_t)stream;
  return forward_host_sample(st, n_batches, cfg, s->logits, token_host, pos_host, force_host, reseed_host, temperature, topp,
                             next_host, [&](const int* tok_d, const int* pos_d, float* lg) {
                               return decode_step_f32(st, n_batches, cfg, w, s, tok_d, pos_d, lg, nullptr, 0);
                             });
}

extern "C" int tha_forward_sample_q8(void* stream, int n_batches, const tha_config_t* cfg, const tha_weights_q8_t* w,
                                     const tha_state_t* s, const int* token_host, const int* pos_host, const int* force_host,
                                     const unsigned long long* reseed_host, float temperature, float topp, int* next_host) {
  if (!cfg || !s || !s->logits) return THA_STATUS_INVALID_ARG;
  cudaStream_t st = (cudaStream_t)stream;
  return forward_host_sample(st, n_batches, cfg, s->logits, token_host, pos_host, force_host, reseed_host, temperature, topp,
                             next_host, [&](const int* tok_d, const int* pos_d, float* lg) {
                               return decode_step_q8(st, n_batches, cfg, w, s, tok_d, pos_d, lg, nullptr);
                             });
}

extern "C" int tha_convert_bf16(void* stream, const float* src_d, uint16_t* dst_d, size_t n) {
  if (!src_d || !dst_d) return THA_STATUS_INVALID_ARG;
  return bf16_convert_launch((cudaStream_t)stream, src_d, dst_d, n);
}

extern "C" int tha_decode_step_bf16(void* stream, int n_batches, const tha_config_t* cfg, const tha_weights_bf16_t* w,
                                    const tha_state_t* s, const int* token_d, const int* pos_d, float* logits_d, int* next_d) {
  return decode_step_bf16((cudaStream_t)stream, n_batches, cfg, w, s, token_d, pos_d, logits_d, next_d);
}

extern "C" int tha_forward_batch_bf16(void* stream, int n_batches, const tha_config_t* cfg, const tha_weights_bf16_t* w,
                                      const tha_state_t* s, const int* token_host, const int* pos_host, float* logits_out) {
  if (!cfg || !s) return THA_STATUS_INVALID_ARG;
  cudaStream_t st = (cudaStream_t)stream;
  return forward_host(st, n_batches, cfg->vocab_size, cfg->seq_len, s->logits, token_host, pos_host, logits_out,
                      [&](const int* tok_d, const int* pos_d, float* target) {
                        return decode_step_bf16(st, n_batches, cfg, w, s, tok_d, pos_d, target, nullptr);
                      });
}

extern "C" int tha_forward_greedy_bf16(void* stream, int n_batches, const tha_config_t* cfg, const tha_weights_bf16_t* w,
                                       const tha_state_t* s, const int* token_host, const int* pos_host, int* next_host) {
  cudaStream_t st = (cudaStream_t)stream;
  return forward_host_greedy(st, n_batches, cfg, token_host, pos_host, next_host,
                             [&](const int* tok_d, const int* pos_d, int* next_d) {
                               return decode_step_bf16(st, n_batches, cfg, w, s, tok_d, pos_d, nullptr, next_d);
                             });
}

extern "C" int tha_forward_sample_bf16(void* stream, int n_batches, const tha_config_t* cfg, const tha_weights_bf16_t* w,
                                       const tha_state_t* s, const int* token_host, const int* pos_host, const int* force_host,
                                       const unsigned long long* reseed_host, float temperature, float topp, int* next_host) {
  if (!cfg || !s || !s->logits) return THA_STATUS_INVALID_ARG;
  cudaStream_t st = (cudaStream_t)stream;
  return forward_host_sample(st, n_batches, cfg, s->logits, token_host, pos_host, force_host, reseed_host, temperature, topp,
                             next_host, [&](const int* tok_d, const int* pos_d, float* lg) {
                               return decode_step_bf16(st, n_batches, cfg, w, s, tok_d, pos_d, lg, nullptr);
                             });
}

extern "C" int tha_prefill_bf16(void* stream, const tha_config_t* cfg, const tha_weights_bf16_t* w, const tha_state_t* s, int n_seqs,
                                const int* slot_host, const int* pos0_host, const int* len_host, const int* tokens_host,
                                float* logits_d, int* next_d) {
  if (n_seqs < 1 || !slot_host || !pos0_host || !len_host || !tokens_host) return THA_STATUS_INVALID_ARG;
  THA_TRY(check_bf16(cfg, w, s));
  Workspace* ws = workspace();
  if (!ws) return THA_STATUS_EXECUTION_FAILED;
  const float *rc, *rs;
  THA_TRY(ws_rope(ws, cfg->dim / cfg->n_heads, cfg->seq_len, &rc, &rs));
  launches_reset();
  const int rcode = step_bf16_prefill((cudaStream_t)stream, ws, *cfg, *w, *s, n_seqs, slot_host, pos0_host, len_host, tokens_host,
                                      logits_d, next_d, rc, rs);
  last_step_launches_set(launches_get());
  return rcode;
}

extern "C" int tha_forward_greedy_f32(void* stream, int n_batches, const tha_config_t* cfg, const tha_weights_f32_t* w,
                                      const tha_state_t* s, const int* token_host, const int* pos_host, int* next_host) {
  cudaStream_t st = (cudaStream_t)stream;
  return forward_host_greedy(st, n_batches, cfg, token_host, pos_host, next_host,
                             [&](const int* tok_d, const int* pos_d, int* next_d) {
                               return decode_step_f32(st, n_batches, cfg, w, s, tok_d, pos_d, nullptr, next_d, 0);
                             });
}

extern "C" int tha_forward_greedy_q8(void* stream, int n_batches, const tha_config_t* cfg, const tha_weights_q8_t* w,
                                     const tha_state_t* s, const int* token_host, const int* pos_host, int* next_host) {
  cudaStream_t st = (cudaStream_t)stream;
  return forward_host_greedy(st, n_batches, cfg, token_host, pos_host, next_host,
                             [&](const int* tok_d, const int* pos_d, int* next_d) {
                               return decode_step_q8(st, n_batches, cfg, w, s, tok_d, pos_d, nullptr, next_d);
                             });
}

// Batched prompt ingestion: what the reference does one forced-prompt step at a time
// (src/llama.cpp:1029-1030) in one pass of tensor-core matmuls over all prompt positions.
extern "C" int tha_prefill_f32(void* stream, const tha_config_t* cfg, const tha_weights_f32_t* w, const tha_state_t* s,
                               int n_seqs, const int* slot_host, const int* pos0_host, const int* len_host,
                               const int* tokens_host, float* logits_d, int* next_d) {
  if (!cfg || !w || !s || n_seqs < 1 || !slot_host || !pos0_host || !len_host || !tokens_host) return THA_STATUS_INVALID_ARG;
  THA_TRY(validate_f32(1, cfg, w, s));
  if (!prefill_tc_supported(cfg)) return THA_STATUS_NOT_SUPPORTED;
  Workspace* ws = workspace();
  if (!ws) return THA_STATUS_EXECUTION_FAILED;
  const float *rc, *rs;
  THA_TRY(ws_rope(ws, cfg->dim / cfg->n_heads, cfg->seq_len, &rc, &rs));
  launches_reset();
  const int rcode = step_tc_prefill((cudaStream_t)stream, ws, *cfg, w, nullptr, *s, n_seqs, slot_host, pos0_host, len_host,
                                    tokens_host, logits_d, next_d, rc, rs);
  last_step_launches_set(launches_get());
  return rcode;
}

extern "C" int tha_prefill_q8(void* stream, const tha_config_t* cfg, const tha_weights_q8_t* w, const tha_state_t* s, int n_seqs,
                              const int* slot_host, const int* pos0_host, const int* len_host, const int* tokens_host,
                              float* logits_d, int* next_d) {
  if (!cfg || !w || !s || n_seqs < 1 || !slot_host || !pos0_host || !len_host || !tokens_host) return THA_STATUS_INVALID_ARG;
  if (cfg->dim <= 0 || cfg->n_heads <= 0 || cfg->dim % cfg->n_heads || cfg->n_kv_heads <= 0 || w->group_size <= 0)
    return THA_STATUS_INVALID_ARG;
  if (!w->token_embedding_table || !w->wq.q || !w->wk.q || !w->wv.q || !w->wo.q || !w->w1.q || !w->w2.q || !w->w3.q ||
      !w->wcls.q || !w->rms_att_weight || !w->rms_ffn_weight || !w->rms_final_weight || !s->key_cache || !s->value_cache)
    return THA_STATUS_INVALID_ARG;
  if (!prefill_tc_supported(cfg) || !step_tc_q8_supported(cfg, 1, w->group_size)) return THA_STATUS_NOT_SUPPORTED;
  Workspace* ws = workspace();
  if (!ws) return THA_STATUS_EXECUTION_FAILED;
  const float *rc, *rs;
  THA_TRY(ws_rope(ws, cfg->dim / cfg->n_heads, cfg->seq_len, &rc, &rs));
  launches_reset();
  const int rcode = step_tc_prefill((cudaStream_t)stream, ws, *cfg, nullptr, w, *s, n_seqs, slot_host, pos0_host, len_host,
                                    tokens_host, logits_d, next_d, rc, rs);
  last_step_launches_set(launches_get());
  return rcode;
}

extern "C" int tha_prefill_supported(const tha_config_t* cfg) { return (cfg && prefill_tc_supported(cfg)) ? 1 : 0; }

extern "C" int tha_last_step_launches(void) { return last_step_launches_get(); }

extern "C" int tha_step_needs_att_scratch(const tha_config_t* cfg) {
  if (!cfg || cfg->n_heads <= 0 || cfg->dim % cfg->n_heads) return 1;
  const int hs = cfg->dim / cfg->n_heads;
  return attn_fused_supported(cfg->n_heads, hs, cfg->dim, hs * cfg->n_kv_heads) ? 0 : 1;
}

extern "C" int tha_decode_step_q8(void* stream, int n_batches, const tha_config_t* cfg, const tha_weights_q8_t* w,
                                  const tha_state_t* s, const int* token_d, const int* pos_d, float* logits_d,
                                  int* next_d) {
  return decode_step_q8((cudaStream_t)stream, n_batches, cfg, w, s, token_d, pos_d, logits_d, next_d);
}

extern "C" int tha_forward_batch_q8(void* stream, int n_batches, const tha_config_t* cfg, const tha_weights_q8_t* w,
                                    const tha_state_t* s, const int* token_host, const int* pos_host,
                                    float* logits_out) {
  if (!cfg || !s) return THA_STATUS_INVALID_ARG;
  cudaStream_t st = (cudaStream_t)stream;
  return forward_host(st, n_batches, cfg->vocab_size, cfg->seq_len, s->logits, token_host, pos_host, logits_out,
                      [&](const int* tok_d, const int* pos_d, float* target) {
                        return decode_step_q8(st, n_batches, cfg, w, s, tok_d, pos_d, target, nullptr);
                      });
}

// ---------------------------------------------------------------------------- tensor parallel ----
// One rank's share of the decode step / prompt ingestion of a model split tp ways (tp.h).  cfg is the
// LOCAL config (n_heads, n_kv_heads, hidden_dim = this rank's share; dim = full width), w / s hold the local shards:
// wq [L, da, dim], wk/wv [L, kvd, dim], wo [L, dim, da], w1/w3 [L, hd, dim], w2 [L, dim, hd]; embedding, norms and
// classifier replicated; KV cache [B][L][S][kvd] of the local heads.  Every rank must make the same sequence of calls.
namespace tha {
struct StepArgsTp { int B; tha_config_t cfg; tha_weights_f32_t w; tha_weights_q8_t wq; tha_state_t s;
                    const int* token_d; const int* pos_d; float* logits_d; int* next_d; int pdl; int q8; int advance; const void* tp; };

static int tp_decode(cudaStream_t st, TpCtx* tp, int B, const tha_config_t* cfg, const tha_weights_f32_t* w,
                     const tha_weights_q8_t* wq, const tha_state_t* s, int* token_d, int* pos_d, float* logits_d,
                     int* next_d, int advance) {
  if (!tp || B <= 0 || !cfg || (!w && !wq) || !s || !token_d || !pos_d || (!logits_d && !s->logits)) return THA_STATUS_INVALID_ARG;
  if (advance && !next_d) return THA_STATUS_INVALID_ARG;
  if (!s->x || !s->xb || !s->q || !s->key_cache || !s->value_cache) return THA_STATUS_INVALID_ARG;
  if (wq ? !step_tc_q8_supported(cfg, B, wq->group_size, tp) : !step_tc_tp_supported(cfg, B, tp)) return THA_STATUS_NOT_SUPPORTED;
  int dev = -1;
  THA_CUDA_OK(cudaGetDevice(&dev));
  if (dev != tp->device) return THA_STATUS_INVALID_ARG;
  Workspace* ws = &tp->ws;
  const float *rc, *rs;
  THA_TRY(ws_rope(ws, tp->head_size, cfg->seq_len, &rc, &rs));
  if (wq) THA_TRY(step_tc_q8_prepare(ws, cfg, B, wq->group_size, tp));
  else THA_TRY(step_tc_prepare(ws, cfg, B, tp));
  StepArgsTp a;
  memset(&a, 0, sizeof(a));
  a.B = B; a.cfg = *cfg; if (w) a.w = *w; if (wq) a.wq = *wq; a.s = *s; a.token_d = token_d; a.pos_d = pos_d;
  a.logits_d = logits_d; a.next_d = next_d; a.pdl = pdl_enabled(); a.q8 = wq != nullptr; a.advance = advance; a.tp = tp;
  float* lg = logits_d ? logits_d : s->logits;
  return run_step(st, ws, &a, sizeof(a), [&]() {
    if (wq) THA_TRY(step_tc_q8_issue(st, ws, B, *cfg, *wq, *s, token_d, pos_d, lg, next_d, rc, rs, a.pdl != 0, tp));
    else THA_TRY(step_tc_issue(st, ws, B, *cfg, *w, *s, token_d, pos_d, lg, next_d, rc, rs, a.pdl != 0, tp));
    if (advance) THA_TRY(advance_launch(st, B, token_d, pos_d, next_d, cfg->seq_len, a.pdl != 0));
    return (int)THA_STATUS_SUCCESS;
  }, tp->mode != 1);
}
}  // namespace tha

extern "C" int tha_tp_decode_step_f32(void* stream, tha_tp_t* tp, int n_batches, const tha_config_t* cfg_local,
                                      const tha_weights_f32_t* w_local, const tha_state_t* s_local, int* token_d, int* pos_d,
                                      float* logits_d, int* next_d, int advance) {
  return tp_decode((cudaStream_t)stream, tp_ctx(tp), n_batches, cfg_local, w_local, nullptr, s_local, token_d, pos_d, logits_d,
                   next_d, advance);
}

extern "C" int tha_tp_decode_step_q8(void* stream, tha_tp_t* tp, int n_batches, const tha_config_t* cfg_local,
                                     const tha_weights_q8_t* w_local, const tha_state_t* s_local, int* token_d, int* pos_d,
                                     float* logits_d, int* next_d, int advance) {
  return tp_decode((cudaStream_t)stream, tp_ctx(tp), n_batches, cfg_local, nullptr, w_local, s_local, token_d, pos_d, logits_d,
                   next_d, advance);
}

static int tp_prefill(cudaStream_t st, TpCtx* tp, const tha_config_t* cfg, const tha_weights_f32_t* w, const tha_weights_q8_t* wq,
                      const tha_state_t* s, int n_seqs, const int* slot_host, const int* pos0_host, const int* len_host,
                      const int* tokens_host, float* logits_d, int* next_d) {
  if (!tp || !cfg || (!w && !wq) || !s || n_seqs < 1 || !slot_host || !pos0_host || !len_host || !tokens_host ||
      !s->key_cache || !s->value_cache)
    return THA_STATUS_INVALID_ARG;
  if (!prefill_tc_supported(cfg, tp)) return THA_STATUS_NOT_SUPPORTED;
  Workspace* ws = &tp->ws;
  const float *rc, *rs;
  THA_TRY(ws_rope(ws, tp->head_size, cfg->seq_len, &rc, &rs));
  launches_reset();
  TpCtx solo;   // prepare only: the same launches without the exchange
  if (tp->mode == 1) {
    solo.tp = 1; solo.rank = 0; solo.head_size = tp->head_size; solo.hidden_of[0] = cfg->hidden_dim; solo.device = tp->device;
    solo.connected = true;
  }
  const int rcode = step_tc_prefill(st, ws, *cfg, w, wq, *s, n_seqs, slot_host, pos0_host, len_host, tokens_host, logits_d,
                                    next_d, rc, rs, tp->mode == 1 ? &solo : tp);
  last_step_launches_set(launches_get());
  return rcode;
}

extern "C" int tha_tp_prefill_f32(void* stream, tha_tp_t* tp, const tha_config_t* cfg_local, const tha_weights_f32_t* w_local,
                                  const tha_state_t* s_local, int n_seqs, const int* slot_host, const int* pos0_host,
                                  const int* len_host, const int* tokens_host, float* logits_d, int* next_d) {
  return tp_prefill((cudaStream_t)stream, tp_ctx(tp), cfg_local, w_local, nullptr, s_local, n_seqs, slot_host, pos0_host,
                    len_host, tokens_host, logits_d, next_d);
}

extern "C" int tha_tp_prefill_q8(void* stream, tha_tp_t* tp, const tha_config_t* cfg_local, const tha_weights_q8_t* w_local,
                                 const tha_state_t* s_local, int n_seqs, const int* slot_host, const int* pos0_host,
                                 const int* len_host, const int* tokens_host, float* logits_d, int* next_d) {
  return tp_prefill((cudaStream_t)stream, tp_ctx(tp), cfg_local, nullptr, w_local, s_local, n_seqs, slot_host, pos0_host,
                    len_host, tokens_host, logits_d, next_d);
}

// One greedy decode step of a tensor-parallel replica whose ranks are all driven by THIS host thread (the shipped
// driver with -p N; ranks may share a device): HOST token / pos in, HOST next token per slot out (rank 0's; every rank
// computes the same).  Order of business: stage the inputs on every rank, PREPARE every rank (scratch, graph -- anything
// that may synchronise with a device), then launch all ranks back to back, then wait.
extern "C" int tha_tp_forward_greedy(int tp, tha_tp_t* const* ranks, void* const* streams, int n_batches, int q8,
                                     const tha_config_t* const* cfg_local, const void* const* w_local,
                                     const tha_state_t* const* s_local, const int* token_host, const int* pos_host,
                                     int* next_host) {
  if (tp < 1 || tp > TP_MAX || !ranks || !streams || n_batches < 1 || !cfg_local || !w_local || !s_local || !token_host ||
      !pos_host || !next_host)
    return THA_STATUS_INVALID_ARG;
  const int B = n_batches;
  if (!cfg_local[0]) return THA_STATUS_INVALID_ARG;
  THA_TRY(check_tokpos(B, cfg_local[0]->vocab_size, cfg_local[0]->seq_len, token_host, pos_host));
  for (int r = 0; r < tp; r++) {
    TpCtx* t = tp_ctx(ranks[r]);
    if (!t || !cfg_local[r] || !w_local[r] || !s_local[r]) return THA_STATUS_INVALID_ARG;
    THA_CUDA_OK(cudaSetDevice(t->device));
    Workspace* ws = &t->ws;
    THA_TRY(ws_ensure_tokpos(ws, B));
    const int cap = ws->tokpos_cap;
    memcpy(ws->tokpos_h, token_host, B * sizeof(int));
    memcpy(ws->tokpos_h + cap, pos_host, B * sizeof(int));
    THA_CUDA_OK(cudaMemcpyAsync(ws->tok_d, ws->tokpos_h, 2 * cap * sizeof(int), cudaMemcpyHostToDevice, (cudaStream_t)streams[r]));
  }
  for (int pass = 0; pass < 2; pass++) {
    for (int r = 0; r < tp; r++) {
      TpCtx* t = tp_ctx(ranks[r]);
      THA_CUDA_OK(cudaSetDevice(t->device));
      const int saved = t->mode;
      if (pass == 0) t->mode = 1;
      const int rc = tp_decode((cudaStream_t)streams[r], t, B, cfg_local[r], q8 ? nullptr : (const tha_weights_f32_t*)w_local[r],
                               q8 ? (const tha_weights_q8_t*)w_local[r] : nullptr, s_local[r], t->ws.tok_d, t->ws.pos_d, nullptr,
                               t->ws.next_d, 0);
      t->mode = saved;
      THA_TRY(rc);
    }
  }
  {
    TpCtx* t0 = tp_ctx(ranks[0]);
    THA_CUDA_OK(cudaSetDevice(t0->device));
    THA_CUDA_OK(cudaMemcpyAsync(t0->ws.next_h, t0->ws.next_d, B * sizeof(int), cudaMemcpyDeviceToHost, (cudaStream_t)streams[0]));
  }
  bool tc_err = false;
  for (int r = 0; r < tp; r++) {
    THA_CUDA_OK(cudaSetDevice(tp_ctx(ranks[r])->device));
    THA_CUDA_OK(cudaStreamSynchronize((cudaStream_t)streams[r]));
    if (gemm_tc_check_error()) tc_err = true;   // this rank's device
  }
  for (int r = 0; r < tp; r++) {
    if (tha_tp_check_error(ranks[r])) {
      set_last_error("tensor-parallel exchange: a wait on a peer timed out", cudaErrorLaunchTimeout);
      return THA_STATUS_EXECUTION_FAILED;
    }
  }
  if (tc_err) {
    set_last_error("tensor-core matmul: a pipeline barrier wait timed out", cudaErrorLaunchTimeout);
    return THA_STATUS_EXECUTION_FAILED;
  }
  memcpy(next_host, tp_ctx(ranks[0])->ws.next_h, B * sizeof(int));
  return THA_STATUS_SUCCESS;
}
