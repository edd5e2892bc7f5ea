// Device-side model and state memory (reference: src/models.cpp:60-179).  All CUDA runtime
// work goes through the C-ABI runtime shim (tha_malloc / tha_memcpy_h2d ...).
#include "models.hpp"

#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <vector>

#include "tha_b200.h"

static inline size_t kv_dim_of(const Config* p) { return (size_t)p->dim * p->n_kv_heads / p->n_heads; }

static float* upload(const float* host, size_t count) {
  float* d = nullptr;
  CHECK_THA(tha_malloc((void**)&d, count * sizeof(float)));
  CHECK_THA(tha_memcpy_h2d(d, host, count * sizeof(float)));
  return d;
}

// src/models.cpp:86-125: 12 allocations + 12 blocking uploads.  When the classifier is tied to
// the embedding (wcls == token_embedding_table on the host) the table is uploaded once and
// shared: the step then streams vocab*dim floats once per step from a single copy.
void copy_weight_to_device(Transformer* t_h, TransformerWeights*& w_d) {
  const Config* p = &t_h->config;
  const size_t dim = p->dim, vocab = p->vocab_size, L = p->n_layers, kvd = kv_dim_of(p), hid = p->hidden_dim;
  const TransformerWeights& h = t_h->weights;
  w_d = (TransformerWeights*)malloc(sizeof(TransformerWeights));
  w_d->token_embedding_table = upload(h.token_embedding_table, vocab * dim);
  w_d->rms_att_weight = upload(h.rms_att_weight, L * dim);
  w_d->rms_ffn_weight = upload(h.rms_ffn_weight, L * dim);
  w_d->wq = upload(h.wq, L * dim * dim);
  w_d->wk = upload(h.wk, L * dim * kvd);
  w_d->wv = upload(h.wv, L * dim * kvd);
  w_d->wo = upload(h.wo, L * dim * dim);
  w_d->w1 = upload(h.w1, L * hid * dim);
  w_d->w2 = upload(h.w2, L * dim * hid);
  w_d->w3 = upload(h.w3, L * hid * dim);
  w_d->rms_final_weight = upload(h.rms_final_weight, dim);
  w_d->wcls = (h.wcls == h.token_embedding_table) ? w_d->token_embedding_table : upload(h.wcls, vocab * dim);
}

void free_weight_on_device(TransformerWeights* w_d, int shared_classifier) {
  if (!w_d) return;
  float* all[] = {w_d->rms_att_weight, w_d->rms_ffn_weight, w_d->wq, w_d->wk, w_d->wv, w_d->wo,
                  w_d->w1, w_d->w2, w_d->w3, w_d->rms_final_weight};
  for (float* p : all) if (p) tha_free(p);
  if (!shared_classifier && w_d->wcls && w_d->wcls != w_d->token_embedding_table) tha_free(w_d->wcls);
  if (w_d->token_embedding_table) tha_free(w_d->token_embedding_table);
  free(w_d);
}

static float* dalloc(size_t count) {
  float* d = nullptr;
  CHECK_THA(tha_malloc((void**)&d, count * sizeof(float)));
  return d;
}

// src/models.cpp:155-179, sizes x batch_size; the KV slabs are zeroed so that a parity run
// never reads uninitialised rows.
void alloc_state_to_device_batch_cfg(const Config* p, RunState*& s, int batch_size, int with_att) {
  const size_t B = batch_size, dim = p->dim, vocab = p->vocab_size, L = p->n_layers, kvd = kv_dim_of(p),
               hid = p->hidden_dim, S = p->seq_len, H = p->n_heads;
  s = (RunState*)calloc(1, sizeof(RunState));
  s->x = dalloc(B * dim);
  s->xb = dalloc(B * dim);
  s->xb2 = dalloc(B * dim);
  s->hb = dalloc(B * hid);
  s->hb2 = dalloc(B * hid);
  s->q = dalloc(B * dim);
  s->att = with_att ? dalloc(B * H * S) : nullptr;
  s->logits = dalloc(B * vocab);
  s->key_cache = dalloc(B * L * S * kvd);
  s->value_cache = dalloc(B * L * S * kvd);
  CHECK_THA(tha_memset(s->key_cache, 0, B * L * S * kvd * sizeof(float)));
  CHECK_THA(tha_memset(s->value_cache, 0, B * L * S * kvd * sizeof(float)));
}

// State for a PAGED KV cache: every activation buffer, no K/V slabs (BatchManager::create_pool + bind_paged supply them).
void alloc_state_to_device_batch_paged(const Config* p, RunState*& s, int batch_size, int with_att) {
  const size_t B = batch_size, dim = p->dim, vocab = p->vocab_size, hid = p->hidden_dim, S = p->seq_len, H = p->n_heads;
  s = (RunState*)calloc(1, sizeof(RunState));
  s->x = dalloc(B * dim);
  s->xb = dalloc(B * dim);
  s->xb2 = dalloc(B * dim);
  s->hb = dalloc(B * hid);
  s->hb2 = dalloc(B * hid);
  s->q = dalloc(B * dim);
  s->att = with_att ? dalloc(B * H * S) : nullptr;
  s->logits = dalloc(B * vocab);
}

void alloc_state_to_device_batch(Transformer* t_h, RunState*& s_d_batch, int batch_size) {
  alloc_state_to_device_batch_cfg(&t_h->config, s_d_batch, batch_size, 1);
}
void alloc_state_to_device_batch_noatt(Transformer* t_h, RunState*& s_d_batch, int batch_size) {
  alloc_state_to_device_batch_cfg(&t_h->config, s_d_batch, batch_size, 0);
}
void alloc_state_to_device(Transformer* t_h, RunState*& s_d) { alloc_state_to_device_batch(t_h, s_d, 1); }

void free_state_on_device(RunState* s) {
  if (!s) return;
  float* all[] = {s->x, s->xb, s->xb2, s->hb, s->hb2, s->q, s->att, s->logits, s->key_cache, s->value_cache};
  for (float* p : all) if (p) tha_free(p);
  free(s);
}

// src/models.cpp:60-84
void copy_transformer_to_device(thablasHandle_t handle, Transformer* t_h, Transformer*& t_d) {
  CHECK_THA(tha_set_device(handle.current_gpu_id));
  t_d = (Transformer*)calloc(1, sizeof(Transformer));
  t_d->config = t_h->config;
  TransformerWeights* w = nullptr;
  RunState* s = nullptr;
  copy_weight_to_device(t_h, w);
  alloc_state_to_device(t_h, s);
  t_d->weights = *w;
  t_d->state = *s;
  free(w);
  free(s);
}

// ---- Q8_0 ---------------------------------------------------------------------------------
static void upload_q(const QuantizedTensor& h, size_t n_values, int gs, QuantizedTensor* d) {
  CHECK_THA(tha_malloc((void**)&d->q, n_values));
  CHECK_THA(tha_malloc((void**)&d->s, n_values / gs * sizeof(float)));
  CHECK_THA(tha_memcpy_h2d(d->q, h.q, n_values));
  CHECK_THA(tha_memcpy_h2d(d->s, h.s, n_values / gs * sizeof(float)));
}

void copy_weight_to_device_q8(TransformerQ8* t_h, TransformerWeightsQ8*& w_d) {
  const Config* p = &t_h->config;
  const size_t dim = p->dim, vocab = p->vocab_size, L = p->n_layers, kvd = kv_dim_of(p), hid = p->hidden_dim;
  const TransformerWeightsQ8& h = t_h->weights;
  const int gs = h.group_size;
  w_d = (TransformerWeightsQ8*)calloc(1, sizeof(TransformerWeightsQ8));
  w_d->group_size = gs;
  w_d->token_embedding_table = upload(h.token_embedding_table, vocab * dim);
  w_d->rms_att_weight = upload(h.rms_att_weight, L * dim);
  w_d->rms_ffn_weight = upload(h.rms_ffn_weight, L * dim);
  w_d->rms_final_weight = upload(h.rms_final_weight, dim);
  upload_q(h.wq, L * dim * dim, gs, &w_d->wq);
  upload_q(h.wk, L * dim * kvd, gs, &w_d->wk);
  upload_q(h.wv, L * dim * kvd, gs, &w_d->wv);
  upload_q(h.wo, L * dim * dim, gs, &w_d->wo);
  upload_q(h.w1, L * hid * dim, gs, &w_d->w1);
  upload_q(h.w2, L * dim * hid, gs, &w_d->w2);
  upload_q(h.w3, L * hid * dim, gs, &w_d->w3);
  upload_q(h.wcls, vocab * dim, gs, &w_d->wcls);
}

void free_weight_on_device_q8(TransformerWeightsQ8* w, int shared_classifier) {
  (void)shared_classifier;
  if (!w) return;
  float* f[] = {w->token_embedding_table, w->rms_att_weight, w->rms_ffn_weight, w->rms_final_weight};
  for (float* p : f) if (p) tha_free(p);
  QuantizedTensor* q[] = {&w->wq, &w->wk, &w->wv, &w->wo, &w->w1, &w->w2, &w->w3, &w->wcls};
  for (QuantizedTensor* t : q) { if (t->q) tha_free(t->q); if (t->s) tha_free(t->s); }
  free(w);
}

// ---- tensor-parallel shards (the slicing hip_llama.cpp_b200/tp.py does with numpy) ------------------
// Local hidden_dim of every rank: whole groups of `unit` values, the first (groups % tp) ranks get one more.
int tp_split_hidden(int hidden, int tp, int unit, int* out) {
  if (tp < 1 || unit < 1 || hidden % unit) return -1;
  const int g = hidden / unit;
  if (g < tp) return -1;
  for (int r = 0; r < tp; r++) out[r] = (g / tp + (r < g % tp ? 1 : 0)) * unit;
  return 0;
}

// n_heads, n_kv_heads, hidden_dim of rank `rank`; dim stays the full residual width
int tp_local_config(const Config* cfg, int tp, int rank, int unit, Config* local) {
  int hid[8];
  if (tp > 8 || rank < 0 || rank >= tp || cfg->n_heads % tp || cfg->n_kv_heads % tp || tp_split_hidden(cfg->hidden_dim, tp, unit, hid)) return -1;
  *local = *cfg;
  local->n_heads = cfg->n_heads / tp;
  local->n_kv_heads = cfg->n_kv_heads / tp;
  local->hidden_dim = hid[rank];
  return 0;
}

// rows [r0, r1) x columns [c0, c1) of every layer of a (L, rows, cols) tensor, as one contiguous device tensor
template <typename T>
static T* upload_slice(const T* src, size_t L, size_t rows, size_t cols, size_t r0, size_t r1, size_t c0, size_t c1) {
  const size_t nr = r1 - r0, nc = c1 - c0;
  T* d = nullptr;
  CHECK_THA(tha_malloc((void**)&d, L * nr * nc * sizeof(T)));
  std::vector<T> tmp(nr * nc);
  for (size_t l = 0; l < L; l++) {
    const T* base = src + l * rows * cols;
    for (size_t r = 0; r < nr; r++) memcpy(tmp.data() + r * nc, base + (r0 + r) * cols + c0, nc * sizeof(T));
    CHECK_THA(tha_memcpy_h2d(d + l * nr * nc, tmp.data(), nr * nc * sizeof(T)));
  }
  return d;
}

struct TpRanges { size_t q0, q1, k0, k1, h0, h1; };
static TpRanges tp_ranges(const Config* p, int tp, int rank, int unit) {
  const size_t hs = p->dim / p->n_heads, da = hs * p->n_heads / tp, kvd = hs * p->n_kv_heads / tp;
  int hid[8];
  if (tp_split_hidden(p->hidden_dim, tp, unit, hid)) { fprintf(stderr, "tensor-parallel split: hidden_dim %d is not %d x a whole number >= %d\n", p->hidden_dim, unit, tp); exit(EXIT_FAILURE); }
  size_t h0 = 0;
  for (int r = 0; r < rank; r++) h0 += hid[r];
  return {rank * da, (rank + 1) * da, rank * kvd, (rank + 1) * kvd, h0, h0 + (size_t)hid[rank]};
}

void copy_weight_shard_to_device(Transformer* t_h, int tp, int rank, TransformerWeights*& w_d) {
  const Config* p = &t_h->config;
  const size_t dim = p->dim, vocab = p->vocab_size, L = p->n_layers, kvd = kv_dim_of(p), hid = p->hidden_dim;
  const TransformerWeights& h = t_h->weights;
  const TpRanges g = tp_ranges(p, tp, rank, 64);
  w_d = (TransformerWeights*)malloc(sizeof(TransformerWeights));
  w_d->token_embedding_table = upload(h.token_embedding_table, vocab * dim);
  w_d->rms_att_weight = upload(h.rms_att_weight, L * dim);
  w_d->rms_ffn_weight = upload(h.rms_ffn_weight, L * dim);
  w_d->wq = upload_slice(h.wq, L, dim, dim, g.q0, g.q1, 0, dim);
  w_d->wk = upload_slice(h.wk, L, kvd, dim, g.k0, g.k1, 0, dim);
  w_d->wv = upload_slice(h.wv, L, kvd, dim, g.k0, g.k1, 0, dim);
  w_d->wo = upload_slice(h.wo, L, dim, dim, 0, dim, g.q0, g.q1);
  w_d->w1 = upload_slice(h.w1, L, hid, dim, g.h0, g.h1, 0, dim);
  w_d->w2 = upload_slice(h.w2, L, dim, hid, 0, dim, g.h0, g.h1);
  w_d->w3 = upload_slice(h.w3, L, hid, dim, g.h0, g.h1, 0, dim);
  w_d->rms_final_weight = upload(h.rms_final_weight, dim);
  w_d->wcls = (h.wcls == h.token_embedding_table) ? w_d->token_embedding_table : upload(h.wcls, vocab * dim);
}

// int8 values sliced like the fp32 tensors; scales [rows, cols / gs] sliced by rows / whole groups of columns
static void upload_q_slice(const QuantizedTensor& h, size_t L, size_t rows, size_t cols, size_t r0, size_t r1, size_t c0, size_t c1,
                           int gs, QuantizedTensor* d) {
  if (c0 % gs || c1 % gs) { fprintf(stderr, "tensor-parallel split cuts a Q8_0 group\n"); exit(EXIT_FAILURE); }
  d->q = upload_slice(h.q, L, rows, cols, r0, r1, c0, c1);
  d->s = upload_slice(h.s, L, rows, cols / gs, r0, r1, c0 / gs, c1 / gs);
}

void copy_weight_shard_to_device_q8(TransformerQ8* t_h, int tp, int rank, TransformerWeightsQ8*& w_d) {
  const Config* p = &t_h->config;
  const size_t dim = p->dim, vocab = p->vocab_size, L = p->n_layers, kvd = kv_dim_of(p), hid = p->hidden_dim;
  const TransformerWeightsQ8& h = t_h->weights;
  const int gs = h.group_size;
  const TpRanges g = tp_ranges(p, tp, rank, gs);
  w_d = (TransformerWeightsQ8*)calloc(1, sizeof(TransformerWeightsQ8));
  w_d->group_size = gs;
  w_d->token_embedding_table = upload(h.token_embedding_table, vocab * dim);
  w_d->rms_att_weight = upload(h.rms_att_weight, L * dim);
  w_d->rms_ffn_weight = upload(h.rms_ffn_weight, L * dim);
  w_d->rms_final_weight = upload(h.rms_final_weight, dim);
  upload_q_slice(h.wq, L, dim, dim, g.q0, g.q1, 0, dim, gs, &w_d->wq);
  upload_q_slice(h.wk, L, kvd, dim, g.k0, g.k1, 0, dim, gs, &w_d->wk);
  upload_q_slice(h.wv, L, kvd, dim, g.k0, g.k1, 0, dim, gs, &w_d->wv);
  upload_q_slice(h.wo, L, dim, dim, 0, dim, g.q0, g.q1, gs, &w_d->wo);
  upload_q_slice(h.w1, L, hid, dim, g.h0, g.h1, 0, dim, gs, &w_d->w1);
  upload_q_slice(h.w2, L, dim, hid, 0, dim, g.h0, g.h1, gs, &w_d->w2);
  upload_q_slice(h.w3, L, hid, dim, g.h0, g.h1, 0, dim, gs, &w_d->w3);
  upload_q(h.wcls, vocab * dim, gs, &w_d->wcls);
}

// state of one rank: q / xb as wide as the local heads, hb as wide as the local hidden units, KV of the local kv heads
void alloc_state_to_device_batch_tp(const Config* cfg_global, const Config* cfg_local, RunState*& s, int batch_size) {
  const size_t B = batch_size, dim = cfg_global->dim, vocab = cfg_global->vocab_size, L = cfg_global->n_layers, S = cfg_global->seq_len;
  const size_t hs = cfg_global->dim / cfg_global->n_heads, da = hs * cfg_local->n_heads, kvd = hs * cfg_local->n_kv_heads;
  s = (RunState*)calloc(1, sizeof(RunState));
  s->x = dalloc(B * dim);
  s->xb = dalloc(B * da);
  s->q = dalloc(B * da);
  s->hb = dalloc(B * cfg_local->hidden_dim);
  s->logits = dalloc(B * vocab);
  s->key_cache = dalloc(B * L * S * kvd);
  s->value_cache = dalloc(B * L * S * kvd);
  CHECK_THA(tha_memset(s->key_cache, 0, B * L * S * kvd * sizeof(float)));
  CHECK_THA(tha_memset(s->value_cache, 0, B * L * S * kvd * sizeof(float)));
}

void free_transformer_device() {}  // the reference declares it (include/models.hpp:134) and never defines it
