// Device-side model and state memory (reference: src/models.cpp:60-179).  All CUDA runtime
// work goes through the C-ABI runtime shim (tha_malloc / tha_memcpy_h2d ...).
#include "models.hpp"

#include <string.h>

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

void free_transformer_device() {}  // the reference declares it (include/models.hpp:134) and never defines it
