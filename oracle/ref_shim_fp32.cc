// TEST INFRASTRUCTURE ONLY -- never linked into the product path.
//
// Thin extern "C" shim around the UNMODIFIED reference fp32 CPU path.  The
// reference translation unit is compiled where it lies (no source is copied):
// `#define TESTING` removes its main() (guard at run.cc:1126) and the whole of
// run.cc becomes visible to this file: forward() run.cc:244-374, rmsnorm :196,
// softmax :211, matmul :231, tokenizer :397-600, sampler :619-745.
//
// Built by oracle/Makefile into oracle/_ref/libref_fp32.so (git-ignored; it
// travels to the GPU box with the snapshot).  Consumers: tests/, bench.py's
// cpu_baseline / --impl reference legs, __graft_entry__.smoke().
#define TESTING 1
#include REF_RUN_CC   // -DREF_RUN_CC='"/root/reference/run.cc"'

#include <omp.h>

extern "C" {

// ---- model ---------------------------------------------------------------
void* ref32_open(const char* checkpoint_path) {
  Transformer* t = (Transformer*)calloc(1, sizeof(Transformer));
  // build_transformer prints the model card to stdout; silence it.
  FILE* saved = stdout;
  FILE* devnull = fopen("/dev/null", "w");
  if (devnull) stdout = devnull;
  build_transformer(t, (char*)checkpoint_path);
  if (devnull) { fflush(devnull); stdout = saved; fclose(devnull); }
  return t;
}

// A second sequence on the same weights: shares the mmap, owns its RunState.
void* ref32_clone(void* h) {
  Transformer* src = (Transformer*)h;
  Transformer* t = (Transformer*)calloc(1, sizeof(Transformer));
  t->config = src->config;
  t->weights = src->weights;
  t->fd = -1;
  t->data = (float*)MAP_FAILED;
  t->file_size = 0;
  malloc_run_state(&t->state, &t->config);
  return t;
}

void ref32_close(void* h) {
  Transformer* t = (Transformer*)h;
  free_transformer(t);
  free(t);
}

void ref32_config(void* h, int out[7]) {
  Config* c = &((Transformer*)h)->config;
  out[0] = c->dim; out[1] = c->hidden_dim; out[2] = c->n_layers; out[3] = c->n_heads;
  out[4] = c->n_kv_heads; out[5] = c->vocab_size; out[6] = c->seq_len;
}

int ref32_shared_classifier(void* h) {
  Transformer* t = (Transformer*)h;
  return t->weights.wcls == t->weights.token_embedding_table;
}

// which: 0 x,1 xb,2 xb2,3 hb,4 hb2,5 q,6 att,7 logits,8 key_cache,9 value_cache
float* ref32_state_ptr(void* h, int which) {
  RunState* s = &((Transformer*)h)->state;
  switch (which) {
    case 0: return s->x;   case 1: return s->xb;  case 2: return s->xb2;
    case 3: return s->hb;  case 4: return s->hb2; case 5: return s->q;
    case 6: return s->att; case 7: return s->logits;
    case 8: return s->key_cache; case 9: return s->value_cache;
  }
  return nullptr;
}

// which: 0 tok_emb,1 rms_att,2 rms_ffn,3 wq,4 wk,5 wv,6 wo,7 w1,8 w2,9 w3,10 rms_final,11 wcls
float* ref32_weight_ptr(void* h, int which) {
  TransformerWeights* w = &((Transformer*)h)->weights;
  switch (which) {
    case 0: return w->token_embedding_table; case 1: return w->rms_att_weight;
    case 2: return w->rms_ffn_weight; case 3: return w->wq; case 4: return w->wk;
    case 5: return w->wv; case 6: return w->wo; case 7: return w->w1; case 8: return w->w2;
    case 9: return w->w3; case 10: return w->rms_final_weight; case 11: return w->wcls;
  }
  return nullptr;
}

float* ref32_forward(void* h, int token, int pos) {
  return forward((Transformer*)h, token, pos);
}

// One decode step for n independent sequences (one Transformer clone each),
// spread over host threads.  Each sequence runs the reference forward()
// untouched, so per-sequence results are bit-identical to the serial call.
void ref32_forward_many(void** hs, const int* token, const int* pos, int n, int n_threads) {
  if (n_threads < 1) n_threads = 1;
#pragma omp parallel for num_threads(n_threads) schedule(dynamic, 1)
  for (int i = 0; i < n; i++) {
    forward((Transformer*)hs[i], token[i], pos[i]);
  }
}

// ---- single ops (run.cc:196-242) -------------------------------------------
void ref32_rmsnorm(float* o, float* x, float* w, int size) { rmsnorm(o, x, w, size); }
void ref32_softmax(float* x, int size) { softmax(x, size); }
void ref32_matmul(float* xout, float* x, float* w, int n, int d) { matmul(xout, x, w, n, d); }
int  ref32_argmax(float* p, int n) { return sample_argmax(p, n); }

// ---- tokenizer (run.cc:397-600) --------------------------------------------
void* ref32_tok_open(const char* path, int vocab_size) {
  Tokenizer* t = (Tokenizer*)calloc(1, sizeof(Tokenizer));
  build_tokenizer(t, (char*)path, vocab_size);
  return t;
}
void ref32_tok_close(void* th) { free_tokenizer((Tokenizer*)th); free(th); }
int ref32_encode(void* th, const char* text, int bos, int eos, int* tokens) {
  int n = 0;
  encode((Tokenizer*)th, (char*)text, (int8_t)bos, (int8_t)eos, tokens, &n);
  return n;
}
const char* ref32_decode(void* th, int prev, int tok) {
  return decode((Tokenizer*)th, prev, tok);
}

// ---- sampler (run.cc:696-743) ----------------------------------------------
void* ref32_sampler_new(int vocab, float temperature, float topp, unsigned long long seed) {
  Sampler* s = (Sampler*)calloc(1, sizeof(Sampler));
  build_sampler(s, vocab, temperature, topp, seed);
  return s;
}
void ref32_sampler_free(void* sh) { free_sampler((Sampler*)sh); free(sh); }
int ref32_sample(void* sh, float* logits) { return sample((Sampler*)sh, logits); }

int ref32_max_threads() { return omp_get_max_threads(); }

}  // extern "C"
