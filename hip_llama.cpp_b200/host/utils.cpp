// Host utilities (reference: src/utils.cpp): test-matrix helpers and the checkpoint loaders.
#include "utils.hpp"

#include <fcntl.h>
#include <stdio.h>
#include <string.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include <fstream>
#include <iterator>

void alloc_mat(float** m, int R, int C) {
  *m = (float*)aligned_alloc(32, (((size_t)R * C * sizeof(float)) + 31) / 32 * 32);
  if (*m == nullptr) { fprintf(stderr, "Failed to allocate memory for mat.\n"); exit(EXIT_FAILURE); }
}
void alloc_vec(float** m, int N) { alloc_mat(m, N, 1); }
void util_free(void* m) { free(m); }
void rand_mat(float* m, int R, int C) {
  for (size_t i = 0; i < (size_t)R * C; i++) m[i] = (float)rand() / (float)RAND_MAX - 0.5f;
}
void rand_vec(float* m, int N) { rand_mat(m, N, 1); }
void zero_mat(float* m, int R, int C) { memset(m, 0, sizeof(float) * (size_t)R * C); }
void zero_vec(float* m, int N) { zero_mat(m, N, 1); }

bool compareFiles(const std::string& filePath1, const std::string& filePath2) {
  std::ifstream a(filePath1, std::ios::binary), b(filePath2, std::ios::binary);
  if (!a.is_open() || !b.is_open()) return false;
  return std::equal(std::istreambuf_iterator<char>(a), std::istreambuf_iterator<char>(), std::istreambuf_iterator<char>(b),
                    std::istreambuf_iterator<char>());
}

// ---- v0 checkpoint -----------------------------------------------------------------------
// Layout after the 7-int header (run.cc:110-137): tok_emb, rms_att, wq, wk, wv, wo, rms_ffn,
// w1, w2, w3, rms_final, [freq_cis_real, freq_cis_imag: seq_len*head_size/2 each, skipped],
// [wcls when not shared].
void memory_map_weights(TransformerWeights* w, Config* p, float* ptr, int shared_weights) {
  const size_t head_size = p->dim / p->n_heads, L = p->n_layers, dim = p->dim, hid = p->hidden_dim;
  const size_t kvd = (size_t)p->n_kv_heads * head_size;
  w->token_embedding_table = ptr; ptr += (size_t)p->vocab_size * dim;
  w->rms_att_weight = ptr;        ptr += L * dim;
  w->wq = ptr;                    ptr += L * dim * (p->n_heads * head_size);
  w->wk = ptr;                    ptr += L * dim * kvd;
  w->wv = ptr;                    ptr += L * dim * kvd;
  w->wo = ptr;                    ptr += L * (p->n_heads * head_size) * dim;
  w->rms_ffn_weight = ptr;        ptr += L * dim;
  w->w1 = ptr;                    ptr += L * dim * hid;
  w->w2 = ptr;                    ptr += L * hid * dim;
  w->w3 = ptr;                    ptr += L * dim * hid;
  w->rms_final_weight = ptr;      ptr += dim;
  ptr += (size_t)p->seq_len * head_size / 2;  // RoPE tables of old exporters
  ptr += (size_t)p->seq_len * head_size / 2;
  w->wcls = shared_weights ? w->token_embedding_table : ptr;
}

void read_checkpoint(char* checkpoint, Config* config, TransformerWeights* weights, int* fd, float** data,
                     ssize_t* file_size) {
  FILE* file = fopen(checkpoint, "rb");
  if (!file) { fprintf(stderr, "Couldn't open file %s\n", checkpoint); exit(EXIT_FAILURE); }
  if (fread(config, sizeof(Config), 1, file) != 1) { fprintf(stderr, "Couldn't read the header of %s\n", checkpoint); exit(EXIT_FAILURE); }
  const int shared_weights = config->vocab_size > 0 ? 1 : 0;  // a negative vocab size flags an unshared classifier
  config->vocab_size = abs(config->vocab_size);
  fseek(file, 0, SEEK_END);
  *file_size = ftell(file);
  fclose(file);
  *fd = open(checkpoint, O_RDONLY);
  if (*fd == -1) { fprintf(stderr, "open failed!\n"); exit(EXIT_FAILURE); }
  *data = (float*)mmap(NULL, *file_size, PROT_READ, MAP_PRIVATE, *fd, 0);
  if (*data == MAP_FAILED) { fprintf(stderr, "mmap failed!\n"); exit(EXIT_FAILURE); }
  memory_map_weights(weights, config, *data + sizeof(Config) / sizeof(float), shared_weights);
}

// Host RunState of the reference's CPU path (src/utils.cpp:86-117).  The product has no CPU
// forward, so nothing reads these buffers; kept because callers free them.
void malloc_run_state(RunState* s, Config* p) { (void)p; memset(s, 0, sizeof(RunState)); }
void free_run_state(RunState* s) { (void)s; }

void build_transformer(Transformer* t, char* checkpoint_path) {
  read_checkpoint(checkpoint_path, &t->config, &t->weights, &t->fd, &t->data, &t->file_size);
  malloc_run_state(&t->state, &t->config);
}

void free_transformer(Transformer* t) {
  if (t->data && t->data != MAP_FAILED) munmap(t->data, t->file_size);
  if (t->fd != -1) close(t->fd);
  free_run_state(&t->state);
}

void print_transformer(Transformer* t) {
  const Config& c = t->config;
  printf("---------Model Information----------\n");
  printf("dim: %d\nhidden_dim: %d\nn_layers: %d\nn_heads: %d\nn_kv_heads: %d\nvocab_size: %d\nseq_len: %d\n", c.dim,
         c.hidden_dim, c.n_layers, c.n_heads, c.n_kv_heads, c.vocab_size, c.seq_len);
  printf("------------------------------------\n");
}

int checkpoint_version(const char* path) {
  FILE* f = fopen(path, "rb");
  if (!f) return -1;
  uint32_t head[2] = {0, 0};
  const size_t n = fread(head, sizeof(uint32_t), 2, f);
  fclose(f);
  if (n != 2) return -1;
  return (head[0] == 0x616b3432u) ? (int)head[1] : 0;
}

// ---- v2 (Q8_0) checkpoint (runq.c:174-251) -------------------------------------------------
// Header: magic "ak42", version 2, Config, shared_classifier byte, group_size int; payload at
// byte 256: fp32 norms, then per tensor and per layer [int8 values][fp32 scales] interleaved.
// The device layout wants every tensor layer-major with values and scales in two contiguous
// arrays, so the loader de-interleaves into host buffers it owns.
static void carve(const char** cur, size_t n_layers, size_t each, int gs, QuantizedTensor* out) {
  out->q = (int8_t*)malloc(n_layers * each);
  out->s = (float*)malloc(n_layers * each / gs * sizeof(float));
  for (size_t l = 0; l < n_layers; l++) {
    memcpy(out->q + l * each, *cur, each);
    *cur += each;
    memcpy(out->s + l * each / gs, *cur, each / gs * sizeof(float));
    *cur += each / gs * sizeof(float);
  }
}

void build_transformer_q8(TransformerQ8* t, char* checkpoint_path) {
  memset(t, 0, sizeof(*t));
  FILE* file = fopen(checkpoint_path, "rb");
  if (!file) { fprintf(stderr, "Couldn't open file %s\n", checkpoint_path); exit(EXIT_FAILURE); }
  uint32_t magic = 0;
  int version = 0, group_size = 0;
  uint8_t shared = 0;
  if (fread(&magic, 4, 1, file) != 1 || magic != 0x616b3432u) { fprintf(stderr, "Bad magic number\n"); exit(EXIT_FAILURE); }
  if (fread(&version, 4, 1, file) != 1 || version != 2) { fprintf(stderr, "Bad version %d, need version 2\n", version); exit(EXIT_FAILURE); }
  if (fread(&t->config, sizeof(Config), 1, file) != 1 || fread(&shared, 1, 1, file) != 1 ||
      fread(&group_size, 4, 1, file) != 1) { fprintf(stderr, "Truncated header\n"); exit(EXIT_FAILURE); }
  fseek(file, 0, SEEK_END);
  t->file_size = ftell(file);
  fclose(file);
  t->fd = open(checkpoint_path, O_RDONLY);
  if (t->fd == -1) { fprintf(stderr, "open failed!\n"); exit(EXIT_FAILURE); }
  t->data = mmap(NULL, t->file_size, PROT_READ, MAP_PRIVATE, t->fd, 0);
  if (t->data == MAP_FAILED) { fprintf(stderr, "mmap failed!\n"); exit(EXIT_FAILURE); }
  t->shared_classifier = shared;
  const Config* p = &t->config;
  TransformerWeightsQ8* w = &t->weights;
  w->group_size = group_size;
  const size_t dim = p->dim, L = p->n_layers, hid = p->hidden_dim, vocab = p->vocab_size;
  const size_t hs = dim / p->n_heads, kvd = hs * p->n_kv_heads;
  const char* cur = (const char*)t->data + 256;
  w->rms_att_weight = (float*)cur; cur += L * dim * sizeof(float);
  w->rms_ffn_weight = (float*)cur; cur += L * dim * sizeof(float);
  w->rms_final_weight = (float*)cur; cur += dim * sizeof(float);
  QuantizedTensor qtok;
  carve(&cur, 1, vocab * dim, group_size, &qtok);
  // runq.c:139-143 dequantize: x[i] = q[i] * s[i / GS]
  t->dequant_embedding = (float*)malloc(vocab * dim * sizeof(float));
  for (size_t i = 0; i < vocab * dim; i++) t->dequant_embedding[i] = qtok.q[i] * qtok.s[i / group_size];
  w->token_embedding_table = t->dequant_embedding;
  carve(&cur, L, dim * dim, group_size, &w->wq);
  carve(&cur, L, dim * kvd, group_size, &w->wk);
  carve(&cur, L, dim * kvd, group_size, &w->wv);
  carve(&cur, L, dim * dim, group_size, &w->wo);
  carve(&cur, L, dim * hid, group_size, &w->w1);
  carve(&cur, L, hid * dim, group_size, &w->w2);
  carve(&cur, L, dim * hid, group_size, &w->w3);
  if (shared) w->wcls = qtok;
  else carve(&cur, 1, dim * vocab, group_size, &w->wcls);
  if (!shared) { free(qtok.q); free(qtok.s); }
}

void free_transformer_q8(TransformerQ8* t) {
  TransformerWeightsQ8* w = &t->weights;
  QuantizedTensor* q[] = {&w->wq, &w->wk, &w->wv, &w->wo, &w->w1, &w->w2, &w->w3, &w->wcls};
  for (QuantizedTensor* x : q) { free(x->q); free(x->s); x->q = nullptr; x->s = nullptr; }
  free(t->dequant_embedding);
  if (t->data && t->data != MAP_FAILED) munmap(t->data, t->file_size);
  if (t->fd != -1) close(t->fd);
}
