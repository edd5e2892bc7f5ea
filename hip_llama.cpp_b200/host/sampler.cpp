// Sampler of the driver (reference: src/llama.cpp:262-422 == run.cc:566-720): greedy argmax,
// multinomial and top-p over CPU softmax, xorshift* coin.  Bit-for-bit the same arithmetic
// order (float accumulations in index order, libc qsort with the same comparator).
#include <math.h>
#include <stdlib.h>

#include "llama_driver.hpp"

void softmax_host(float* x, int size) {
  float max_val = x[0];
  for (int i = 1; i < size; i++) if (x[i] > max_val) max_val = x[i];
  float sum = 0.0f;
  for (int i = 0; i < size; i++) { x[i] = expf(x[i] - max_val); sum += x[i]; }
  for (int i = 0; i < size; i++) x[i] /= sum;
}

int sample_argmax(const float* p, int n) {
  int best = 0;
  float best_p = p[0];
  for (int i = 1; i < n; i++) if (p[i] > best_p) { best = i; best_p = p[i]; }
  return best;
}

static int sample_mult(const float* p, int n, float coin) {
  float cdf = 0.0f;
  for (int i = 0; i < n; i++) { cdf += p[i]; if (coin < cdf) return i; }
  return n - 1;
}

static int by_prob_desc(const void* a, const void* b) {
  const ProbIndex* x = (const ProbIndex*)a;
  const ProbIndex* y = (const ProbIndex*)b;
  if (x->prob > y->prob) return -1;
  if (x->prob < y->prob) return 1;
  return 0;
}

static int sample_topp(const float* p, int n, float topp, ProbIndex* probindex, float coin) {
  // candidates below (1 - topp) / (n - 1) cannot be part of the nucleus
  int n0 = 0;
  const float cutoff = (1.0f - topp) / (n - 1);
  for (int i = 0; i < n; i++) if (p[i] >= cutoff) { probindex[n0].index = i; probindex[n0].prob = p[i]; n0++; }
  qsort(probindex, n0, sizeof(ProbIndex), by_prob_desc);
  float cumulative = 0.0f;
  int last = n0 - 1;
  for (int i = 0; i < n0; i++) { cumulative += probindex[i].prob; if (cumulative > topp) { last = i; break; } }
  const float r = coin * cumulative;
  float cdf = 0.0f;
  for (int i = 0; i <= last; i++) { cdf += probindex[i].prob; if (r < cdf) return probindex[i].index; }
  return probindex[last].index;
}

void build_sampler(Sampler* s, int vocab_size, float temperature, float topp, unsigned long long rng_seed) {
  s->vocab_size = vocab_size;
  s->temperature = temperature;
  s->topp = topp;
  s->rng_state = rng_seed;
  s->probindex = (ProbIndex*)malloc((size_t)vocab_size * sizeof(ProbIndex));
}
void free_sampler(Sampler* s) { free(s->probindex); s->probindex = nullptr; }

unsigned int random_u32(unsigned long long* state) {  // xorshift*
  *state ^= *state >> 12;
  *state ^= *state << 25;
  *state ^= *state >> 27;
  return (unsigned int)((*state * 0x2545F4914F6CDD1Dull) >> 32);
}
float random_f32(unsigned long long* state) { return (random_u32(state) >> 8) / 16777216.0f; }

int sample(Sampler* s, float* logits) {
  if (s->temperature == 0.0f) return sample_argmax(logits, s->vocab_size);
  for (int q = 0; q < s->vocab_size; q++) logits[q] /= s->temperature;
  softmax_host(logits, s->vocab_size);
  const float coin = random_f32(&s->rng_state);
  if (s->topp <= 0 || s->topp >= 1) return sample_mult(logits, s->vocab_size, coin);
  return sample_topp(logits, s->vocab_size, s->topp, s->probindex, coin);
}
