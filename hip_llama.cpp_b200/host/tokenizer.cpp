// Byte-pair tokenizer of llama2.c over assets/tokenizer.bin (reference: src/llama.cpp:34-256 ==
// run.cc:380-560).  Same ids for the same text (pinned by tests/golden/tokenizer_golden.json);
// the implementation is a hash map + std::string rather than a sorted array + sprintf.
#include <ctype.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "llama_driver.hpp"

void build_tokenizer(Tokenizer* t, const char* tokenizer_path, int vocab_size) {
  t->vocab_size = vocab_size;
  t->vocab.assign(vocab_size, std::string());
  t->vocab_scores.assign(vocab_size, 0.0f);
  for (int i = 0; i < 256; i++) {
    t->byte_pieces[i * 2] = (unsigned char)i;
    t->byte_pieces[i * 2 + 1] = '\0';
  }
  FILE* file = fopen(tokenizer_path, "rb");
  if (!file) { fprintf(stderr, "couldn't load %s\n", tokenizer_path); exit(EXIT_FAILURE); }
  // file: int max_token_length, then per token: float score, int len, len bytes
  if (fread(&t->max_token_length, sizeof(int), 1, file) != 1) { fprintf(stderr, "failed read\n"); exit(EXIT_FAILURE); }
  std::string buf;
  for (int i = 0; i < vocab_size; i++) {
    int len = 0;
    if (fread(&t->vocab_scores[i], sizeof(float), 1, file) != 1 || fread(&len, sizeof(int), 1, file) != 1 || len < 0) {
      fprintf(stderr, "failed read\n"); exit(EXIT_FAILURE);
    }
    buf.resize(len);
    if (len && fread(&buf[0], len, 1, file) != 1) { fprintf(stderr, "failed read\n"); exit(EXIT_FAILURE); }
    t->vocab[i] = buf;
  }
  fclose(file);
  t->lookup.reserve(vocab_size * 2);
  // duplicates: the reference's bsearch may land on any of them; the lowest id is the
  // deterministic choice and no duplicate exists in assets/tokenizer.bin
  for (int i = vocab_size - 1; i >= 0; i--) t->lookup[t->vocab[i]] = i;
}

void free_tokenizer(Tokenizer* t) {
  t->vocab.clear();
  t->vocab_scores.clear();
  t->lookup.clear();
}

const char* decode(Tokenizer* t, int prev_token, int token) {
  const char* piece = t->vocab[token].c_str();
  if (prev_token == 1 && piece[0] == ' ') piece++;  // sentencepiece drops the space after BOS
  unsigned char byte_val;
  if (sscanf(piece, "<0x%02hhX>", &byte_val) == 1) piece = (const char*)t->byte_pieces + byte_val * 2;
  return piece;
}

void append_str(const char* piece, std::string& str) {
  if (piece == nullptr || piece[0] == '\0') return;
  if (piece[1] == '\0') {
    const unsigned char b = piece[0];
    if (!(isprint(b) || isspace(b))) return;
  }
  str += piece;
}

static int str_lookup(const Tokenizer* t, const std::string& s) {
  auto it = t->lookup.find(s);
  return it == t->lookup.end() ? -1 : it->second;
}

void encode(Tokenizer* t, const char* text, int8_t bos, int8_t eos, int* tokens, int* n_tokens) {
  if (text == nullptr) { fprintf(stderr, "cannot encode NULL text\n"); exit(EXIT_FAILURE); }
  int n = 0;
  if (bos) tokens[n++] = 1;
  // dummy prefix: sentencepiece prepends a space to any non-empty input
  if (text[0] != '\0') tokens[n++] = str_lookup(t, " ");
  // one token per UTF-8 code point; unknown code points fall back to their bytes (+3: <unk>, <s>, </s>)
  std::string cp;
  for (const char* c = text; *c != '\0'; c++) {
    if ((*c & 0xC0) != 0x80) cp.clear();  // not a continuation byte: a new code point starts
    cp.push_back(*c);
    if ((*(c + 1) & 0xC0) == 0x80 && cp.size() < 4) continue;  // more continuation bytes follow
    const int id = str_lookup(t, cp);
    if (id != -1) tokens[n++] = id;
    else for (unsigned char b : cp) tokens[n++] = (int)b + 3;
    cp.clear();
  }
  // greedy merges: repeatedly fuse the adjacent pair whose concatenation has the best score
  std::string pair;
  for (;;) {
    float best_score = -1e10f;
    int best_id = -1, best_idx = -1;
    for (int i = 0; i < n - 1; i++) {
      pair = t->vocab[tokens[i]];
      pair += t->vocab[tokens[i + 1]];
      const int id = str_lookup(t, pair);
      if (id != -1 && t->vocab_scores[id] > best_score) { best_score = t->vocab_scores[id]; best_id = id; best_idx = i; }
    }
    if (best_idx == -1) break;
    tokens[best_idx] = best_id;
    for (int i = best_idx + 1; i < n - 1; i++) tokens[i] = tokens[i + 1];
    n--;
  }
  if (eos) tokens[n++] = 2;
  *n_tokens = n;
}
