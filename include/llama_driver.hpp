// llama_driver.hpp -- the host side of ./build/apps/llama: tokenizer, sampler, request file
// IO and the batching driver.  In the reference all of this lives inside src/llama.cpp
// (tokenizer :34-256, sampler :262-422, Requests :424-505, test_data_parallelism :891-1083);
// here it is a small library so that the app, the C++ tests and other hosts share it.
#pragma once
#include <string>
#include <unordered_map>
#include <vector>

#include "BatchManager.hpp"
#include "models.hpp"
#include "thaDNN.hpp"

// ---- tokenizer (llama2.c byte-pair tokenizer over assets/tokenizer.bin) ------------------
struct Tokenizer {
  std::vector<std::string> vocab;
  std::vector<float> vocab_scores;
  std::unordered_map<std::string, int> lookup;  // the reference bsearches a sorted copy; same answers
  int vocab_size = 0;
  unsigned int max_token_length = 0;
  unsigned char byte_pieces[512];
};
void build_tokenizer(Tokenizer* t, const char* tokenizer_path, int vocab_size);
void free_tokenizer(Tokenizer* t);
// text -> ids; bos/eos != 0 add token 1 / 2.  tokens must hold strlen(text)+3 ints.
void encode(Tokenizer* t, const char* text, int8_t bos, int8_t eos, int* tokens, int* n_tokens);
// the text piece of `token` after `prev_token` (leading space dropped after BOS, <0xXX> bytes decoded)
const char* decode(Tokenizer* t, int prev_token, int token);
// append a piece unless it is a single unprintable byte (src/llama.cpp:112-125)
void append_str(const char* piece, std::string& str);

// ---- sampler (src/llama.cpp:262-422) -----------------------------------------------------
struct ProbIndex { float prob; int index; };
struct Sampler {
  int vocab_size;
  ProbIndex* probindex;
  float temperature;
  float topp;
  unsigned long long rng_state;
};
void build_sampler(Sampler* sampler, int vocab_size, float temperature, float topp, unsigned long long rng_seed);
void free_sampler(Sampler* sampler);
int sample_argmax(const float* probabilities, int n);  // first maximum wins
int sample(Sampler* sampler, float* logits);            // mutates logits (softmax in place) when temperature != 0
unsigned int random_u32(unsigned long long* state);
float random_f32(unsigned long long* state);
void softmax_host(float* x, int size);

// ---- request file IO (src/llama.cpp:424-505) ---------------------------------------------
// input:  line 1 = number of requests N, then one prompt per line (missing lines = empty prompts)
// output: line 1 = N, then the generated text of each request followed by a blank line
struct Requests {
  int num_reqs = 0;
  int max_token_len = 0;
  int max_seq_len = 0;
  std::vector<std::string> str_reqs;
  std::vector<std::string> str_gens;
  std::vector<std::vector<int>> tok_gens;  // B200 addition: the token ids behind str_gens (parity checks compare these)
};
int read_inputfile(const char* input_filename, int max_token_len, int max_seq_len, Requests* reqs);
int write_outputfile(const char* output_filename, Requests* reqs);
int write_tokenfile(const char* output_filename, Requests* reqs);  // one line of ids per request

// ---- the batching driver -----------------------------------------------------------------
struct DriverOptions {
  int batch_size = 1;          // slots per GPU (-b)
  int n_devices = 0;           // 0 = all visible
  float temperature = 1.0f;    // test mode of the reference: 1.0 / 0.9 / seed 314028 per request (src/llama.cpp:897-900)
  float topp = 0.9f;
  unsigned long long seed = 314028;
  int max_new_positions = 0;   // 0 = run to seq_len; otherwise stop a request after this many positions
  int ignore_stop = 0;         // keep going through BOS/EOS (fixed-length timing runs)
  int resident_greedy = 1;     // temperature == 0: keep token/pos/argmax on the device, read back token ids only
  int device_offset = 0, device_stride = 1;  // request sharding across processes: this process serves requests r with r % stride == offset
  int prefill = 1;             // fp32 models: ingest a new request's prompt in one tensor-core pass (tha_prefill_f32) instead of
                               // one forced-prompt decode step per token (src/llama.cpp:1029-1030); 0 = the reference's scheme
  int tp = 1;                  // > 1: ONE replica split tensor-parallel over `tp` ranks (app flag -w); rank r runs on device r % n_devices,
                               // all ranks are driven by one host thread; greedy decode only (temperature 0), prompts fed token by token
  int kv_blocks = 0;           // 0: dense KV slabs [B][L][seq_len][kv_dim] as the reference; > 0: PAGED cache, the B slots share this many
                               // blocks of 16 positions (BatchManager::create_pool); -1: as many as fit in 80 % of the free memory
  int host_sampler = 0;        // temperature != 0: 1 = the reference's scheme (B * vocab logits to the host, sample() on the host cores);
                               // 0 = softmax / top-p / coin on the device (tha_forward_sample_*), B token ids back per step
  int verbose = 0;
};
struct DriverStats {
  long long gen_tokens = 0;    // sum over requests of (pos - 1), the reference's count (src/llama.cpp:1062)
  long long steps = 0;         // forward calls issued, all devices
  long long prefill_rows = 0;  // prompt positions ingested through tha_prefill_f32
  double load_seconds = 0;     // weight upload, max over devices
  double loop_seconds = 0;     // step loop, max over devices
};
// src/llama.cpp:891-1083: one host thread per GPU, replicated weights, requests pulled from a
// shared queue, B slots per GPU refilled as requests finish.  fp32 (v0) and Q8_0 (v2) models.
int test_data_parallelism(Transformer* transformer, Tokenizer* tokenizer, const char* tokenizer_path, Requests* requests,
                          int batch_size);
int run_data_parallel(Transformer* t32, TransformerQ8* tq8, const char* tokenizer_path, Requests* requests,
                      const DriverOptions& opt, DriverStats* stats);
// One replica, tensor parallel: same slot state machine, the step is tha_tp_forward_greedy.
int run_tensor_parallel(Transformer* t32, TransformerQ8* tq8, const char* tokenizer_path, Requests* requests,
                        const DriverOptions& opt, DriverStats* stats);
long time_in_ms();
