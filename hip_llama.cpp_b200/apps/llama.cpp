// ./build/apps/llama -- the inference app of the reference (src/llama.cpp main(), :1514-1638),
// same command line:
//
//   llama <checkpoint> -m test -f <input file> -o <output file> -b <slots per GPU>
//   llama <checkpoint> [-m generate] -i "prompt" -n 256 -t 1.0 -p 0.9 -s seed
//
// Options (reference :1490-1510): -t temperature, -p top-p, -s seed, -n steps, -i prompt,
// -z tokenizer path, -m mode (generate|test), -f/-o files, -b batch size.  Additions: -g <n>
// number of GPUs, -k <file> write the generated token ids, -x 1 ignore stop tokens (timing), -r 0 prompt token by token,
// -q 1 host-side sampler, -c <blocks> paged KV cache,
// and in test mode -t/-p/-s are honoured when given (-t 0 = greedy).  The checkpoint may be a
// v0 fp32 file or a v2 Q8_0 file; the version is detected from the header.
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

#include "llama_driver.hpp"

static void error_usage() {
  fprintf(stderr, "Usage:   llama <checkpoint> [options]\n");
  fprintf(stderr, "Example: llama model.bin -n 256 -i \"Once upon a time\"\n");
  fprintf(stderr, "Example: llama model.bin -m test -f <input_filename> -o <output_filename> -b <batch>\n");
  fprintf(stderr, "Options:\n");
  fprintf(stderr, "  -t <float>  temperature in [0,inf], default 1.0\n");
  fprintf(stderr, "  -p <float>  p value in top-p (nucleus) sampling in [0,1] default 0.9\n");
  fprintf(stderr, "  -s <int>    random seed, default time(NULL) (test mode: 314028)\n");
  fprintf(stderr, "  -n <int>    number of steps to run for, default 256. 0 = max_seq_len\n");
  fprintf(stderr, "  -i <string> input prompt\n");
  fprintf(stderr, "  -z <string> optional path to custom tokenizer\n");
  fprintf(stderr, "  -m <string> mode: generate|test, default: generate\n");
  fprintf(stderr, "  -f <string> (only for test mode) input filename\n");
  fprintf(stderr, "  -o <string> (only for test mode) output filename\n");
  fprintf(stderr, "  -b <int>    batch size (slots per GPU)\n");
  fprintf(stderr, "  -g <int>    number of GPUs (default: all)\n");
  fprintf(stderr, "  -w <int>    tensor-parallel width: ONE replica split over this many ranks (rank r on GPU r %% n_gpus), greedy only (-t 0)\n");
  fprintf(stderr, "  -k <string> also write the generated token ids to this file\n");
  fprintf(stderr, "  -x <0|1>    ignore stop tokens (fixed-length timing runs)\n");
  fprintf(stderr, "  -c <int>    paged KV cache: the slots of a GPU share this many blocks of 16 positions (-1: what fits; default 0: dense slabs)\n");
  fprintf(stderr, "  -q <0|1>    sampled runs: 1 = sample on the host from B*vocab logits as the reference does (default 0: sampler on the GPU)\n");
  fprintf(stderr, "  -r <0|1>    fp32 models: ingest prompts in one tensor-core pass (default 1); 0 = one decode step per prompt token\n");
  exit(EXIT_FAILURE);
}

int main(int argc, char* argv[]) {
  const long total_start = time_in_ms();
  char* checkpoint_path = nullptr;
  const char* tokenizer_path = "./assets/tokenizer.bin";
  float temperature = 1.0f, topp = 0.9f;
  int steps = 256, batch = 1, n_gpus = 0, ignore_stop = 0, prefill = 1, tp = 1, host_sampler = 0, kv_blocks = 0;
  const char* prompt = nullptr;
  unsigned long long rng_seed = 0;
  const char* mode = "generate";
  const char *input_filename = nullptr, *output_filename = nullptr, *token_filename = nullptr;
  bool t_given = false, s_given = false, n_given = false;

  if (argc >= 2) checkpoint_path = argv[1]; else error_usage();
  for (int i = 2; i < argc; i += 2) {
    if (i + 1 >= argc || argv[i][0] != '-' || strlen(argv[i]) != 2) error_usage();
    const char f = argv[i][1];
    const char* v = argv[i + 1];
    if (f == 't') { temperature = atof(v); t_given = true; }
    else if (f == 'p') topp = atof(v);
    else if (f == 's') { rng_seed = strtoull(v, nullptr, 10); s_given = true; }
    else if (f == 'n') { steps = atoi(v); n_given = true; }
    else if (f == 'i') prompt = v;
    else if (f == 'z') tokenizer_path = v;
    else if (f == 'm') mode = v;
    else if (f == 'y') { /* chat system prompt: chat mode is an empty branch in the reference (:1583-1584) */ }
    else if (f == 'f') input_filename = v;
    else if (f == 'o') output_filename = v;
    else if (f == 'b') batch = atoi(v);
    else if (f == 'g') n_gpus = atoi(v);
    else if (f == 'k') token_filename = v;
    else if (f == 'x') ignore_stop = atoi(v);
    else if (f == 'r') prefill = atoi(v);
    else if (f == 'w') tp = atoi(v);
    else if (f == 'q') host_sampler = atoi(v);
    else if (f == 'c') kv_blocks = atoi(v);
    else error_usage();
  }
  if (rng_seed == 0) rng_seed = (unsigned int)time(nullptr);
  if (temperature < 0.0f) temperature = 0.0f;
  if (topp < 0.0f || 1.0f < topp) topp = 0.9f;
  if (steps < 0) steps = 0;
  if (batch < 1) batch = 1;

  const int version = checkpoint_version(checkpoint_path);
  if (version != 0 && version != 2) { fprintf(stderr, "unreadable checkpoint %s\n", checkpoint_path); return EXIT_FAILURE; }
  Transformer t32;
  TransformerQ8 tq8;
  if (version == 0) build_transformer(&t32, checkpoint_path); else build_transformer_q8(&tq8, checkpoint_path);
  const Config cfg = version == 0 ? t32.config : tq8.config;
  if (steps == 0 || steps > cfg.seq_len) steps = cfg.seq_len;

  Requests requests;
  DriverOptions opt;
  opt.batch_size = batch;
  opt.n_devices = n_gpus;
  opt.ignore_stop = ignore_stop;
  opt.prefill = prefill;
  opt.host_sampler = host_sampler;
  opt.kv_blocks = kv_blocks;
  opt.tp = tp < 1 ? 1 : tp;
  // ranks of one replica that share a GPU hand-shake between concurrent kernels: every rank's stream needs its own
  // hardware queue (set before the CUDA context exists)
  if (opt.tp > 1) setenv("CUDA_DEVICE_MAX_CONNECTIONS", "32", 0);

  if (strcmp(mode, "generate") == 0) {
    // one request through the same driver (the reference's generate() is the single-sequence loop, :1087-1147 of run.cc)
    requests.num_reqs = 1;
    requests.max_seq_len = steps;
    requests.str_reqs.assign(1, prompt ? prompt : "");
    requests.str_gens.assign(1, "");
    requests.tok_gens.assign(1, std::vector<int>());
    opt.batch_size = 1;
    opt.n_devices = opt.tp > 1 ? n_gpus : 1;
    opt.temperature = temperature; opt.topp = topp; opt.seed = rng_seed;
    DriverStats st;
    const long start = time_in_ms();
    run_data_parallel(version == 0 ? &t32 : nullptr, version == 2 ? &tq8 : nullptr, tokenizer_path, &requests, opt, &st);
    const long end = time_in_ms();
    printf("%s", requests.str_gens[0].c_str());
    fprintf(stderr, "achieved tok/s: %f\n", st.gen_tokens / (double)(end - start) * 1000);
  } else if (strcmp(mode, "test") == 0) {
    if (input_filename == nullptr || output_filename == nullptr) error_usage();
    Tokenizer probe;   // max_token_length sizes the request buffers in the reference (:1590)
    build_tokenizer(&probe, tokenizer_path, cfg.vocab_size);
    // the reference always runs a request to seq_len; an explicit -n caps the positions per request (short test runs)
    read_inputfile(input_filename, probe.max_token_length, n_given ? steps : cfg.seq_len, &requests);
    free_tokenizer(&probe);
    if (t_given) { opt.temperature = temperature; opt.topp = topp; }
    if (s_given) opt.seed = rng_seed;
    DriverStats st;
    const long start = time_in_ms();
    const int num_gen_tokens = run_data_parallel(version == 0 ? &t32 : nullptr, version == 2 ? &tq8 : nullptr, tokenizer_path,
                                                 &requests, opt, &st);
    const long end = time_in_ms();
    fprintf(stdout, "Total achieved token: %d\n", num_gen_tokens);
    fprintf(stdout, "elapsed time(s): %f, achieved throughput(tok/s): %f\n", (double)(end - start) / 1000,
            num_gen_tokens / (double)(end - start) * 1000);
    fprintf(stdout, "step loop(s): %f (%lld steps), weight upload(s): %f, loop throughput(tok/s): %f\n", st.loop_seconds,
            st.steps, st.load_seconds, st.loop_seconds > 0 ? num_gen_tokens / st.loop_seconds : 0.0);
    write_outputfile(output_filename, &requests);
    if (token_filename) write_tokenfile(token_filename, &requests);
  } else {
    fprintf(stderr, "unknown mode: %s\n", mode);
    error_usage();
  }

  if (version == 0) free_transformer(&t32); else free_transformer_q8(&tq8);
  fprintf(stdout, "total elapsed time(s): %lf\n", (double)(time_in_ms() - total_start) / 1000);
  return 0;
}
