// The batching driver (reference: test_data_parallelism, src/llama.cpp:891-1083) and request
// file IO (:424-505).
//
// Same control flow as the reference -- one host thread per GPU, replicated weights, requests
// pulled from one shared queue, B slots per GPU refilled as requests finish, prompts forced
// token by token, stop on BOS/EOS or seq_len, gen_cnt += pos - 1 -- with these differences:
//   * slot state lives in a BatchManager (slot table + KV block accounting);
//   * idle slots carry pos = -1 and are skipped by the step (the reference computes them);
//   * greedy runs (temperature 0) read back B token ids per step instead of B*vocab logits;
//   * fp32 (v0) and Q8_0 (v2) checkpoints both run on the GPU.
#include <omp.h>
#include <stdio.h>
#include <string.h>
#include <time.h>

#include <algorithm>
#include <atomic>
#include <fstream>

#include "llama_driver.hpp"
#include "tha_b200.h"

long time_in_ms() {
  struct timespec t;
  clock_gettime(CLOCK_REALTIME, &t);
  return t.tv_sec * 1000 + t.tv_nsec / 1000000;
}

// ---- request files ----------------------------------------------------------------------
int read_inputfile(const char* input_filename, int max_token_len, int max_seq_len, Requests* reqs) {
  printf("max_token_len: %d, max_seq_len: %d\n", max_token_len, max_seq_len);
  std::ifstream in(input_filename);
  if (!in.is_open()) { fprintf(stderr, "cannot open the file: %s\n", input_filename); exit(EXIT_FAILURE); }
  std::string line;
  std::getline(in, line);
  const int n = atoi(line.c_str());
  reqs->num_reqs = n;
  reqs->max_token_len = max_token_len;
  reqs->max_seq_len = max_seq_len;
  reqs->str_reqs.assign(n, std::string());   // fewer body lines than N: the rest are empty prompts (BOS only)
  reqs->str_gens.assign(n, std::string());
  reqs->tok_gens.assign(n, std::vector<int>());
  for (int idx = 0; idx < n && std::getline(in, line); idx++) reqs->str_reqs[idx] = line;
  return 0;
}

int write_outputfile(const char* output_filename, Requests* reqs) {
  std::ofstream out(output_filename);
  if (!out.is_open()) { fprintf(stderr, "cannot write the file: %s\n", output_filename); exit(EXIT_FAILURE); }
  out << reqs->num_reqs << "\n";
  for (int i = 0; i < reqs->num_reqs; i++) out << reqs->str_gens[i] << "\n";
  return 0;
}

int write_tokenfile(const char* output_filename, Requests* reqs) {
  std::ofstream out(output_filename);
  if (!out.is_open()) { fprintf(stderr, "cannot write the file: %s\n", output_filename); exit(EXIT_FAILURE); }
  out << reqs->num_reqs << "\n";
  for (int i = 0; i < reqs->num_reqs; i++) {
    for (size_t j = 0; j < reqs->tok_gens[i].size(); j++) out << (j ? " " : "") << reqs->tok_gens[i][j];
    out << "\n";
  }
  return 0;
}

// ---- the driver ---------------------------------------------------------------------------
namespace {
struct Slot {
  std::vector<int> prompt;   // prompt token ids (BOS first)
  std::string text;          // generated text so far
  std::vector<int> ids;      // every token fed or produced, in order
  int steps = 0;             // position limit
};
}  // namespace

int run_data_parallel(Transformer* t32, TransformerQ8* tq8, const char* tokenizer_path, Requests* requests,
                      const DriverOptions& opt, DriverStats* stats) {
  if (opt.tp > 1) return run_tensor_parallel(t32, tq8, tokenizer_path, requests, opt, stats);
  const Config cfg = t32 ? t32->config : tq8->config;
  const int vocab = cfg.vocab_size, B = opt.batch_size;
  int n_devices = 0;
  CHECK_THA(tha_device_count(&n_devices));
  if (opt.n_devices > 0 && opt.n_devices < n_devices) n_devices = opt.n_devices;
  if (n_devices < 1) { fprintf(stderr, "no CUDA device\n"); exit(EXIT_FAILURE); }
  fprintf(stderr, "\n DATA PARALLELISM \n\n Num Devices %d\n\n Batch Size %d\n", n_devices, B);

  // one sampler per REQUEST so that the output does not depend on slot or device (src/llama.cpp:895-900)
  std::vector<Sampler> samplers(requests->num_reqs);
  for (auto& s : samplers) build_sampler(&s, vocab, opt.temperature, opt.topp, opt.seed);

  std::atomic<int> next_idx(opt.device_offset);
  long long gen_cnt = 0, n_steps = 0, n_prefill_rows = 0;
  double load_s = 0, loop_s = 0;
  const bool greedy = opt.temperature == 0.0f && opt.resident_greedy;

#pragma omp parallel for num_threads(n_devices) reduction(+ : gen_cnt, n_steps, n_prefill_rows) reduction(max : load_s, loop_s)
  for (int gid = 0; gid < n_devices; gid++) {
    CHECK_THA(tha_set_device(gid));
    Tokenizer tok;
    build_tokenizer(&tok, tokenizer_path, vocab);
    thablasHandle_t handle;
    thablasCreate(&handle);

    const long t_load0 = time_in_ms();
    TransformerWeights* w_d = nullptr;
    TransformerWeightsQ8* wq_d = nullptr;
    if (t32) copy_weight_to_device(t32, w_d);
    else copy_weight_to_device_q8(tq8, wq_d);
    CHECK_THA(tha_device_sync());
    load_s = (time_in_ms() - t_load0) / 1000.0;
    fprintf(stdout, "\nLoad model time thread %d: %f\n", gid, load_s);

    const int kvd = cfg.dim * cfg.n_kv_heads / cfg.n_heads;
    BatchManager bm(B, cfg.seq_len, cfg.n_layers, kvd);
    RunState* s_d = nullptr;
    // opt.kv_blocks != 0: PAGED KV cache -- the slots share a pool of blocks of 16 positions (K and V rows of all layers)
    // handed out as sequences grow and returned when a request finishes, instead of B slabs of seq_len positions each
    const bool paged = opt.kv_blocks != 0;
    if (paged) {
      alloc_state_to_device_batch_paged(&cfg, s_d, B, 0);
      size_t n_blocks = opt.kv_blocks > 0 ? (size_t)opt.kv_blocks : (size_t)B * bm.nrow_blocks;
      if (opt.kv_blocks < 0) {   // auto: what fits in 80 % of the free memory, at most one full context per slot
        const size_t fit = (size_t)(0.8 * (double)bm.get_gpu_memory(gid)) / bm.block_bytes();
        if (fit < n_blocks) n_blocks = fit;
      }
      if (!bm.create_pool(n_blocks)) { fprintf(stderr, "cannot allocate a KV pool of %zu blocks: %s\n", n_blocks, tha_last_error()); exit(EXIT_FAILURE); }
      bm.bind_paged(s_d);
      fprintf(stdout, "paged KV cache: %zu blocks of %d positions (%.1f MB) for %d slots\n", n_blocks, bm.block_size,
              n_blocks * bm.block_bytes() / 1e6, B);
    } else {
      alloc_state_to_device_batch_cfg(&cfg, s_d, B, tha_step_needs_att_scratch((const tha_config_t*)&cfg));
      bm.bind(s_d);
    }
    int pending_req = -1;   // paged: a request that did not fit the pool yet
    // temperature != 0: softmax / top-p / coin run on the device (tha_forward_sample_*, csrc/sampler.cu), one xorshift*
    // stream per request as in the reference (src/llama.cpp:895-900); opt.host_sampler = 1 keeps the reference's scheme
    // (B * vocab logits to pinned host memory, sample() on the host cores)
    const bool device_sampler = !greedy && !opt.host_sampler;
    float* logits_host = nullptr;
    if (!greedy && !device_sampler) CHECK_THA(tha_host_alloc((void**)&logits_host, (size_t)B * vocab * sizeof(float)));
    std::vector<int> force(B, -1);
    std::vector<unsigned long long> reseed(B, 0ull);

    std::vector<Slot> slots(B);
    std::vector<int> token(B, 0), pos(B, -1), next(B, 0);
    Config cfg_local = cfg;

    const bool use_prefill = opt.prefill && tha_prefill_supported((const tha_config_t*)&cfg_local);
    std::vector<int> pf_slot, pf_pos0, pf_len, pf_tokens;

    const long t_loop0 = time_in_ms();
    for (;;) {
      // ---- refill free slots from the shared queue
      bool queue_empty = false;
      pf_slot.clear(); pf_pos0.clear(); pf_len.clear(); pf_tokens.clear();
      for (int s = 0; s < B && !queue_empty; s++) {
        if (bm.request_of(s) >= 0) continue;
        const int req = pending_req >= 0 ? pending_req : next_idx.fetch_add(opt.device_stride);
        pending_req = -1;
        if (req >= requests->num_reqs) { queue_empty = true; break; }
        const std::string& prompt = requests->str_reqs[req];
        std::vector<int> ptoks(prompt.size() + 3, 0);
        int n_prompt = 0;
        encode(&tok, prompt.c_str(), 1, 0, ptoks.data(), &n_prompt);
        if (n_prompt < 1) { fprintf(stderr, "something is wrong, expected at least 1 prompt token\n"); exit(EXIT_FAILURE); }
        if (paged) {
          // admission control: a request enters only if the pool can still hold every admitted request grown to its
          // position limit -- a running sequence then never finds the pool empty
          int limit = requests->max_seq_len;
          if (opt.max_new_positions > 0 && opt.max_new_positions < limit) limit = opt.max_new_positions;
          size_t reserved = 0;
          for (int t = 0; t < B; t++)
            if (bm.request_of(t) >= 0) reserved += bm.blocks_for(slots[t].steps) - (size_t)bm.n_block_table_filled[t];
          if (bm.free_blocks() < reserved + bm.blocks_for(limit)) {
            if (bm.active_slots() == 0) { fprintf(stderr, "KV pool too small for one request of %d positions\n", limit); exit(EXIT_FAILURE); }
            pending_req = req;   // wait for a finishing request
            break;
          }
        }
        const int got = bm.acquire_slot(req);   // lowest free slot == s
        Slot& sl = slots[got];
        sl.prompt.assign(ptoks.begin(), ptoks.begin() + n_prompt);
        sl.text.clear();
        sl.ids.assign(1, sl.prompt[0]);
        sl.steps = requests->max_seq_len;
        if (opt.max_new_positions > 0 && opt.max_new_positions < sl.steps) sl.steps = opt.max_new_positions;
        token[got] = sl.prompt[0];
        pos[got] = 0;
        reseed[got] = opt.seed;   // the request's own sampler stream starts here
        // Prompt ingestion in one pass: positions 0..m-1 go through tha_prefill_f32 (cache rows only), the slot
        // then joins the step loop at position m with the prompt's token m -- the state the reference reaches
        // after m forced-prompt steps (src/llama.cpp:1029-1030), text and ids included.
        int m = use_prefill ? n_prompt - 1 : 0;
        if (m > sl.steps - 1) m = sl.steps - 1;
        if (m >= 1) {
          if (!bm.ensure_position(got, m - 1)) { fprintf(stderr, "KV pool exhausted\n"); exit(EXIT_FAILURE); }
          pf_slot.push_back(got); pf_pos0.push_back(0); pf_len.push_back(m);
          pf_tokens.insert(pf_tokens.end(), sl.prompt.begin(), sl.prompt.begin() + m);
          for (int j = 1; j <= m; j++) {
            append_str(decode(&tok, sl.prompt[j - 1], sl.prompt[j]), sl.text);
            sl.ids.push_back(sl.prompt[j]);
            bm.advance(got);
          }
          token[got] = sl.prompt[m];
          pos[got] = m;
        }
        if (opt.verbose) fprintf(stderr, "\nDevice %d - seq id %d in batch, Request %d\n", gid, got, req);
      }
      if (bm.active_slots() == 0) {
        if (pending_req >= 0) continue;   // cannot happen (checked above), kept for clarity
        break;
      }
      if (paged) {   // the block that receives this step's K/V row of every active slot exists before the step runs
        for (int s = 0; s < B; s++)
          if (bm.request_of(s) >= 0 && !bm.ensure_position(s, pos[s])) {
            fprintf(stderr, "KV pool exhausted (%zu blocks): raise -c or lower -b\n", bm.num_physical_blocks);
            exit(EXIT_FAILURE);
          }
        bm.sync_table((void*)handle.calc_stream);
      }
      if (!pf_slot.empty()) {
        if (t32)
          CHECK_THA(tha_prefill_f32((void*)handle.calc_stream, (const tha_config_t*)&cfg_local, (const tha_weights_f32_t*)w_d,
                                    (const tha_state_t*)s_d, (int)pf_slot.size(), pf_slot.data(), pf_pos0.data(), pf_len.data(),
                                    pf_tokens.data(), nullptr, nullptr));
        else
          CHECK_THA(tha_prefill_q8((void*)handle.calc_stream, (const tha_config_t*)&cfg_local, (const tha_weights_q8_t*)wq_d,
                                   (const tha_state_t*)s_d, (int)pf_slot.size(), pf_slot.data(), pf_pos0.data(), pf_len.data(),
                                   pf_tokens.data(), nullptr, nullptr));
        n_prefill_rows += (long long)pf_tokens.size();
      }

      // ---- one decode step for every active slot
      int rc;
      if (greedy) {
        rc = t32 ? tha_forward_greedy_f32((void*)handle.calc_stream, B, (const tha_config_t*)&cfg_local, (const tha_weights_f32_t*)w_d,
                                          (const tha_state_t*)s_d, token.data(), pos.data(), next.data())
                 : tha_forward_greedy_q8((void*)handle.calc_stream, B, (const tha_config_t*)&cfg_local, (const tha_weights_q8_t*)wq_d,
                                         (const tha_state_t*)s_d, token.data(), pos.data(), next.data());
      } else if (device_sampler) {
        for (int s = 0; s < B; s++) {
          const int req = bm.request_of(s);
          force[s] = (req >= 0 && pos[s] < (int)slots[s].prompt.size() - 1) ? slots[s].prompt[pos[s] + 1] : -1;
        }
        rc = t32 ? tha_forward_sample_f32((void*)handle.calc_stream, B, (const tha_config_t*)&cfg_local, (const tha_weights_f32_t*)w_d,
                                          (const tha_state_t*)s_d, token.data(), pos.data(), force.data(), reseed.data(),
                                          opt.temperature, opt.topp, next.data())
                 : tha_forward_sample_q8((void*)handle.calc_stream, B, (const tha_config_t*)&cfg_local, (const tha_weights_q8_t*)wq_d,
                                         (const tha_state_t*)s_d, token.data(), pos.data(), force.data(), reseed.data(),
                                         opt.temperature, opt.topp, next.data());
        std::fill(reseed.begin(), reseed.end(), 0ull);
      } else {
        rc = t32 ? (int)thaDNN_s_forward_batch(handle, handle, handle, B, &cfg_local, w_d, s_d, token.data(), pos.data(), logits_host)
                 : (int)thaDNN_q8_forward_batch(handle, B, &cfg_local, wq_d, s_d, token.data(), pos.data(), logits_host);
      }
      CHECK_THA(rc);
      n_steps++;

      // ---- advance the state machine of every slot (src/llama.cpp:1025-1050)
      for (int s = 0; s < B; s++) {
        const int req = bm.request_of(s);
        if (req < 0) continue;
        Slot& sl = slots[s];
        int nx;
        if (pos[s] < (int)sl.prompt.size() - 1) nx = sl.prompt[pos[s] + 1];            // still inside the prompt
        else if (greedy || device_sampler) nx = next[s];
        else nx = sample(&samplers[req], logits_host + (size_t)s * vocab);
        bm.advance(s);
        pos[s] += 1;
        bool done = false;
        if ((nx == 1 || nx == 2) && !opt.ignore_stop) {
          done = true;
        } else {
          append_str(decode(&tok, token[s], nx), sl.text);
          sl.ids.push_back(nx);
          token[s] = nx;
          if (pos[s] >= sl.steps) done = true;
        }
        if (done) {
          requests->str_gens[req] = sl.text + "\n";
          requests->tok_gens[req] = sl.ids;
          gen_cnt += pos[s] - 1;
          bm.release_slot(s);
          pos[s] = -1;
          token[s] = 0;
          if (opt.verbose) fprintf(stderr, "\nThread %d DONE Request %d \n", gid, req);
        }
      }
    }
    loop_s = (time_in_ms() - t_loop0) / 1000.0;

    if (logits_host) tha_host_free(logits_host);
    bm.destroy_pool();   // paged: the pools are the BatchManager's (s_d's cache pointers are cleared with them)
    free_state_on_device(s_d);
    if (w_d) free_weight_on_device(w_d, w_d->wcls == w_d->token_embedding_table);
    if (wq_d) free_weight_on_device_q8(wq_d, 1);
    thablasDestroy(handle);
    free_tokenizer(&tok);
  }

  for (auto& s : samplers) free_sampler(&s);
  if (stats) { stats->gen_tokens = gen_cnt; stats->steps = n_steps; stats->prefill_rows = n_prefill_rows; stats->load_seconds = load_s; stats->loop_seconds = loop_s; }
  return (int)gen_cnt;
}

// ---- one replica, tensor parallel -------------------------------------------------------------
// The slot state machine of run_data_parallel with ONE set of slots; the model is split over opt.tp ranks (rank r on
// device r % n_devices: with fewer devices than ranks several ranks share a device, which is how the single-GPU tests
// run it), all driven by this thread through tha_tp_forward_greedy.  Greedy decode; prompts are fed token by token as in
// the reference (src/llama.cpp:1029-1030).
int run_tensor_parallel(Transformer* t32, TransformerQ8* tq8, const char* tokenizer_path, Requests* requests,
                        const DriverOptions& opt, DriverStats* stats) {
  const Config cfg = t32 ? t32->config : tq8->config;
  const int vocab = cfg.vocab_size, B = opt.batch_size, tp = opt.tp;
  if (opt.temperature != 0.0f) { fprintf(stderr, "tensor parallel (-w): greedy decode only, use -t 0\n"); exit(EXIT_FAILURE); }
  if (tp > 8) { fprintf(stderr, "tensor parallel (-w): at most 8 ranks\n"); exit(EXIT_FAILURE); }
  int n_devices = 0;
  CHECK_THA(tha_device_count(&n_devices));
  if (opt.n_devices > 0 && opt.n_devices < n_devices) n_devices = opt.n_devices;
  if (n_devices < 1) { fprintf(stderr, "no CUDA device\n"); exit(EXIT_FAILURE); }
  const int unit = t32 ? 64 : tq8->weights.group_size;
  int hidden_of[8];
  Config cfg_local[8];
  for (int r = 0; r < tp; r++)
    if (tp_local_config(&cfg, tp, r, unit, &cfg_local[r])) {
      fprintf(stderr, "tensor parallel (-w %d): heads %d / kv heads %d / hidden %d cannot be split\n", tp, cfg.n_heads, cfg.n_kv_heads, cfg.hidden_dim);
      exit(EXIT_FAILURE);
    }
  tp_split_hidden(cfg.hidden_dim, tp, unit, hidden_of);
  fprintf(stderr, "\n TENSOR PARALLELISM \n\n Ranks %d on %d device(s)\n\n Batch Size %d\n", tp, n_devices, B);
  // ranks sharing a device hand-shake between concurrent kernels of that device: an early-launched (PDL) persistent
  // matmul of one rank must not hold the SMs its peer needs
  if (n_devices < tp) tha_set_option("pdl", 0);

  Tokenizer tok;
  build_tokenizer(&tok, tokenizer_path, vocab);
  thablasHandle_t handle[8];
  thaTP_t ctx[8] = {nullptr};
  TransformerWeights* w_d[8] = {nullptr};
  TransformerWeightsQ8* wq_d[8] = {nullptr};
  RunState* s_d[8] = {nullptr};
  const int head_size = cfg.dim / cfg.n_heads;
  const long t_load0 = time_in_ms();
  for (int r = 0; r < tp; r++) {
    CHECK_THA(tha_set_device(r % n_devices));
    thablasCreate(&handle[r]);
    if (t32) copy_weight_shard_to_device(t32, tp, r, w_d[r]);
    else copy_weight_shard_to_device_q8(tq8, tp, r, wq_d[r]);
    alloc_state_to_device_batch_tp(&cfg, &cfg_local[r], s_d[r], B);
    if (thaTPCreate(&ctx[r], tp, r, head_size, hidden_of, (size_t)16 * B * cfg.dim) != THABLAS_STATUS_SUCCESS) {
      fprintf(stderr, "thaTPCreate failed: %s\n", tha_last_error());
      exit(EXIT_FAILURE);
    }
    CHECK_THA(tha_device_sync());
  }
  if (thaTPConnectLocal(ctx, tp) != THABLAS_STATUS_SUCCESS) {
    fprintf(stderr, "thaTPConnectLocal failed (no peer access between the devices?): %s\n", tha_last_error());
    exit(EXIT_FAILURE);
  }
  const double load_s = (time_in_ms() - t_load0) / 1000.0;
  fprintf(stdout, "\nLoad model time (%d shards): %f\n", tp, load_s);

  void* streams[8];
  const tha_config_t* cfgs[8];
  const void* ws[8];
  const tha_state_t* ss[8];
  for (int r = 0; r < tp; r++) {
    streams[r] = (void*)handle[r].calc_stream;
    cfgs[r] = (const tha_config_t*)&cfg_local[r];
    ws[r] = t32 ? (const void*)w_d[r] : (const void*)wq_d[r];
    ss[r] = (const tha_state_t*)s_d[r];
  }

  const int kvd = cfg.dim * cfg.n_kv_heads / cfg.n_heads;
  BatchManager bm(B, cfg.seq_len, cfg.n_layers, kvd);   // slot table only: the KV slabs are the ranks' own
  std::vector<Slot> slots(B);
  std::vector<int> token(B, 0), pos(B, -1), next(B, 0);
  int next_idx = opt.device_offset;
  long long gen_cnt = 0, n_steps = 0;
  const long t_loop0 = time_in_ms();
  for (;;) {
    for (int s = 0; s < B; s++) {   // refill free slots
      if (bm.request_of(s) >= 0) continue;
      const int req = next_idx;
      if (req >= requests->num_reqs) break;
      next_idx += opt.device_stride;
      const int got = bm.acquire_slot(req);
      Slot& sl = slots[got];
      const std::string& prompt = requests->str_reqs[req];
      sl.prompt.assign(prompt.size() + 3, 0);
      int n_prompt = 0;
      encode(&tok, prompt.c_str(), 1, 0, sl.prompt.data(), &n_prompt);
      if (n_prompt < 1) { fprintf(stderr, "something is wrong, expected at least 1 prompt token\n"); exit(EXIT_FAILURE); }
      sl.prompt.resize(n_prompt);
      sl.text.clear();
      sl.ids.assign(1, sl.prompt[0]);
      sl.steps = requests->max_seq_len;
      if (opt.max_new_positions > 0 && opt.max_new_positions < sl.steps) sl.steps = opt.max_new_positions;
      token[got] = sl.prompt[0];
      pos[got] = 0;
    }
    if (bm.active_slots() == 0) break;
    const int rc = tha_tp_forward_greedy(tp, ctx, streams, B, t32 ? 0 : 1, cfgs, ws, ss, token.data(), pos.data(), next.data());
    if (rc) { fprintf(stderr, "tha_tp_forward_greedy failed (%d): %s\n", rc, tha_last_error()); exit(EXIT_FAILURE); }
    n_steps++;
    for (int s = 0; s < B; s++) {   // src/llama.cpp:1025-1050
      const int req = bm.request_of(s);
      if (req < 0) continue;
      Slot& sl = slots[s];
      const int nx = (pos[s] < (int)sl.prompt.size() - 1) ? sl.prompt[pos[s] + 1] : next[s];
      bm.advance(s);
      pos[s] += 1;
      bool done = false;
      if ((nx == 1 || nx == 2) && !opt.ignore_stop) {
        done = true;
      } else {
        append_str(decode(&tok, token[s], nx), sl.text);
        sl.ids.push_back(nx);
        token[s] = nx;
        if (pos[s] >= sl.steps) done = true;
      }
      if (done) {
        requests->str_gens[req] = sl.text + "\n";
        requests->tok_gens[req] = sl.ids;
        gen_cnt += pos[s] - 1;
        bm.release_slot(s);
        pos[s] = -1;
        token[s] = 0;
      }
    }
  }
  const double loop_s = (time_in_ms() - t_loop0) / 1000.0;

  for (int r = 0; r < tp; r++) {
    CHECK_THA(tha_set_device(r % n_devices));
    CHECK_THA(tha_device_sync());
  }
  for (int r = 0; r < tp; r++) {
    CHECK_THA(tha_set_device(r % n_devices));
    thaTPDestroy(ctx[r]);
    free_state_on_device(s_d[r]);
    if (w_d[r]) free_weight_on_device(w_d[r], w_d[r]->wcls == w_d[r]->token_embedding_table);
    if (wq_d[r]) free_weight_on_device_q8(wq_d[r], 1);
    thablasDestroy(handle[r]);
  }
  free_tokenizer(&tok);
  if (stats) { stats->gen_tokens = gen_cnt; stats->steps = n_steps; stats->prefill_rows = 0; stats->load_seconds = load_s; stats->loop_seconds = loop_s; }
  return (int)gen_cnt;
}

// The reference's entry point, same signature (src/llama.cpp:891): test-mode sampling
// (temperature 1.0, top-p 0.9, seed 314028 for every request).
int test_data_parallelism(Transformer* transformer, Tokenizer* tokenizer, const char* tokenizer_path, Requests* requests,
                          int batch_size) {
  (void)tokenizer;  // every device thread builds its own, as in the reference (:930-931)
  DriverOptions opt;
  opt.batch_size = batch_size;
  return run_data_parallel(transformer, nullptr, tokenizer_path, requests, opt, nullptr);
}
