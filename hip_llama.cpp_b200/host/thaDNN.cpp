// thaDNN host layer: the reference's operator names (include/thaDNN.hpp, include/thaDNN/*.hpp)
// forwarded to the C-ABI.  No CUDA in this file.
#include "thaDNN.hpp"

#include <stdio.h>
#include <string.h>

#include "tha_b200.h"

static_assert(sizeof(Config) == sizeof(tha_config_t), "Config must mirror tha_config_t");
static_assert(sizeof(TransformerWeights) == sizeof(tha_weights_f32_t), "TransformerWeights must mirror tha_weights_f32_t");
static_assert(sizeof(RunState) == sizeof(tha_state_t), "RunState must mirror tha_state_t");
static_assert(sizeof(TransformerWeightsQ8) == sizeof(tha_weights_q8_t), "TransformerWeightsQ8 must mirror tha_weights_q8_t");

static inline thablasStatus_t st(int rc) { return (thablasStatus_t)rc; }
static inline void* calc(thablasHandle_t* h) { return (void*)h->calc_stream; }

thablasStatus_t thaDNN_s_rmsnorm_v2_batch(thablasHandle_t* handle, int n_batches, float* o_batch, float* x_batch,
                                          float* weight, int size, int dim) {
  if (!handle) return THABLAS_STATUS_HANDLE_IS_NULLPTR;
  return st(tha_rmsnorm_batch_f32(calc(handle), n_batches, o_batch, x_batch, weight, size, dim));
}

thablasStatus_t thaDNN_s_rope(thablasHandle_t* handle, int dim, int head_size, int kv_dim, int pos, float* q, float* k) {
  if (!handle) return THABLAS_STATUS_HANDLE_IS_NULLPTR;
  return st(tha_rope_f32(calc(handle), dim, head_size, kv_dim, pos, q, k));
}

// src/thaDNN/thaDNN_softmax.cpp:102-119: argument check, device of the handle, default stream
thablasStatus_t thaDNN_s_softmax_v2(thablasHandle_t handle, float* x, int size) {
  if (size <= 0 || x == nullptr) {
    printf("THABLAS ERROR: INVALID ARGUMENT\n");
    return THABLAS_STATUS_ALLOC_FAILED;
  }
  tha_set_device(handle.current_gpu_id);
  return st(tha_softmax_f32(nullptr, x, size));
}
thablasStatus_t thaDNN_s_softmax_v2(thablasHandle_t* handle, float* x, int size) {
  if (!handle) return THABLAS_STATUS_HANDLE_IS_NULLPTR;
  return thaDNN_s_softmax_v2(*handle, x, size);
}

thablasStatus_t thaDNN_s_swiglu(thablasHandle_t* handle, float* hb, float* hb2, int hidden_dim) {
  if (!handle) return THABLAS_STATUS_HANDLE_IS_NULLPTR;
  return st(tha_swiglu_f32(calc(handle), hb, hb2, hidden_dim));
}

// ---- attention stages (src/thaDNN/thaDNN_mha.cpp) --------------------------------------
thablasStatus_t thaDNN_s_multiheads_1_v1_batch(thablasHandle_t* handle, int n_batches, int pos[], int pos_d[],
                                               int n_heads, int n_layers, float* s_q_batch, float* s_att_batch,
                                               float* s_key_cache_batch, int head_size, int seq_len, int loff,
                                               int kv_dim, int dim, int kv_mul) {
  (void)pos;
  if (!handle) return THABLAS_STATUS_HANDLE_IS_NULLPTR;
  return st(tha_mha_scores_v1_f32(calc(handle), n_batches, pos_d, n_heads, n_layers, s_q_batch, s_att_batch,
                                  s_key_cache_batch, head_size, seq_len, loff, kv_dim, dim, kv_mul));
}
// size_batch[] is the device pos array: the kernel normalises pos+1 entries per row (:306-374)
thablasStatus_t thaDNN_s_multiheads_2_v1_batch(thablasHandle_t* handle, int n_batches, float* s_att_batch,
                                               int size_batch[], int seq_len, int n_heads) {
  if (!handle) return THABLAS_STATUS_HANDLE_IS_NULLPTR;
  return st(tha_mha_softmax_f32(calc(handle), n_batches, s_att_batch, size_batch, seq_len, n_heads));
}
thablasStatus_t thaDNN_s_multiheads_2_batch(thablasHandle_t* handle, int n_batches, float* s_att_batch,
                                            int size_batch[], int seq_len, int n_heads) {
  return thaDNN_s_multiheads_2_v1_batch(handle, n_batches, s_att_batch, size_batch, seq_len, n_heads);
}
thablasStatus_t thaDNN_s_multiheads_3_v1_batch(thablasHandle_t* handle, int n_batches, int pos_d[], int n_heads,
                                               float* s_xb_batch, float* s_att_batch, float* s_value_cache_batch,
                                               int head_size, int seq_len, int loff, int kv_dim, int kv_mul,
                                               int dim, int n_layers) {
  if (!handle) return THABLAS_STATUS_HANDLE_IS_NULLPTR;
  return st(tha_mha_weighted_v1_f32(calc(handle), n_batches, pos_d, n_heads, s_xb_batch, s_att_batch,
                                    s_value_cache_batch, head_size, seq_len, loff, kv_dim, kv_mul, dim, n_layers));
}
thablasStatus_t thaDNN_s_multiheads_1_v2_batch(thablasHandle_t* handle, int batch_size, int pipe_size, int pos[],
                                               int pos_d[], int n_heads, float* s_q_batch, float* s_att_batch,
                                               float* s_key_cache_batch, int head_size, int n_words, int kv_dim,
                                               int dim, int kv_mul) {
  (void)pipe_size; (void)pos;
  if (!handle) return THABLAS_STATUS_HANDLE_IS_NULLPTR;
  return st(tha_mha_scores_v2_f32(calc(handle), batch_size, pos_d, n_heads, s_q_batch, s_att_batch,
                                  s_key_cache_batch, head_size, n_words, kv_dim, dim, kv_mul));
}
thablasStatus_t thaDNN_s_multiheads_3_v2_batch(thablasHandle_t* handle, int batch_size, int pos_d[], int n_heads,
                                               float* s_xb_batch, float* s_att_batch, float* s_value_cache_batch,
                                               int head_size, int n_words, int kv_dim, int kv_mul, int dim,
                                               int pipe_size) {
  (void)pipe_size;
  if (!handle) return THABLAS_STATUS_HANDLE_IS_NULLPTR;
  return st(tha_mha_weighted_v2_f32(calc(handle), batch_size, pos_d, n_heads, s_xb_batch, s_att_batch,
                                    s_value_cache_batch, head_size, n_words, kv_dim, kv_mul, dim));
}
thablasStatus_t thaDNN_s_multiheads_fused_v1_batch(thablasHandle_t* handle, int n_batches, int pos_d[], int n_heads,
                                                   int n_kv_heads, int n_layers, float* s_q_batch, float* s_xb_batch,
                                                   float* s_key_cache_batch, float* s_value_cache_batch,
                                                   int head_size, int seq_len, int loff) {
  if (!handle) return THABLAS_STATUS_HANDLE_IS_NULLPTR;
  return st(tha_mha_fused_v1_f32(calc(handle), n_batches, pos_d, n_heads, n_kv_heads, n_layers, s_q_batch, s_xb_batch,
                                 s_key_cache_batch, s_value_cache_batch, head_size, seq_len, loff));
}

// ---- the forward pass ------------------------------------------------------------------
thablasStatus_t thaDNN_s_forward_batch(thablasHandle_t handle1, thablasHandle_t handle2, thablasHandle_t handle3,
                                       int n_batches, Config* p, TransformerWeights* w, RunState* s_batch,
                                       int token[], int pos[], float* logits_host) {
  (void)handle2; (void)handle3;  // the reference uses handle1 only (src/thaDNN.cpp:40)
  return st(tha_forward_batch_f32((void*)handle1.calc_stream, n_batches, (const tha_config_t*)p,
                                  (const tha_weights_f32_t*)w, (const tha_state_t*)s_batch, token, pos, logits_host));
}

thablasStatus_t thaDNN_q8_forward_batch(thablasHandle_t handle, int n_batches, Config* p, TransformerWeightsQ8* w,
                                        RunState* s_batch, int token[], int pos[], float* logits_host) {
  return st(tha_forward_batch_q8((void*)handle.calc_stream, n_batches, (const tha_config_t*)p,
                                 (const tha_weights_q8_t*)w, (const tha_state_t*)s_batch, token, pos, logits_host));
}

thablasStatus_t thaDNN_s_prefill_batch(thablasHandle_t* handle, Config* p, TransformerWeights* w, RunState* s_batch,
                                       int n_seqs, int slot[], int pos0[], int len[], int tokens[], float* logits_d,
                                       int* next_d) {
  if (!handle) return THABLAS_STATUS_HANDLE_IS_NULLPTR;
  return st(tha_prefill_f32(calc(handle), (const tha_config_t*)p, (const tha_weights_f32_t*)w, (const tha_state_t*)s_batch,
                            n_seqs, slot, pos0, len, tokens, logits_d, next_d));
}

thablasStatus_t thaDNN_s_decode_greedy_batch(thablasHandle_t* handle, int n_batches, Config* p, TransformerWeights* w,
                                             RunState* s_batch, int* token_d, int* pos_d, int* next_d,
                                             float* logits_d, int advance) {
  if (!handle) return THABLAS_STATUS_HANDLE_IS_NULLPTR;
  if (advance)
    return st(tha_decode_greedy_f32(calc(handle), n_batches, (const tha_config_t*)p, (const tha_weights_f32_t*)w,
                                    (const tha_state_t*)s_batch, token_d, pos_d, logits_d, next_d));
  return st(tha_decode_step_f32(calc(handle), n_batches, (const tha_config_t*)p, (const tha_weights_f32_t*)w,
                                (const tha_state_t*)s_batch, token_d, pos_d, logits_d, next_d));
}

// ---- tensor parallelism (include/tha_b200.h: tha_tp_*) ------------------------------------------------
thablasStatus_t thaTPCreate(thaTP_t* tp, int tp_size, int rank, int head_size, const int hidden_of_rank[], size_t xchg_floats) {
  return st(tha_tp_create(tp, tp_size, rank, head_size, hidden_of_rank, xchg_floats));
}
thablasStatus_t thaTPConnectLocal(thaTP_t ranks[], int tp_size) {
  if (!ranks || tp_size < 1 || tp_size > 8) return THABLAS_STATUS_ALLOC_FAILED;
  void* bases[8];
  for (int r = 0; r < tp_size; r++) {
    const int rc = tha_tp_local_ptr(ranks[r], &bases[r]);
    if (rc) return st(rc);
  }
  // tha_tp_connect_ptr enables peer access FROM the current device: the caller selects each rank's device in turn,
  // exactly as the reference's per-device threads do (src/llama.cpp:927); here we query the owner of each block.
  int prev = 0;
  tha_get_device(&prev);
  for (int r = 0; r < tp_size; r++) {
    int rc = tha_tp_select_device(ranks[r]);
    if (!rc) rc = tha_tp_connect_ptr(ranks[r], bases);
    if (rc) { tha_set_device(prev); return st(rc); }
  }
  tha_set_device(prev);
  return THABLAS_STATUS_SUCCESS;
}
thablasStatus_t thaTPDestroy(thaTP_t tp) { return st(tha_tp_destroy(tp)); }
thablasStatus_t thaDNN_s_forward_batch_tp(thablasHandle_t* handle, thaTP_t tp, int n_batches, Config* p_local,
                                          TransformerWeights* w_local, RunState* s_local, int* token_d, int* pos_d,
                                          float* logits_d, int* next_d, int advance) {
  if (!handle) return THABLAS_STATUS_HANDLE_IS_NULLPTR;
  return st(tha_tp_decode_step_f32(calc(handle), tp, n_batches, (const tha_config_t*)p_local, (const tha_weights_f32_t*)w_local,
                                   (const tha_state_t*)s_local, token_d, pos_d, logits_d, next_d, advance));
}
thablasStatus_t thaDNN_q8_forward_batch_tp(thablasHandle_t* handle, thaTP_t tp, int n_batches, Config* p_local,
                                           TransformerWeightsQ8* w_local, RunState* s_local, int* token_d, int* pos_d,
                                           float* logits_d, int* next_d, int advance) {
  if (!handle) return THABLAS_STATUS_HANDLE_IS_NULLPTR;
  return st(tha_tp_decode_step_q8(calc(handle), tp, n_batches, (const tha_config_t*)p_local, (const tha_weights_q8_t*)w_local,
                                  (const tha_state_t*)s_local, token_d, pos_d, logits_d, next_d, advance));
}
