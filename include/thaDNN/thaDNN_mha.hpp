// reference: include/thaDNN/thaDNN_mha.hpp:1-47, src/thaDNN/thaDNN_mha.cpp:60-426
//
// The three attention stages of the reference, same names and arguments:
//   1  att[b,h,t] = q[b,h,:] . K[b,t,h/kv_mul,:] / sqrt(head_size),   t <= pos[b]
//   2  softmax over t of every (b,h) row, size_batch[b] = pos[b] + 1 entries
//   3  xb[b,h,:] = sum_t att[b,h,t] * V[b,t,h/kv_mul,:]
// v1: KV cache batch-major [B][n_layers][seq_len][kv_dim], `loff` = l*seq_len*kv_dim.
// v2: one layer's KV, time-major [t][B][kv_dim]; att rows are n_words long.
// `pos[]` is the host copy (unused by the kernels), `pos_d[]` the device copy.
//
// The decode step itself (thaDNN_s_forward_batch) does not call these: it runs ONE fused
// QK^T / online-softmax / PV kernel (thaDNN_s_multiheads_fused_v1_batch below).
#pragma once
#include "thaBLAS.hpp"

thablasStatus_t thaDNN_s_multiheads_1_v2_batch(thablasHandle_t* handle, int batch_size, int pipe_size, int pos[],
                                               int pos_d[], int n_heads, float* s_q_batch, float* s_att_batch,
                                               float* s_key_cache_batch, int head_size, int n_words, int kv_dim,
                                               int dim, int kv_mul);

thablasStatus_t thaDNN_s_multiheads_2_batch(thablasHandle_t* handle, int n_batches, float* s_att_batch,
                                            int size_batch[], int seq_len, int n_heads);

thablasStatus_t thaDNN_s_multiheads_3_v2_batch(thablasHandle_t* handle, int batch_size, int pos_d[], int n_heads,
                                               float* s_xb_batch, float* s_att_batch, float* s_value_cache_batch,
                                               int head_size, int n_words, int kv_dim, int kv_mul, int dim,
                                               int pipe_size);

thablasStatus_t thaDNN_s_multiheads_1_v1_batch(thablasHandle_t* handle, int n_batches, int pos[], int pos_d[],
                                               int n_heads, int n_layers, float* s_q_batch, float* s_att_batch,
                                               float* s_key_cache_batch, int head_size, int seq_len, int loff,
                                               int kv_dim, int dim, int kv_mul);
thablasStatus_t thaDNN_s_multiheads_2_v1_batch(thablasHandle_t* handle, int n_batches, float* s_att_batch,
                                               int size_batch[], int seq_len, int n_heads);
thablasStatus_t thaDNN_s_multiheads_3_v1_batch(thablasHandle_t* handle, int n_batches, int pos_d[], int n_heads,
                                               float* s_xb_batch, float* s_att_batch, float* s_value_cache_batch,
                                               int head_size, int seq_len, int loff, int kv_dim, int kv_mul,
                                               int dim, int n_layers);

// B200 addition: the three stages in one kernel (v1 layout); slots with pos_d[b] < 0 are skipped.
thablasStatus_t thaDNN_s_multiheads_fused_v1_batch(thablasHandle_t* handle, int n_batches, int pos_d[], int n_heads,
                                                   int n_kv_heads, int n_layers, float* s_q_batch, float* s_xb_batch,
                                                   float* s_key_cache_batch, float* s_value_cache_batch,
                                                   int head_size, int seq_len, int loff);
