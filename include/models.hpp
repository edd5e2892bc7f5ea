// models.hpp -- model / state structs and the device-memory helpers of the reference
// (include/models.hpp:10-134, src/models.cpp:86-179), same names and field order.
//
// `Config` is read straight from the checkpoint header (src/utils.cpp:155), so its field order
// is part of the file format.  The structs are layout-identical to the C-ABI mirrors in
// tha_b200.h (static_asserts in host/models.cpp), which is how they cross the boundary
// without copies.
#pragma once
#include <stdint.h>
#include <stdlib.h>
#include <sys/types.h>

#include "tha_helper.hpp"
#include "thaBLAS.hpp"

typedef struct {
  int dim;         // transformer dimension
  int hidden_dim;  // ffn width
  int n_layers;
  int n_heads;     // query heads
  int n_kv_heads;  // key/value heads (<= n_heads)
  int vocab_size;
  int seq_len;     // max sequence length
} Config;

typedef struct {
  float* token_embedding_table;  // (vocab_size, dim)
  float* rms_att_weight;         // (layer, dim)
  float* rms_ffn_weight;         // (layer, dim)
  float* wq;                     // (layer, dim, dim)          row-major [out, in]
  float* wk;                     // (layer, kv_dim, dim)
  float* wv;                     // (layer, kv_dim, dim)
  float* wo;                     // (layer, dim, dim)
  float* w1;                     // (layer, hidden_dim, dim)
  float* w2;                     // (layer, dim, hidden_dim)
  float* w3;                     // (layer, hidden_dim, dim)
  float* rms_final_weight;       // (dim,)
  float* wcls;                   // (vocab_size, dim); == token_embedding_table when shared
} TransformerWeights;

typedef struct {
  float* x;    // (B, dim) residual stream
  float* xb;   // (B, dim)
  float* xb2;  // (B, dim)
  float* hb;   // (B, hidden_dim)
  float* hb2;  // (B, hidden_dim)
  float* q;    // (B, dim)
  float* k;    // unused on the device: k/v rows are written straight into the cache
  float* v;
  float* att;     // (B, n_heads, seq_len) scratch of the 3-stage attention ops (NULL in the fused step)
  float* logits;  // (B, vocab_size)
  float* key_cache;    // (B, layer, seq_len, kv_dim)
  float* value_cache;  // (B, layer, seq_len, kv_dim)
  float* key_matmul;   // pipeline / swap variants of the reference (include/models.hpp:56-59); unused here
  float* value_matmul;
  float* key_layer_cache;
  float* value_layer_cache;
} RunState;

typedef struct {
  Config config;
  TransformerWeights weights;  // host pointers into the mapping (build_transformer) or device pointers
  RunState state;
  int fd;             // file descriptor of the mapping
  float* data;        // mapped checkpoint
  ssize_t file_size;  // bytes
} Transformer;

typedef struct {
  float* addr;
  int occupied;
} KVBlock;

// ---- Q8_0 (runq.c:34-37, :174-251): the reference's GPU app cannot load these; this one can.
typedef struct {
  int8_t* q;  // quantised values
  float* s;   // one scale per group
} QuantizedTensor;

typedef struct {
  float* token_embedding_table;  // dequantised (vocab, dim) fp32 (runq.c:205-206)
  float* rms_att_weight;
  float* rms_ffn_weight;
  float* rms_final_weight;
  QuantizedTensor wq, wk, wv, wo, w1, w2, w3, wcls;  // layer-major, contiguous
  int group_size;
} TransformerWeightsQ8;

typedef struct {
  Config config;
  TransformerWeightsQ8 weights;  // host pointers (build_transformer_q8) or device pointers
  int shared_classifier;
  int fd;
  void* data;
  ssize_t file_size;
  float* dequant_embedding;  // host buffer owned by the loader
} TransformerQ8;

// ---- device helpers (src/models.cpp) ---------------------------------------------------
// Upload the 12 weight tensors of a v0 checkpoint to the current device (src/models.cpp:86-125).
void copy_weight_to_device(Transformer* t_h, TransformerWeights*& w_d);
// Batched decode state: every activation buffer is [batch_size, ...] (src/models.cpp:155-179).
void alloc_state_to_device(Transformer* t_h, RunState*& s_d);
void alloc_state_to_device_batch(Transformer* t_h, RunState*& s_d_batch, int batch_size);
// weights + single-slot state as one device Transformer (src/models.cpp:60-84)
void copy_transformer_to_device(thablasHandle_t handle, Transformer* t_h, Transformer*& t_d);
// B200 additions: release what the helpers above allocated (the reference never frees), and
// a state allocator that skips the att scratch the fused step does not need.
void free_weight_on_device(TransformerWeights* w_d, int shared_classifier);
void free_state_on_device(RunState* s_d);
void alloc_state_to_device_batch_noatt(Transformer* t_h, RunState*& s_d_batch, int batch_size);
void copy_weight_to_device_q8(TransformerQ8* t_h, TransformerWeightsQ8*& w_d);
void free_weight_on_device_q8(TransformerWeightsQ8* w_d, int shared_classifier);
void alloc_state_to_device_batch_cfg(const Config* cfg, RunState*& s_d_batch, int batch_size, int with_att);
// activation buffers only: the K/V rows live in a BatchManager block pool (BatchManager::create_pool / bind_paged)
void alloc_state_to_device_batch_paged(const Config* cfg, RunState*& s_d_batch, int batch_size, int with_att);
// Tensor-parallel shards (B200 addition): rows of wq/wk/wv/w1/w3, columns of wo/w2, everything
// else replicated.  Hidden units are dealt in whole groups of `unit` (the Q8_0 group size; 64 for fp32 models).
int tp_split_hidden(int hidden, int tp, int unit, int* hidden_of_rank);
int tp_local_config(const Config* cfg, int tp, int rank, int unit, Config* local);
void copy_weight_shard_to_device(Transformer* t_h, int tp, int rank, TransformerWeights*& w_d);
void copy_weight_shard_to_device_q8(TransformerQ8* t_h, int tp, int rank, TransformerWeightsQ8*& w_d);
void alloc_state_to_device_batch_tp(const Config* cfg_global, const Config* cfg_local, RunState*& s_d_batch, int batch_size);
void free_transformer_device();
