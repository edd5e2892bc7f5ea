/* tha_b200.h -- the C-ABI boundary of the B200 (sm_100a) llama2 forward pass.
 *
 * Everything that touches the GPU goes through the `extern "C"` entry points
 * declared here: plain pointers, ints and an opaque `void* stream`
 * (a cudaStream_t; NULL = the legacy default stream).  No C++ or torch types.
 * The C++ headers that keep the reference's names (include/thaBLAS.hpp,
 * include/thaDNN.hpp, include/models.hpp, ...) are thin wrappers over these
 * symbols; a Python / cgo / JNI caller binds the same symbols directly
 * (see INTEGRATION.md).
 *
 * Citations are to the reference tree (/root/reference): each entry point names
 * the reference interface it replaces.
 *
 * Return value: a thablasStatus_t value (include/thaBLAS.hpp:5-19 in the
 * reference): 0 = success, 2 = ALLOC_FAILED (the reference's "invalid argument"
 * code, src/thaBLAS.cpp:80-84), 5 = EXECUTION_FAILED (a CUDA launch error),
 * 7 = NOT_SUPPORTED.  All pointers are DEVICE pointers unless named *_host.
 * All ops are asynchronous on `stream` unless stated otherwise.
 */
#ifndef THA_B200_H
#define THA_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define THA_STATUS_SUCCESS 0
#define THA_STATUS_INVALID_ARG 2   /* THABLAS_STATUS_ALLOC_FAILED, what the reference returns for bad args */
#define THA_STATUS_EXECUTION_FAILED 5
#define THA_STATUS_NOT_SUPPORTED 7

/* ---- POD mirrors of include/models.hpp:10-60 (identical field order) ------ */
typedef struct {
  int dim, hidden_dim, n_layers, n_heads, n_kv_heads, vocab_size, seq_len;
} tha_config_t; /* == Config; fread straight from the checkpoint (src/utils.cpp:155) */

typedef struct {
  float* token_embedding_table; /* (vocab, dim) */
  float* rms_att_weight;        /* (layer, dim) */
  float* rms_ffn_weight;        /* (layer, dim) */
  float* wq;                    /* (layer, dim, dim) row-major [out, in] */
  float* wk;                    /* (layer, kv_dim, dim) */
  float* wv;                    /* (layer, kv_dim, dim) */
  float* wo;                    /* (layer, dim, dim) */
  float* w1;                    /* (layer, hidden, dim) */
  float* w2;                    /* (layer, dim, hidden) */
  float* w3;                    /* (layer, hidden, dim) */
  float* rms_final_weight;      /* (dim) */
  float* wcls;                  /* (vocab, dim) */
} tha_weights_f32_t; /* == TransformerWeights */

typedef struct {
  float *x, *xb, *xb2, *hb, *hb2, *q, *k, *v, *att, *logits;
  float *key_cache, *value_cache; /* [B][n_layers][seq_len][kv_dim] (src/models.cpp:176-177) */
  float *key_matmul, *value_matmul, *key_layer_cache, *value_layer_cache;
} tha_state_t; /* == RunState (batched: every activation buffer is [B, ...]) */

/* Q8_0 weights, runq.c:34-37 format: int8 values [out,in] + one fp32 scale per
 * `group_size` consecutive values.  Per-layer tensors are stored layer-major
 * and contiguous (q: [L, out*in], s: [L, out*in/GS]). */
typedef struct { int8_t* q; float* s; } tha_qtensor_t;
typedef struct {
  float* token_embedding_table; /* dequantised fp32 (runq.c:205-206) */
  float *rms_att_weight, *rms_ffn_weight, *rms_final_weight;
  tha_qtensor_t wq, wk, wv, wo, w1, w2, w3, wcls;
  int group_size;
} tha_weights_q8_t;

/* bf16 weights (the reference declares 16-bit weight structs at include/models.hpp:77-117 and never fills them;
 * BASELINE.json configs[3] "fp32/bf16"): the tensors of tha_weights_f32_t, matrices and embedding table as raw bf16 bit
 * patterns (round-to-nearest-even of the fp32 checkpoint, tha_convert_bf16), norm gains fp32.  Residual stream, KV cache,
 * attention and logits stay fp32; every vector entering a weight matmul is rounded to bf16. */
typedef struct {
  uint16_t* token_embedding_table;
  float *rms_att_weight, *rms_ffn_weight, *rms_final_weight;
  uint16_t *wq, *wk, *wv, *wo, *w1, *w2, *w3, *wcls;
} tha_weights_bf16_t;

/* ---- library ---------------------------------------------------------------- */
int tha_version(void);
const char* tha_last_error(void);       /* text of the last CUDA error seen by this library */
int tha_set_option(const char* name, int value); /* "pdl", "graph", "mega" (0/1); "tc": -1 auto (tensor-core
                                                    matmuls for more than 4 slots), 0 never, 1 whenever supported;
                                                    "dual": slots from which a step runs as two concurrent half-batch
                                                    branches (default 256, 0 never); "ropefuse" (RoPE + KV write inside
                                                    the attention kernel), "sample_resident", "chain", "skip": A/B and
                                                    diagnostic switches, see DESIGN.md section 4 */

/* ---- runtime shim: the only way the C++ host layer touches the CUDA runtime ---------
 * (reference: hipSetDevice / hipMalloc / hipMemcpy / hipHostMalloc / hipStreamCreate /
 *  hipDeviceSynchronize calls in src/models.cpp:86-179, src/llama.cpp:919-947, src/thaBLAS.cpp:49-58) */
int tha_device_count(int* n);
int tha_get_device(int* dev);
int tha_set_device(int dev);
int tha_device_sync(void);
int tha_stream_create(void** stream);
int tha_stream_destroy(void* stream);
int tha_stream_sync(void* stream);
int tha_malloc(void** dptr, size_t bytes);
int tha_free(void* dptr);
int tha_host_alloc(void** hptr, size_t bytes);   /* pinned */
int tha_host_free(void* hptr);
int tha_memset(void* dptr, int value, size_t bytes);
int tha_memcpy_h2d(void* dst, const void* src, size_t bytes);                 /* blocking */
int tha_memcpy_d2h(void* dst, const void* src, size_t bytes);                 /* blocking */
int tha_memcpy_d2d(void* dst, const void* src, size_t bytes);                 /* blocking */
int tha_memcpy_async(void* dst, const void* src, size_t bytes, void* stream); /* direction inferred */
int tha_mem_info(size_t* free_bytes, size_t* total_bytes);

/* ---- thaBLAS level 1/2 -------------------------------------------------------- */
/* thablas_Svds, src/thaBLAS.cpp:72-95: B[i] = A[i] / val */
int tha_svds_f32(void* stream, int n, const float* A, float* B, float val);
/* thaBLAS_s_vecaddvec, src/thaBLAS.cpp:104-126: a[i] += b[i] */
int tha_vecaddvec_f32(void* stream, float* a, const float* b, int size);
/* thaBLAS_s_matmul_batch, src/thaBLAS.cpp:191-228 (the weight matmuls):
 *   C_b[0:M] = A[M,K] . B_b[K];  C_b = C + Coff + has_pos*pos_d[b] + b*C_batch_size,
 *   B_b = B + b*B_batch_size.  pos_d may be NULL when has_pos == 0. */
int tha_matmul_batch_f32(void* stream, int n_batches, float* C_batch, const float* B_batch,
                         const float* A, int K, int M, int Coff, int has_pos, const int* pos_d,
                         int C_batch_size, int B_batch_size);
/* thaDNN_s_matmulvec_v2 / thaBLAS_s_matmulvec, src/thaBLAS.cpp:241-277: C[M] = A[M,K] . B[K] */
int tha_matmulvec_f32(void* stream, float* C, const float* B, const float* A, int K, int M);
/* thaBLAS_s_matmul_reduction, src/thaBLAS.cpp:281-319: A row-major [M,K]; B holds N
 * rows of length K; C holds N rows of length M: C[j*M+i] = sum_k A[i,k] B[j,k] */
int tha_matmul_reduction_f32(void* stream, const float* A, const float* B, float* C, int M, int N, int K);
/* thaBLAS_s_matmul, src/thaBLAS.cpp:134-170: C[M,N] = A[M,K] . B[K,N], all row-major */
int tha_matmul_f32(void* stream, int m, int n, int k, const float* A, const float* B, float* C);

/* Tensor-core form (tcgen05, error-compensated 3xTF32, fp32 in / fp32 out) of the skinny matmul
 * thaBLAS_s_matmul_reduction, src/thaBLAS.cpp:281-319: out[n*ldo + i] = sum_k W[i,k] X[n*ldx + k]
 * for N activation rows against an [M, K] row-major weight matrix.  NOT_SUPPORTED unless
 * K % 4 == 0, K >= 32 and W is 16-byte aligned. */
int tha_gemm_tf32x3(void* stream, float* out, long long ldo, const float* W, int M, int K, const float* X, int N,
                    long long ldx);
/* 1 if a tensor-core matmul on the current device gave up on a pipeline barrier since the last call */
int tha_gemm_check_error(void);
/* bf16 weights x bf16 activations on the tensor cores (tcgen05.mma.kind::f16, fp32 accumulate), fp32 out:
 * out[n*ldo + row] = sum_k W[row,k] X[n,k]; W [M,K], X [N,ldx] as raw bf16 bit patterns; K % 8 == 0, K >= 64.
 * The matmul of the bf16 model path (tha_*_bf16 below). */
int tha_gemm_bf16(void* stream, float* out, long long ldo, const uint16_t* W, int M, int K, const uint16_t* X, int N,
                  long long ldx);
/* runq.c:317-342 on the INT8 tensor cores: out[n*ldo + m] = sum_g (float)(sum_{k in g} Wq[m,k] Xq[n,k]) * Ws[m,g] * XsT[g*ldxs + n],
 * groups in ascending order, no FMA contraction -- bit-identical to the reference loop.  Xq int8 [N, K]; XsT the activation
 * group scales TRANSPOSED [K/gs, ldxs] with ldxs >= N rounded up to 16/32/64 (the column tile) and a multiple of 4.
 * gs in {32, 64, 128}, K % gs == 0, K >= 128. */
int tha_matmul_q8_tc(void* stream, float* out, long long ldo, const int8_t* Wq, const float* Ws, int gs, int M, int K,
                     const int8_t* Xq, const float* XsT, int ldxs, int N);

/* ---- thaDNN ops ----------------------------------------------------------------- */
/* thaDNN_s_rmsnorm_v2_batch, src/thaDNN/thaDNN_rmsnorm.cpp:35-83; rows `dim` apart, o may alias x */
int tha_rmsnorm_batch_f32(void* stream, int n_batches, float* o, const float* x, const float* weight,
                          int size, int dim);
/* thaDNN_s_rope, src/thaDNN/thaDNN_rope.cpp:25-63 (powf/cosf/sinf on the device) */
int tha_rope_f32(void* stream, int dim, int head_size, int kv_dim, int pos, float* q, float* k);
/* thaDNN_s_softmax_v2, src/thaDNN/thaDNN_softmax.cpp:62-119: in place, one vector */
int tha_softmax_f32(void* stream, float* x, int size);
/* thaDNN_s_swiglu, src/thaDNN/thaDNN_swiglu.cpp:5-35: hb = hb*sigmoid(hb)*hb2 */
int tha_swiglu_f32(void* stream, float* hb, const float* hb2, int hidden_dim);
/* thaDNN_s_multiheads_{1,2,3}_v1_batch, src/thaDNN/thaDNN_mha.cpp:246-426:
 * KV batch-major [B][n_layers][seq_len][kv_dim]; att scratch [B][n_heads][seq_len]. */
int tha_mha_scores_v1_f32(void* stream, int n_batches, const int* pos_d, int n_heads, int n_layers,
                          const float* q, float* att, const float* key_cache, int head_size,
                          int seq_len, int loff, int kv_dim, int dim, int kv_mul);
int tha_mha_softmax_f32(void* stream, int n_batches, float* att, const int* pos_d, int seq_len, int n_heads);
int tha_mha_weighted_v1_f32(void* stream, int n_batches, const int* pos_d, int n_heads, float* xb,
                            const float* att, const float* value_cache, int head_size, int seq_len,
                            int loff, int kv_dim, int kv_mul, int dim, int n_layers);
/* thaDNN_s_multiheads_{1,3}_v2_batch, src/thaDNN/thaDNN_mha.cpp:60-240: one layer's KV,
 * time-major [t][B][kv_dim]; att rows `n_words` long. */
int tha_mha_scores_v2_f32(void* stream, int n_batches, const int* pos_d, int n_heads, const float* q,
                          float* att, const float* key_cache, int head_size, int n_words, int kv_dim,
                          int dim, int kv_mul);
int tha_mha_weighted_v2_f32(void* stream, int n_batches, const int* pos_d, int n_heads, float* xb,
                            const float* att, const float* value_cache, int head_size, int n_words,
                            int kv_dim, int kv_mul, int dim);
/* Fused replacement of the three stages (QK^T / online softmax / PV in one kernel),
 * v1 layout.  Slots with pos_d[b] < 0 are skipped. */
int tha_mha_fused_v1_f32(void* stream, int n_batches, const int* pos_d, int n_heads, int n_kv_heads,
                         int n_layers, const float* q, float* xb, const float* key_cache,
                         const float* value_cache, int head_size, int seq_len, int loff);
/* The same kernel and arguments for callers that may not allocate.  No counterpart in the reference: added
 * for code under CUDA graph capture, tight serving loops and harnesses that forbid
 * cudaMalloc.  The call allocates nothing (no cudaMalloc / cudaFree, no allocating library routine): the split
 * scratch is a file-scope device array of 787,456 bytes (32768 partial slots of 24 bytes + 256 arrival counters)
 * that every GPU loading the library holds until exit.  A launch whose time split needs more
 * (n_batches * segments * splits * 32 J > 32768 slots, or n_batches * segments > 256 with a split) returns
 * THA_STATUS_INVALID_ARG and tha_last_error() names the sizes; use tha_mha_fused_v1_f32 then.  All calls on a device
 * share the array: the counters a launch uses are cleared on its stream first, and calls that overlap on the GPU
 * (two streams) race -- serialise them.  Results are bit-identical to tha_mha_fused_v1_f32. */
int tha_mha_fused_v1_f32_static(void* stream, int n_batches, const int* pos_d, int n_heads, int n_kv_heads,
                                int n_layers, const float* q, float* xb, const float* key_cache,
                                const float* value_cache, int head_size, int seq_len, int loff);

/* ---- Q8_0 (runq.c) ---------------------------------------------------------------- */
/* quantize, runq.c:145-171, bit-exact: n = rows*cols values, groups of `gs` */
int tha_quantize_q8(void* stream, int8_t* q, float* s, const float* x, int n, int gs);
/* matmul, runq.c:317-342 for n_batches activation rows:
 *   out[b*out_stride + i] = sum_g (float)(sum_k xq.wq) * ws[i,g] * xs[b,g] */
int tha_matmul_batch_q8(void* stream, int n_batches, float* out, int out_stride, const int8_t* xq,
                        const float* xs, const int8_t* wq, const float* ws, int K, int M, int gs);

/* ---- the forward pass --------------------------------------------------------------- */
/* thaDNN_s_forward_batch, src/thaDNN.cpp:13-81: one decode step for n_batches slots,
 * slot b at its own pos_host[b] (pos < 0: idle slot, skipped).  token_host / pos_host
 * are HOST arrays; logits_out receives [n_batches, vocab] fp32 and may be pinned host
 * memory (as in the reference, src/llama.cpp:935-936) or device memory.  Synchronises
 * `stream` before returning, like the reference's hipDeviceSynchronize (:78). */
int tha_forward_batch_f32(void* stream, int n_batches, const tha_config_t* cfg,
                          const tha_weights_f32_t* w, const tha_state_t* s, const int* token_host,
                          const int* pos_host, float* logits_out);
/* Same step with everything resident: token_d / pos_d are DEVICE arrays, nothing is
 * copied and nothing synchronises.  logits_d (nullable) gets [B, vocab]; next_d
 * (nullable) gets the greedy argmax per slot (first max wins, src/llama.cpp:275-286). */
int tha_decode_step_f32(void* stream, int n_batches, const tha_config_t* cfg,
                        const tha_weights_f32_t* w, const tha_state_t* s, const int* token_d,
                        const int* pos_d, float* logits_d, int* next_d);
/* The greedy loop of the batching driver kept on the device: the step above with
 * next_d, then token_d[b] = next_d[b] and pos_d[b] += 1 (src/llama.cpp:1027-1050); a slot
 * that reaches seq_len goes idle (pos = -1).  One kernel launch per step for models the
 * persistent step kernel covers (B <= 8, dim <= 1024, hidden_dim <= 2048). */
int tha_decode_greedy_f32(void* stream, int n_batches, const tha_config_t* cfg,
                          const tha_weights_f32_t* w, const tha_state_t* s, int* token_d,
                          int* pos_d, float* logits_d, int* next_d);
/* Q8_0 counterparts (runq.c:344-481 semantics; fp32 KV cache, state as above) */
int tha_forward_batch_q8(void* stream, int n_batches, const tha_config_t* cfg,
                         const tha_weights_q8_t* w, const tha_state_t* s, const int* token_host,
                         const int* pos_host, float* logits_out);
int tha_decode_step_q8(void* stream, int n_batches, const tha_config_t* cfg,
                       const tha_weights_q8_t* w, const tha_state_t* s, const int* token_d,
                       const int* pos_d, float* logits_d, int* next_d);
/* thaDNN_s_forward_batch + sample_argmax (src/llama.cpp:275-286) in one call: HOST token / pos
 * in, HOST next[] out (first maximum per active slot; idle slots are left untouched on the
 * device and read back as whatever they held).  B ints cross PCIe instead of B*vocab floats. */
int tha_forward_greedy_f32(void* stream, int n_batches, const tha_config_t* cfg, const tha_weights_f32_t* w,
                           const tha_state_t* s, const int* token_host, const int* pos_host, int* next_host);
int tha_forward_greedy_q8(void* stream, int n_batches, const tha_config_t* cfg, const tha_weights_q8_t* w,
                          const tha_state_t* s, const int* token_host, const int* pos_host, int* next_host);
/* The driver's sampler on the device (replaces sample() / sample_topp() / sample_mult() / random_f32() of
 * src/llama.cpp:262-411 for B slots at once).  logits_d [B, vocab] is overwritten with the
 * probabilities (as the reference's sample() does).  rng_d [B]: xorshift* state of the request in each slot, advanced once
 * per SAMPLED token; reseed_d (nullable) [B]: non-zero = start the slot's stream from this seed first; force_d (nullable)
 * [B]: >= 0 = the next token is this prompt token, no coin is drawn (src/llama.cpp:1029-1030); pos_d [B]: < 0 = idle.
 * temperature 0 = greedy (first maximum). */
int tha_sample(void* stream, int n_batches, float* logits_d, int vocab, float temperature, float topp,
               unsigned long long* rng_d, const unsigned long long* reseed_d, const int* force_d, const int* pos_d, int* next_d);
/* thaDNN_s_forward_batch + sample() per slot in one call (src/llama.cpp:1025-1050): HOST token / pos / forced prompt
 * token (-1 = sample) / sampler reseed (0 = keep the slot's stream) in, HOST next token per slot out.  The sampler
 * state of every slot lives on the device between calls. */
int tha_forward_sample_f32(void* stream, int n_batches, const tha_config_t* cfg, const tha_weights_f32_t* w,
                           const tha_state_t* s, const int* token_host, const int* pos_host, const int* force_host,
                           const unsigned long long* reseed_host, float temperature, float topp, int* next_host);
int tha_forward_sample_q8(void* stream, int n_batches, const tha_config_t* cfg, const tha_weights_q8_t* w,
                          const tha_state_t* s, const int* token_host, const int* pos_host, const int* force_host,
                          const unsigned long long* reseed_host, float temperature, float topp, int* next_host);
/* bf16 model path: same calls as the fp32 ones (thaDNN_s_forward_batch semantics, state as tha_state_t) */
int tha_convert_bf16(void* stream, const float* src_d, uint16_t* dst_d, size_t n);   /* fp32 -> bf16, round to nearest even */
int tha_forward_batch_bf16(void* stream, int n_batches, const tha_config_t* cfg, const tha_weights_bf16_t* w,
                           const tha_state_t* s, const int* token_host, const int* pos_host, float* logits_out);
int tha_decode_step_bf16(void* stream, int n_batches, const tha_config_t* cfg, const tha_weights_bf16_t* w,
                         const tha_state_t* s, const int* token_d, const int* pos_d, float* logits_d, int* next_d);
int tha_forward_greedy_bf16(void* stream, int n_batches, const tha_config_t* cfg, const tha_weights_bf16_t* w,
                            const tha_state_t* s, const int* token_host, const int* pos_host, int* next_host);
int tha_forward_sample_bf16(void* stream, int n_batches, const tha_config_t* cfg, const tha_weights_bf16_t* w,
                            const tha_state_t* s, const int* token_host, const int* pos_host, const int* force_host,
                            const unsigned long long* reseed_host, float temperature, float topp, int* next_host);
int tha_prefill_bf16(void* stream, const tha_config_t* cfg, const tha_weights_bf16_t* w, const tha_state_t* s, int n_seqs,
                     const int* slot_host, const int* pos0_host, const int* len_host, const int* tokens_host,
                     float* logits_d, int* next_d);
/* Paged KV cache (fills in include/BatchManager.hpp:11-52 of the reference, whose alloc_new_block /
 * convert_block_ids_to_block_addrs / push_kv_data_to_kv_cache are stubs).  key_pool / value_pool: device arrays
 * [n_blocks][n_layers][block_size][kv_dim] fp32 owned by the caller; block_table_d: device array [slots][table_stride] of
 * block ids, row s listing in order the blocks that hold slot s's positions (block_size, a power of two, positions each).
 * After registration a tha_state_t whose key_cache == key_pool and value_cache == value_pool is PAGED: the attention,
 * prefill-attention and RoPE/scatter kernels find every cache row through the table (read at run time: update it,
 * stream-ordered, before the step that first touches a new block).  Paged states run the tensor-core row pipeline at
 * every batch size (fp32, Q8_0, bf16; decode and prefill). */
int tha_kv_pool_register(float* key_pool, float* value_pool, const int* block_table_d, int table_stride, int block_size,
                         int n_blocks);
int tha_kv_pool_unregister(float* key_pool);
/* greedy sampler on the device: next[b] = first argmax of logits[b, :] */
int tha_argmax_batch_f32(void* stream, int n_batches, const float* logits, int vocab, int* next_d);
/* greedy bookkeeping of the batching driver (src/llama.cpp:1027-1050) kept on the device:
 * token_d[b] = next_d[b], pos_d[b] += 1 (a slot reaching seq_len goes idle: pos = -1) */
int tha_advance_slots(void* stream, int n_batches, int* token_d, int* pos_d, const int* next_d, int seq_len);
/* 1 when the step for this shape falls back to the three attention stages and therefore needs
 * tha_state_t.att ([B, n_heads, seq_len]); 0 when the fused attention kernel covers it */
/* Batched prompt ingestion ("prefill").  The reference feeds a prompt one forced token per
 * decode step (src/llama.cpp:1029-1030); this runs all prompt positions of n_seqs sequences through the tcgen05
 * matmuls in one pass.  HOST arrays: sequence i lives in cache slab slot_host[i]; its len_host[i] tokens (concatenated
 * in tokens_host, sequence after sequence) go to positions pos0_host[i] .. pos0_host[i]+len_host[i]-1.  Every row's K/V
 * is written to the cache exactly where the decode step would have put it.  logits_d (device, [n_seqs, vocab], nullable):
 * logits of the LAST position of each sequence; next_d (device, [n_seqs], nullable): their argmax.  With logits_d == NULL
 * the call only fills the cache.  Asynchronous on `stream`; the host arrays may be reused when it returns. */
int tha_prefill_f32(void* stream, const tha_config_t* cfg, const tha_weights_f32_t* w, const tha_state_t* s, int n_seqs,
                    const int* slot_host, const int* pos0_host, const int* len_host, const int* tokens_host,
                    float* logits_d, int* next_d);
/* Same for a Q8_0 model (runq.c semantics: every matmul input quantised per group first). */
int tha_prefill_q8(void* stream, const tha_config_t* cfg, const tha_weights_q8_t* w, const tha_state_t* s, int n_seqs,
                   const int* slot_host, const int* pos0_host, const int* len_host, const int* tokens_host,
                   float* logits_d, int* next_d);
int tha_prefill_supported(const tha_config_t* cfg); /* 1 when tha_prefill_f32 handles this architecture */

/* ---- tensor parallelism (Llama-2-7B split 2/4/8 ways over NVLink; SURVEY.md 8e "Optional TP") ------------------
 * One tha_tp_t per rank (= per GPU; one process per GPU, or one host thread per GPU in one process as the reference's
 * driver runs, src/llama.cpp:913).  wq/wk/wv/w1/w3 are split by output rows (heads, hidden units), wo/w2 by input
 * columns (the reference has no TP; its closest relative is the pipeline split of src/thaDNN.cpp:83-427).  The
 * all-reduce after wo and w2 is FUSED into its consumer: each rank's matmul leaves its partial sums in an exchange
 * buffer in its own HBM, mapped by its peers (cudaIpc across processes, peer access inside one), and the residual-add
 * + RMSNorm kernel of every rank reads all ranks' partials over NVLink in a fixed order after a flag handshake -- no
 * separate collective kernel, bit-identical residual stream on every rank.
 *   create:  hidden_of_rank[tp] = every rank's local hidden_dim (a multiple of the Q8_0 group size; may differ between
 *            ranks: 11008 / 8 is not a multiple of 64); xchg_floats = capacity of each of the two exchange buffers,
 *            >= max over ranks of (split-K factor <= 16) * rows * dim.
 *   connect: across processes, gather every rank's 64-byte handle (tha_tp_export) and pass the [tp][64] array to
 *            tha_tp_connect_ipc; inside one process pass the tp block pointers (tha_tp_local_ptr) to tha_tp_connect_ptr.
 *   step:    cfg_local has n_heads / n_kv_heads / hidden_dim of THIS rank and the full dim; weights / state are the
 *            local shards (wq [L,da,dim], wk,wv [L,kvd,dim], wo [L,dim,da], w1,w3 [L,hd,dim], w2 [L,dim,hd]; embedding,
 *            norms, classifier replicated; KV cache [B][L][S][kvd] of the local heads; q, xb >= [B, da]).  Every rank
 *            makes the same sequence of calls; advance != 0 adds the greedy slot bookkeeping (tha_advance_slots).
 *   A peer that never shows up ends in tha_tp_check_error() == 1 after ~20 s, not in a hang. */
typedef struct tha_tp tha_tp_t;
int tha_tp_create(tha_tp_t** out, int tp, int rank, int head_size, const int* hidden_of_rank, size_t xchg_floats);
int tha_tp_export(tha_tp_t* t, void* handle64_out);
int tha_tp_local_ptr(tha_tp_t* t, void** base);
int tha_tp_connect_ipc(tha_tp_t* t, const void* handles);
int tha_tp_connect_ptr(tha_tp_t* t, void* const* bases);
int tha_tp_check_error(tha_tp_t* t);
int tha_tp_set_mode(tha_tp_t* t, int mode); /* 1: the next step / prefill call only prepares (scratch, graph, module loads) */
int tha_tp_select_device(tha_tp_t* t);   /* cudaSetDevice(the rank's device) */
int tha_tp_destroy(tha_tp_t* t);
int tha_tp_decode_step_f32(void* stream, tha_tp_t* tp, int n_batches, const tha_config_t* cfg_local,
                           const tha_weights_f32_t* w_local, const tha_state_t* s_local, int* token_d, int* pos_d,
                           float* logits_d, int* next_d, int advance);
int tha_tp_decode_step_q8(void* stream, tha_tp_t* tp, int n_batches, const tha_config_t* cfg_local,
                          const tha_weights_q8_t* w_local, const tha_state_t* s_local, int* token_d, int* pos_d,
                          float* logits_d, int* next_d, int advance);
/* One greedy step of a replica whose `tp` ranks are all driven by the calling host thread (ranks may share a device):
 * HOST token / pos in, HOST next token per slot out.  streams[r], cfg_local[r], w_local[r] (tha_weights_f32_t* or
 * tha_weights_q8_t* by `q8`), s_local[r] belong to rank r.  Prepares every rank before the first launch. */
int tha_tp_forward_greedy(int tp, tha_tp_t* const* ranks, void* const* streams, int n_batches, int q8,
                          const tha_config_t* const* cfg_local, const void* const* w_local,
                          const tha_state_t* const* s_local, const int* token_host, const int* pos_host, int* next_host);
int tha_tp_prefill_f32(void* stream, tha_tp_t* tp, const tha_config_t* cfg_local, const tha_weights_f32_t* w_local,
                       const tha_state_t* s_local, int n_seqs, const int* slot_host, const int* pos0_host,
                       const int* len_host, const int* tokens_host, float* logits_d, int* next_d);
int tha_tp_prefill_q8(void* stream, tha_tp_t* tp, const tha_config_t* cfg_local, const tha_weights_q8_t* w_local,
                      const tha_state_t* s_local, int n_seqs, const int* slot_host, const int* pos0_host,
                      const int* len_host, const int* tokens_host, float* logits_d, int* next_d);

int tha_trace_dump(int skip);   /* development aid: THA_TRACE=1 + graphs off -> live per-kernel times (event to event) */
int tha_step_needs_att_scratch(const tha_config_t* cfg);
/* number of kernels the last tha_decode_step_* call on this thread launched */
int tha_last_step_launches(void);

#ifdef __cplusplus
}
#endif
#endif /* THA_B200_H */
