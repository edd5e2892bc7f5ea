// thaDNN.hpp -- the forward pass of the reference (include/thaDNN.hpp:1-79, src/thaDNN.cpp:13-81).
#pragma once
#include "thaBLAS.hpp"
#include "utils.hpp"
#include "thaDNN/thaDNN_rmsnorm.hpp"
#include "thaDNN/thaDNN_rope.hpp"
#include "thaDNN/thaDNN_mha.hpp"
#include "thaDNN/thaDNN_softmax.hpp"
#include "thaDNN/thaDNN_swiglu.hpp"

// include/thaDNN.hpp:15-28 (never returned by any function of the reference; kept for source compatibility)
typedef enum {
  thaDNNStatusSuccess = 0,
  thaDNNStatusNotInitialized = 1,
  thaDNNStatusInvalidValue = 2,
  thaDNNStatusBadParm = 3,
  thaDNNStatusAllocFailed = 4,
  thaDNNStatusInternalError = 5,
  thaDNNStatusNotImplemented = 6,
  thaDNNStatusUnknownError = 7,
  thaDNNStatusUnsupportedOp = 8,
  thaDNNStatusGpuOperationsSkipped = 9,
  thaDNNStatusVersionMismatch = 10,
} thaDNNStatus_t;

// One decode step for n_batches independent slots, slot b at its own pos[b]
// (src/thaDNN.cpp:13-81).  token[] / pos[] are HOST arrays; logits_host receives
// [n_batches, vocab_size] fp32 and is pinned host memory allocated by the caller
// (src/llama.cpp:935-936).  Only handle1 is used, as in the reference (:40).  Returns after
// the logits have landed (the reference ends with hipDeviceSynchronize, :78).
// Differences, all on the safe side: a slot with pos[b] < 0 is idle and skipped (the reference
// computes garbage rows for idle slots); nothing is allocated per call.
thablasStatus_t thaDNN_s_forward_batch(thablasHandle_t handle1, thablasHandle_t handle2, thablasHandle_t handle3,
                                       int n_batches, Config* p, TransformerWeights* w, RunState* s_batch,
                                       int token[], int pos[], float* logits_host);

// Same step for a Q8_0 model (runq.c:344-481 semantics, which the reference runs on the CPU only).
thablasStatus_t thaDNN_q8_forward_batch(thablasHandle_t handle, int n_batches, Config* p, TransformerWeightsQ8* w,
                                        RunState* s_batch, int token[], int pos[], float* logits_host);

// Batched prompt ingestion (additive).  HOST arrays: sequence i lives in cache slab slot[i];
// its len[i] tokens (concatenated in tokens[]) go to positions pos0[i]..pos0[i]+len[i]-1 in one pass of tensor-core
// matmuls, leaving the cache as len[i] forced-prompt steps of thaDNN_s_forward_batch would (src/llama.cpp:1029-1030).
// logits_d / next_d (DEVICE, nullable): last-position logits [n_seqs, vocab] and their argmax.  Asynchronous.
thablasStatus_t thaDNN_s_prefill_batch(thablasHandle_t* handle, Config* p, TransformerWeights* w, RunState* s_batch,
                                       int n_seqs, int slot[], int pos0[], int len[], int tokens[], float* logits_d,
                                       int* next_d);

// Resident greedy step of the batching driver: token_d / pos_d are DEVICE arrays; the step
// picks the first-maximum token per slot on the device (sample_argmax, src/llama.cpp:275-286),
// stores it in next_d, and (advance != 0) moves every active slot on: token = next, pos += 1.
// Nothing is copied and nothing synchronises.  logits_d may be NULL.
thablasStatus_t thaDNN_s_decode_greedy_batch(thablasHandle_t* handle, int n_batches, Config* p, TransformerWeights* w,
                                             RunState* s_batch, int* token_d, int* pos_d, int* next_d,
                                             float* logits_d, int advance);

// Tensor-parallel step (additive; SURVEY.md 8e "Optional TP"): ONE replica of a model split over the GPUs of a box,
// one rank per GPU (one host thread per device as in src/llama.cpp:913, or one process per device).  p / w / s_batch
// are the rank's LOCAL config, weight shards and state (include/tha_b200.h, "tensor parallelism"); the all-reduce after
// the wo and w2 matmuls is fused into the kernel that consumes it, over NVLink peer memory.  Every rank makes the same
// calls.  token_d / pos_d / next_d / logits_d are DEVICE arrays (next_d, logits_d nullable); advance != 0 moves every
// active slot on (token = next, pos += 1).  Asynchronous on handle->calc_stream.
struct tha_tp;
typedef struct tha_tp* thaTP_t;
thablasStatus_t thaTPCreate(thaTP_t* tp, int tp_size, int rank, int head_size, const int hidden_of_rank[], size_t xchg_floats);
thablasStatus_t thaTPConnectLocal(thaTP_t ranks[], int tp_size);   // all ranks in this process: enables peer access, wires pointers
thablasStatus_t thaTPDestroy(thaTP_t tp);
thablasStatus_t thaDNN_s_forward_batch_tp(thablasHandle_t* handle, thaTP_t tp, int n_batches, Config* p_local,
                                          TransformerWeights* w_local, RunState* s_local, int* token_d, int* pos_d,
                                          float* logits_d, int* next_d, int advance);
thablasStatus_t thaDNN_q8_forward_batch_tp(thablasHandle_t* handle, thaTP_t tp, int n_batches, Config* p_local,
                                           TransformerWeightsQ8* w_local, RunState* s_local, int* token_d, int* pos_d,
                                           float* logits_d, int* next_d, int advance);
