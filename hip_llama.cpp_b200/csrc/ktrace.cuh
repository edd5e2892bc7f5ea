// Development aid: a timeline of the kernels of a step, taken on the device (%globaltimer), that works under CUDA-graph
// replay and PDL where host-side events do not.  tha_ktrace(nullptr, 0, 1) arms it; every traced kernel then appends one
// record for its first and one for its last block: [0] block entered, [1] griddepcontrol.wait returned, [2..5], [7..14] kernel
// specific, [6] block done, [15] kind | flags.  tha_ktrace(out, max, 0) copies the records out and disarms.
// Off (pointer null): one predictable load per stamp by one thread.  tools/step_timeline.py prints it.
#pragma once
#include <cuda_runtime.h>

namespace tha {

struct KtSlots { unsigned long long* buf; unsigned int* ctr; };
constexpr int KT_STAMPS = 16, KT_MAX = 8192;
void kt_register(int (*setter)(const KtSlots&));

static __device__ KtSlots d_kt;   // this translation unit's copy of the pointers (no relocatable device code in this library)
static int kt_set_tu(const KtSlots& s) { return cudaMemcpyToSymbol(d_kt, &s, sizeof(s)) == cudaSuccess ? 0 : 5; }
namespace { struct KtReg { KtReg() { kt_register(&kt_set_tu); } }; static KtReg kt_reg_instance; }

enum { KT_GEMM = 1, KT_PREP = 2, KT_ROPE = 3, KT_SWIGLU = 4, KT_ATTN = 5, KT_ARGMAX = 6, KT_GEMM_Q8 = 7, KT_OTHER = 8 };

__device__ __forceinline__ unsigned long long kt_now() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
__device__ __forceinline__ bool kt_first_block() { return (blockIdx.x | blockIdx.y | blockIdx.z) == 0; }
__device__ __forceinline__ bool kt_last_block() {
  return blockIdx.x == gridDim.x - 1 && blockIdx.y == gridDim.y - 1 && blockIdx.z == gridDim.z - 1;
}
// ONE thread of the first / last block calls this on entry; -1 = not traced
__device__ __forceinline__ int kt_begin(unsigned long long kind) {
  if (!d_kt.buf || !(kt_first_block() || kt_last_block())) return -1;
  const unsigned id = atomicAdd(d_kt.ctr, 1u);
  if (id >= (unsigned)KT_MAX) return -1;
  d_kt.buf[(size_t)id * KT_STAMPS + 15] = kind | (kt_first_block() ? 0ull : 0x100ull);
  d_kt.buf[(size_t)id * KT_STAMPS] = kt_now();
  return (int)id;
}
__device__ __forceinline__ void kt_stamp(int id, int k) {
  if (id >= 0) d_kt.buf[(size_t)id * KT_STAMPS + k] = kt_now();
}

}  // namespace tha
