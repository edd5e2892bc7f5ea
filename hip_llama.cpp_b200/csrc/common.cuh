// Shared device/host helpers for the sm_100a kernels.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "tha_b200.h"

#define THA_WARP 32
#define THA_FULL_MASK 0xffffffffu

namespace tha {

// ---- error plumbing --------------------------------------------------------------
void set_last_error(const char* where, cudaError_t e);
int pdl_enabled();     // programmatic dependent launch between the step's kernels
int graph_enabled();   // capture the decode step into a CUDA graph

#define THA_CUDA_OK(expr)                                  \
  do {                                                     \
    cudaError_t _e = (expr);                               \
    if (_e != cudaSuccess) {                               \
      tha::set_last_error(#expr, _e);                      \
      return THA_STATUS_EXECUTION_FAILED;                  \
    }                                                      \
  } while (0)

#define THA_LAUNCH_OK()                                    \
  do {                                                     \
    cudaError_t _e = cudaGetLastError();                   \
    if (_e != cudaSuccess) {                               \
      tha::set_last_error("kernel launch", _e);            \
      return THA_STATUS_EXECUTION_FAILED;                  \
    }                                                      \
  } while (0)

// Thread-local count of kernels launched (reported by tha_last_step_launches()).
void count_launch(int n = 1);
// Development aid (THA_TRACE=1, graphs off): an event after every launch; tha_trace_dump() prints the time between
// consecutive events per kernel (= its live duration including the gap before it).
void trace_launch(const void* fn, cudaStream_t st);
// THA_CARVEOUT=1 (experiment, off by default): every kernel of a step asks for the maximum shared-memory carve-out, so an SM
// that alternates between the ~200 KB tensor-core matmuls and small-footprint glue kernels never reconfigures its L1 / shared
// split.  Measured slower (0.883 -> 0.928 ms/step at 256 slots): the glue kernels lose their L1.
void prefer_max_smem(const void* fn);

// Launch with optional programmatic-dependent-launch attribute.  Kernels launched
// this way call pdl_wait() before touching anything a previous kernel wrote and
// pdl_trigger() as early as they like.
template <typename... KArgs, typename... Args>
inline cudaError_t launch(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem,
                          cudaStream_t stream, bool pdl, Args... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl ? 1 : 0;
  count_launch();
  prefer_max_smem(reinterpret_cast<const void*>(kernel));
  const cudaError_t e = cudaLaunchKernelEx(&cfg, kernel, KArgs(args)...);
  trace_launch(reinterpret_cast<const void*>(kernel), stream);
  return e;
}

// ---- paged KV cache (include/BatchManager.hpp) ---------------------------------------------
// The cache rows of a slot either sit in the dense slab [slot][layer][seq_len][kv_dim] (table == nullptr; the layout of
// src/models.cpp:176-177) or in BLOCKS of (1 << bs_shift) positions handed out by a BatchManager: pool
// [block][layer][position in block][kv_dim], one table row of block ids per slot.
struct KvPaged {
  const int* table;
  int stride;               // table row length (blocks per slot)
  int bs_shift, bs_mask;
  long long blk_stride;     // floats per block: n_layers * block_size * kv_dim
  long long layer_off;      // floats: layer * block_size * kv_dim
};
#if defined(__CUDACC__)
// offset of cache row (slot, t) from the pool base, layer included
__device__ __forceinline__ long long kv_paged_off(const KvPaged& g, int slot, int t, int kvd) {
  const int blk = __ldg(g.table + (size_t)slot * g.stride + (t >> g.bs_shift));
  return (long long)blk * g.blk_stride + g.layer_off + (long long)(t & g.bs_mask) * kvd;
}
#endif

// ---- device helpers ------------------------------------------------------------------
#ifdef __CUDACC__

// Wait until the kernels this launch depends on have completed and their writes are
// visible (no-op when the launch carried no programmatic attribute).
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
// Allow the next kernel in the stream to start launching.
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;"); }

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(THA_FULL_MASK, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(THA_FULL_MASK, v, o));
  return v;
}

// Block-wide sum for blockDim.x <= 1024 (multiple of 32); result valid in all threads.
// `red` is a 33-float shared scratch.
__device__ __forceinline__ float block_sum(float v, float* red) {
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
  v = warp_sum(v);
  __syncthreads();  // protect `red` against a previous use
  if (lane == 0) red[wid] = v;
  __syncthreads();
  if (wid == 0) {
    float t = lane < nw ? red[lane] : 0.0f;
    t = warp_sum(t);
    if (lane == 0) red[32] = t;
  }
  __syncthreads();
  return red[32];
}
__device__ __forceinline__ float block_max(float v, float* red) {
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
  v = warp_max(v);
  __syncthreads();
  if (lane == 0) red[wid] = v;
  __syncthreads();
  if (wid == 0) {
    float t = lane < nw ? red[lane] : -INFINITY;
    t = warp_max(t);
    if (lane == 0) red[32] = t;
  }
  __syncthreads();
  return red[32];
}

// Streaming 128-bit load of read-once data (weights): read-only path, do not
// allocate in L1.
__device__ __forceinline__ float4 ldg_stream(const float4* p) {
  float4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w)
               : "l"(p));
  return r;
}
__device__ __forceinline__ int4 ldg_stream_i4(const int4* p) {
  int4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.s32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}

__device__ __forceinline__ float dot4(const float4& a, const float4& b, float acc) {
  acc = fmaf(a.x, b.x, acc);
  acc = fmaf(a.y, b.y, acc);
  acc = fmaf(a.z, b.z, acc);
  acc = fmaf(a.w, b.w, acc);
  return acc;
}

// The runq.c:145-171 quantiser on float4 lanes (scale = max|x| / 127 per group, round half away from zero; bit-exact).
// Four consecutive values per thread, a group = gs/4 consecutive lanes (8, 16 or 32): the group maximum is a
// segmented shuffle reduction, the four int8 results leave as one 32-bit store.  v = the thread's values (index i4 in
// units of float4 within the row).  scale -> xsT[g * ldxs + row] (transposed: the matmul's epilogue wants the NT scales
// of a group side by side).
// Called by whole warps (lanes past the row pass valid = false and zeros).
__device__ __forceinline__ void q8_quant4(float4 v, int gs, int i4, int8_t* __restrict__ xq_row, float* __restrict__ xsT,
                                          int ldxs, int row, bool valid) {
  float amax = fmaxf(fmaxf(fabsf(v.x), fabsf(v.y)), fmaxf(fabsf(v.z), fabsf(v.w)));
  for (int o = gs >> 3; o > 0; o >>= 1) amax = fmaxf(amax, __shfl_xor_sync(THA_FULL_MASK, amax, o));
  const float scale = __fdiv_rn(amax, 127.0f);
  if (!valid) return;
  const int g = (4 * i4) / gs;
  if (((4 * i4) % gs) == 0) xsT[(size_t)g * ldxs + row] = scale;
  const float in[4] = {v.x, v.y, v.z, v.w};
  unsigned int packed = 0;
#pragma unroll
  for (int e = 0; e < 4; e++) {
    const float r = __fdiv_rn(in[e], scale);
    const int q = (r != r) ? 0 : __float2int_rn(roundf(r));
    packed |= ((unsigned int)q & 0xFFu) << (8 * e);
  }
  *reinterpret_cast<unsigned int*>(xq_row + 4 * i4) = packed;
}

#endif  // __CUDACC__

inline int ceil_div(int a, int b) { return (a + b - 1) / b; }
inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

}  // namespace tha
