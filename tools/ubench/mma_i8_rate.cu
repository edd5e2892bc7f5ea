// Rate of tcgen05.mma.cta_group::1.kind::i8 (M = 128, N = NT, K = 32 per instruction), both operands from shared memory,
// back-to-back from one thread, alternating between two accumulators; also kind::f16 (K = 16) for comparison.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I../../hip_llama.cpp_b200/csrc -I../../include -o mma_i8_rate mma_i8_rate.cu -lcuda
#include <cstdio>
#include <cstdint>
#include <cuda.h>
#include <cuda_runtime.h>
#include "tc_common.cuh"
using namespace tha;

__host__ __device__ constexpr uint32_t idesc_i8(int nt) { return (2u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(nt >> 3) << 17) | ((uint32_t)(128 >> 4) << 24); }
__host__ __device__ constexpr uint32_t idesc_bf16(int nt) { return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(nt >> 3) << 17) | ((uint32_t)(128 >> 4) << 24); }

template <int KIND>
__device__ __forceinline__ void mma(uint32_t d, uint64_t a, uint64_t b, uint32_t idesc, uint32_t acc) {
  if (KIND == 0)
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t}" ::"r"(d), "l"(a), "l"(b), "r"(idesc), "r"(acc) : "memory");
  else
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d), "l"(a), "l"(b), "r"(idesc), "r"(acc) : "memory");
}

__device__ __forceinline__ void mma_ts_bf16(uint32_t d, uint32_t a_tmem, uint64_t b, uint32_t idesc, uint32_t acc) {
  asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}" ::"r"(d), "r"(a_tmem), "l"(b), "r"(idesc), "r"(acc) : "memory");
}
// variant: 0 = two accumulators (as above), 1 = eight accumulators, 2 = never accumulate, 3 = A from TMEM (bf16), 4 = two issuing warps
template <int KIND>
__global__ void kv(long long* out, int nt, int n_mma, int variant) {
  extern __shared__ unsigned char raw[];
  unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(raw) + 1023) & ~(uintptr_t)1023);
  __shared__ uint64_t bar[2];
  __shared__ uint32_t slot;
  __shared__ int abortf;
  const int warp = threadIdx.x >> 5;
  for (int i = threadIdx.x; i < 48 * 1024 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0x01010101u;
  if (threadIdx.x == 0) { tc_mbar_init(&bar[0], 1); tc_mbar_init(&bar[1], 1); abortf = 0; asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tc_smem_u32(&slot)), "r"(512));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::);
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const int nissue = variant == 4 ? 2 : 1;
  if ((threadIdx.x & 31) == 0 && warp < nissue) {
    const uint64_t a = tc_smem_desc(tc_smem_u32(smem)), b = tc_smem_desc(tc_smem_u32(smem + 16384));
    const uint32_t idesc = KIND == 0 ? idesc_i8(nt) : idesc_bf16(nt);
    const int nacc = variant == 1 ? 8 : 2;
    const int ntc = nt > 32 ? 32 : nt;     // accumulator column stride so that 8 fit beside each other for small nt
    const int dstride = variant == 1 ? ntc : (nt > 128 ? 0 : nt);
    const long long t0 = clock64();
#pragma unroll 8
    for (int i = 0; i < n_mma; i++) {
      const uint32_t d = slot + (uint32_t)(256 * warp + (i & (nacc - 1)) * dstride);
      const uint32_t accf = variant == 2 ? 0u : (uint32_t)(i >= nacc);
      if (variant == 3) mma_ts_bf16(d, slot + 448, b + (uint64_t)((i & 3) * 2), idesc, accf);
      else mma<KIND>(d, a + (uint64_t)((i & 3) * 2), b + (uint64_t)((i & 3) * 2), idesc, accf);
    }
    tc_commit(&bar[warp]);
    tc_mbar_wait(&bar[warp], 0, &abortf);
    const long long t1 = clock64();
    out[warp] = t1 - t0;
  }
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(slot), "r"(512));
}

template <int KIND>
__global__ void k(long long* out, int nt, int n_mma, int per_commit) {
  extern __shared__ unsigned char raw[];
  unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(raw) + 1023) & ~(uintptr_t)1023);
  __shared__ uint64_t bar;
  __shared__ uint32_t slot;
  __shared__ int abortf;
  const int warp = threadIdx.x >> 5;
  for (int i = threadIdx.x; i < 48 * 1024 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0x01010101u;
  if (threadIdx.x == 0) { tc_mbar_init(&bar, 1); abortf = 0; asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tc_smem_u32(&slot)), "r"(512));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::);
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  if (threadIdx.x == 0) {
    const uint64_t a = tc_smem_desc(tc_smem_u32(smem)), b = tc_smem_desc(tc_smem_u32(smem + 16384));
    const uint32_t idesc = KIND == 0 ? idesc_i8(nt) : idesc_bf16(nt);
    const long long t0 = clock64();
    unsigned ph = 0;
    for (int i = 0; i < n_mma; i++) {
      mma<KIND>(slot + (uint32_t)((i & 1) * nt), a + (uint64_t)((i & 3) * 2), b + (uint64_t)((i & 3) * 2), idesc, (uint32_t)(i >= 2));
      if ((i + 1) % per_commit == 0) { tc_commit(&bar); tc_mbar_wait(&bar, ph, &abortf); ph ^= 1u; }
    }
    const long long t1 = clock64();
    out[0] = t1 - t0;
  }
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(slot), "r"(512));
}

int main() {
  long long* d; cudaMalloc(&d, 16);
  cudaFuncSetAttribute(k<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
  cudaFuncSetAttribute(k<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
  const int n = 4096;
  for (int kind : {0, 1})
    for (int nt : {16, 32, 64, 128, 256})
      for (int pc : {4, 64, 4096}) {
        if (kind == 0) k<0><<<1, 128, 64 * 1024>>>(d, nt, n, pc); else k<1><<<1, 128, 64 * 1024>>>(d, nt, n, pc);
        cudaError_t e = cudaDeviceSynchronize();
        long long h = 0; cudaMemcpy(&h, d, 8, cudaMemcpyDeviceToHost);
        printf("%s N=%3d commit+wait every %4d: %7.1f cycles per MMA (%s)\n", kind == 0 ? "i8  K=32" : "bf16 K=16", nt, pc, (double)h / n, cudaGetErrorString(e));
      }
  cudaFuncSetAttribute(kv<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
  cudaFuncSetAttribute(kv<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
  const char* vn[5] = {"2 accumulators", "8 accumulators", "no accumulate", "A from TMEM", "two issuing warps"};
  for (int kind : {0, 1})
    for (int nt : {16, 64, 128})
      for (int v = 0; v < 5; v++) {
        if (kind == 0 && v == 3) continue;
        if (kind == 0) kv<0><<<1, 128, 64 * 1024>>>(d, nt, n, v); else kv<1><<<1, 128, 64 * 1024>>>(d, nt, n, v);
        cudaError_t e = cudaDeviceSynchronize();
        long long h[2] = {0, 0}; cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost);
        printf("%s N=%3d %-18s: %7.1f cycles per MMA per issuer (%s)\n", kind == 0 ? "i8  K=32" : "bf16 K=16", nt, vn[v], (double)h[0] / n, cudaGetErrorString(e));
      }
  return 0;
}
