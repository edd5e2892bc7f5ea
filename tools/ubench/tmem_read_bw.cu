// TMEM read bandwidth per SM: W warps (W / 4 per 32-lane quarter) each issue tcgen05.ld.32x32b.xN over 512 columns, R rounds.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tmem_read_bw tmem_read_bw.cu && ./tmem_read_bw
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

template <int N> __device__ __forceinline__ void ld(uint32_t a, uint32_t* v);
template <> __device__ __forceinline__ void ld<16>(uint32_t a, uint32_t* v) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
               : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]),
                 "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]) : "r"(a));
}
template <> __device__ __forceinline__ void ld<32>(uint32_t a, uint32_t* v) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
               : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]),
                 "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]),
                 "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]),
                 "=r"(v[30]), "=r"(v[31]) : "r"(a));
}

template <int N>
__global__ void k(long long* out, int rounds, int wait_every) {
  __shared__ uint32_t slot;
  const int warp = threadIdx.x >> 5;
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"((uint32_t)__cvta_generic_to_shared(&slot)), "r"(512));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::);
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t base = slot + ((uint32_t)((warp & 3) * 32) << 16);
  uint32_t v[N];
  uint32_t acc = 0;
  __syncthreads();
  const long long t0 = clock64();
  for (int r = 0; r < rounds; r++) {
    for (int c = 0; c < 512; c += N) {
      ld<N>(base + c, v);
      if (((c / N) + 1) % wait_every == 0) {
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
        for (int i = 0; i < N; i++) acc += v[i];
      }
    }
  }
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
  __syncthreads();
  const long long t1 = clock64();
  if (threadIdx.x == 0) out[0] = t1 - t0;
  if (acc == 0x12345678u) out[1] = acc;
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(slot), "r"(512));
}

int main() {
  long long* d; cudaMalloc(&d, 16);
  const int rounds = 64;
  for (int warps : {4, 8, 16}) {
    for (int n : {16, 32}) {
      for (int we : {1, 4}) {
        if (n == 16) k<16><<<1, warps * 32>>>(d, rounds, we); else k<32><<<1, warps * 32>>>(d, rounds, we);
        cudaDeviceSynchronize();
        long long h = 0; cudaMemcpy(&h, d, 8, cudaMemcpyDeviceToHost);
        const double bytes = (double)warps * 32 * 512 * 4 * rounds;
        printf("warps %2d  x%-2d  wait every %d loads: %lld cycles, %.1f B/clk/SM (%s)\n", warps, n, we, h, bytes / h, cudaGetErrorString(cudaGetLastError()));
      }
    }
  }
  return 0;
}
