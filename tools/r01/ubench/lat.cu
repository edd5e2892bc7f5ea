// Development aid: latency calibration on the B200 (L2 pointer chase, grid barrier, bar.sync, fence).
#include <cstdio>
#include <cuda_runtime.h>
#include <vector>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); return 1; } } while (0)

__global__ void chase(const unsigned* __restrict__ next, int iters, unsigned* out, long long* cyc, int cg) {
  unsigned i = threadIdx.x + blockIdx.x * 977;
  long long t0 = clock64();
  for (int k = 0; k < iters; k++) {
    unsigned v;
    if (cg) asm volatile("ld.global.cg.u32 %0, [%1];" : "=r"(v) : "l"(next + i));
    else asm volatile("ld.global.ca.u32 %0, [%1];" : "=r"(v) : "l"(next + i));
    i = v;
  }
  long long t1 = clock64();
  out[blockIdx.x] = i;
  cyc[blockIdx.x] = (t1 - t0) / iters;
}

__device__ __forceinline__ unsigned ld_acq(const unsigned* p) { unsigned v; asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory"); return v; }
__device__ __forceinline__ unsigned ld_rlx(const unsigned* p) { unsigned v; asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory"); return v; }

// mode 0: bar.sync + red.release + acquire poll + bar.sync; mode 1: + per-thread __threadfence (old); mode 2: relaxed red/poll (no fence)
__global__ void gridbar(unsigned* ctr, int rounds, int mode, long long* cyc, float* data) {
  long long t0 = clock64();
  for (int r = 1; r <= rounds; r++) {
    data[blockIdx.x * blockDim.x + threadIdx.x] = (float)r;   // a store the barrier must publish
    if (mode == 1) __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) {
      if (mode == 2) asm volatile("red.relaxed.gpu.global.add.u32 [%0], 1;" ::"l"(ctr) : "memory");
      else asm volatile("red.release.gpu.global.add.u32 [%0], 1;" ::"l"(ctr) : "memory");
      unsigned target = r * gridDim.x;
      if (mode == 2) while (ld_rlx(ctr) < target) {}
      else while (ld_acq(ctr) < target) {}
    }
    __syncthreads();
  }
  long long t1 = clock64();
  if (threadIdx.x == 0) cyc[blockIdx.x] = (t1 - t0) / rounds;
}

__global__ void barsync(int rounds, long long* cyc) {
  long long t0 = clock64();
  for (int r = 0; r < rounds; r++) __syncthreads();
  long long t1 = clock64();
  if (threadIdx.x == 0) cyc[blockIdx.x] = (t1 - t0) / rounds;
}

// one CTA writes a line, another reads it right after seeing a flag: ping-pong latency between two SMs
__global__ void pingpong(volatile unsigned* flag, int rounds, long long* cyc) {
  long long t0 = clock64();
  if (threadIdx.x == 0) {
    for (int r = 1; r <= rounds; r++) {
      if (blockIdx.x == 0) { asm volatile("st.relaxed.gpu.global.u32 [%0], %1;" ::"l"(flag), "r"(2 * r - 1) : "memory"); while (ld_rlx((const unsigned*)flag) != 2u * r) {} }
      else { while (ld_rlx((const unsigned*)flag) != 2u * r - 1) {} asm volatile("st.relaxed.gpu.global.u32 [%0], %1;" ::"l"(flag), "r"(2 * r) : "memory"); }
    }
  }
  long long t1 = clock64();
  if (threadIdx.x == 0) cyc[blockIdx.x] = (t1 - t0) / rounds;
}

int main() {
  int dev = 0; cudaDeviceProp prop; CK(cudaGetDeviceProperties(&prop, dev));
  int khz = 0; cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, dev);
  printf("%s SMs %d clock %d MHz\n", prop.name, prop.multiProcessorCount, khz / 1000);
  long long* cyc; CK(cudaMallocManaged(&cyc, 1024 * sizeof(long long)));
  unsigned* out; CK(cudaMalloc(&out, 4096));
  // pointer chase over 32 MB (L2 resident) and 2 GB (DRAM), stride pattern
  for (size_t mb : {32ul, 2048ul}) {
    size_t n = mb * 1024 * 1024 / 4;
    std::vector<unsigned> h(n);
    size_t stride = 128 / 4 * 1031;   // jump lines pseudo-randomly
    for (size_t i = 0; i < n; i++) h[i] = (unsigned)((i + stride) % n);
    unsigned* d; CK(cudaMalloc(&d, n * 4)); CK(cudaMemcpy(d, h.data(), n * 4, cudaMemcpyHostToDevice));
    for (int cg = 1; cg >= 0; cg--) {
      chase<<<1, 1>>>(d, 20000, out, cyc, cg); CK(cudaDeviceSynchronize());
      chase<<<1, 1>>>(d, 20000, out, cyc, cg); CK(cudaDeviceSynchronize());
      printf("chase %zu MB %s: %lld cycles/load\n", mb, cg ? "cg" : "ca", cyc[0]);
    }
    cudaFree(d);
  }
  unsigned* ctr; CK(cudaMalloc(&ctr, 256)); float* data; CK(cudaMalloc(&data, 148 * 1024 * 4));
  for (int mode = 0; mode < 3; mode++) for (int threads : {32, 512}) {
    CK(cudaMemset(ctr, 0, 256));
    void* args[] = {&ctr, nullptr, &mode, &cyc, &data};
    int rounds = 2000; args[1] = &rounds;
    CK(cudaLaunchCooperativeKernel((void*)gridbar, dim3(prop.multiProcessorCount), dim3(threads), args, 0, 0));
    CK(cudaDeviceSynchronize());
    long long mx = 0; for (int i = 0; i < prop.multiProcessorCount; i++) mx = cyc[i] > mx ? cyc[i] : mx;
    printf("grid barrier mode %d threads %d: %lld cycles/round\n", mode, threads, mx);
  }
  barsync<<<1, 512>>>(10000, cyc); CK(cudaDeviceSynchronize()); printf("bar.sync 512 threads: %lld cycles\n", cyc[0]);
  CK(cudaMemset(ctr, 0, 256));
  { void* args[] = {&ctr, nullptr, &cyc}; int rounds = 5000; args[1] = &rounds;
    CK(cudaLaunchCooperativeKernel((void*)pingpong, dim3(2), dim3(32), args, 0, 0)); CK(cudaDeviceSynchronize());
    printf("ping-pong between 2 SMs: %lld cycles per round trip (2 one-way hops)\n", cyc[0]); }
  return 0;
}
