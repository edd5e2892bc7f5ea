// Standalone element-wise / reduction ops of the thaBLAS + thaDNN surface.
// In the fused decode step these are folded into the GEMV kernels (gemv.cu); the
// stand-alone forms exist because every reference entry point must exist
//  and they are what the per-op parity tests call.
#include "common.cuh"

namespace tha {

// B = A / val  (src/thaBLAS.cpp:72-77).  IEEE division, as the reference.
__global__ void svds_kernel(int n, const float* __restrict__ A, float* __restrict__ B, float val) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  int stride = gridDim.x * blockDim.x;
  for (; i < n; i += stride) B[i] = A[i] / val;
}

// a += b (src/thaBLAS.cpp:104-108)
__global__ void vecadd_kernel(float* __restrict__ a, const float* __restrict__ b, int n) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  int stride = gridDim.x * blockDim.x;
  for (; i < n; i += stride) a[i] += b[i];
}

// hb = hb * (1/(1+exp(-hb))) * hb2  (src/thaDNN/thaDNN_swiglu.cpp:5-14), same op order.
__device__ __forceinline__ float swiglu_scalar(float a, float c) {
  float v = a;
  v *= (1.0f / (1.0f + expf(-v)));
  v *= c;
  return v;
}
__global__ void swiglu_kernel(float* __restrict__ hb, const float* __restrict__ hb2, int n) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  int stride = gridDim.x * blockDim.x;
  for (; i < n; i += stride) hb[i] = swiglu_scalar(hb[i], hb2[i]);
}

// RoPE on one q (all dim) and one k (first kv_dim) vector, src/thaDNN/thaDNN_rope.cpp:25-42.
__global__ void rope_kernel(int dim, int head_size, int kv_dim, int pos, float* __restrict__ q,
                            float* __restrict__ k) {
  int i = (blockIdx.x * blockDim.x + threadIdx.x) * 2;
  if (i >= dim) return;
  int head_dim = i % head_size;
  float freq = 1.0f / powf(10000.0f, head_dim / (float)head_size);
  float val = pos * freq;
  float fcr = cosf(val), fci = sinf(val);
  float v0 = q[i], v1 = q[i + 1];
  q[i] = v0 * fcr - v1 * fci;
  q[i + 1] = v0 * fci + v1 * fcr;
  if (i < kv_dim) {
    v0 = k[i]; v1 = k[i + 1];
    k[i] = v0 * fcr - v1 * fci;
    k[i + 1] = v0 * fci + v1 * fcr;
  }
}

// One block per row: o = w * (x * rsqrt(mean(x^2)+1e-5)), src/thaDNN/thaDNN_rmsnorm.cpp:35-65.
// o may alias x (the reference's final norm is in place, src/thaDNN.cpp:75).
__global__ void rmsnorm_batch_kernel(float* o_batch, const float* x_batch, const float* __restrict__ weight,
                                     int size, int dim) {
  __shared__ float red[33];
  const float* x = x_batch + (size_t)blockIdx.x * dim;
  float* o = o_batch + (size_t)blockIdx.x * dim;
  float ss = 0.0f;
  for (int i = threadIdx.x; i < size; i += blockDim.x) { float t = x[i]; ss = fmaf(t, t, ss); }
  ss = block_sum(ss, red);
  ss /= size;
  ss += 1e-5f;
  ss = 1.0f / sqrtf(ss);
  for (int i = threadIdx.x; i < size; i += blockDim.x) o[i] = weight[i] * (ss * x[i]);
}

// In-place softmax of one vector by one block (src/thaDNN/thaDNN_softmax.cpp:62-97).
__global__ void softmax_kernel(float* x, int size) {
  __shared__ float red[33];
  float m = -INFINITY;
  for (int i = threadIdx.x; i < size; i += blockDim.x) m = fmaxf(m, x[i]);
  m = block_max(m, red);
  float s = 0.0f;
  for (int i = threadIdx.x; i < size; i += blockDim.x) { float e = expf(x[i] - m); x[i] = e; s += e; }
  s = block_sum(s, red);
  for (int i = threadIdx.x; i < size; i += blockDim.x) x[i] /= s;
}

// x[b,:] = table[token[b],:]  (src/thaDNN.cpp:31-36, B blocking D2D copies there).
__global__ void embed_kernel(float* __restrict__ x, const float* __restrict__ table,
                             const int* __restrict__ token, const int* __restrict__ pos, int dim) {
  pdl_trigger();   // the next kernel may start its launch (it waits for this grid to complete before touching memory)
  pdl_wait();
  int b = blockIdx.x;
  if (pos[b] < 0) return;
  const float4* src = reinterpret_cast<const float4*>(table + (size_t)token[b] * dim);
  float4* dst = reinterpret_cast<float4*>(x + (size_t)b * dim);
  for (int i = threadIdx.x; i < dim / 4; i += blockDim.x) dst[i] = src[i];
}

// Greedy sampler: first maximum wins (strict >, src/llama.cpp:275-286).  One block per row.
__global__ void argmax_kernel(const float* __restrict__ logits, int vocab, int* __restrict__ next) {
  pdl_trigger();   // the next kernel may start its launch (it waits for this grid to complete before touching memory)
  pdl_wait();
  __shared__ float sv[32];
  __shared__ int si[32];
  const float* p = logits + (size_t)blockIdx.x * vocab;
  float bv = -INFINITY;
  int bi = 0x7fffffff;
  for (int i = threadIdx.x; i < vocab; i += blockDim.x) {
    float v = p[i];
    if (v > bv) { bv = v; bi = i; }  // ascending i per thread: first max kept
  }
  // NaN-free inputs assumed (as the CPU code); ties resolved to the lowest index
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    float ov = __shfl_xor_sync(THA_FULL_MASK, bv, o);
    int oi = __shfl_xor_sync(THA_FULL_MASK, bi, o);
    if (ov > bv || (ov == bv && oi < bi)) { bv = ov; bi = oi; }
  }
  int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  if (lane == 0) { sv[wid] = bv; si[wid] = bi; }
  __syncthreads();
  if (wid == 0) {
    int nw = blockDim.x >> 5;
    bv = lane < nw ? sv[lane] : -INFINITY;
    bi = lane < nw ? si[lane] : 0x7fffffff;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      float ov = __shfl_xor_sync(THA_FULL_MASK, bv, o);
      int oi = __shfl_xor_sync(THA_FULL_MASK, bi, o);
      if (ov > bv || (ov == bv && oi < bi)) { bv = ov; bi = oi; }
    }
    if (lane == 0) next[blockIdx.x] = (bi == 0x7fffffff) ? 0 : bi;
  }
}

// Q8_0 activation quantiser, bit-exact with runq.c:145-171.  One warp per group.
//   scale = max|x| / 127.0f   (IEEE fp32 division)
//   q     = (int8) round(x / scale)   (round half away from zero; NaN -> 0, the x86 result)
__global__ void quantize_q8_kernel(int8_t* __restrict__ q, float* __restrict__ s,
                                   const float* __restrict__ x, int n_groups, int gs) {
  int g = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (g >= n_groups) return;
  int lane = threadIdx.x & 31;
  const float* xg = x + (size_t)g * gs;
  float amax = 0.0f;
  for (int i = lane; i < gs; i += 32) amax = fmaxf(amax, fabsf(xg[i]));
  amax = warp_max(amax);
  float scale = __fdiv_rn(amax, 127.0f);
  if (lane == 0) s[g] = scale;
  for (int i = lane; i < gs; i += 32) {
    float r = __fdiv_rn(xg[i], scale);
    q[(size_t)g * gs + i] = (r != r) ? (int8_t)0 : (int8_t)__float2int_rn(roundf(r));
  }
}

// Greedy decode bookkeeping on the device (the reference does this on the host,
// src/llama.cpp:1027-1050): token[b] = next[b]; pos[b] += 1 for active slots.
__global__ void advance_slots_kernel(int B, int* __restrict__ token, int* __restrict__ pos,
                                     const int* __restrict__ next, int seq_len) {
  pdl_trigger();   // the next kernel may start its launch (it waits for this grid to complete before touching memory)
  pdl_wait();
  int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= B) return;
  int p = pos[b];
  if (p < 0) return;
  token[b] = next[b];
  pos[b] = (p + 1 < seq_len) ? p + 1 : -1;   // a slot that ran out of context goes idle
}

int advance_launch(cudaStream_t st, int B, int* token, int* pos, const int* next, int seq_len, bool pdl) {
  THA_CUDA_OK(launch(advance_slots_kernel, dim3(ceil_div(B, 128)), dim3(128), 0, st, pdl, B, token, pos, next, seq_len));
  return THA_STATUS_SUCCESS;
}
int embed_launch(cudaStream_t st, int B, float* x, const float* table, const int* token, const int* pos, int dim, bool pdl) {
  THA_CUDA_OK(launch(embed_kernel, dim3(B), dim3(256), 0, st, pdl, x, table, token, pos, dim));
  return THA_STATUS_SUCCESS;
}
int argmax_launch(cudaStream_t st, int B, const float* logits, int vocab, int* next, bool pdl) {
  THA_CUDA_OK(launch(argmax_kernel, dim3(B), dim3(1024), 0, st, pdl, logits, vocab, next));
  return THA_STATUS_SUCCESS;
}

}  // namespace tha

using namespace tha;

static inline int grid_for(int n, int block) { int g = ceil_div(n, block); return g > 4096 ? 4096 : (g < 1 ? 1 : g); }

extern "C" int tha_svds_f32(void* stream, int n, const float* A, float* B, float val) {
  if (n <= 0 || !A || !B || val == 0.0f) { printf("THABLAS ERROR: INVALID ARGUMENT\n"); return THA_STATUS_INVALID_ARG; }
  svds_kernel<<<grid_for(n, 256), 256, 0, (cudaStream_t)stream>>>(n, A, B, val);
  count_launch();
  THA_LAUNCH_OK();
  return THA_STATUS_SUCCESS;
}

extern "C" int tha_vecaddvec_f32(void* stream, float* a, const float* b, int size) {
  if (size <= 0 || !a || !b) return THA_STATUS_INVALID_ARG;
  vecadd_kernel<<<grid_for(size, 256), 256, 0, (cudaStream_t)stream>>>(a, b, size);
  count_launch();
  THA_LAUNCH_OK();
  return THA_STATUS_SUCCESS;
}

extern "C" int tha_swiglu_f32(void* stream, float* hb, const float* hb2, int hidden_dim) {
  if (hidden_dim <= 0 || !hb || !hb2) return THA_STATUS_INVALID_ARG;
  swiglu_kernel<<<grid_for(hidden_dim, 256), 256, 0, (cudaStream_t)stream>>>(hb, hb2, hidden_dim);
  count_launch();
  THA_LAUNCH_OK();
  return THA_STATUS_SUCCESS;
}

extern "C" int tha_rope_f32(void* stream, int dim, int head_size, int kv_dim, int pos, float* q, float* k) {
  if (dim <= 0 || head_size <= 0 || !q || (kv_dim > 0 && !k)) return THA_STATUS_INVALID_ARG;
  int pairs = (dim + 1) / 2;
  rope_kernel<<<ceil_div(pairs, 128), 128, 0, (cudaStream_t)stream>>>(dim, head_size, kv_dim, pos, q, k);
  count_launch();
  THA_LAUNCH_OK();
  return THA_STATUS_SUCCESS;
}

extern "C" int tha_rmsnorm_batch_f32(void* stream, int n_batches, float* o, const float* x,
                                     const float* weight, int size, int dim) {
  if (n_batches <= 0 || size <= 0 || !o || !x || !weight) return THA_STATUS_INVALID_ARG;
  int block = size >= 1024 ? 1024 : (size >= 256 ? 256 : 64);
  rmsnorm_batch_kernel<<<n_batches, block, 0, (cudaStream_t)stream>>>(o, x, weight, size, dim);
  count_launch();
  THA_LAUNCH_OK();
  return THA_STATUS_SUCCESS;
}

extern "C" int tha_softmax_f32(void* stream, float* x, int size) {
  if (size <= 0 || !x) { fprintf(stderr, "THABLAS SOFTMAX ERROR: INVALID ARGUMENT\n"); return THA_STATUS_INVALID_ARG; }
  softmax_kernel<<<1, 1024, 0, (cudaStream_t)stream>>>(x, size);
  count_launch();
  THA_LAUNCH_OK();
  return THA_STATUS_SUCCESS;
}

extern "C" int tha_argmax_batch_f32(void* stream, int n_batches, const float* logits, int vocab, int* next_d) {
  if (n_batches <= 0 || vocab <= 0 || !logits || !next_d) return THA_STATUS_INVALID_ARG;
  return argmax_launch((cudaStream_t)stream, n_batches, logits, vocab, next_d, false);
}

extern "C" int tha_advance_slots(void* stream, int n_batches, int* token_d, int* pos_d, const int* next_d, int seq_len) {
  if (n_batches <= 0 || !token_d || !pos_d || !next_d) return THA_STATUS_INVALID_ARG;
  THA_CUDA_OK(launch(advance_slots_kernel, dim3(ceil_div(n_batches, 128)), dim3(128), 0, (cudaStream_t)stream, false,
                     n_batches, token_d, pos_d, next_d, seq_len));
  return THA_STATUS_SUCCESS;
}

extern "C" int tha_quantize_q8(void* stream, int8_t* q, float* s, const float* x, int n, int gs) {
  if (n <= 0 || gs <= 0 || n % gs != 0 || !q || !s || !x) return THA_STATUS_INVALID_ARG;
  int groups = n / gs;
  quantize_q8_kernel<<<ceil_div(groups, 8), 256, 0, (cudaStream_t)stream>>>(q, s, x, groups, gs);
  count_launch();
  THA_LAUNCH_OK();
  return THA_STATUS_SUCCESS;
}
