// Host-side interface of the tcgen05 weight matmul (gemm_tc.cu).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace tha {

#define THA_TRY_RC(expr) do { int _st = (expr); if (_st != THA_STATUS_SUCCESS) return _st; } while (0)

struct TcArgs {
  int n_mats;                 // 1..3 stacked weight matrices with the same K
  const float* W[3]; int M[3]; int out_off[3];   // rows of matrix i land at out[n*ldo + out_off[i] + row]
  int K;
  const float* Xhi; const float* Xlo; int N; long long ldx;   // activations [N, ldx] split: hi = tf32(x), lo = x - hi
  float* out; long long ldo; long long zstride;  // split z writes out + z*zstride (partials, summed by the consumer)
  int splits;                 // split-K factor (>= 1); must divide the k-blocks into exactly `splits` non-empty parts
  const float* colscale;      // nullable [N]: out *= colscale[n]  (the 1/rms factor of a folded RMSNorm)
};

struct TcP {                  // kernel parameter block
  int M[3]; int tiles[3]; int out_off[3];
  int K, N, ny, kblocks_total, kblocks_per_split;   // ny = column tiles
  float* out; long long ldo; long long zstride;
  const float* colscale;
  int* error_flag;
  int splits, n_items;          // work items = (tiles of all matrices) x column tiles x splits, walked by persistent CTAs
  long long* trace;           // development aid (tha_gemm_trace)
};

// Q8_0 weights x Q8_0 activations on the int8 tensor cores (gemm_q8i_kernel)
struct Q8iArgs {
  int n_mats; const int8_t* Wq[3]; const float* Ws[3]; int M[3]; int out_off[3];
  int K, gs;
  const int8_t* Xq; const float* xsT; int ldxs; int N;   // activations int8 [N, K]; scales transposed [K/gs, ldxs]
  float* out; long long ldo; long long zstride; int splits;   // split over 128-value slots
};
bool gemm_q8i_ok(const Q8iArgs& a);
int gemm_q8i_pick_nt(int N, int gs);
int gemm_q8i_launch(cudaStream_t st, const Q8iArgs& a, bool pdl);

// bf16 weights x bf16 activations (gemm_bf16.cu): both operands straight from TMA-written shared memory
struct BfArgs {
  int n_mats; const uint16_t* W[3]; int M[3]; int out_off[3];
  int K;
  const uint16_t* X; int N; long long ldx;       // activations [N, ldx] bf16
  float* out; long long ldo; long long zstride; int splits;
  const float* colscale;
  // SwiGLU in the read-out (prompt-sized N, two matrices w1 | w3 of equal height, no split-K): a work item multiplies 128 rows
  // of w1 AND the same 128 rows of w3; its read-out writes silu(w1 x) * (w3 x), rounded to bf16, to swiglu_out[n * ld_swiglu + row]
  // (the input of the w2 matmul) and `out` is not written at all.
  uint16_t* swiglu_out; long long ld_swiglu;
};
bool gemm_bf16_swiglu_ok(const BfArgs& a);   // the launch can take the SwiGLU read-out
bool gemm_bf16_ok(const BfArgs& a);
int gemm_bf16_pick_nt(int N);
int gemm_bf16_launch(cudaStream_t st, const BfArgs& a, bool pdl);

bool gemm_tc_ok(const TcArgs& a);
int gemm_tc_pick_nt(int N);
int gemm_tc_launch(cudaStream_t st, const TcArgs& a, bool pdl);
int gemm_tc_check_error();
int gemm_tc_init(int** flag_out);   // allocate the per-device error flag (outside any stream capture)
void gemm_tc_forget();

}  // namespace tha
