// tcgen05 / TMEM / TMA building blocks shared by the tensor-core weight matmuls (gemm_tc.cu: one matmul per launch;
// chain_tc.cu: a chain of dependent matmuls with fused epilogues in one persistent launch).
#pragma once
#include <cuda.h>

#include "common.cuh"

namespace tha {

constexpr int TC_BM = 128;          // weight rows per tile = UMMA M
constexpr int TC_BK = 32;           // floats per k-block: 128 bytes, the swizzle width
constexpr int TC_CGROUPS = 3;       // converter/epilogue warp groups (4 warps each), k-blocks dealt round-robin
constexpr int TC_THREADS = 128 + 128 * TC_CGROUPS;
constexpr int TC_SMEM_LIMIT = 227 * 1024;

// ------------------------------------------------------------------ PTX wrappers ----
__device__ __forceinline__ uint32_t tc_smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void tc_mbar_init(uint64_t* bar, unsigned count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(tc_smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void tc_mbar_expect_tx(uint64_t* bar, unsigned bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(tc_smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void tc_mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(tc_smem_u32(bar)) : "memory");
}
__device__ __forceinline__ bool tc_mbar_try(uint64_t* bar, unsigned parity) {
  unsigned ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok) : "r"(tc_smem_u32(bar)), "r"(parity) : "memory");
  return ok != 0;
}
// Bounded wait: a broken pipeline must end the kernel (and fail the parity test), never hang the GPU.
__device__ __forceinline__ bool tc_mbar_wait(uint64_t* bar, unsigned parity, int* abort_flag) {
  for (unsigned n = 0; !tc_mbar_try(bar, parity); n++) {
    if (n > (1u << 22)) { *abort_flag = 1; return false; }
    if ((n & 1023u) == 1023u && *((volatile int*)abort_flag)) return false;
  }
  return true;
}
__device__ __forceinline__ void tc_tma_2d(void* dst, const CUtensorMap* map, int c0, int c1, uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
      ::"r"(tc_smem_u32(dst)), "l"(map), "r"(tc_smem_u32(bar)), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void tc_tma_2d_hint(void* dst, const CUtensorMap* map, int c0, int c1, uint64_t* bar, uint64_t pol) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1, {%3, %4}], [%2], %5;"
      ::"r"(tc_smem_u32(dst)), "l"(map), "r"(tc_smem_u32(bar)), "r"(c0), "r"(c1), "l"(pol) : "memory");
}
// K-major operand tile, 128-byte rows, SWIZZLE_128B: 8-row groups 1024 bytes apart (cute::UMMA::SmemDescriptor).
__device__ __forceinline__ uint64_t tc_smem_desc(uint32_t saddr) {
  uint64_t d = 0;
  d |= (uint64_t)((saddr & 0x3FFFFu) >> 4);          // start address, 16-byte units
  d |= (uint64_t)1 << 16;                            // leading byte offset: unused for swizzled K-major
  d |= (uint64_t)(1024 >> 4) << 32;                  // stride byte offset between 8-row groups
  d |= (uint64_t)1 << 46;                            // descriptor version (Blackwell)
  d |= (uint64_t)2 << 61;                            // SWIZZLE_128B
  return d;
}
// kind::tf32, fp32 accumulate, both operands K-major, M = 128, N = nt (cute::UMMA::InstrDescriptor)
__host__ __device__ constexpr uint32_t tc_idesc(int nt) {
  return (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(nt >> 3) << 17) | ((uint32_t)(TC_BM >> 4) << 24);
}
__device__ __forceinline__ void tc_mma_tf32(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
// A operand in TMEM (lane = row of the 128-row tile, one tf32 per 32-bit column), B through a shared-memory descriptor
__device__ __forceinline__ void tc_mma_tf32_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n\t}"
      ::"r"(tmem_d), "r"(tmem_a), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void tc_tmem_st32(uint32_t taddr, const uint32_t (&v)[32]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, "
      "%17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32};"
      ::"r"(taddr), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]), "r"(v[8]), "r"(v[9]),
        "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15]), "r"(v[16]), "r"(v[17]), "r"(v[18]), "r"(v[19]),
        "r"(v[20]), "r"(v[21]), "r"(v[22]), "r"(v[23]), "r"(v[24]), "r"(v[25]), "r"(v[26]), "r"(v[27]), "r"(v[28]), "r"(v[29]),
        "r"(v[30]), "r"(v[31]) : "memory");
}
__device__ __forceinline__ void tc_tmem_ld16(uint32_t taddr, uint32_t (&v)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
        "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
      : "r"(taddr));
}
__device__ __forceinline__ void tc_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(tc_smem_u32(bar)) : "memory");
}
// round to nearest (ties away from zero) onto the 10-bit tf32 mantissa == cvt.rna.tf32.f32, with two
// full-rate integer instructions instead of one quarter-rate conversion
__device__ __forceinline__ float tc_tf32(float v) {
  return __uint_as_float((__float_as_uint(v) + 0x1000u) & 0xFFFFE000u);
}

__device__ __forceinline__ long long tc_clock() { long long t; asm volatile("mov.u64 %0, %%clock64;" : "=l"(t)); return t; }
#define TC_TRACE(slot, i) do { if (p.trace && blockIdx.x == 0 && blockIdx.z == 0 && (i) < 64) p.trace[(slot) * 64 + (i)] = tc_clock(); } while (0)

template <int NT>
struct TcSmem {
  static constexpr int W_BYTES = TC_BM * 128;                        // one weight k-block, raw fp32, swizzled rows
  static constexpr int X_BYTES = NT * 128;
  static constexpr int XS_BYTES = 2 * X_BYTES;                       // Xhi | Xlo of one k-block (one 2*NT-row MMA operand)
  static constexpr bool SPLIT_ACC = true;                            // lo-terms in their own accumulator columns [NT, 2NT)
  static constexpr int ACC_COLS = 2 * NT;
  static constexpr int A_COLS = 2 * TC_BK;                           // per k-block: W hi (32 columns) | W lo (32 columns)
  // A pipeline STAGE is KG k-blocks: the single MMA-issuing thread pays its fixed per-stage cost (barrier wait,
  // fence, commit: ~380 cycles measured) once per KG x 8 MMAs (~60 cycles each).
  static constexpr int KG = NT <= 64 ? 2 : 1;
  static constexpr int W_STAGE = KG * W_BYTES, XS_STAGE = KG * XS_BYTES, A_STAGE = KG * A_COLS;
  // Three rings.  W (shared memory): filled by TMA, freed as soon as a converter has read the rows -- it only has
  // to cover the HBM latency.  A (TMEM, converter -> MMA) and X (shared memory, TMA -> MMA) share one index, one
  // full barrier (4 converter warps + the X bytes) and one empty barrier (the MMAs' commit).  Ring lengths are
  // multiples of the number of converter groups so that a slot is always served by the same group (a waiter must
  // never be two barrier phases ahead of the barrier).
  static constexpr int SA_RAW = (512 - ACC_COLS) / A_STAGE;
  static constexpr int SA = SA_RAW >= 6 ? 6 : 3;
  static constexpr int SX = SA;
  static constexpr int SW_RAW = (TC_SMEM_LIMIT - 2048 - SX * XS_STAGE) / W_STAGE;
  static constexpr int SW = SW_RAW >= 12 ? 12 : (SW_RAW / TC_CGROUPS) * TC_CGROUPS;
  static constexpr int NBARS = 2 * SW + 2 * SA + 2;
  static constexpr int TOTAL = SW * W_STAGE + SX * XS_STAGE + 1024 /*alignment slack*/ + NBARS * 8 + 64;
  static constexpr int TMEM_COLS = 512;
  static_assert(SW >= TC_CGROUPS && SW % TC_CGROUPS == 0 && SA % TC_CGROUPS == 0, "ring / group mismatch");
  static_assert(ACC_COLS + SA * A_STAGE <= 512, "TMEM overflow");
  static_assert(TOTAL <= TC_SMEM_LIMIT, "shared memory overflow");
};

// Tensor map of a row-major [rows, K] fp32 matrix (elem_bytes = 1: int8) with boxes of box_rows x 128 bytes, SWIZZLE_128B;
// cached by (pointer, shape) -- gemm_tc.cu.
int tc_map(const void* ptr, long long rows, int K, long long ld, int box_rows, CUtensorMap* out, int elem_bytes = 4);

}  // namespace tha
