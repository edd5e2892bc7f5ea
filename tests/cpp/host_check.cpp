// host_check -- command-line probe of the C++ host layer, driven by tests/test_host_cpp.py.
//   host_check encode <tokenizer.bin> <vocab>            prompts on stdin (one per line) -> ids per line
//   host_check decode <tokenizer.bin> <vocab>            ids per line on stdin -> text per line (decode + append_str)
//   host_check sample <vocab> <temp> <topp> <seed> <n>   n*vocab fp32 logits on stdin -> n sampled ids
//   host_check bm                                        BatchManager bookkeeping self-check
//   host_check api                                       (GPU) every thaBLAS / thaDNN entry point against plain loops,
//                                                        with the shapes and tolerances of the reference's own API tests
//                                                        (scripts/test/thaBLAS.test.cpp:148-152, thaDNN.test.cpp)
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <iostream>
#include <sstream>
#include <string>
#include <vector>

#include "BatchManager.hpp"
#include "llama_driver.hpp"
#include "tha_b200.h"
#include "thaDNN.hpp"

static int cmd_encode(const char* path, int vocab) {
  Tokenizer t;
  build_tokenizer(&t, path, vocab);
  std::string line;
  while (std::getline(std::cin, line)) {
    std::vector<int> ids(line.size() + 3);
    int n = 0;
    encode(&t, line.c_str(), 1, 0, ids.data(), &n);
    for (int i = 0; i < n; i++) printf("%s%d", i ? " " : "", ids[i]);
    printf("\n");
  }
  return 0;
}

static int cmd_decode(const char* path, int vocab) {
  Tokenizer t;
  build_tokenizer(&t, path, vocab);
  std::string line;
  while (std::getline(std::cin, line)) {
    std::istringstream is(line);
    std::string out;
    int prev = -1, id;
    while (is >> id) {
      if (prev >= 0) append_str(decode(&t, prev, id), out);
      prev = id;
    }
    printf("%s\n", out.c_str());
  }
  return 0;
}

static int cmd_sample(int vocab, float temp, float topp, unsigned long long seed, int n) {
  Sampler s;
  build_sampler(&s, vocab, temp, topp, seed);
  std::vector<float> lg(vocab);
  for (int i = 0; i < n; i++) {
    if (fread(lg.data(), sizeof(float), vocab, stdin) != (size_t)vocab) return 2;
    printf("%d\n", sample(&s, lg.data()));
  }
  free_sampler(&s);
  return 0;
}

#define EXPECT(c) do { if (!(c)) { printf("FAIL %s:%d %s\n", __FILE__, __LINE__, #c); return 1; } } while (0)

static int cmd_bm() {
  BatchManager bm(3, 40, 2, 8);   // 3 slots, 40 positions, block = 16 positions -> 3 blocks per slot
  EXPECT(bm.nrow_blocks == 3 && bm.num_physical_blocks == 9 && bm.cur_physical_blocks == 0);
  EXPECT(bm.acquire_slot(10) == 0 && bm.acquire_slot(11) == 1 && bm.acquire_slot(12) == 2 && bm.acquire_slot(13) == -1);
  EXPECT(bm.active_slots() == 3 && bm.request_of(1) == 11);
  for (int i = 0; i < 17; i++) bm.advance(1);          // positions 0..16 -> blocks 0 and 1
  EXPECT(bm.n_block_table_filled[1] == 2 && bm.cur_physical_blocks == 2);
  bm.advance(0);
  EXPECT(bm.cur_physical_blocks == 3);
  EXPECT(bm.get_block_by_id(nullptr, 1 * 3 + 1).occupied == 1 && bm.get_block_by_id(nullptr, 2 * 3).occupied == 0);
  bm.release_slot(1);
  EXPECT(bm.cur_physical_blocks == 1 && bm.request_of(1) == -1 && bm.active_slots() == 2);
  EXPECT(bm.acquire_slot(14) == 1 && bm.logical_KV_table[1].second == 0);
  const size_t id = bm.alloc_new_block(nullptr);       // lowest free: slot 0 block 1
  EXPECT(id == 1 && bm.cur_physical_blocks == 2);
  EXPECT(bm.block_bytes() == (size_t)2 * 2 * 16 * 8 * 4);
  printf("OK\n");
  return 0;
}

// ---- GPU API check -------------------------------------------------------------------------
static bool close_enough(float got, float want, float eps) {  // the reference's rule: abs <= eps OR rel <= eps
  const float a = fabsf(got - want);
  return a <= eps || a <= eps * fabsf(want) || (isnan(got) && isnan(want));
}
static int compare(const char* what, const std::vector<float>& got, const std::vector<float>& want, float eps) {
  int bad = 0;
  for (size_t i = 0; i < want.size(); i++)
    if (!close_enough(got[i], want[i], eps)) { if (bad++ < 3) printf("  %s[%zu] = %g, want %g\n", what, i, got[i], want[i]); }
  printf("%s %s (%zu values, eps %g)\n", bad ? "FAIL" : "ok  ", what, want.size(), eps);
  return bad != 0;
}
struct Dev {
  float* p = nullptr; size_t n;
  explicit Dev(const std::vector<float>& h) : n(h.size()) { CHECK_THA(tha_malloc((void**)&p, n * 4)); CHECK_THA(tha_memcpy_h2d(p, h.data(), n * 4)); }
  explicit Dev(size_t n_) : n(n_) { CHECK_THA(tha_malloc((void**)&p, n * 4)); CHECK_THA(tha_memset(p, 0, n * 4)); }
  std::vector<float> host() const { std::vector<float> h(n); CHECK_THA(tha_device_sync()); CHECK_THA(tha_memcpy_d2h(h.data(), p, n * 4)); return h; }
  ~Dev() { tha_free(p); }
};
static std::vector<float> rnd(size_t n) { std::vector<float> v(n); rand_vec(v.data(), (int)n); return v; }

static int cmd_api() {
  int fails = 0;
  thablasHandle_t h;
  if (thablasCreate(&h) != THABLAS_STATUS_SUCCESS) { printf("FAIL thablasCreate\n"); return 1; }
  srand(7);
  // argument validation (src/thaBLAS.cpp:80-84, :155-158, thaDNN_softmax.cpp:103-106)
  fails += thablas_Svds(h, 0, nullptr, nullptr, 1.f) != THABLAS_STATUS_ALLOC_FAILED;
  fails += thaBLAS_s_matmul(h, 0, 1, 1, nullptr, nullptr, nullptr) != THABLAS_STATUS_ALLOC_FAILED;
  fails += thaDNN_s_softmax_v2(&h, nullptr, 0) != THABLAS_STATUS_ALLOC_FAILED;
  printf("%s argument validation\n", fails ? "FAIL" : "ok  ");
  for (int n : {1, 111, 1000, 100000}) {  // vec / scalar, vec + vec
    auto a = rnd(n), b = rnd(n);
    Dev da(a), db(n), dc(b);
    fails += thablas_Svds(h, n, da.p, db.p, 3.25f) != 0;
    std::vector<float> want(n);
    for (int i = 0; i < n; i++) want[i] = a[i] / 3.25f;
    fails += compare("thablas_Svds", db.host(), want, 1e-6f);
    fails += thaBLAS_s_vecaddvec(&h, da.p, dc.p, n) != 0;
    CHECK_THA(tha_stream_sync((void*)h.calc_stream));
    for (int i = 0; i < n; i++) want[i] = a[i] + b[i];
    fails += compare("thaBLAS_s_vecaddvec", da.host(), want, 1e-6f);
  }
  for (int s : {3, 100, 1000}) {  // C = A.B row-major (scripts/test/thaBLAS.test.cpp:148-152), eps 1e-3
    auto A = rnd((size_t)s * s), Bm = rnd((size_t)s * s);
    Dev dA(A), dB(Bm), dC((size_t)s * s);
    fails += thaBLAS_s_matmul(h, s, s, s, dA.p, dB.p, dC.p) != 0;
    std::vector<float> want((size_t)s * s, 0.f);
    for (int i = 0; i < s; i++) for (int k = 0; k < s; k++) { const float a = A[(size_t)i * s + k]; for (int j = 0; j < s; j++) want[(size_t)i * s + j] += a * Bm[(size_t)k * s + j]; }
    fails += compare("thaBLAS_s_matmul", dC.host(), want, 1e-3f);
  }
  {  // GEMV forms and the batched weight matmul with cache-row addressing
    const int M = 288, K = 768, B = 3, S = 8, L = 2;
    auto A = rnd((size_t)M * K), X = rnd((size_t)B * K);
    Dev dA(A), dX(X), dC((size_t)M), dCb((size_t)B * L * S * M), dR((size_t)B * M);
    std::vector<float> want(M);
    for (int i = 0; i < M; i++) { float v = 0; for (int k = 0; k < K; k++) v += A[(size_t)i * K + k] * X[k]; want[i] = v; }
    fails += thaBLAS_s_matmulvec(h, dC.p, dX.p, dA.p, K, M) != 0;
    CHECK_THA(tha_stream_sync((void*)h.calc_stream));
    fails += compare("thaBLAS_s_matmulvec", dC.host(), want, 1e-3f);
    fails += thaDNN_s_matmulvec_v2(&h, dC.p, dX.p, dA.p, K, M) != 0;
    CHECK_THA(tha_stream_sync((void*)h.calc_stream));
    fails += compare("thaDNN_s_matmulvec_v2", dC.host(), want, 1e-3f);
    int pos_h[B] = {5, 0, 3};
    int* pos_d = nullptr;
    CHECK_THA(tha_malloc((void**)&pos_d, sizeof(pos_h)));
    CHECK_THA(tha_memcpy_h2d(pos_d, pos_h, sizeof(pos_h)));
    const int l = 1, Coff = l * S * M;
    fails += thaBLAS_s_matmul_batch(&h, B, dCb.p, dX.p, dA.p, K, M, Coff, M, pos_d, L * S * M, K) != 0;
    CHECK_THA(tha_stream_sync((void*)h.calc_stream));
    std::vector<float> wantb((size_t)B * L * S * M, 0.f);
    for (int b = 0; b < B; b++) for (int i = 0; i < M; i++) {
      float v = 0; for (int k = 0; k < K; k++) v += A[(size_t)i * K + k] * X[(size_t)b * K + k];
      wantb[(size_t)b * L * S * M + Coff + (size_t)pos_h[b] * M + i] = v;
    }
    fails += compare("thaBLAS_s_matmul_batch", dCb.host(), wantb, 1e-3f);
    fails += thaBLAS_s_matmul_reduction(&h, dA.p, dX.p, dR.p, M, B, K) != 0;
    CHECK_THA(tha_stream_sync((void*)h.calc_stream));
    std::vector<float> wantr((size_t)B * M);
    for (int b = 0; b < B; b++) for (int i = 0; i < M; i++) wantr[(size_t)b * M + i] = wantb[(size_t)b * L * S * M + Coff + (size_t)pos_h[b] * M + i];
    fails += compare("thaBLAS_s_matmul_reduction", dR.host(), wantr, 1e-3f);
    tha_free(pos_d);
  }
  for (int n : {768, 4096, 16384}) {  // rmsnorm, swiglu, softmax (thaDNN.test.cpp shapes)
    const int B = 2;
    auto x = rnd((size_t)B * n), w = rnd(n), g = rnd(n);
    Dev dx(x), dw(w), dOut((size_t)B * n), dg(g);
    fails += thaDNN_s_rmsnorm_v2_batch(&h, B, dOut.p, dx.p, dw.p, n, n) != 0;
    CHECK_THA(tha_stream_sync((void*)h.calc_stream));
    std::vector<float> want((size_t)B * n);
    for (int b = 0; b < B; b++) {
      float ss = 0; for (int i = 0; i < n; i++) ss += x[(size_t)b * n + i] * x[(size_t)b * n + i];
      ss = 1.0f / sqrtf(ss / n + 1e-5f);
      for (int i = 0; i < n; i++) want[(size_t)b * n + i] = w[i] * (ss * x[(size_t)b * n + i]);
    }
    fails += compare("thaDNN_s_rmsnorm_v2_batch", dOut.host(), want, 1e-4f);
    Dev dh(std::vector<float>(x.begin(), x.begin() + n));
    fails += thaDNN_s_swiglu(&h, dh.p, dg.p, n) != 0;
    CHECK_THA(tha_stream_sync((void*)h.calc_stream));
    std::vector<float> wants(n);
    for (int i = 0; i < n; i++) wants[i] = x[i] * (1.0f / (1.0f + expf(-x[i]))) * g[i];
    fails += compare("thaDNN_s_swiglu", dh.host(), wants, 1e-5f);
    Dev dsm(std::vector<float>(x.begin(), x.begin() + n));
    fails += thaDNN_s_softmax_v2(&h, dsm.p, n) != 0;
    std::vector<float> sm(x.begin(), x.begin() + n);
    softmax_host(sm.data(), n);
    fails += compare("thaDNN_s_softmax_v2", dsm.host(), sm, 1e-5f);
  }
  {  // rope
    const int dim = 768, hs = 64, kvd = 768, pos = 37;
    auto q = rnd(dim), k = rnd(kvd);
    Dev dq(q), dk(k);
    fails += thaDNN_s_rope(&h, dim, hs, kvd, pos, dq.p, dk.p) != 0;
    CHECK_THA(tha_stream_sync((void*)h.calc_stream));
    for (int i = 0; i < dim; i += 2) {
      const float freq = 1.0f / powf(10000.0f, (i % hs) / (float)hs), val = pos * freq, fcr = cosf(val), fci = sinf(val);
      float v0 = q[i], v1 = q[i + 1]; q[i] = v0 * fcr - v1 * fci; q[i + 1] = v0 * fci + v1 * fcr;
      v0 = k[i]; v1 = k[i + 1]; k[i] = v0 * fcr - v1 * fci; k[i + 1] = v0 * fci + v1 * fcr;
    }
    fails += compare("thaDNN_s_rope q", dq.host(), q, 1e-4f);
    fails += compare("thaDNN_s_rope k", dk.host(), k, 1e-4f);
  }
  thablasDestroy(h);
  printf(fails ? "API CHECK FAILED (%d)\n" : "API CHECK OK\n", fails);
  return fails != 0;
}

int main(int argc, char** argv) {
  if (argc < 2) return 2;
  const std::string c = argv[1];
  if (c == "encode" && argc == 4) return cmd_encode(argv[2], atoi(argv[3]));
  if (c == "decode" && argc == 4) return cmd_decode(argv[2], atoi(argv[3]));
  if (c == "sample" && argc == 7) return cmd_sample(atoi(argv[2]), atof(argv[3]), atof(argv[4]), strtoull(argv[5], nullptr, 10), atoi(argv[6]));
  if (c == "bm") return cmd_bm();
  if (c == "api") return cmd_api();
  return 2;
}
