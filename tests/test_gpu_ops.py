"""Per-op parity on the B200: every thaBLAS / thaDNN entry point, called through the C-ABI,
against the oracle, with the shapes and tolerances of the reference's own comparers
(scripts/test/thaBLAS.test.cpp, scripts/test/thaDNN.test.cpp)."""
import ctypes as C

import numpy as np
import pytest

from gpu_util import assert_close_scaled, P, assert_close, stream_ptr, tdev
from pkgload import load_pkg

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def L():
    import torch
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    return load_pkg("cabi").lib()


def sync():
    import torch
    torch.cuda.synchronize()


@pytest.mark.parametrize("n", [10, 1000, 100000])
def test_svds(L, n):
    rng = np.random.default_rng(n)
    a = rng.random(n, dtype=np.float32) - 0.5
    A, B = tdev(a), tdev(np.zeros(n, np.float32))
    L.check(L.c.tha_svds_f32(stream_ptr(), n, P(A), P(B), 0.37))
    sync()
    assert np.array_equal(B.cpu().numpy(), a / np.float32(0.37))  # IEEE division: bit-exact
    assert L.c.tha_svds_f32(stream_ptr(), 0, P(A), P(B), 1.0) == 2       # reference: ALLOC_FAILED on bad args
    assert L.c.tha_svds_f32(stream_ptr(), n, P(A), P(B), 0.0) == 2


@pytest.mark.parametrize("n", [1, 768, 33333])
def test_vecaddvec(L, n):
    rng = np.random.default_rng(n)
    a, b = rng.standard_normal(n).astype(np.float32), rng.standard_normal(n).astype(np.float32)
    A, B = tdev(a), tdev(b)
    L.check(L.c.tha_vecaddvec_f32(stream_ptr(), P(A), P(B), n))
    sync()
    assert np.array_equal(A.cpu().numpy(), a + b)


@pytest.mark.parametrize("size", [1, 111, 11111, 65536])
def test_rmsnorm(L, oracle, size):
    rng = np.random.default_rng(size)
    nb = 3
    x = (rng.random((nb, size), dtype=np.float32) - 0.5)
    w = (rng.random(size, dtype=np.float32) - 0.5)
    X, W, O = tdev(x), tdev(w), tdev(np.zeros_like(x))
    L.check(L.c.tha_rmsnorm_batch_f32(stream_ptr(), nb, P(O), P(X), P(W), size, size))
    sync()
    got = O.cpu().numpy()
    for b in range(nb):
        assert_close(got[b], oracle.rmsnorm(x[b], w), 1e-5, f"rmsnorm size {size}")
    # in place, as the reference's final norm (src/thaDNN.cpp:75)
    L.check(L.c.tha_rmsnorm_batch_f32(stream_ptr(), nb, P(X), P(X), P(W), size, size))
    sync()
    assert np.array_equal(X.cpu().numpy(), got)


@pytest.mark.parametrize("size", [1, 111, 11111, 32000])
def test_softmax(L, oracle, size):
    rng = np.random.default_rng(size)
    x = (rng.random(size, dtype=np.float32) - 0.5) * 8
    X = tdev(x)
    L.check(L.c.tha_softmax_f32(stream_ptr(), P(X), size))
    sync()
    assert_close(X.cpu().numpy(), oracle.softmax(x), 1e-5, f"softmax {size}")
    assert L.c.tha_softmax_f32(stream_ptr(), P(X), 0) == 2


@pytest.mark.parametrize("dim,hs,kv_dim,pos", [(256, 16, 64, 0), (2, 2, 2, 1), (16384, 256, 11112, 512), (2220, 222, 2110, 111)])
def test_rope(L, oracle, dim, hs, kv_dim, pos):
    # shapes after scripts/test/thaDNN.test.cpp:510-516 (made even where the original was odd)
    rng = np.random.default_rng(dim)
    q, k = rng.random(dim, dtype=np.float32) - 0.5, rng.random(kv_dim, dtype=np.float32) - 0.5
    Q, K = tdev(q), tdev(k)
    L.check(L.c.tha_rope_f32(stream_ptr(), dim, hs, kv_dim, pos, P(Q), P(K)))
    sync()
    wq, wk = oracle.rope(q, k, hs, pos)
    assert_close(Q.cpu().numpy(), wq, 1e-4, "rope q")
    assert_close(K.cpu().numpy(), wk, 1e-4, "rope k")


@pytest.mark.parametrize("n", [256, 1, 1000000, 33333])
def test_swiglu(L, oracle, n):
    rng = np.random.default_rng(n)
    a, b = (rng.random(n, dtype=np.float32) - 0.5) * 6, rng.random(n, dtype=np.float32) - 0.5
    A, B = tdev(a), tdev(b)
    L.check(L.c.tha_swiglu_f32(stream_ptr(), P(A), P(B), n))
    sync()
    assert_close(A.cpu().numpy(), oracle.swiglu(a, b), 1e-4, "swiglu")


@pytest.mark.parametrize("m,n,k", [(3, 3, 3), (100, 100, 100), (1000, 1000, 1000), (65, 130, 17)])
def test_matmul(L, m, n, k):
    rng = np.random.default_rng(m * n)
    a = rng.random((m, k), dtype=np.float32) - 0.5
    b = rng.random((k, n), dtype=np.float32) - 0.5
    A, B, Cc = tdev(a), tdev(b), tdev(np.zeros((m, n), np.float32))
    L.check(L.c.tha_matmul_f32(stream_ptr(), m, n, k, P(A), P(B), P(Cc)))
    sync()
    want = (a.astype(np.float64) @ b.astype(np.float64))
    assert_close(Cc.cpu().numpy(), want, 1e-3, "matmul")
    assert L.c.tha_matmul_f32(stream_ptr(), 0, n, k, P(A), P(B), P(Cc)) == 2


# thaBLAS_s_matmul at shapes that take the tcgen05 route (gemm_tc.cu matmul_tc_route: B transposed once, A split hi / lo,
# 3xTF32): fp32-level accuracy (norm-wise 2e-5 as test_gemm_tf32x3), ragged tiles in every dimension
@pytest.mark.parametrize("m,n,k", [(1000, 1000, 1000), (256, 4096, 768), (33, 65, 68), (2048, 2048, 2048), (130, 11008, 4096)])
def test_matmul_tensor_cores(L, m, n, k):
    rng = np.random.default_rng(m + n + k)
    a = rng.standard_normal((m, k)).astype(np.float32)
    b = rng.standard_normal((k, n)).astype(np.float32)
    A, B, Cc = tdev(a), tdev(b), tdev(np.full((m, n), 3.0, np.float32))
    for _ in range(2):   # twice: the scratch and the tensor maps are reused
        L.check(L.c.tha_matmul_f32(stream_ptr(), m, n, k, P(A), P(B), P(Cc)))
    sync()
    assert L.c.tha_gemm_check_error() == 0
    import torch
    want = (torch.from_numpy(a).double() @ torch.from_numpy(b).double()).numpy()
    assert_close_scaled(Cc.cpu().numpy(), want, 2e-5, f"matmul {m}x{n}x{k} on the tensor cores")


# the weight matmul: stories110M and 7B shapes, odd shapes through the scalar path
@pytest.mark.parametrize("M,K", [(768, 768), (2048, 768), (768, 2048), (32000, 768), (4096, 11008), (100, 100), (3, 3), (31, 20)])
@pytest.mark.parametrize("B", [1, 3, 8, 13])
def test_matmul_batch(L, oracle, M, K, B):
    if M * K > 4e7 and B > 3:
        pytest.skip("large shape covered at small B")
    rng = np.random.default_rng(M + K + B)
    w = (rng.random((M, K), dtype=np.float32) - 0.5)
    x = (rng.random((B, K), dtype=np.float32) - 0.5)
    W, X, O = tdev(w), tdev(x), tdev(np.zeros((B, M), np.float32))
    L.check(L.c.tha_matmul_batch_f32(stream_ptr(), B, P(O), P(X), P(W), K, M, 0, 0, None, M, K))
    sync()
    got = O.cpu().numpy()
    for b in range(B):
        assert_close(got[b], oracle.matmul(w, x[b]), 1e-3 if K > 100 else 1e-5, f"matmul_batch M{M} K{K} b{b}")


def test_matmul_batch_kv_cache_addressing(L, oracle):
    """Coff / has_pos / pos_d / C_batch_size: rows land straight in the KV cache (src/thaDNN.cpp:45-46)."""
    B, K, M, n_layers, seq = 4, 64, 32, 2, 8
    rng = np.random.default_rng(5)
    w = rng.standard_normal((M, K)).astype(np.float32); x = rng.standard_normal((B, K)).astype(np.float32)
    pos = np.array([0, 3, 7, 2], np.int32)
    cache = np.zeros((B, n_layers, seq, M), np.float32)
    l = 1
    W, X, Cc, Pd = tdev(w), tdev(x), tdev(cache), tdev(pos)
    L.check(L.c.tha_matmul_batch_f32(stream_ptr(), B, P(Cc), P(X), P(W), K, M, l * seq * M, M, P(Pd), n_layers * seq * M, K))
    sync()
    got = Cc.cpu().numpy()
    want = cache.copy()
    for b in range(B):
        want[b, l, pos[b]] = oracle.matmul(w, x[b])
    assert_close(got, want, 1e-5, "kv addressing")
    assert np.count_nonzero(got) == B * M


def test_matmulvec_and_reduction(L, oracle):
    rng = np.random.default_rng(9)
    M, K, N = 96, 160, 5
    a = rng.standard_normal((M, K)).astype(np.float32); b = rng.standard_normal((N, K)).astype(np.float32)
    A, Bm, Cc = tdev(a), tdev(b), tdev(np.zeros((N, M), np.float32))
    L.check(L.c.tha_matmul_reduction_f32(stream_ptr(), P(A), P(Bm), P(Cc), M, N, K))
    L.check(L.c.tha_matmulvec_f32(stream_ptr(), P(Cc), P(Bm), P(A), K, M))   # row 0 again
    sync()
    got = Cc.cpu().numpy()
    for j in range(N):
        assert_close(got[j], oracle.matmul(a, b[j]), 1e-5, "matmul_reduction")


@pytest.mark.parametrize("layout", ["v1", "v2"])
@pytest.mark.parametrize("n_heads,n_kv,hs", [(4, 4, 16), (6, 2, 16), (12, 12, 64)])
def test_mha_three_stage(L, oracle, layout, n_heads, n_kv, hs):
    B, n_layers, seq = 3, 2, 40
    dim, kvd, kv_mul = n_heads * hs, n_kv * hs, n_heads // n_kv
    rng = np.random.default_rng(hs + n_heads)
    pos = np.array([0, 17, 39], np.int32)
    q = rng.standard_normal((B, dim)).astype(np.float32)
    l = 1
    kc = rng.standard_normal((B, n_layers, seq, kvd)).astype(np.float32)
    vc = rng.standard_normal((B, n_layers, seq, kvd)).astype(np.float32)
    Q, Pd = tdev(q), tdev(pos)
    ATT, XB = tdev(np.zeros((B, n_heads, seq), np.float32)), tdev(np.zeros((B, dim), np.float32))
    if layout == "v1":
        KC, VC = tdev(kc), tdev(vc)
        loff = l * seq * kvd
        L.check(L.c.tha_mha_scores_v1_f32(stream_ptr(), B, P(Pd), n_heads, n_layers, P(Q), P(ATT), P(KC), hs, seq, loff, kvd, dim, kv_mul))
        L.check(L.c.tha_mha_softmax_f32(stream_ptr(), B, P(ATT), P(Pd), seq, n_heads))
        L.check(L.c.tha_mha_weighted_v1_f32(stream_ptr(), B, P(Pd), n_heads, P(XB), P(ATT), P(VC), hs, seq, loff, kvd, kv_mul, dim, n_layers))
    else:  # time-major [t][B][kvd], one layer
        KC = tdev(np.ascontiguousarray(kc[:, l].transpose(1, 0, 2)))
        VC = tdev(np.ascontiguousarray(vc[:, l].transpose(1, 0, 2)))
        L.check(L.c.tha_mha_scores_v2_f32(stream_ptr(), B, P(Pd), n_heads, P(Q), P(ATT), P(KC), hs, seq, kvd, dim, kv_mul))
        L.check(L.c.tha_mha_softmax_f32(stream_ptr(), B, P(ATT), P(Pd), seq, n_heads))
        L.check(L.c.tha_mha_weighted_v2_f32(stream_ptr(), B, P(Pd), n_heads, P(XB), P(ATT), P(VC), hs, seq, kvd, kv_mul, dim))
    sync()
    got, gatt = XB.cpu().numpy(), ATT.cpu().numpy()
    for b in range(B):
        want, watt = oracle.attention(q[b], kc[b, l], vc[b, l], int(pos[b]), n_heads, kv_mul, hs)
        assert_close(got[b], want, 1e-5, f"mha out b{b}")
        assert_close(gatt[b, :, :pos[b] + 1], watt[:, :pos[b] + 1], 1e-5, f"mha att b{b}")


@pytest.mark.parametrize("n_heads,n_kv,hs,seq,B", [(12, 12, 64, 1024, 2), (8, 8, 32, 64, 5), (8, 2, 32, 96, 3),
                                                     (32, 32, 128, 256, 1), (4, 4, 32, 2048, 1), (12, 12, 64, 128, 40)])
def test_mha_fused(L, oracle, n_heads, n_kv, hs, seq, B):
    n_layers = 2
    dim, kvd, kv_mul = n_heads * hs, n_kv * hs, n_heads // n_kv
    rng = np.random.default_rng(hs * seq + B)
    pos = rng.integers(0, seq, size=B).astype(np.int32)
    pos[0] = seq - 1
    if B > 1:
        pos[1] = 0
    if B > 2:
        pos[2] = -1  # idle slot: must be left untouched
    q = rng.standard_normal((B, dim)).astype(np.float32)
    kc = rng.standard_normal((B, n_layers, seq, kvd)).astype(np.float32)
    vc = rng.standard_normal((B, n_layers, seq, kvd)).astype(np.float32)
    l = 1
    Q, Pd, KC, VC = tdev(q), tdev(pos), tdev(kc), tdev(vc)
    XB = tdev(np.full((B, dim), 7.0, np.float32))
    for _ in range(2):  # twice: the split counters must reset themselves
        L.check(L.c.tha_mha_fused_v1_f32(stream_ptr(), B, P(Pd), n_heads, n_kv, n_layers, P(Q), P(XB), P(KC), P(VC), hs, seq, l * seq * kvd))
    sync()
    got = XB.cpu().numpy()
    for b in range(B):
        if pos[b] < 0:
            assert np.all(got[b] == 7.0)
            continue
        want, _ = oracle.attention(q[b], kc[b, l], vc[b, l], int(pos[b]), n_heads, kv_mul, hs)
        assert_close(got[b], want, 1e-5, f"fused attention b{b} pos{pos[b]}")


def test_argmax_first_max_wins(L, oracle):
    rng = np.random.default_rng(2)
    B, V = 5, 32000
    lg = rng.standard_normal((B, V)).astype(np.float32)
    lg[1, 77] = lg[1, 31000] = 9.0   # tie: lowest index
    lg[2, 0] = 50.0
    lg[3, V - 1] = 50.0
    LG, N = tdev(lg), tdev(np.zeros(B, np.int32))
    L.check(L.c.tha_argmax_batch_f32(stream_ptr(), B, P(LG), V, P(N)))
    sync()
    assert N.cpu().numpy().tolist() == [oracle.argmax(lg[b]) for b in range(B)]


def test_quantize_q8_bit_exact(L, oracle, golden_dir):
    import os
    g = np.load(os.path.join(golden_dir, "quantize_golden.npz"))
    for k in sorted(k[:-2] for k in g.files if k.endswith("_x")):
        gs = int(k.split("_")[0][2:])
        x = g[k + "_x"]
        X, Q, S = tdev(x), tdev(np.zeros(x.size, np.int8)), tdev(np.zeros(x.size // gs, np.float32))
        L.check(L.c.tha_quantize_q8(stream_ptr(), P(Q), P(S), P(X), x.size, gs))
        sync()
        assert np.array_equal(Q.cpu().numpy(), g[k + "_q"]), k
        assert np.array_equal(S.cpu().numpy().view(np.uint32), g[k + "_s"].view(np.uint32)), k
    rng = np.random.default_rng(4)
    for n, gs in ((768, 64), (2048, 64), (11008, 64), (96, 32)):
        x = rng.standard_normal(n).astype(np.float32)
        x[:gs] = 0.0  # all-zero group
        X, Q, S = tdev(x), tdev(np.zeros(n, np.int8)), tdev(np.zeros(n // gs, np.float32))
        L.check(L.c.tha_quantize_q8(stream_ptr(), P(Q), P(S), P(X), n, gs))
        sync()
        wq, ws = oracle.quantize(x, gs)
        assert np.array_equal(Q.cpu().numpy(), wq) and np.array_equal(S.cpu().numpy(), ws)


@pytest.mark.parametrize("M,K,gs", [(768, 768, 64), (2048, 768, 64), (768, 2048, 64), (4096, 11008, 64), (96, 96, 32), (33, 64, 16)])
@pytest.mark.parametrize("B", [1, 4, 11])
def test_matmul_batch_q8(L, oracle, M, K, gs, B):
    if M * K > 4e7 and B > 4:
        pytest.skip("large shape covered at small B")
    rng = np.random.default_rng(M + K + B)
    w = rng.standard_normal((M, K)).astype(np.float32)
    wq, ws = oracle.quantize(w.reshape(-1), gs)
    wq = wq.reshape(M, K)
    x = rng.standard_normal((B, K)).astype(np.float32)
    xq = np.zeros((B, K), np.int8); xs = np.zeros((B, K // gs), np.float32)
    for b in range(B):
        xq[b], xs[b] = oracle.quantize(x[b], gs)
    WQ, WS, XQ, XS, O = tdev(wq), tdev(ws), tdev(xq), tdev(xs), tdev(np.zeros((B, M), np.float32))
    L.check(L.c.tha_matmul_batch_q8(stream_ptr(), B, P(O), M, P(XQ), P(XS), P(WQ), P(WS), K, M, gs))
    sync()
    got = O.cpu().numpy()
    for b in range(B):
        # integer group sums are exact; only the fp32 accumulation order over the K/gs group terms
        # differs (two rows' lanes add in a tree, the CPU adds left to right): a few ulps of the
        # largest partial sum, so the bound is stated relative to the output scale
        assert_close_scaled(got[b], oracle.matmul_q8(wq, ws, xq[b], xs[b], gs), 4e-6, f"q8 matmul b{b}")


# ---- tensor-core weight matmul (tcgen05, 3xTF32) ------------------------------------------
@pytest.mark.parametrize("M,K,N", [(128, 32, 1), (100, 132, 5), (768, 768, 9), (2304, 768, 33), (768, 2048, 64),
                                   (4096, 768, 128), (1000, 1000, 130), (512, 4096, 257)])
def test_gemm_tf32x3_vs_oracle(L, oracle, M, K, N):
    """out[n, i] = sum_k W[i, k] X[n, k] on the tensor cores against the CPU matmul (run.cc:217-228) at the
    reference's matmul tolerance 1e-3 abs-or-rel (scripts/test/thaBLAS.test.cpp) -- and, tighter, within 5e-5 of the
    largest output (3xTF32 keeps fp32-level accuracy; plain TF32 would sit at ~1e-3)."""
    import torch
    rng = np.random.default_rng(M + K + N)
    w = (rng.random((M, K), dtype=np.float32) - 0.5)
    x = (rng.random((N, K), dtype=np.float32) - 0.5)
    W, X, O = tdev(w), tdev(x), tdev(np.full((N, M), np.nan, np.float32))
    sync()
    L.check(L.c.tha_gemm_tf32x3(stream_ptr(), P(O), M, P(W), M, K, P(X), N, K), "tha_gemm_tf32x3")
    sync()
    assert L.c.tha_gemm_check_error() == 0
    got = O.cpu().numpy()
    want = np.stack([oracle.matmul(w, x[n]) for n in range(min(N, 6))])
    assert_close(got[: want.shape[0]], want, 1e-3, f"gemm {M}x{K}x{N}")
    full = x.astype(np.float64) @ w.astype(np.float64).T
    assert np.abs(got - full).max() <= 5e-5 * np.abs(full).max()


@pytest.mark.parametrize("M,K,N,gs", [(128, 128, 1, 64), (100, 192, 5, 64), (768, 768, 9, 64), (2304, 704, 33, 64), (768, 2048, 64, 64),
                                      (4096, 768, 100, 128), (1000, 1024, 130, 32), (512, 4096, 70, 64), (4096, 11008, 3, 64)])
def test_matmul_q8_int8_tensor_cores_bit_exact(L, oracle, M, K, N, gs):
    """runq.c's matmul (317-342) on the int8 tensor cores: exact int32 group sums, scaled and accumulated in the
    reference's order -- every output BIT-IDENTICAL to the oracle's loop (built with -ffp-contract=off)."""
    rng = np.random.default_rng(M + K + N)
    wq, ws = oracle.quantize(rng.standard_normal(M * K).astype(np.float32), gs)
    wq = wq.reshape(M, K)
    x = rng.standard_normal((N, K)).astype(np.float32) * rng.uniform(0.1, 4.0, size=(N, 1)).astype(np.float32)
    xq = np.zeros((N, K), np.int8); xs = np.zeros((N, K // gs), np.float32)
    for n in range(N):
        xq[n], xs[n] = oracle.quantize(x[n], gs)
    ldxs = ((N + 63) // 64) * 64
    xsT = np.zeros((K // gs, ldxs), np.float32)
    xsT[:, :N] = xs.T
    WQ, WS, XQ, XST, O = tdev(wq), tdev(ws), tdev(xq), tdev(xsT), tdev(np.full((N, M), np.nan, np.float32))
    sync()
    L.check(L.c.tha_matmul_q8_tc(stream_ptr(), P(O), M, P(WQ), P(WS), gs, M, K, P(XQ), P(XST), ldxs, N), "tha_matmul_q8_tc")
    sync()
    assert L.c.tha_gemm_check_error() == 0
    got = O.cpu().numpy()
    for n in range(min(N, 5)):
        want = oracle.matmul_q8(wq, ws, xq[n], xs[n], gs)
        assert np.array_equal(got[n].view(np.uint32), want.view(np.uint32)), \
            f"row {n}: {np.count_nonzero(got[n] != want)} of {M} outputs differ, max |d| {np.abs(got[n] - want).max()}"
    # all rows: against the exact group arithmetic in fp64
    full = np.einsum("mgk,ngk->nmg", wq.reshape(M, -1, gs).astype(np.float64), xq.reshape(N, -1, gs).astype(np.float64))
    full = (full * ws.reshape(M, -1).astype(np.float64)[None] * xs.astype(np.float64)[:, None, :]).sum(-1)
    assert np.abs(got - full).max() <= 2e-6 * np.abs(full).max()


def test_gemm_tf32x3_rejects_bad_shapes(L):
    W, X, O = tdev(np.zeros((8, 30), np.float32)), tdev(np.zeros((2, 30), np.float32)), tdev(np.zeros((2, 8), np.float32))
    assert L.c.tha_gemm_tf32x3(stream_ptr(), P(O), 8, P(W), 8, 30, P(X), 2, 30) == 7      # K % 4 != 0: NOT_SUPPORTED
    assert L.c.tha_gemm_tf32x3(stream_ptr(), P(O), 8, P(W), 0, 32, P(X), 2, 32) == 2      # M < 1: invalid argument


@pytest.mark.parametrize("M,K,N", [(128, 64, 1), (300, 768, 5), (768, 2048, 16), (2304, 768, 40), (4096, 4096, 64),
                                   (1000, 11008, 130), (4096, 4096, 300), (32000, 768, 256), (640, 4096, 2048), (1100, 768, 600),
                                   (4096, 4096, 1027)])   # the last three: 256-row work items (N >= 512), ragged in rows and columns
def test_gemm_bf16_tensor_cores(M, K, N):
    """tha_gemm_bf16 (tcgen05.mma.kind::f16, both operands from TMA-written shared memory, double-buffered TMEM
    accumulator) against a float64 product of the same bf16 values: only fp32 accumulation order differs."""
    import torch
    lib = load_pkg("cabi").lib()
    g = torch.Generator(device="cuda"); g.manual_seed(M + K + N)
    W = (torch.randn(M, K, device="cuda", generator=g) * 0.05).to(torch.bfloat16)
    X = torch.randn(N, K, device="cuda", generator=g).to(torch.bfloat16)
    out = torch.zeros(N, M, device="cuda")
    lib.check(lib.c.tha_gemm_bf16(stream_ptr(), P(out), M, P(W), M, K, P(X), N, K), "tha_gemm_bf16")
    torch.cuda.synchronize()
    assert lib.c.tha_gemm_check_error() == 0
    want = (X.double() @ W.double().t())
    err = (out.double() - want).abs().max().item()
    scale = want.abs().max().item()
    assert err <= 2e-5 * scale, f"max err {err:.3e} vs scale {scale:.3e}"   # fp32 accumulation in the tensor core over up to 11008 terms
