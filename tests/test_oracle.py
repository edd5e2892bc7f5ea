"""Pin the oracle (oracle/llama2_oracle.c): bit-exact against (a) the committed golden
vectors, which were produced by the reference's own exporter + run.cc / runq.c
(tests/golden/make_golden.py), and (b) oracle/_ref (reference sources compiled
unmodified) when those binaries are present."""
import os

import numpy as np
import pytest

from oracle_lib import Oracle, Ref32, RefQ8, ref_available

ARCHS = ["tinyA", "tinyB_gqa"]


@pytest.mark.parametrize("name", ARCHS)
def test_fp32_forward_matches_golden(oracle, golden_dir, name):
    g = np.load(os.path.join(golden_dir, f"{name}_golden.npz"))
    m = oracle.open(os.path.join(golden_dir, f"{name}_v0.bin"))
    s = m.new_seq()
    for pos, tok in enumerate(g["tokens"]):
        logits = s.forward(tok, pos)
        assert np.array_equal(logits, g["logits_fp32"][pos]), f"step {pos}"
    assert np.array_equal(s.buf("key_cache"), g["key_cache_fp32"])
    assert np.array_equal(s.buf("value_cache"), g["value_cache_fp32"])


@pytest.mark.parametrize("name", ARCHS)
def test_q8_forward_matches_golden(oracle, golden_dir, name):
    g = np.load(os.path.join(golden_dir, f"{name}_golden.npz"))
    m = oracle.open(os.path.join(golden_dir, f"{name}_v2.bin"), quantized=True)
    assert m.group_size == int(g["group_size"])
    s = m.new_seq()
    for pos, tok in enumerate(g["tokens"]):
        logits = s.forward(tok, pos)
        assert np.array_equal(logits, g["logits_q8"][pos]), f"step {pos}"
    assert np.array_equal(s.buf("key_cache"), g["key_cache_q8"])
    assert np.array_equal(s.buf("value_cache"), g["value_cache_q8"])
    xq, xs = s.qbuf(0)
    assert np.array_equal(xq, g["last_xq"]) and np.array_equal(xs, g["last_xs"])
    hq, hs = s.qbuf(1)
    assert np.array_equal(hq, g["last_hq"]) and np.array_equal(hs, g["last_hs"])
    # the quantiser alone, on the fp32 vector the reference quantised last
    q2, s2 = oracle.quantize(g["last_hb"], m.group_size)
    assert np.array_equal(q2, g["last_hq"]) and np.array_equal(s2, g["last_hs"])


def test_quantize_known_answers(oracle, golden_dir):
    g = np.load(os.path.join(golden_dir, "quantize_golden.npz"))
    keys = sorted(k[:-2] for k in g.files if k.endswith("_x"))
    assert len(keys) == 12
    for k in keys:
        gs = int(k.split("_")[0][2:])
        q, s = oracle.quantize(g[k + "_x"], gs)
        assert np.array_equal(q, g[k + "_q"]), k
        assert np.array_equal(s.view(np.uint32), g[k + "_s"].view(np.uint32)), k


def test_quantize_zero_group_is_pinned_to_zero(oracle):
    x = np.zeros(128, np.float32)
    x[64:] = np.linspace(-1, 1, 64, dtype=np.float32)
    q, s = oracle.quantize(x, 64)
    assert s[0] == 0.0 and not q[:64].any()
    assert q[64] == -127 and q[127] == 127


def test_argmax_first_max_wins(oracle):
    p = np.array([0.1, 0.7, 0.7, 0.2], np.float32)
    assert oracle.argmax(p) == 1


@pytest.mark.skipif(not ref_available("fp32"), reason="oracle/_ref not built")
def test_ops_bit_exact_vs_reference_build(oracle):
    ref = Ref32()
    rng = np.random.default_rng(0)
    for n in (1, 111, 768, 11111):
        x = rng.standard_normal(n).astype(np.float32); w = rng.standard_normal(n).astype(np.float32)
        o = np.empty_like(x)
        ref.lib.ref32_rmsnorm(o.ctypes.data_as(ref.lib.ref32_rmsnorm.argtypes[0]), x.ctypes.data_as(ref.lib.ref32_rmsnorm.argtypes[0]),
                              w.ctypes.data_as(ref.lib.ref32_rmsnorm.argtypes[0]), n)
        assert np.array_equal(o, oracle.rmsnorm(x, w))
        y = x.copy()
        ref.lib.ref32_softmax(y.ctypes.data_as(ref.lib.ref32_softmax.argtypes[0]), n)
        assert np.array_equal(y, oracle.softmax(x))
    for d, n in ((3, 3), (100, 100), (768, 2048)):
        w = rng.standard_normal((d, n)).astype(np.float32); x = rng.standard_normal(n).astype(np.float32)
        o = np.empty(d, np.float32)
        P = ref.lib.ref32_matmul.argtypes[0]
        ref.lib.ref32_matmul(o.ctypes.data_as(P), x.ctypes.data_as(P), w.ctypes.data_as(P), n, d)
        assert np.array_equal(o, oracle.matmul(w, x))


@pytest.mark.skipif(not ref_available("q8"), reason="oracle/_ref not built")
def test_q8_ops_bit_exact_vs_reference_build(oracle):
    ref = RefQ8()
    rng = np.random.default_rng(1)
    for gs in (32, 64):
        x = rng.standard_normal(gs * 24).astype(np.float32)
        q, s = ref.quantize(x, gs)
        q2, s2 = oracle.quantize(x, gs)
        assert np.array_equal(q, q2) and np.array_equal(s, s2)
        w = rng.standard_normal((40, x.size)).astype(np.float32)
        wq, ws = oracle.quantize(w.reshape(-1), gs)
        wq = wq.reshape(w.shape)
        assert np.array_equal(ref.matmul(wq, ws, q, s, gs), oracle.matmul_q8(wq, ws, q, s, gs))


@pytest.mark.skipif(not ref_available("fp32"), reason="oracle/_ref not built")
@pytest.mark.parametrize("name", ARCHS)
def test_forward_bit_exact_vs_reference_build(oracle, golden_dir, name):
    """Live run of the compiled reference on a token stream the fixtures do not contain."""
    ref = Ref32()
    path = os.path.join(golden_dir, f"{name}_v0.bin")
    h, cfg = ref.open(path)
    m = oracle.open(path)
    assert cfg == [m.dim, m.hidden_dim, m.n_layers, m.n_heads, m.n_kv_heads, m.vocab_size, m.seq_len]
    s = m.new_seq()
    rng = np.random.default_rng(99)
    for pos, tok in enumerate(rng.integers(0, m.vocab_size, size=m.seq_len)):
        assert np.array_equal(ref.forward(h, tok, pos, m.vocab_size), s.forward(tok, pos)), pos
