"""Host logic of the tensor-parallel split (hip_llama.cpp_b200/tp.py): the shards of every rank, put back together,
are the unsplit tensors; partial wo / w2 products sum to the unsplit product; Q8_0 column splits never cut a group."""
import numpy as np
import pytest

from pkgload import load_pkg


def test_split_hidden_whole_groups():
    tpm = load_pkg("tp")
    assert tpm.split_hidden(11008, 8, 64) == [1408] * 4 + [1344] * 4      # 172 groups -> 22,22,22,22,21,21,21,21
    assert tpm.split_hidden(11008, 4, 64) == [2752] * 4
    assert tpm.split_hidden(11008, 2, 64) == [5504] * 2
    assert sum(tpm.split_hidden(704, 2, 64)) == 704
    with pytest.raises(ValueError):
        tpm.split_hidden(700, 2, 64)


@pytest.mark.parametrize("tp", [2, 4])
def test_f32_shards_reassemble_and_partials_sum(tp):
    ck, tpm = load_pkg("checkpoint"), load_pkg("tp")
    cfg = ck.Config(dim=256, hidden_dim=704 if tp == 2 else 1024, n_layers=2, n_heads=8, n_kv_heads=4, vocab_size=64, seq_len=16)
    w = ck.random_weights(cfg, seed=1, std=0.1)
    shards = [tpm.shard_f32(cfg, w, tp, r) for r in range(tp)]
    assert np.array_equal(np.concatenate([s[1]["wq"] for s in shards], axis=1), w["wq"])
    assert np.array_equal(np.concatenate([s[1]["wk"] for s in shards], axis=1), w["wk"])
    assert np.array_equal(np.concatenate([s[1]["wo"] for s in shards], axis=2), w["wo"])
    assert np.array_equal(np.concatenate([s[1]["w1"] for s in shards], axis=1), w["w1"])
    assert np.array_equal(np.concatenate([s[1]["w2"] for s in shards], axis=2), w["w2"])
    for lcfg, lw in shards:
        assert lcfg.dim == cfg.dim and lcfg.n_heads == cfg.n_heads // tp and lcfg.n_kv_heads == cfg.n_kv_heads // tp
        assert lw["w1"].shape[1] == lcfg.hidden_dim == lw["w2"].shape[2]
    # what the exchange sums: per-rank partial products of the column-split matrices
    rng = np.random.default_rng(0)
    xb = rng.standard_normal(cfg.dim).astype(np.float32)
    hs = cfg.dim // cfg.n_heads
    da = hs * cfg.n_heads // tp
    part = sum(shards[r][1]["wo"][0].astype(np.float64) @ xb[r * da:(r + 1) * da] for r in range(tp))
    assert np.allclose(part, w["wo"][0].astype(np.float64) @ xb, rtol=1e-12, atol=1e-12)


def test_q8_shards_keep_groups():
    ck, tpm = load_pkg("checkpoint"), load_pkg("tp")
    cfg = ck.Config(dim=256, hidden_dim=704, n_layers=2, n_heads=8, n_kv_heads=8, vocab_size=64, seq_len=16)
    w = ck.random_weights(cfg, seed=2, std=0.1)
    gs = 64
    qt = {}
    for n in ck.V2_QORDER:
        t = w[n]
        if n == "tok_emb":
            q, s = ck.quantize_q80(t, gs)
            qt[n] = (q.reshape(t.shape), s)
        else:
            qs, ss = zip(*[ck.quantize_q80(t[l], gs) for l in range(cfg.n_layers)])
            qt[n] = (np.stack(qs).reshape(t.shape), np.stack(ss))
    qt["wcls"] = qt["tok_emb"]
    tp = 2
    sh = [tpm.shard_q8(cfg, gs, qt, tp, r) for r in range(tp)]
    # dequantised shards, reassembled, equal the dequantised unsplit tensor
    def deq(q, s):
        L = q.shape[0]
        return (q.reshape(L, -1, gs).astype(np.float32) * np.asarray(s).reshape(L, -1, 1)).reshape(q.shape)
    for name, axis in (("wq", 1), ("wo", 2), ("w1", 1), ("w2", 2)):
        full = deq(*qt[name])
        got = np.concatenate([deq(*sh[r][1][name]) for r in range(tp)], axis=axis)
        assert np.array_equal(got, full), name
    assert [s[0].hidden_dim for s in sh] == [384, 320]


def _tp_forward_numpy(cfg, shards, tokens):
    """The arithmetic of the tensor-parallel step in numpy float64: every rank works on its
    shard with the LOCAL config, the wo / w2 partial products of all ranks are summed into one residual stream.
    Returns the logits after each token (run.cc:244-374 semantics)."""
    tp = len(shards)
    d, L, S = cfg.dim, cfg.n_layers, cfg.seq_len
    hs = cfg.dim // cfg.n_heads
    w0 = shards[0][1]
    kc = [np.zeros((L, S, hs * sh[0].n_kv_heads)) for sh in shards]
    vc = [np.zeros((L, S, hs * sh[0].n_kv_heads)) for sh in shards]
    out = []

    def rms(x, g):
        return g * (x / np.sqrt((x * x).mean() + 1e-5))

    def rope(v, pos):
        v = v.copy()
        for i in range(0, v.size, 2):
            f = 1.0 / (10000.0 ** ((i % hs) / hs))
            c, s = np.cos(pos * f), np.sin(pos * f)
            v[i], v[i + 1] = v[i] * c - v[i + 1] * s, v[i] * s + v[i + 1] * c
        return v

    for pos, tok in enumerate(tokens):
        x = w0["tok_emb"][tok].astype(np.float64)
        for l in range(L):
            xb = rms(x, w0["rms_att"][l])
            part = np.zeros(d)
            for r, (lc, lw) in enumerate(shards):
                q = rope(lw["wq"][l].astype(np.float64) @ xb, pos)
                kc[r][l, pos] = rope(lw["wk"][l].astype(np.float64) @ xb, pos)
                vc[r][l, pos] = lw["wv"][l].astype(np.float64) @ xb
                kv_mul = lc.n_heads // lc.n_kv_heads
                att_out = np.zeros(hs * lc.n_heads)
                for h in range(lc.n_heads):
                    kh = kc[r][l, :pos + 1, (h // kv_mul) * hs:(h // kv_mul + 1) * hs]
                    vh = vc[r][l, :pos + 1, (h // kv_mul) * hs:(h // kv_mul + 1) * hs]
                    sc = kh @ q[h * hs:(h + 1) * hs] / np.sqrt(hs)
                    p = np.exp(sc - sc.max()); p /= p.sum()
                    att_out[h * hs:(h + 1) * hs] = p @ vh
                part += lw["wo"][l].astype(np.float64) @ att_out        # this rank's partial of the residual update
            x = x + part                                                # the exchange: every rank adds all partials
            xb = rms(x, w0["rms_ffn"][l])
            part = np.zeros(d)
            for lc, lw in shards:
                a = lw["w1"][l].astype(np.float64) @ xb
                c = lw["w3"][l].astype(np.float64) @ xb
                part += lw["w2"][l].astype(np.float64) @ (a / (1.0 + np.exp(-a)) * c)
            x = x + part
        out.append(w0["tok_emb"].astype(np.float64) @ rms(x, w0["rms_final"]))
    return out


@pytest.mark.parametrize("tp,kv", [(2, 8), (2, 4), (4, 4)])
def test_tp_arithmetic_matches_unsplit_oracle(oracle, tmp_path, tp, kv):
    """The split itself (which rows / columns / heads / hidden units go where, grouped-query heads, uneven hidden
    widths) pinned on the CPU: the sharded arithmetic reproduces the oracle's logits of the UNSPLIT model."""
    import os
    ck, tpm = load_pkg("checkpoint"), load_pkg("tp")
    cfg = ck.Config(dim=128, hidden_dim=448, n_layers=2, n_heads=8, n_kv_heads=kv, vocab_size=96, seq_len=12)
    w = ck.random_weights(cfg, seed=11, std=0.1)
    rng = np.random.default_rng(3)
    for k in ("rms_att", "rms_ffn", "rms_final"):
        w[k] = (1.0 + 0.1 * rng.standard_normal(w[k].shape)).astype(np.float32)
    path = os.path.join(str(tmp_path), "m.bin")
    ck.write_v0(path, cfg, w)
    seq = oracle.open(path).new_seq()
    shards = [tpm.shard_f32(cfg, w, tp, r) for r in range(tp)]
    assert sum(s[0].hidden_dim for s in shards) == cfg.hidden_dim       # 448 = 7 groups of 64: uneven at tp = 2 and 4
    tokens = rng.integers(0, cfg.vocab_size, size=6)
    got = _tp_forward_numpy(cfg, shards, tokens)
    for pos, tok in enumerate(tokens):
        want = seq.forward(int(tok), pos)
        assert np.allclose(got[pos], want, rtol=2e-4, atol=2e-4), f"tp={tp} kv={kv} pos {pos}: {np.abs(got[pos] - want).max()}"
