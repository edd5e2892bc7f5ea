"""Paged KV cache (what the reference's BatchManager stub -- include/BatchManager.hpp:11-52 --
gestures at): K/V rows in blocks of 16 positions, found through a per-slot block table by the attention, prefill-attention
and RoPE/scatter kernels.  With a deliberately FRAGMENTED table (blocks of one slot scattered over the pool, interleaved
with other slots', out of order) the results must be what the dense slab gives: logits against the oracle at the usual
tolerance, and every cache row found through the table equal to the oracle's row."""
import os

import numpy as np
import pytest

from gpu_util import assert_close, assert_close_scaled, small_model_files
from pkgload import load_pkg

pytestmark = pytest.mark.gpu
MODEL = dict(dim=256, hidden_dim=704, n_layers=3, n_heads=8, n_kv_heads=4, vocab_size=512, seq_len=48)


def fragmented_table(B, nrow, n_blocks, seed):
    """Every (slot, row block) gets its own block, drawn from a shuffled pool larger than needed."""
    rng = np.random.default_rng(seed)
    ids = rng.permutation(n_blocks)[: B * nrow]
    return ids.reshape(nrow, B).T.copy()          # consecutive blocks of a slot are B apart in the shuffle


@pytest.mark.parametrize("kind", ["f32", "q8", "bf16"])
@pytest.mark.parametrize("B", [3, 12])
def test_paged_decode_and_prefill_vs_oracle(oracle, tmp_path, kind, B):
    import torch
    ck, eng = load_pkg("checkpoint"), load_pkg("engine")
    cfg, w, v0, v2 = small_model_files(str(tmp_path), "paged", MODEL)
    if kind == "f32":
        dw = eng.DeviceWeights(cfg, ck.read_v0(v0)[1])
        m = oracle.open(v0)
    elif kind == "q8":
        c, gs, norms, qt, shared = ck.read_v2(v2)
        dw = eng.DeviceWeightsQ8(c, gs, norms, qt)
        m = oracle.open(v2, quantized=True)
    else:
        wb = ck.bf16_weights(w)
        pb = str(tmp_path / "paged_bf16.bin")
        ck.write_v0(pb, cfg, wb)
        dw = eng.DeviceWeightsBF16(cfg, wb)
        m = oracle.open(pb, bf16=True)
    paged = eng.PagedKV(cfg, B, n_blocks=B * 3 + 11)
    paged.set_table(fragmented_table(B, paged.nrow, paged.n_blocks, B))
    e = eng.Engine(dw, B, paged=paged)
    check = assert_close if kind == "f32" else assert_close_scaled
    eps = 1e-4 if kind == "f32" else 5e-2
    rng = np.random.default_rng(B)
    seqs = [m.new_seq() for _ in range(B)]
    # slots 0.. start with prompts of different lengths ingested in one pass (crossing block boundaries), then all decode
    lens = [(5 + 7 * b) % 30 + 1 for b in range(B)]
    prompts = [rng.integers(0, cfg.vocab_size, size=n).astype(np.int32) for n in lens]
    for b in range(B):
        for p_, t in enumerate(prompts[b]):
            seqs[b].forward(int(t), p_)
    logits = torch.zeros(B, cfg.vocab_size, device="cuda")
    e.prefill([(b, 0, prompts[b]) for b in range(B)], logits, None)
    torch.cuda.synchronize()
    got = logits.cpu().numpy()
    for b in range(B):
        check(got[b], seqs[b].buf("logits"), eps, f"{kind} paged prefill logits slot {b}")
    for step in range(14):
        tok = rng.integers(0, cfg.vocab_size, size=B).astype(np.int32)
        pos = np.array([lens[b] + step for b in range(B)], np.int32)
        lg = e.forward_batch(tok, pos).numpy()
        for b in range(B):
            check(lg[b], seqs[b].forward(int(tok[b]), int(pos[b])), eps, f"{kind} paged decode slot {b} step {step}")
    assert load_pkg("cabi").lib().c.tha_gemm_check_error() == 0
    for b in range(B):
        n = lens[b] + 14
        for cache in ("key_cache", "value_cache"):
            want = seqs[b].buf(cache).reshape(m.n_layers, m.seq_len, m.kv_dim)[:, :n]
            check(paged.rows(cache, b, n), want, eps, f"{kind} {cache} slot {b} through the block table")
    paged.close()


def test_llama_app_paged_matches_dense(tmp_path):
    """./build/apps/llama -c <blocks>: BatchManager::create_pool + the driver's admission control; greedy token streams
    must equal the dense run's, also when the pool is too small for all slots at once (requests wait for blocks)."""
    import test_host_cpp as T
    load_pkg("build").build_all()
    cfg, w, v0, v2 = small_model_files(str(tmp_path), "app", dict(dim=256, hidden_dim=704, n_layers=2, n_heads=8, n_kv_heads=4,
                                                                  vocab_size=32000, seq_len=48), std=0.1)

    def go(tag, path, extra):
        out_txt, out_tok = str(tmp_path / f"{tag}.txt"), str(tmp_path / f"{tag}.tok")
        stdout = T.run([T.LLAMA, path, "-m", "test", "-f", os.path.join(T.ASSETS, "gen_in_8.txt"), "-o", out_txt, "-k", out_tok, "-b", "4",
                        "-t", "0", "-z", T.TOKENIZER, "-g", "1"] + extra)
        return open(out_tok).read(), stdout

    for path in (v0, v2):
        dense, _ = go("dense", path, [])
        paged, so = go("paged", path, ["-c", "12"])     # 4 slots x 3 blocks: everything fits
        assert "paged KV cache: 12 blocks" in so
        assert paged == dense
        tight, _ = go("tight", path, ["-c", "7"])       # fewer blocks than 4 full contexts: admission control at work
        assert tight == dense
