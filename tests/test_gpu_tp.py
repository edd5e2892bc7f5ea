"""Tensor-parallel step (tha_tp_*, SURVEY.md 8e "Optional TP") against the oracle of the UNSPLIT model.

The ranks of one replica normally sit on different GPUs; the exchange only needs every rank to be able to address its
peers' exchange blocks, so on a single-GPU box the same code runs with all ranks on cuda:0, one stream per rank (the
kernels of different ranks overlap and hand-shake through the same flags).  PDL is switched off for that layout: an
early-launched persistent matmul of one rank would otherwise hold every SM's shared memory while it waits for a peer
that needs an SM to make progress.  Multi-GPU runs (tools/tp_check.py under torchrun, bench.py --tp) keep PDL on."""
import os

import numpy as np
import pytest

from gpu_util import assert_close, assert_close_scaled, small_model_files
from pkgload import load_pkg

pytestmark = pytest.mark.gpu

TP_MODELS = {
    # hs 32: tp=2 -> 4 local heads (attention width 128); hidden 704 = 11 groups of 64 -> 6 + 5 (uneven)
    "tpA": dict(dim=256, hidden_dim=704, n_layers=3, n_heads=8, n_kv_heads=8, vocab_size=512, seq_len=48),
    # GQA: 4 kv heads -> 2 per rank
    "tpGQA": dict(dim=256, hidden_dim=640, n_layers=2, n_heads=8, n_kv_heads=4, vocab_size=320, seq_len=40),
    # hs 64, tp=4 -> 2 local heads each; hidden 1408 = 22 groups -> 6,6,5,5
    "tp4": dict(dim=512, hidden_dim=1408, n_layers=2, n_heads=8, n_kv_heads=8, vocab_size=384, seq_len=40),
}


def make_ranks(cfg, w, tp, B, q8, max_rows=None):
    ck, eng, tpm = load_pkg("checkpoint"), load_pkg("engine"), load_pkg("tp")
    ranks = []
    if q8:
        gs = ck.v2_group_size(cfg)
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
        norms = {k: w[k] for k in ("rms_att", "rms_ffn", "rms_final")}
    for r in range(tp):
        if q8:
            lcfg, lqt = tpm.shard_q8(cfg, gs, qt, tp, r)
            dw = eng.DeviceWeightsQ8(lcfg, gs, norms, lqt)
        else:
            lcfg, lw = tpm.shard_f32(cfg, w, tp, r)
            dw = eng.DeviceWeights(lcfg, lw)
        ranks.append(tpm.TpRank(cfg, dw, tp, r, B, max_rows=max_rows))
    tpm.connect_in_process(ranks)
    return ranks


_STREAMS = []


def rank_streams(tp):
    """One stream per rank, created once (back to back: distinct hardware queues, see conftest.py) and reused."""
    import torch
    while len(_STREAMS) < tp:
        _STREAMS.append(torch.cuda.Stream())
    return _STREAMS[:tp]


def step_all(ranks, streams, call):
    """One call on every rank of a replica whose ranks share this process (and here: one device).  First every rank
    PREPARES (scratch, module loads, graph instantiation -- anything that may synchronise with the device), then the
    real calls go out back to back: while a rank's kernels wait for a peer, the host does nothing but enqueue."""
    import torch
    for r, rk in enumerate(ranks):
        rk.set_mode(True)
        call(r, streams[r])
        rk.set_mode(False)
    torch.cuda.synchronize()
    for r in range(len(ranks)):
        call(r, streams[r])
    torch.cuda.synchronize()
    for r, rk in enumerate(ranks):
        assert not rk.check_error(), f"rank {r}: a wait on a peer timed out"


@pytest.fixture
def no_pdl():
    lib = load_pkg("cabi").lib()
    lib.c.tha_set_option(b"pdl", 0)
    yield lib
    lib.c.tha_set_option(b"pdl", 1)


@pytest.mark.parametrize("name,tp", [("tpA", 2), ("tpGQA", 2), ("tp4", 4)])
@pytest.mark.parametrize("q8", [False, True])
@pytest.mark.parametrize("B", [3, 12])
def test_tp_decode_vs_oracle(oracle, tmp_path, no_pdl, name, tp, q8, B):
    import torch
    cfg, w, v0, v2 = small_model_files(str(tmp_path), name, TP_MODELS[name])
    m = oracle.open(v2 if q8 else v0, quantized=q8)
    ranks = make_ranks(cfg, w, tp, B, q8)
    streams = rank_streams(tp)
    seqs = [m.new_seq() for _ in range(B)]
    rng = np.random.default_rng(B + tp)
    n_steps = 10
    logits = [torch.zeros(B, cfg.vocab_size, device="cuda") for _ in range(tp)]
    nxt = [torch.zeros(B, dtype=torch.int32, device="cuda") for _ in range(tp)]
    try:
        for step in range(n_steps):
            tok = rng.integers(0, cfg.vocab_size, size=B).astype(np.int32)
            pos = np.array([step - (b % 3) for b in range(B)], np.int32)   # ragged: slots join at steps 0, 1, 2
            pos[pos < 0] = -1
            tok_d = [torch.from_numpy(tok).cuda() for _ in range(tp)]
            pos_d = [torch.from_numpy(pos).cuda() for _ in range(tp)]
            step_all(ranks, streams, lambda r, st: ranks[r].decode_step(tok_d[r], pos_d[r], logits[r], nxt[r], stream=st))
            got = [l.cpu().numpy() for l in logits]
            for r in range(1, tp):   # every rank sums the same partials in the same order
                act = pos >= 0
                assert np.array_equal(got[0][act], got[r][act]), f"step {step}: rank {r} differs from rank 0"
            for b in range(B):
                if pos[b] < 0:
                    continue
                want = seqs[b].forward(int(tok[b]), int(pos[b]))
                if q8:
                    assert_close_scaled(got[0][b], want, 5e-2, f"tp{tp} q8 step {step} slot {b}")
                else:
                    assert_close(got[0][b], want, 1e-4, f"tp{tp} step {step} slot {b}")
                    assert int(nxt[0][b]) == int(want.argmax())
        assert load_pkg("cabi").lib().c.tha_gemm_check_error() == 0
        # the local KV shards, side by side, are the unsplit cache (fp32: 1e-4)
        if not q8:
            hs = cfg.dim // cfg.n_heads
            kvd_l = hs * cfg.n_kv_heads // tp
            for b in range(min(B, 3)):
                n = n_steps - (b % 3)
                want = seqs[b].buf("key_cache").reshape(m.n_layers, m.seq_len, m.kv_dim)[:, :n]
                have = np.concatenate([ranks[r].key_cache[b].cpu().numpy()[:, :n] for r in range(tp)], axis=2)
                assert have.shape[2] == kvd_l * tp
                assert_close(have, want, 1e-4, f"key cache slot {b}")
    finally:
        for r in ranks:
            r.close()


@pytest.mark.parametrize("q8", [False, True])
def test_tp_prefill_then_decode(oracle, tmp_path, no_pdl, q8):
    """Prompt ingestion through the split model, then greedy decode steps with the device-side slot bookkeeping."""
    import torch
    name, tp, B = "tpA", 2, 3
    cfg, w, v0, v2 = small_model_files(str(tmp_path), name, TP_MODELS[name])
    m = oracle.open(v2 if q8 else v0, quantized=q8)
    ranks = make_ranks(cfg, w, tp, B, q8, max_rows=64)
    streams = rank_streams(tp)
    rng = np.random.default_rng(5)
    lens = [5, 17, 30]
    prompts = [rng.integers(0, cfg.vocab_size, size=n).astype(np.int32) for n in lens]
    seqs = [m.new_seq() for _ in range(B)]
    want_last = []
    for b in range(B):
        for p_, t in enumerate(prompts[b]):
            lg = seqs[b].forward(int(t), p_)
        want_last.append(lg.copy())
    logits = [torch.zeros(B, cfg.vocab_size, device="cuda") for _ in range(tp)]
    nxt = [torch.zeros(B, dtype=torch.int32, device="cuda") for _ in range(tp)]
    try:
        step_all(ranks, streams, lambda r, st: ranks[r].prefill([(b, 0, prompts[b]) for b in range(B)], logits[r], nxt[r], stream=st))
        got = logits[0].cpu().numpy()
        assert np.array_equal(got, logits[1].cpu().numpy())
        for b in range(B):
            (assert_close_scaled if q8 else assert_close)(got[b], want_last[b], 5e-2 if q8 else 1e-4, f"prefill slot {b}")
        if q8:
            return
        # greedy continuation on the device: token = argmax, pos += 1 (tha_advance_slots inside the step)
        tok_d = [n.clone() for n in nxt]
        pos_d = [torch.tensor(lens, dtype=torch.int32, device="cuda") for _ in range(tp)]
        for step in range(4):
            want_tok = [int(want_last[b].argmax()) for b in range(B)]
            assert tok_d[0].cpu().tolist() == want_tok
            for b in range(B):
                want_last[b] = seqs[b].forward(want_tok[b], lens[b] + step).copy()
            step_all(ranks, streams, lambda r, st: ranks[r].decode_step(tok_d[r], pos_d[r], logits[r], nxt[r], advance=True, stream=st))
            for b in range(B):
                assert_close(logits[0][b].cpu().numpy(), want_last[b], 1e-4, f"decode after prefill step {step} slot {b}")
        assert pos_d[1].cpu().tolist() == [n + 4 for n in lens]
    finally:
        for r in ranks:
            r.close()


def test_tp_on_two_real_devices():
    """The same parity check with one PROCESS PER GPU (tools/tp_check.py under torchrun: IPC-mapped exchange blocks, PDL on,
    ranks on different devices), fp32 and Q8_0, 3 and 20 slots, against the oracle of the unsplit model; every rank must
    also agree with rank 0 bit for bit.  Needs two devices: skipped on a one-GPU box."""
    import subprocess
    import sys
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", "29533", os.path.join(root, "tools", "tp_check.py")]
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=600)
    assert r.returncode == 0 and "TP_CHECK PASS" in r.stdout, r.stdout[-3000:]
