"""The bf16 model path (csrc/step_bf16.cu, gemm_bf16.cu; BASELINE.json configs[3] "fp32/bf16") against the oracle's
restatement of the same rule: bf16-rounded weights, every vector entering a weight matmul rounded to bf16 (nearest even),
everything else fp32 as run.cc:244-374.  GPU and oracle then differ by fp32 accumulation order (~1e-7), plus -- now and
then -- a value that sits on a bf16 rounding boundary and lands on the other side: that element moves by 2^-8 of its
magnitude, which on these small random networks is worth a few 1e-3 of the largest logit, and the K/V rows written after it
keep the difference for the rest of the sequence (the Q8_0 quantiser shows the same effect, tests/test_gpu_baseline_shapes.py).
How often that happens depends on how far the value being rounded is from the oracle's: ~1e-7 of its scale on the CUDA-core
GEMV path (1-8 slots), ~1e-6 .. 5e-6 after a tensor-core matmul (tcgen05 accumulates the 16 products of an instruction with
less than fp32 alignment: tests/test_gpu_ops.py::test_gemm_bf16_tensor_cores measures 5e-6 at K = 4096), i.e. a flipped
rounding every few tokens.  Stated here: every row within CAP (not rounding at all costs 5e-3 .. 7e-3, tools/bf16_debug.py;
CAP is three such steps) with the median within half of it; on one-layer models, where few roundings can flip and no
earlier layer's K/V rows carry a flip forward, most rows agree to fp32 noise; the oracle's own sensitivity (every norm gain
moved by -1 / 0 / +1 ulp: it produces the same 3e-3 .. 7e-3 jumps) is printed beside each result."""
import numpy as np
import pytest

from gpu_util import assert_close_scaled, small_model_files
from pkgload import load_pkg

pytestmark = pytest.mark.gpu
TIGHT = 1e-5
CAP = 2e-2
BF16_EPS = CAP


def floor_model(oracle, tmp_path, cfg, wb, tag):
    """The same bf16 model with every norm gain moved by -1 / 0 / +1 ulp."""
    ck = load_pkg("checkpoint")
    rng = np.random.default_rng(99)
    w2 = dict(wb)
    for k in ("rms_att", "rms_ffn", "rms_final"):
        bits = np.ascontiguousarray(wb[k], np.float32).copy().view(np.int32)
        bits += rng.integers(-1, 2, size=bits.shape).astype(np.int32)
        w2[k] = bits.view(np.float32)
    path = str(tmp_path / f"{tag}_floor.bin")
    ck.write_v0(path, cfg, w2)
    return oracle.open(path, bf16=True)


def ragged_ratios(e, m, n_steps, B, rng, stagger=0, floor=None, sample=None):
    """run_ragged of test_gpu_forward.py, returning |err| / max|logit| of every (step, slot) instead of asserting; with
    `floor` (a second oracle model) also that model's spread against the first one on the slots in `sample`."""
    join = [(b % stagger if stagger else b) for b in range(B)]
    seqs = [m.new_seq() for _ in range(B)]
    sample = list(range(B)) if sample is None else sample
    fseqs = {b: floor.new_seq() for b in sample} if floor else {}
    fl = []
    toks = rng.integers(0, m.vocab_size, size=(B, n_steps)).astype(np.int32)
    ratios = []
    for step in range(n_steps):
        tok = np.zeros(B, np.int32); pos = np.full(B, -1, np.int32)
        for b in range(B):
            if step >= join[b]:
                pos[b] = step - join[b]
                tok[b] = toks[b, step - join[b]]
        logits = e.forward_batch(tok, pos).numpy()
        live = [b for b in range(B) if pos[b] >= 0]
        m.forward_many([seqs[b] for b in live], tok[live], pos[live], 8)
        fl_live = [b for b in live if b in fseqs]
        if fl_live:
            floor.forward_many([fseqs[b] for b in fl_live], tok[fl_live], pos[fl_live], 8)
        for b in live:
            want = seqs[b].buf("logits")
            ratios.append(float(np.abs(logits[b] - want).max() / max(np.abs(want).max(), 1e-30)))
            if b in fseqs:
                fl.append(float(np.abs(fseqs[b].buf("logits") - want).max() / max(np.abs(want).max(), 1e-30)))
    return np.array(ratios), seqs, np.array(fl)


def check_ratios(r, fl, what, min_tight=None):
    tight = float((r <= TIGHT).mean())
    print(f"{what}: {r.size} rows, {100 * tight:.1f}% within {TIGHT} of the largest logit, 90% {np.quantile(r, 0.9):.3e}, worst {r.max():.3e}"
          + (f"; the oracle against itself (gains +-1 ulp): {100 * float((fl <= TIGHT).mean()):.1f}% / {np.quantile(fl, 0.9):.3e} / {fl.max():.3e}"
             if fl.size else ""))
    assert r.max() <= CAP, f"{what}: worst row {r.max():.3e}"
    assert np.median(r) <= CAP / 2, f"{what}: median {np.median(r):.3e}"
    if min_tight is not None:
        assert tight >= min_tight, f"{what}: only {100 * tight:.1f}% of the rows agree to fp32 noise"

MODELS = {
    "fusedA": dict(dim=256, hidden_dim=704, n_layers=3, n_heads=8, n_kv_heads=8, vocab_size=512, seq_len=48),
    "fusedGQA": dict(dim=256, hidden_dim=704, n_layers=2, n_heads=8, n_kv_heads=2, vocab_size=320, seq_len=40),
    "hs64": dict(dim=256, hidden_dim=512, n_layers=2, n_heads=4, n_kv_heads=2, vocab_size=300, seq_len=200),
    "hs128": dict(dim=512, hidden_dim=1024, n_layers=2, n_heads=4, n_kv_heads=2, vocab_size=300, seq_len=420),
}


def bf16_model(tmp_path, name, kwargs, std=0.08):
    ck, eng = load_pkg("checkpoint"), load_pkg("engine")
    cfg, w, v0, v2 = small_model_files(str(tmp_path), name, kwargs, std=std)
    wb = ck.bf16_weights(w)
    path = str(tmp_path / f"{name}_bf16_v0.bin")
    ck.write_v0(path, cfg, wb)
    return cfg, wb, path


@pytest.mark.parametrize("name", ["fusedA", "fusedGQA"])
@pytest.mark.parametrize("B", [1, 5, 16, 70, 130, 260])
def test_bf16_batched_ragged_vs_oracle(oracle, tmp_path, name, B):
    eng = load_pkg("engine")
    cfg, wb, path = bf16_model(tmp_path, name, MODELS[name])
    e = eng.Engine(eng.DeviceWeightsBF16(cfg, wb), B)
    m = oracle.open(path, bf16=True)
    fm = floor_model(oracle, tmp_path, cfg, wb, name)
    r, _, fl = ragged_ratios(e, m, min(cfg.seq_len, 8), B, np.random.default_rng(B), stagger=5 if B > 16 else 0, floor=fm,
                             sample=list(range(min(B, 24))))
    assert load_pkg("cabi").lib().c.tha_gemm_check_error() == 0
    check_ratios(r, fl, f"bf16 {name} B={B}")


@pytest.mark.parametrize("B", [1, 2, 7])
def test_bf16_gemv_step_vs_oracle(oracle, tmp_path, B):
    """tha_set_option("tc", 0): the fused GEMV step over bf16 rows (gemv_bf16.cu), no longer the default at any slot count."""
    eng = load_pkg("engine")
    lib = load_pkg("cabi").lib()
    cfg, wb, path = bf16_model(tmp_path, "fusedA", MODELS["fusedA"])
    m = oracle.open(path, bf16=True)
    fm = floor_model(oracle, tmp_path, cfg, wb, "fusedA")
    try:
        lib.check(lib.c.tha_set_option(b"tc", 0))
        e = eng.Engine(eng.DeviceWeightsBF16(cfg, wb), B)
        r, _, fl = ragged_ratios(e, m, min(cfg.seq_len, 8), B, np.random.default_rng(B), floor=fm, sample=list(range(B)))
        launches = e.last_step_launches()
    finally:
        lib.check(lib.c.tha_set_option(b"tc", -1))
    assert launches == 5 * cfg.n_layers + 2, launches   # embed + 5 fused kernels per layer + classifier (final norm folded in)
    check_ratios(r, fl, f"bf16 GEMV step B={B}")


@pytest.mark.parametrize("B", [1, 40])
def test_bf16_one_layer_agrees_to_fp32_noise(oracle, tmp_path, B):
    """One layer: five vectors are rounded per token, so a flipped rounding is rare and most rows must match to ~1e-7."""
    eng = load_pkg("engine")
    kw = dict(MODELS["fusedA"]); kw["n_layers"] = 1
    cfg, wb, path = bf16_model(tmp_path, "one", kw)
    e = eng.Engine(eng.DeviceWeightsBF16(cfg, wb), B)
    r, _, fl = ragged_ratios(e, oracle.open(path, bf16=True), 6, B, np.random.default_rng(B), stagger=3)
    check_ratios(r, fl, f"bf16 one layer B={B}", min_tight=0.6)


@pytest.mark.parametrize("name", ["fusedA", "hs64", "hs128"])
def test_bf16_prefill_vs_oracle(oracle, tmp_path, name):
    """Head sizes 64 and 128 take the tensor-core prompt attention (attn_prefill_tc.cu: q, k, v, p rounded to bf16, keys in
    tiles of 128 -- the oracle restates that rule for the prompt positions); head size 32 the CUDA-core kernel."""
    import torch
    eng = load_pkg("engine")
    cfg, wb, path = bf16_model(tmp_path, name, MODELS[name])
    m = oracle.open(path, bf16=True)
    tc_attention = cfg.head_size in (64, 128)
    m.set_bf16_attention(tc_attention)
    B = 4
    e = eng.Engine(eng.DeviceWeightsBF16(cfg, wb), B)
    rng = np.random.default_rng(3)
    lens = [1, 9, min(cfg.seq_len - 6, 70), min(cfg.seq_len - 6, 130 if name != "hs128" else 400)]
    prompts = [rng.integers(0, cfg.vocab_size, size=n).astype(np.int32) for n in lens]
    seqs = [m.new_seq() for _ in range(B)]
    want_last = []
    for b in range(B):
        for p_, t in enumerate(prompts[b]):
            lg = seqs[b].forward(int(t), p_)
        want_last.append(lg.copy())
    logits = torch.zeros(B, cfg.vocab_size, device="cuda")
    nxt = torch.zeros(B, dtype=torch.int32, device="cuda")
    e.prefill([(b, 0, prompts[b]) for b in range(B)], logits, nxt)
    torch.cuda.synchronize()
    assert load_pkg("cabi").lib().c.tha_gemm_check_error() == 0
    got = logits.cpu().numpy()
    rr = np.array([np.abs(got[b] - want_last[b]).max() / np.abs(want_last[b]).max() for b in range(B)])
    print(f"bf16 prefill {name}: |err| / max|logit| per sequence {rr}")
    for b in range(B):
        assert_close_scaled(got[b], want_last[b], BF16_EPS, f"bf16 prefill logits slot {b} ({lens[b]} positions)")
        n = lens[b]
        for cache in ("key_cache", "value_cache"):
            have = getattr(e.state, cache)[b].cpu().numpy()[:, :n]
            want = seqs[b].buf(cache).reshape(m.n_layers, m.seq_len, m.kv_dim)[:, :n]
            assert_close_scaled(have, want, BF16_EPS, f"{cache} slot {b}")
    m.set_bf16_attention(False)   # decode positions: run.cc's fp32 attention
    for step in range(3):   # decode continues from the prefilled cache
        tok = rng.integers(0, cfg.vocab_size, size=B).astype(np.int32)
        pos = np.array([lens[b] + step for b in range(B)], np.int32)
        lg = e.forward_batch(tok, pos).numpy()
        for b in range(B):
            assert_close_scaled(lg[b], seqs[b].forward(int(tok[b]), int(pos[b])), BF16_EPS, f"decode after prefill, slot {b} step {step}")


def test_bf16_stories110m_shape(oracle, tmp_path):
    ck, eng = load_pkg("checkpoint"), load_pkg("engine")
    cfg = ck.STORIES110M
    w = ck.bf16_weights(ck.random_weights(cfg, seed=1337, std=0.05))
    path = str(tmp_path / "s110m_bf16.bin")
    ck.write_v0(path, cfg, w)
    for B in (2, 40):
        e = eng.Engine(eng.DeviceWeightsBF16(cfg, w), B)
        fm = floor_model(oracle, tmp_path, cfg, w, f"s110m{B}")
        r, _, fl = ragged_ratios(e, oracle.open(path, bf16=True), 4, B, np.random.default_rng(B), stagger=3, floor=fm,
                                 sample=list(range(min(B, 6))))
        check_ratios(r, fl, f"bf16 stories110M shape B={B}")
        del e
