"""Parity at the shapes BASELINE.json names, through the same entry points bench.py times.

  * stories110M fp32 at 256 slots (configs[1], the headline line): the tensor-core row pipeline with two 128-column
    tiles and the planner's split-K, teacher-forced from position 0 past position 64, sampled slots against the
    oracle at the reference's own 1e-4 (scripts/test/thaDNN.test.cpp:224, full-forward check :541-548);
  * stories110M Q8_0 at 64 slots (configs[2]);
  * a 2-layer truncation at Llama-2-7B widths (dim 4096, hidden 11008, 32 heads, vocab 32000), fp32 and Q8_0, at
    1 / 8 / 64 slots (configs[3], configs[4]);
  * Q8_0 layer by layer with EXACT inputs: every layer of a model as its own 1-layer model whose embedding rows are
    realistic activations, so no rounding flip of an earlier layer reaches the comparison (runq.c:344-481).
The oracle runs a handful of sampled slots only (it needs ~70 ms per stories110M position and core)."""
import os

import numpy as np
import pytest

from gpu_util import assert_close, assert_close_scaled, close, engine_from_v0, engine_from_v2, small_model_files
from pkgload import load_pkg

pytestmark = pytest.mark.gpu

FP32_EPS = 1e-4
# Q8_0 end to end.  runq.c rounds every matmul input to int8 (runq.c:145-171); a 1-ulp difference in the fp32 value
# being rounded (a reordered RMSNorm sum is enough) now and then moves one int8 by one step.  On N(0, 0.02)
# random-init networks the residual stream is as small as the weights, so ONE such step is worth ~1e-2 of the largest
# logit, and K/V rows computed after it stay off for the rest of the sequence.  That is a property of the reference
# arithmetic, and it is measured here instead of assumed: the oracle runs a second time with every norm gain moved by
# -1 / 0 / +1 ulp ("floor"), and the GPU has to sit within twice that spread.  The kernels themselves are pinned
# without this noise: bit-exact quantiser and int8 matmul (test_gpu_ops.py), one layer at a time below.
Q8_E2E_CAP = 5e-2          # hard cap whatever the floor says (norm-wise)
Q8_LAYER_EPS = 1e-4        # one layer with exact inputs: fp32 noise unless an activation rounding flipped
Q8_LAYER_CAP = 5e-2


def n_threads():
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:
        return os.cpu_count() or 1


def perturbed_norms(w, seed):
    rng = np.random.default_rng(seed)
    for k in ("rms_att", "rms_ffn", "rms_final"):
        w[k] = (1.0 + 0.1 * rng.standard_normal(w[k].shape)).astype(np.float32)
    return w


@pytest.fixture(scope="module")
def s110m_files(tmp_path_factory):
    ck = load_pkg("checkpoint")
    cfg = ck.STORIES110M
    w = perturbed_norms(ck.random_weights(cfg, seed=1337, std=0.05), 5)
    d = tmp_path_factory.mktemp("s110m")
    v0, v2 = str(d / "s110m_v0.bin"), str(d / "s110m_v2.bin")
    ck.write_v0(v0, cfg, w)
    # Q8_0 end to end is compared on the reference's own initialisation law (train/model.py:232-247, N(0, 0.02)): with
    # 2.5x larger weights every matmul amplifies (gain 0.05 * sqrt(768) = 1.4 per projection) and a single flipped
    # activation rounding grows layer by layer into percents of the logits -- a statement about that random network,
    # not about the kernels (those are pinned bit-exactly by the op tests and layer by layer below).
    wq = perturbed_norms(ck.random_weights(cfg, seed=1337, std=0.02), 5)
    ck.write_v2(v2, cfg, wq)
    return cfg, v0, v2, wq


def ulp_perturbed_copy(oracle, ck, cfg, w, path):
    """The same model with every norm gain moved by -1 / 0 / +1 ulp: what a reordered fp32 sum does to the values
    the activation quantiser rounds."""
    rng = np.random.default_rng(99)
    w2 = dict(w)
    for k in ("rms_att", "rms_ffn", "rms_final"):
        bits = w[k].copy().view(np.int32)
        bits += rng.integers(-1, 2, size=bits.shape).astype(np.int32)
        w2[k] = bits.view(np.float32)
    ck.write_v2(path, cfg, w2)
    return oracle.open(path, quantized=True)


def teacher_forced(e, m, B, n_steps, check_steps, sample, eps, q8, seed, what, floor_model=None, scaled=False):
    """All B slots decode positions 0..n_steps-1 with their own random token stream; the sampled slots are replayed
    through the oracle (every position: it needs the history) and compared on the steps in check_steps."""
    rng = np.random.default_rng(seed)
    toks = rng.integers(3, m.vocab_size, size=(n_steps, B)).astype(np.int32)
    seqs = [m.new_seq() for _ in sample]
    fseqs = [floor_model.new_seq() for _ in sample] if floor_model else None
    ratios, floor = [], []
    flips = 0
    for step in range(n_steps):
        pos = np.full(B, step, np.int32)
        logits = e.forward_batch(toks[step], pos).numpy()
        m.forward_many(seqs, toks[step, sample], pos[sample], min(n_threads(), len(sample)))
        if fseqs:
            floor_model.forward_many(fseqs, toks[step, sample], pos[sample], min(n_threads(), len(sample)))
        if step not in check_steps:
            continue
        for i, b in enumerate(sample):
            want = seqs[i].buf("logits")
            if fseqs:
                floor.append(float(np.abs(fseqs[i].buf("logits") - want).max() / max(np.abs(want).max(), 1e-30)))
            if not q8 and scaled:
                assert_close_scaled(logits[b], want, eps, f"{what}: step {step} slot {b}")
            elif not q8:
                assert_close(logits[b], want, eps, f"{what}: step {step} slot {b}")
            ratios.append(float(np.abs(logits[b] - want).max() / max(np.abs(want).max(), 1e-30)))
            flips += int(logits[b].argmax()) != int(want.argmax())
    r = np.array(ratios)
    print(f"{what}: |err| / max|logit| over {r.size} rows: median {np.median(r):.3e}, 90% {np.quantile(r, 0.9):.3e}, "
          f"worst {r.max():.3e}; argmax flips {flips}")
    if q8:
        assert r.max() <= eps, f"{what}: worst row {r.max():.3e} > {eps}"
        if floor:
            fl = np.array(floor)
            print(f"{what}: the oracle against itself with norm gains moved by <= 1 ulp: median {np.median(fl):.3e}, "
                  f"90% {np.quantile(fl, 0.9):.3e}, worst {fl.max():.3e}")
            # the spread is bimodal (a row either saw a flipped rounding upstream or not): compare the upper end
            assert np.quantile(r, 0.9) <= 2 * np.quantile(fl, 0.9) + 1e-3, f"{what}: 90% {np.quantile(r, 0.9):.3e} vs floor {np.quantile(fl, 0.9):.3e}"
            assert r.max() <= 2 * fl.max() + 1e-3, f"{what}: worst {r.max():.3e} vs floor {fl.max():.3e}"
    return seqs, flips


def check_kv(e, m, seqs, sample, n, eps, q8, what):
    for i, b in enumerate(sample):
        for cache in ("key_cache", "value_cache"):
            have = getattr(e.state, cache)[b].cpu().numpy()[:, :n]
            want = seqs[i].buf(cache).reshape(m.n_layers, m.seq_len, m.kv_dim)[:, :n]
            (assert_close_scaled if q8 else assert_close)(have, want, eps, f"{what}: {cache} slot {b}")


@pytest.mark.parametrize("chain", [0, 1])
def test_stories110m_fp32_256_slots_row_pipeline(oracle, s110m_files, chain):
    """The configuration bench.py times (gemm_tf32x3 at dim 768 / hidden 2048 / vocab 32000 / 256 slots); chain = 1: the
    same step as chained matmuls with fused epilogues (chain_tc.cu)."""
    cfg, v0, _, _ = s110m_files
    B = 256
    lib = load_pkg("cabi").lib()
    lib.check(lib.c.tha_set_option(b"chain", chain))
    try:
        _s110m_256(oracle, v0, B)
    finally:
        lib.check(lib.c.tha_set_option(b"chain", 0))


def _s110m_256(oracle, v0, B):
    e = engine_from_v0(v0, B)
    m = oracle.open(v0)
    sample = [0, 1, 17, 63, 64, 100, 127, 128, 129, 130, 191, 192, 200, 254, 255, 31]
    n_steps = 74
    check = set(range(0, 5)) | set(range(64, 74))
    seqs, flips = teacher_forced(e, m, B, n_steps, check, sample, FP32_EPS, False, 11, "s110m fp32 256 slots")
    assert e.last_step_launches() > 1          # not the <= 8-slot persistent kernel
    assert load_pkg("cabi").lib().c.tha_gemm_check_error() == 0
    assert flips == 0
    check_kv(e, m, seqs, sample[:4], n_steps, FP32_EPS, False, "s110m fp32 256 slots")


def test_stories110m_q8_64_slots(oracle, s110m_files, tmp_path):
    cfg, _, v2, wq = s110m_files
    B = 64
    e = engine_from_v2(v2, B)
    m = oracle.open(v2, quantized=True)
    m2 = ulp_perturbed_copy(oracle, load_pkg("checkpoint"), cfg, wq, str(tmp_path / "floor_v2.bin"))
    sample = [0, 1, 15, 16, 31, 32, 47, 63]
    n_steps = 70
    check = set(range(0, 4)) | set(range(64, 70))
    seqs, _ = teacher_forced(e, m, B, n_steps, check, sample, Q8_E2E_CAP, True, 12, "s110m q8 64 slots", floor_model=m2)
    assert load_pkg("cabi").lib().c.tha_gemm_check_error() == 0
    check_kv(e, m, seqs, sample[:2], n_steps, Q8_E2E_CAP, True, "s110m q8 64 slots")


# ---- Llama-2-7B widths, two layers ------------------------------------------------------------------------------
WIDE = dict(dim=4096, hidden_dim=11008, n_layers=2, n_heads=32, n_kv_heads=32, vocab_size=32000, seq_len=64)


@pytest.fixture(scope="module")
def wide_files(tmp_path_factory):
    ck = load_pkg("checkpoint")
    cfg = ck.Config(**WIDE)
    w = perturbed_norms(ck.random_weights(cfg, seed=77, std=0.02), 6)
    d = tmp_path_factory.mktemp("wide")
    v0, v2 = str(d / "wide_v0.bin"), str(d / "wide_v2.bin")
    ck.write_v0(v0, cfg, w)
    ck.write_v2(v2, cfg, w)
    f2 = str(d / "wide_floor_v2.bin")
    return cfg, v0, v2, w, f2


@pytest.mark.parametrize("B", [1, 8, 64])
def test_7b_widths_fp32(oracle, wide_files, B):
    cfg, v0 = wide_files[:2]
    e = engine_from_v0(v0, B)
    m = oracle.open(v0)
    sample = sorted(set([0, B // 2, B - 1] + ([7, 33] if B == 64 else [])))
    # norm-wise at these widths: run.cc's own sequential fp32 sums over K = 4096 / 11008 carry ~1e-5 of the row's scale
    # (logits reach |5|), so "1e-4 absolute on every element" would measure the reference's rounding, not the kernel
    seqs, flips = teacher_forced(e, m, B, 6, set(range(6)), sample, FP32_EPS, False, 20 + B, f"7B widths fp32 B={B}", scaled=True)
    assert load_pkg("cabi").lib().c.tha_gemm_check_error() == 0
    assert flips == 0
    check_kv(e, m, seqs, sample[:2], 6, FP32_EPS, False, f"7B widths fp32 B={B}")


@pytest.mark.parametrize("B", [1, 8, 64])
def test_7b_widths_q8(oracle, wide_files, B):
    cfg, _, v2, w, f2 = wide_files
    e = engine_from_v2(v2, B)
    m = oracle.open(v2, quantized=True)
    m2 = ulp_perturbed_copy(oracle, load_pkg("checkpoint"), cfg, w, f2)
    sample = sorted(set([0, B // 2, B - 1] + ([7, 33] if B == 64 else [])))
    seqs, _ = teacher_forced(e, m, B, 6, set(range(6)), sample, Q8_E2E_CAP, True, 30 + B, f"7B widths q8 B={B}", floor_model=m2)
    assert load_pkg("cabi").lib().c.tha_gemm_check_error() == 0
    check_kv(e, m, seqs, sample[:2], 6, Q8_E2E_CAP, True, f"7B widths q8 B={B}")


# ---- Q8_0, one layer at a time with exact inputs ----------------------------------------------------------------
LAYERED = dict(dim=256, hidden_dim=704, n_layers=4, n_heads=8, n_kv_heads=8, vocab_size=512, seq_len=24)


@pytest.mark.parametrize("B", [3, 24])
def test_q8_layer_by_layer_exact_inputs(oracle, tmp_path, B):
    """Layer l of a 4-layer model as its own 1-layer Q8_0 model.  Its embedding table holds rows with the spread of a
    residual stream entering layer l (embedding rows plus Gaussian updates that grow with l); GPU and runq.c-oracle
    dequantise the same table, so both start the layer from bit-identical rows and whatever differs afterwards was
    produced inside this one layer.  Rows are right to fp32 noise (1e-4 of the row's scale) unless one activation
    rounding flipped (then the row moves by a step of the int8 grid, percents of the logits on these small networks);
    a flipped K/V row stays in the cache, so later positions of that slot inherit it.  Most rows must be flip-free."""
    ck = load_pkg("checkpoint")
    cfg = ck.Config(**LAYERED)
    w = perturbed_norms(ck.random_weights(cfg, seed=9, std=0.08), 10)
    rng = np.random.default_rng(B)
    n_pos = 10
    V = cfg.vocab_size
    tight = total = k_tight = k_total = 0
    for l in range(cfg.n_layers):
        one = ck.Config(cfg.dim, cfg.hidden_dim, 1, cfg.n_heads, cfg.n_kv_heads, V, cfg.seq_len)
        w1 = {k: (v[l:l + 1] if k not in ("tok_emb", "rms_final") else v) for k, v in w.items()}
        w1 = dict(w1)
        w1["tok_emb"] = (w["tok_emb"] + (0.05 * l) * rng.standard_normal(w["tok_emb"].shape)).astype(np.float32)
        v2 = str(tmp_path / f"layer{l}_v2.bin")
        ck.write_v2(v2, one, w1)
        e = engine_from_v2(v2, B)
        m = oracle.open(v2, quantized=True)
        seqs = [m.new_seq() for _ in range(B)]
        toks = rng.integers(0, V, size=(n_pos, B)).astype(np.int32)
        for p in range(n_pos):
            pos = np.full(B, p, np.int32)
            got = e.forward_batch(toks[p], pos).numpy()
            m.forward_many(seqs, toks[p], pos, min(n_threads(), B))
            for b in range(B):
                want = seqs[b].buf("logits")
                err = float(np.abs(got[b] - want).max() / np.abs(want).max())
                total += 1
                tight += err <= Q8_LAYER_EPS
                assert err <= Q8_LAYER_CAP, f"layer {l} pos {p} slot {b}: {err:.3e}"
        for b in range(min(B, 4)):
            have = e.state.key_cache[b].cpu().numpy()[0, :n_pos]
            want = seqs[b].buf("key_cache").reshape(1, m.seq_len, m.kv_dim)[0, :n_pos]
            scale = np.abs(want).max()
            row_err = np.abs(have - want).max(axis=1) / scale
            assert row_err.max() <= Q8_LAYER_CAP, f"layer {l} K rows slot {b}: {row_err}"
            k_tight += int((row_err <= Q8_LAYER_EPS).sum()); k_total += row_err.size
    print(f"q8 layer-by-layer B={B}: {tight}/{total} logit rows and {k_tight}/{k_total} K rows within {Q8_LAYER_EPS} of the row scale")
    assert tight >= 0.7 * total and k_tight >= 0.9 * k_total
