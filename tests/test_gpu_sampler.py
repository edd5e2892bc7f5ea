"""The driver's sampler on the device (csrc/sampler.cu, tha_sample) against the reference's sample() compiled unmodified
(oracle/_ref, run.cc:566-720 == src/llama.cpp:262-411) and against an exact restatement of its rule.

The device kernel keeps the reference's random stream bit for bit and its selection RULE; it cannot keep the fp32
rounding pattern of 32000 sequential additions over host-libm exponentials (sampler.cu header).  So for every draw:
the device token must be the reference's token, or the coin must sit within CDF_EPS (in units of probability mass) of
the boundary between the two tokens in the exact (float64) cumulative distribution -- and that may only happen rarely on
peaked (trained-model-like) distributions."""
import ctypes as C

import numpy as np
import pytest

from gpu_util import P, stream_ptr, tdev
from pkgload import load_pkg

pytestmark = pytest.mark.gpu
CDF_EPS = 2e-5


def xorshift_coin(state):
    """random_f32 (src/llama.cpp:401-411) on a Python int; returns (coin, new state)."""
    M = (1 << 64) - 1
    state ^= state >> 12
    state ^= (state << 25) & M
    state ^= state >> 27
    u = ((state * 0x2545F4914F6CDD1D) & M) >> 32
    return np.float32((u >> 8) / 16777216.0), state


def exact_rule(logits, temperature, topp, coin):
    """The reference's rule in float64: returns (token, sorted candidate ids, their cdf, r)."""
    x = logits.astype(np.float64) / temperature
    p = np.exp(x - x.max()); p /= p.sum()
    n = p.size
    if topp <= 0 or topp >= 1:
        cdf = np.cumsum(p)
        k = int(np.searchsorted(cdf, coin, side="right"))
        return min(k, n - 1), np.arange(n), cdf, float(coin)
    cutoff = (1.0 - topp) / (n - 1)
    cand = np.nonzero(p >= cutoff)[0]
    order = cand[np.argsort(-p[cand], kind="stable")]
    cdf = np.cumsum(p[order])
    last = int(np.searchsorted(cdf, topp, side="right"))
    last = min(last, order.size - 1)
    r = coin * cdf[last]
    k = int(np.searchsorted(cdf[:last + 1], r, side="right"))
    return int(order[min(k, last)]), order, cdf, float(r)


def run_device(logits, temperature, topp, seeds, force=None, pos=None, draws=1):
    import torch
    lib = load_pkg("cabi").lib()
    B, V = logits.shape
    rng = torch.zeros(B, dtype=torch.int64, device="cuda")
    reseed = tdev(np.asarray(seeds, np.uint64).view(np.int64))
    zero = torch.zeros(B, dtype=torch.int64, device="cuda")
    pos_d = tdev(np.zeros(B, np.int32) if pos is None else np.asarray(pos, np.int32))
    force_d = None if force is None else tdev(np.asarray(force, np.int32))
    nxt = torch.zeros(B, dtype=torch.int32, device="cuda")
    out = []
    for d in range(draws):
        lg = tdev(logits)
        lib.check(lib.c.tha_sample(stream_ptr(), B, P(lg), V, C.c_float(temperature), C.c_float(topp), P(rng),
                                   P(reseed if d == 0 else zero), P(force_d), P(pos_d), P(nxt)), "tha_sample")
        torch.cuda.synchronize()
        out.append(nxt.cpu().numpy().copy())
    return np.stack(out), rng.cpu().numpy().view(np.uint64)


def check_draws(logits, temperature, topp, seeds, got, max_boundary_frac):
    """got [draws, B]."""
    from oracle_lib import Ref32, ref_available
    B, V = logits.shape
    ref = Ref32() if ref_available("fp32") else None
    boundary = total = 0
    for b in range(B):
        state = int(seeds[b])
        smp = ref.lib.ref32_sampler_new(V, C.c_float(temperature), C.c_float(topp), C.c_ulonglong(state)) if ref else None
        for d in range(got.shape[0]):
            coin, state = xorshift_coin(state)
            want, order, cdf, r = exact_rule(logits[b], temperature, topp, coin)
            if ref:   # the reference itself (its own softmax, qsort, fp32 sums)
                lg = logits[b].copy()
                want_ref = ref.lib.ref32_sample(smp, lg.ctypes.data_as(C.POINTER(C.c_float)))
            else:
                want_ref = want
            have = int(got[d, b])
            total += 1
            if have == want_ref:
                continue
            # both tokens must be neighbours of the coin in the exact cumulative distribution
            boundary += 1
            pos_in = {int(t): i for i, t in enumerate(order)}
            assert have in pos_in, f"slot {b} draw {d}: device token {have} is not even a candidate"
            i = pos_in[have]
            lo = cdf[i - 1] if i > 0 else 0.0
            assert lo - CDF_EPS <= r <= cdf[i] + CDF_EPS, \
                f"slot {b} draw {d}: device {have} covers [{lo}, {cdf[i]}], target {r}, reference {want_ref}"
    assert boundary <= max_boundary_frac * total, f"{boundary}/{total} draws differ from the reference"
    return boundary, total


@pytest.mark.parametrize("topp", [0.9, 0.5, 1.0, 0.0])
@pytest.mark.parametrize("temperature", [1.0, 0.7])
def test_sampler_peaked_distributions(topp, temperature):
    """Logits with the spread of a trained model (std 3): a handful to a few hundred candidates in the nucleus."""
    rng = np.random.default_rng(5)
    B, V, draws = 48, 32000, 6
    logits = (3.0 * rng.standard_normal((B, V))).astype(np.float32)
    seeds = rng.integers(1, 2 ** 62, size=B).astype(np.uint64)
    got, _ = run_device(logits, temperature, topp, seeds, draws=draws)
    # how often the reference's own fp32 rounding decides between two neighbours: rarely inside a nucleus of a few hundred
    # tokens, a few percent of the draws when all 32000 tokens take part (multinomial: 32000 sequential fp32 additions)
    boundary, total = check_draws(logits, temperature, topp, seeds, got, 0.06 if 0.0 < topp < 1.0 else 0.12)
    print(f"peaked topp={topp} T={temperature}: {total - boundary}/{total} draws identical to the reference")


def test_sampler_near_uniform_distribution():
    """Random-init models: every token is a candidate and neighbours in sorted order are 3e-5 of mass apart, so the
    reference's own fp32 cdf decides between neighbours by rounding; the rule (candidate set, order, target) still holds."""
    rng = np.random.default_rng(6)
    B, V = 16, 32000
    logits = (0.05 * rng.standard_normal((B, V))).astype(np.float32)
    seeds = rng.integers(1, 2 ** 62, size=B).astype(np.uint64)
    got, _ = run_device(logits, 1.0, 0.9, seeds, draws=4)
    boundary, total = check_draws(logits, 1.0, 0.9, seeds, got, 1.0)
    print(f"near-uniform: {total - boundary}/{total} draws identical to the reference")


def test_sampler_ties_small_vocab_and_stream_state():
    """Equal probabilities keep index order (what a stable sort leaves); forced prompt tokens draw no coin; idle slots are
    skipped; temperature 0 is the first maximum; the random stream continues across calls."""
    B, V = 6, 8
    logits = np.zeros((B, V), np.float32)
    logits[:, [2, 5, 6]] = 2.0                      # three equal leaders
    seeds = np.arange(1, B + 1, dtype=np.uint64) * 7919
    got, state = run_device(logits, 1.0, 0.9, seeds, force=[-1, -1, 4, -1, -1, -1], pos=[0, 0, 0, -1, 0, 0], draws=3)
    for d in range(3):
        assert got[d, 2] == 4 and got[d, 3] == 0
    for b in (0, 1, 4, 5):
        st = int(seeds[b])
        for d in range(3):
            coin, st = xorshift_coin(st)
            want, *_ = exact_rule(logits[b], 1.0, 0.9, coin)
            assert got[d, b] == want, f"slot {b} draw {d}"
        assert int(state[b]) == st                   # three coins consumed
    assert int(state[2]) == int(seeds[2])            # forced: reseeded, no coin drawn
    g, _ = run_device(np.array([[0.1, 3.0, 3.0, -1.0]] * 2, np.float32), 0.0, 0.9, [1, 2])
    assert g[0].tolist() == [1, 1]


def test_forward_sample_equals_forward_then_sample(tmp_path):
    """tha_forward_sample_f32 == thaDNN_s_forward_batch followed by tha_sample on the same logits."""
    import torch
    from gpu_util import engine_from_v0, small_model_files
    cfg, w, v0, v2 = small_model_files(str(tmp_path), "smp", dict(dim=256, hidden_dim=704, n_layers=2, n_heads=8, n_kv_heads=8,
                                                                   vocab_size=512, seq_len=32), std=0.3)
    B = 5
    e1, e2 = engine_from_v0(v0, B), engine_from_v0(v0, B)
    rng = np.random.default_rng(1)
    seeds = rng.integers(1, 2 ** 62, size=B).astype(np.uint64)
    tok = rng.integers(0, cfg.vocab_size, size=B).astype(np.int32)
    lib = load_pkg("cabi").lib()
    state = torch.zeros(B, dtype=torch.int64, device="cuda")
    for step in range(5):
        pos = np.full(B, step, np.int32)
        force = np.array([-1, 7 if step < 2 else -1, -1, -1, -1], np.int32)
        nx = e1.forward_sample(tok, pos, force, seeds if step == 0 else None, 1.0, 0.9).copy()
        lg = tdev(e2.forward_batch(tok, pos).numpy())
        nxt = torch.zeros(B, dtype=torch.int32, device="cuda")
        reseed = tdev((seeds if step == 0 else np.zeros(B, np.uint64)).view(np.int64))
        force_d, pos_d = tdev(force), tdev(pos)      # named: the tensors must outlive the launch
        lib.check(lib.c.tha_sample(stream_ptr(), B, P(lg), cfg.vocab_size, C.c_float(1.0), C.c_float(0.9), P(state), P(reseed),
                                   P(force_d), P(pos_d), P(nxt)), "tha_sample")
        torch.cuda.synchronize()
        assert nx.tolist() == nxt.cpu().numpy().tolist(), f"step {step}"
        tok = nx.astype(np.int32)


def test_sampler_resident_and_streaming_forms_pick_the_same_tokens():
    """The two forms of the kernel (row resident in shared memory: integer keys, four key bits per level, no atomics; row in
    HBM / L2: three histogram levels of shared atomics) accumulate the same integer masses over the same candidates, so
    they must agree on EVERY token and leave the same probabilities and stream states -- whatever the distribution."""
    import torch
    lib = load_pkg("cabi").lib()
    rng = np.random.default_rng(77)
    B, V = 48, 32000
    cases = [(rng.standard_normal((B, V)) * 2.0, 1.0, 0.9), (rng.standard_normal((B, V)) * 0.02, 1.0, 0.9),
             (rng.standard_normal((B, V)) * 6.0, 0.7, 0.5), (rng.standard_normal((B, 5000)) * 3.0, 1.3, 0.95),
             (rng.standard_normal((B, V)) * 2.0, 1.0, 1.0)]           # the last: multinomial over the whole row
    seeds = np.arange(1, B + 1, dtype=np.uint64) * np.uint64(2654435761)
    try:
        for logits, temperature, topp in cases:
            logits = logits.astype(np.float32)
            logits[3, 100:200] = logits[3, 100]                        # a run of equal probabilities: index order decides
            got = {}
            for resident in (1, 0):
                lib.check(lib.c.tha_set_option(b"sample_resident", resident))
                toks, state = run_device(logits, temperature, topp, seeds, draws=4)
                got[resident] = (toks, state)
            assert np.array_equal(got[1][0], got[0][0]), f"tokens differ (temperature {temperature}, top-p {topp}, vocab {logits.shape[1]})"
            assert np.array_equal(got[1][1], got[0][1]), "stream states differ"
        # the probabilities the kernel leaves in the logits buffer are the same bits in both forms
        lg = {}
        for resident in (1, 0):
            lib.check(lib.c.tha_set_option(b"sample_resident", resident))
            t = tdev(cases[0][0].astype(np.float32))
            st = tdev(seeds.view(np.int64)); nxt = torch.zeros(B, dtype=torch.int32, device="cuda")
            pos = torch.zeros(B, dtype=torch.int32, device="cuda")
            lib.check(lib.c.tha_sample(stream_ptr(), B, P(t), V, C.c_float(1.0), C.c_float(0.9), P(st), None, None, P(pos), P(nxt)))
            torch.cuda.synchronize()
            lg[resident] = t.cpu().numpy()
        assert np.array_equal(lg[1], lg[0]), "probabilities differ between the two forms"
    finally:
        lib.check(lib.c.tha_set_option(b"sample_resident", 1))
