"""tha_mha_fused_v1_f32_static: the no-allocation variant of the fused attention (split scratch in a file-scope device
array, hip_llama.cpp_b200/csrc/attention.cu) against the existing entry point tha_mha_fused_v1_f32, through the C-ABI.

The entry point has no counterpart in the reference (include/tha_b200.h says why it exists).  This file is named so that
pytest collects it LAST: a failure here must not hide any other test.

What is checked, on two calls in a row that share the static array (different slot counts and positions, so the second
call meets the first one's leftovers):
  * output == the existing entry point's on the same inputs -- within the 1e-5 rule of tests/test_gpu_ops.py
    (test_mha_fused), and in fact bit for bit (same kernel, same fixed-order merge of the splits);
  * output against the oracle's attention (the same bar as test_mha_fused);
  * the device's free memory does not move across the calls;
  * a launch beyond the array's stated size is refused with a message that names the sizes, and the next call still works.
"""
import ctypes as C

import numpy as np
import pytest

from gpu_util import P, assert_close, stream_ptr, tdev
from pkgload import load_pkg

pytestmark = pytest.mark.gpu

TOL = 1e-5   # tests/test_gpu_ops.py::test_mha_fused


@pytest.fixture(scope="module")
def L():
    import torch
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    return load_pkg("cabi").lib()


def sync():
    import torch
    torch.cuda.synchronize()


def free_bytes(L):
    f, t = C.c_size_t(0), C.c_size_t(0)
    L.check(L.c.tha_mem_info(C.byref(f), C.byref(t)))
    return f.value


class Case:
    """Inputs of one attention call: B slots of a [B, n_layers, seq, kv_dim] cache, layer 1."""

    def __init__(self, n_heads, n_kv, hs, seq, B, seed):
        self.n_heads, self.n_kv, self.hs, self.seq, self.B = n_heads, n_kv, hs, seq, B
        self.n_layers, self.layer = 2, 1
        self.dim, self.kvd = n_heads * hs, n_kv * hs
        rng = np.random.default_rng(seed)
        self.pos = rng.integers(0, seq, size=B).astype(np.int32)
        self.pos[0] = seq - 1
        if B > 1:
            self.pos[1] = 0
        if B > 2:
            self.pos[2] = -1   # idle slot: left untouched
        self.q = rng.standard_normal((B, self.dim)).astype(np.float32)
        self.kc = rng.standard_normal((B, self.n_layers, seq, self.kvd)).astype(np.float32)
        self.vc = rng.standard_normal((B, self.n_layers, seq, self.kvd)).astype(np.float32)
        self.Q, self.Pd, self.KC, self.VC = tdev(self.q), tdev(self.pos), tdev(self.kc), tdev(self.vc)

    def run(self, L, fn):
        XB = tdev(np.full((self.B, self.dim), 7.0, np.float32))
        st = fn(stream_ptr(), self.B, P(self.Pd), self.n_heads, self.n_kv, self.n_layers, P(self.Q), P(XB), P(self.KC),
                P(self.VC), self.hs, self.seq, self.layer * self.seq * self.kvd)
        return st, XB

    def check_oracle(self, oracle, got, what):
        kv_mul = self.n_heads // self.n_kv
        for b in range(self.B):
            if self.pos[b] < 0:
                assert np.all(got[b] == 7.0), f"{what}: idle slot {b} was written"
                continue
            want, _ = oracle.attention(self.q[b], self.kc[b, self.layer], self.vc[b, self.layer], int(self.pos[b]),
                                       self.n_heads, kv_mul, self.hs)
            assert_close(got[b], want, TOL, f"{what} slot {b} pos {self.pos[b]}")


def test_static_attention_matches_existing_entry_point(L, oracle):
    # (heads, kv heads, head size, seq, slots): calls 0, 1 and 3 split the time axis (slots * segments well under 148 CTAs),
    # so they use the scratch -- with different slot and split counts, over each other's leftovers; call 2 is a large
    # unsplit launch (no scratch) in between
    cases = [Case(12, 12, 64, 1024, 2, 11), Case(12, 12, 64, 1024, 5, 12), Case(12, 12, 64, 128, 40, 13),
             Case(8, 2, 32, 96, 3, 14)]
    # warm-up of both entry points on every shape: module loads and the growable workspace of the EXISTING entry point
    # happen here, outside the no-allocation window
    for c in cases:
        for fn in (L.c.tha_mha_fused_v1_f32, L.c.tha_mha_fused_v1_f32_static):
            st, _ = c.run(L, fn)
            L.check(st, "warm-up")
    sync()
    want = []
    for c in cases:
        st, XB = c.run(L, L.c.tha_mha_fused_v1_f32)
        L.check(st, "tha_mha_fused_v1_f32")
        want.append(XB)
    sync()
    want = [w.cpu().numpy() for w in want]
    outs = [tdev(np.full((c.B, c.dim), 7.0, np.float32)) for c in cases]
    sync()
    before = free_bytes(L)
    for c, XB in zip(cases, outs):   # calls in a row on one stream, sharing the static array
        L.check(L.c.tha_mha_fused_v1_f32_static(stream_ptr(), c.B, P(c.Pd), c.n_heads, c.n_kv, c.n_layers, P(c.Q), P(XB),
                                                P(c.KC), P(c.VC), c.hs, c.seq, c.layer * c.seq * c.kvd),
                "tha_mha_fused_v1_f32_static")
    sync()
    after = free_bytes(L)
    assert after == before, f"free device memory moved across the static calls: {before} -> {after}"
    for i, (c, XB) in enumerate(zip(cases, outs)):
        got = XB.cpu().numpy()
        assert_close(got, want[i], TOL, f"static vs existing, call {i}")
        assert np.array_equal(got, want[i]), f"static vs existing, call {i}: not bit-identical"
        c.check_oracle(oracle, got, f"static attention, call {i}")


def test_static_attention_rejects_oversized_launch(L, oracle):
    # 8 query heads of 128 on ONE K/V head: a CTA covers the whole group (J = 8, 256 float4 per CTA); 9 slots split 32 ways
    # need 9 * 32 * 256 = 73728 scratch slots > 32768.  (Without grouped heads J = 1 and a split launch needs <= 9472.)
    big = Case(8, 1, 128, 1024, 9, 21)
    st, XB = big.run(L, L.c.tha_mha_fused_v1_f32_static)
    sync()
    assert st == 2, f"oversized launch returned {st}, expected THA_STATUS_INVALID_ARG"
    msg = L.c.tha_last_error().decode()
    assert "73728" in msg and "32768" in msg and "static" in msg, msg
    assert np.all(XB.cpu().numpy() == 7.0), "a refused call wrote its output"
    st, XB = big.run(L, L.c.tha_mha_fused_v1_f32)   # the allocating entry point takes the same launch
    L.check(st, "tha_mha_fused_v1_f32")
    sync()
    big.check_oracle(oracle, XB.cpu().numpy(), "existing entry point, large launch")
    # bad arguments: the reference's status for them
    assert L.c.tha_mha_fused_v1_f32_static(stream_ptr(), 0, P(big.Pd), 8, 1, 2, P(big.Q), P(XB), P(big.KC), P(big.VC), 128, 1024, 0) == 2
    # and the array still serves the next call that fits (same wide segments, fewer slots: 2 * 32 * 256 = 16384)
    ok = Case(8, 1, 128, 1024, 2, 22)
    st, want = ok.run(L, L.c.tha_mha_fused_v1_f32)
    L.check(st)
    st, got = ok.run(L, L.c.tha_mha_fused_v1_f32_static)
    L.check(st)
    sync()
    assert np.array_equal(got.cpu().numpy(), want.cpu().numpy())
    ok.check_oracle(oracle, got.cpu().numpy(), "static attention after a refused call")
