"""Generate the committed golden fixtures from the REFERENCE ITSELF.

Runs only in the build container (needs /root/reference and oracle/_ref); the
GPU box never runs it -- tests read the files it leaves in tests/golden/.

For each tiny architecture:
  1. build `train/model.py:Transformer` (the reference's own torch model) with a fixed seed,
  2. export it with the reference's own `train/export.py` (`legacy_export`, `version2_export`),
  3. run the reference's compiled CPU code (run.cc / runq.c via oracle/_ref) token by token
     and record logits, the KV cache, and (Q8_0) the last quantised activation.
Also: tokenizer known answers -- the vectors of /root/reference/test.c:49-75 are re-checked
against run.cc's encode(), and the encodings of the gen_in prompts are recorded.

usage: python tests/golden/make_golden.py
"""
import ctypes as C
import io
import json
import os
import sys
from contextlib import redirect_stdout

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF = "/root/reference"
sys.path.insert(0, os.path.join(REF, "train"))

from model import ModelArgs, Transformer  # noqa: E402  (reference code)
from export import legacy_export, version2_export  # noqa: E402  (reference code)

ARCHS = {
    # name: (ModelArgs kwargs, token ids fed in order)
    "tinyA": dict(dim=64, n_layers=2, n_heads=4, n_kv_heads=4, vocab_size=256, multiple_of=32, max_seq_len=32),
    "tinyB_gqa": dict(dim=96, n_layers=3, n_heads=6, n_kv_heads=2, vocab_size=320, multiple_of=32, max_seq_len=24),
}
N_STEPS = 12


def ref_fp32():
    lib = C.CDLL(os.path.join(ROOT, "oracle/_ref/libref_fp32.so"))
    lib.ref32_open.restype = C.c_void_p
    lib.ref32_open.argtypes = [C.c_char_p]
    lib.ref32_forward.restype = C.POINTER(C.c_float)
    lib.ref32_forward.argtypes = [C.c_void_p, C.c_int, C.c_int]
    lib.ref32_state_ptr.restype = C.POINTER(C.c_float)
    lib.ref32_state_ptr.argtypes = [C.c_void_p, C.c_int]
    lib.ref32_tok_open.restype = C.c_void_p
    lib.ref32_tok_open.argtypes = [C.c_char_p, C.c_int]
    lib.ref32_encode.restype = C.c_int
    lib.ref32_encode.argtypes = [C.c_void_p, C.c_char_p, C.c_int, C.c_int, C.POINTER(C.c_int)]
    return lib


def ref_q8():
    lib = C.CDLL(os.path.join(ROOT, "oracle/_ref/libref_q8.so"))
    lib.refq8_open.restype = C.c_void_p
    lib.refq8_open.argtypes = [C.c_char_p]
    lib.refq8_forward.restype = C.POINTER(C.c_float)
    lib.refq8_forward.argtypes = [C.c_void_p, C.c_int, C.c_int]
    lib.refq8_state_ptr.restype = C.POINTER(C.c_float)
    lib.refq8_state_ptr.argtypes = [C.c_void_p, C.c_int]
    lib.refq8_state_q.restype = C.POINTER(C.c_int8)
    lib.refq8_state_q.argtypes = [C.c_void_p, C.c_int]
    lib.refq8_state_s.restype = C.POINTER(C.c_float)
    lib.refq8_state_s.argtypes = [C.c_void_p, C.c_int]
    lib.refq8_group_size.restype = C.c_int
    lib.refq8_quantize.argtypes = [C.POINTER(C.c_int8), C.POINTER(C.c_float), C.POINTER(C.c_float), C.c_int]
    return lib


def arr(ptr, n, dtype=np.float32):
    return np.ctypeslib.as_array(ptr, shape=(n,)).astype(dtype, copy=True)


def main():
    l32, lq8 = ref_fp32(), ref_q8()
    for name, kw in ARCHS.items():
        torch.manual_seed(1234)
        model = Transformer(ModelArgs(**kw))
        # make the norms non-trivial so the fixtures exercise the rmsnorm weights
        with torch.no_grad():
            for p_name, p in model.named_parameters():
                if p_name.endswith("norm.weight"):
                    p.copy_(1.0 + 0.1 * torch.randn_like(p))
                else:
                    p.mul_(4.0)  # std 0.08: logits with usable top-1 margins
        v0 = os.path.join(HERE, f"{name}_v0.bin")
        v2 = os.path.join(HERE, f"{name}_v2.bin")
        with redirect_stdout(io.StringIO()):
            legacy_export(model, v0)
            version2_export(model, v2, group_size=64)
        rng = np.random.default_rng(7)
        tokens = rng.integers(3, kw["vocab_size"], size=N_STEPS).astype(np.int32)
        L, S = kw["n_layers"], kw["max_seq_len"]
        kvd = kw["dim"] * kw["n_kv_heads"] // kw["n_heads"]
        V = kw["vocab_size"]

        h = l32.ref32_open(v0.encode())
        logits32 = np.stack([arr(l32.ref32_forward(h, int(t), p), V) for p, t in enumerate(tokens)])
        kc32 = arr(l32.ref32_state_ptr(h, 8), L * S * kvd)
        vc32 = arr(l32.ref32_state_ptr(h, 9), L * S * kvd)

        hq = lq8.refq8_open(v2.encode())
        gs = lq8.refq8_group_size()
        logitsq8 = np.stack([arr(lq8.refq8_forward(hq, int(t), p), V) for p, t in enumerate(tokens)])
        kcq = arr(lq8.refq8_state_ptr(hq, 8), L * S * kvd)
        vcq = arr(lq8.refq8_state_ptr(hq, 9), L * S * kvd)
        d, hd = kw["dim"], model.layers[0].feed_forward.w1.weight.shape[0]
        xq = arr(lq8.refq8_state_q(hq, 0), d, np.int8)
        xs = arr(lq8.refq8_state_s(hq, 0), d // gs)
        hq_q = arr(lq8.refq8_state_q(hq, 1), hd, np.int8)
        hq_s = arr(lq8.refq8_state_s(hq, 1), hd // gs)
        hb = arr(lq8.refq8_state_ptr(hq, 3), hd)  # the fp32 vector hq was quantised from

        np.savez_compressed(
            os.path.join(HERE, f"{name}_golden.npz"),
            tokens=tokens, logits_fp32=logits32, key_cache_fp32=kc32, value_cache_fp32=vc32,
            logits_q8=logitsq8, key_cache_q8=kcq, value_cache_q8=vcq, group_size=np.int32(gs),
            last_xq=xq, last_xs=xs, last_hb=hb, last_hq=hq_q, last_hs=hq_s,
        )
        print(name, "v0", os.path.getsize(v0), "v2", os.path.getsize(v2), "GS", gs,
              "argmax fp32", logits32.argmax(1).tolist(), "q8", logitsq8.argmax(1).tolist())

    # ---- quantiser known answers straight from runq.c:145-171 (GS = 64 and 32) ----
    rng = np.random.default_rng(11)
    cases = {}
    for gs in (64, 32):
        lq8.refq8_set_group_size(gs)
        n = gs * 12
        xs_in = {
            "normal": rng.standard_normal(n).astype(np.float32),
            "uniform": (rng.random(n, dtype=np.float32) - np.float32(0.5)),
            "ties": (np.tile(np.arange(-63, 65, dtype=np.float32), n // 128 + 1)[:n] * np.float32(0.5)),
            "allequal": np.full(n, 0.37, np.float32),
            "outlier": np.concatenate([np.full(n - 1, 1e-3, np.float32), np.array([50.0], np.float32)]),
            "tiny": (rng.standard_normal(n).astype(np.float32) * np.float32(1e-30)),
        }
        # ties: make max|x| = 127*0.5 so x/scale lands on exact .5 values
        xs_in["ties"][::gs] = np.float32(63.5)
        for k, x in xs_in.items():
            q = np.zeros(n, np.int8)
            s = np.zeros(n // gs, np.float32)
            xc = np.ascontiguousarray(x)
            lq8.refq8_quantize(q.ctypes.data_as(C.POINTER(C.c_int8)), s.ctypes.data_as(C.POINTER(C.c_float)),
                               xc.ctypes.data_as(C.POINTER(C.c_float)), n)
            cases[f"gs{gs}_{k}_x"] = xc
            cases[f"gs{gs}_{k}_q"] = q
            cases[f"gs{gs}_{k}_s"] = s
    np.savez_compressed(os.path.join(HERE, "quantize_golden.npz"), **cases)

    # ---- tokenizer known answers ----
    tok = l32.ref32_tok_open(os.path.join(REF, "assets/tokenizer.bin").encode(), 32000)
    kat = {
        "": [1],
        "I believe the meaning of life is": [1, 306, 4658, 278, 6593, 310, 2834, 338],
        "Simply put, the theory of relativity states that ":
            [1, 3439, 17632, 1925, 29892, 278, 6368, 310, 14215, 537, 5922, 393, 29871],
    }
    buf = (C.c_int * 4096)()
    for text, want in kat.items():  # vectors of /root/reference/test.c:49-63
        n = l32.ref32_encode(tok, text.encode(), 1, 0, buf)
        assert list(buf[:n]) == want, (text, list(buf[:n]))
    with open(os.path.join(REF, "assets/in/gen_in_128.txt")) as f:
        lines = f.read().split("\n")
    prompts = lines[1:1 + 138]
    enc = {}
    for p in prompts:
        n = l32.ref32_encode(tok, p.encode(), 1, 0, buf)
        enc[p] = list(buf[:n])
    with open(os.path.join(HERE, "tokenizer_golden.json"), "w") as f:
        json.dump({"test_c_vectors": kat, "gen_in_prompts": enc}, f)
    print("tokenizer: %d prompts encoded" % len(enc))


if __name__ == "__main__":
    main()
