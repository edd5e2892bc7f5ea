#!/usr/bin/env python
"""Device-side timeline of ONE decode step under graph replay + PDL (ktrace.cuh): per kernel, when its first / last block
entered, when griddepcontrol.wait returned, kernel-specific stamps, when the block finished.
    python tools/step_timeline.py [--batch 256] [--pos 10] [--layers 2] [--model s110m]"""
import argparse, ctypes as C, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
from pkgload import load_pkg
import torch
ap = argparse.ArgumentParser()
ap.add_argument("--batch", type=int, default=256)
ap.add_argument("--pos", type=int, default=10)
ap.add_argument("--show", type=int, default=24, help="kernels to print (from the second layer on)")
ap.add_argument("--model", default="s110m", choices=["s110m", "7b"])
ap.add_argument("--quant", default="f32", choices=["f32", "q8"])
ap.add_argument("--seq-len", type=int, default=0)
a = ap.parse_args()
ck, eng = load_pkg("checkpoint"), load_pkg("engine")
import dataclasses
cfg = ck.STORIES110M if a.model == "s110m" else ck.LLAMA2_7B
if a.seq_len:
    cfg = dataclasses.replace(cfg, seq_len=a.seq_len)
dw = eng.device_random_weights(cfg, q8=(a.quant == "q8"), seed=1)
e = eng.Engine(dw, a.batch, with_att=False)
B = a.batch
tok = torch.randint(3, cfg.vocab_size, (B,), dtype=torch.int32, device="cuda")
pos = torch.full((B,), a.pos, dtype=torch.int32, device="cuda")
nxt = torch.zeros(B, dtype=torch.int32, device="cuda")
lg = torch.zeros(B, cfg.vocab_size, device="cuda")
st = torch.cuda.Stream()
lib = e.lib.c
lib.tha_ktrace.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p]
with torch.cuda.stream(st):
    for _ in range(4):
        e.decode_step(tok, pos, lg, nxt)
    st.synchronize()
    assert lib.tha_ktrace(None, 0, 1, None) == 0
    e.decode_step(tok, pos, lg, nxt)
    st.synchronize()
buf = np.zeros((8192, 16), np.uint64)
n = C.c_int(0)
assert lib.tha_ktrace(buf.ctypes.data_as(C.c_void_p), 8192, 0, C.byref(n)) == 0
recs = buf[: n.value].astype(np.int64)
KIND = {1: "gemm", 2: "prep", 3: "rope", 4: "swiglu", 5: "attn", 6: "argmax", 7: "gemm_q8", 8: "other"}
recs = recs[np.argsort(recs[:, 0], kind="stable")]
t0 = int(recs[:, 0].min())
print(f"{n.value} records; step spans {(int(recs[:, :15].max()) - t0) / 1e3:.1f} us")
print("  us from the first stamp.  gemm: s2 = first operand stage ready (MMA starts), s3 = last MMA issued, s4 = accumulator complete, s5 = epilogue stores issued")
# group first/last block records of the same launch: consecutive records with the same kind
def fmt(r, k):
    return f"{(int(r[k]) - t0) / 1e3:8.2f}" if r[k] else "      --"
shown = 0
skip = 9 * 1 + 1   # start at layer 1
prev_end = None
launches = []
i = 0
for r in recs:
    kind = int(r[15]) & 0xFF
    last = bool(int(r[15]) & 0x100)
    launches.append((kind, last, r))
for kind, last, r in launches:
    extra = ""
    if kind in (1, 7):
        extra = f" z={(int(r[15]) >> 16) & 0xFFFF} items={int(r[15]) >> 32}"
    print(f"{KIND.get(kind, '?'):7s} {'last ' if last else 'first'} enter={fmt(r, 0)} dep={fmt(r, 1)} s2={fmt(r, 2)} s3={fmt(r, 3)} s4={fmt(r, 4)} s5={fmt(r, 5)} done={fmt(r, 6)}{extra}" + (" epi: " + " ".join(fmt(r, k) for k in range(7, 15)) if kind in (1, 7) else ""))
    shown += 1
    if shown >= a.show * 2 + 2 * skip:
        break
