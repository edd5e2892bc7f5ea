#!/usr/bin/env python
"""Time tha_sample alone: python tools/sampler_time.py [--batch 256] [--vocab 32000]   (THA_SAMPLE_RESIDENT=0: the HBM / L2 form)"""
import argparse, ctypes as C, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
from pkgload import load_pkg
import torch
ap = argparse.ArgumentParser()
ap.add_argument("--batch", type=int, default=256); ap.add_argument("--vocab", type=int, default=32000)
a = ap.parse_args()
lib = load_pkg("cabi").lib()
B, V = a.batch, a.vocab
src = torch.randn(B, V, device="cuda") * 2.0
lg = src.clone()
rng = torch.arange(1, B + 1, dtype=torch.int64, device="cuda") * 7919
pos = torch.zeros(B, dtype=torch.int32, device="cuda")
nxt = torch.zeros(B, dtype=torch.int32, device="cuda")
st = torch.cuda.current_stream().cuda_stream
P = lambda t: C.c_void_p(t.data_ptr())
def once():
    lib.check(lib.c.tha_sample(C.c_void_p(st), B, P(lg), V, C.c_float(1.0), C.c_float(0.9), P(rng), None, None, P(pos), P(nxt)))
for _ in range(3):
    lg.copy_(src); once()
torch.cuda.synchronize()
ts = []
for _ in range(10):
    lg.copy_(src)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); once(); e1.record(); torch.cuda.synchronize()
    ts.append(e0.elapsed_time(e1) * 1e3)
lib.c.tha_ktrace.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p]
lib.c.tha_ktrace(None, 0, 1, None)
lg.copy_(src); once(); torch.cuda.synchronize()
buf = np.zeros((64, 16), np.uint64); nrec = C.c_int(0)
lib.c.tha_ktrace(buf.ctypes.data_as(C.c_void_p), 64, 0, C.byref(nrec))
for r in buf[: nrec.value].astype(np.int64):
    t0 = r[0]
    print("  block", "last " if r[15] & 0x100 else "first", "us from entry: row staged + max", (r[2] - t0) / 1e3, "exp + sum", (r[3] - t0) / 1e3,
          "normalised", (r[4] - t0) / 1e3, "nucleus found", (r[5] - t0) / 1e3, "token", (r[6] - t0) / 1e3,
          "| first selection, end of level 0..6:", " ".join(f"{(r[7 + l] - t0) / 1e3:.2f}" for l in range(7)))
print(f"tha_sample {B} x {V}, temperature 1.0, top-p 0.9: median {sorted(ts)[len(ts) // 2]:.1f} us (min {min(ts):.1f})  resident={os.environ.get('THA_SAMPLE_RESIDENT', '1')}")
