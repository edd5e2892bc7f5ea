#!/usr/bin/env python
"""Timeline of one chain launch (chain_tc.cu) inside a stories110M fp32 decode step: %globaltimer stamps of CTAs 0 and 97.
    THA_GRAPH=0 python tools/chain_trace.py [--batch 256] [--launch 5] [--pos 10]"""
import argparse, ctypes as C, os, sys
import numpy as np
os.environ.setdefault("THA_GRAPH", "0")
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
from pkgload import load_pkg
import torch
ap = argparse.ArgumentParser()
ap.add_argument("--batch", type=int, default=256)
ap.add_argument("--launch", type=int, default=5, help="which chain launch of the step (1-based; 13 = the one with the classifier)")
ap.add_argument("--pos", type=int, default=10)
a = ap.parse_args()
ck, eng = load_pkg("checkpoint"), load_pkg("engine")
cfg = ck.STORIES110M
dw = eng.device_random_weights(cfg, q8=False, seed=1)
e = eng.Engine(dw, a.batch, with_att=False)
B = a.batch
tok = torch.randint(3, cfg.vocab_size, (B,), dtype=torch.int32, device="cuda")
pos = torch.full((B,), a.pos, dtype=torch.int32, device="cuda")
nxt = torch.zeros(B, dtype=torch.int32, device="cuda")
lg = torch.zeros(B, cfg.vocab_size, device="cuda")
st = torch.cuda.Stream()
lib = e.lib.c
with torch.cuda.stream(st):
    for _ in range(3):
        e.decode_step(tok, pos, lg, nxt)
    st.synchronize()
    lib.tha_chain_trace.argtypes = [C.c_void_p, C.c_int]
    lib.tha_chain_trace(None, a.launch)
    e.decode_step(tok, pos, lg, nxt)
    st.synchronize()
buf = np.zeros((2, 64, 16), np.uint64)
lib.tha_chain_trace(buf.ctypes.data_as(C.c_void_p), 0)
names = ["Xdep", "Xissued", "MMA1st", "MMAdone", "accum", "tmem_out", "splits_in", "share_done", "signalled", "Wissued"]
t0 = min(int(x) for x in buf[:, :, :10].reshape(-1) if x > 0)
for cta in range(2):
    print(f"--- CTA {0 if cta == 0 else 97}  (us from the first stamp)")
    for it in range(64):
        row = buf[cta, it]
        if not row[:10].any():
            continue
        hdr = int(row[10])
        cells = " ".join(f"{names[k]}={(int(row[k]) - t0) / 1e3:8.2f}" if row[k] else f"{names[k]}=   --   " for k in range(10))
        print(f"item {it:2d} phase {hdr // 1000} z={hdr % 1000:2d}: {cells}")
