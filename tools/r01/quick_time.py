"""Quick device-side timing of the resident decode step (development aid, not the bench)."""
import sys, os, time
import numpy as np, torch
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests"))
from pkgload import load_pkg
ck = load_pkg("checkpoint"); eng = load_pkg("engine"); cabi = load_pkg("cabi")

def run(cfg, Bs, steps, q8=False, pos0=0):
    w = ck.random_weights(cfg, seed=1, std=0.05)
    if q8:
        import tempfile
        with tempfile.TemporaryDirectory() as td:
            p = os.path.join(td, "m.bin"); ck.write_v2(p, cfg, w)
            c, gs, norms, qt, sh = ck.read_v2(p)
            dw = eng.DeviceWeightsQ8(c, gs, norms, qt)
    else:
        dw = eng.DeviceWeights(cfg, w)
    for B in Bs:
        e = eng.Engine(dw, B, with_att=False)
        s = torch.cuda.Stream()
        with torch.cuda.stream(s):
            tok = torch.randint(3, cfg.vocab_size, (B,), dtype=torch.int32, device="cuda")
            pos = torch.full((B,), pos0, dtype=torch.int32, device="cuda")
            nxt = torch.zeros(B, dtype=torch.int32, device="cuda")
            lg = torch.zeros(B, cfg.vocab_size, device="cuda")
            for _ in range(3):
                e.decode_step(tok, pos, lg, nxt); pos += 1
            s.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(s)
            for _ in range(steps):
                e.decode_step(tok, pos, lg, nxt); pos += 1
            e1.record(s); s.synchronize()
        ms = e0.elapsed_time(e1) / steps
        wb = dw.nbytes()
        print(f"{'q8' if q8 else 'f32'} d={cfg.dim} B={B:3d} pos~{pos0+steps//2}: {ms*1e3:8.1f} us/step  {B/ms*1e3:10.0f} tok/s  weights {wb/ms/1e6:7.0f} GB/s  launches {e.last_step_launches()}", flush=True)
        del e

if __name__ == "__main__":
    lib = cabi.lib()
    for pdl in (1, 0):
        lib.c.tha_set_option(b"pdl", pdl)
        print("pdl", pdl)
        run(ck.STORIES110M, [1, 4, 8, 16, 64], 64)
    lib.c.tha_set_option(b"pdl", 1)
    lib.c.tha_set_option(b"graph", 0); print("graph off"); run(ck.STORIES110M, [1, 8], 64)
    lib.c.tha_set_option(b"graph", 1)
    run(ck.STORIES110M, [1, 8], 64, q8=True)
    run(ck.STORIES110M, [1, 8], 64, pos0=900)
