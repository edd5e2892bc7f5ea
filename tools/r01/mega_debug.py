"""Development aid: run the persistent step kernel on a small model against the oracle, step by step."""
import os, sys, tempfile
import numpy as np, torch
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests"))
from gpu_util import close, engine_from_v0, small_model_files
from oracle_lib import Oracle
from pkgload import load_pkg

cfgs = {"fusedA": dict(dim=256, hidden_dim=704, n_layers=3, n_heads=8, n_kv_heads=8, vocab_size=512, seq_len=48)}
B = int(sys.argv[1]) if len(sys.argv) > 1 else 1
lib = load_pkg("cabi").lib()
orc = Oracle()
with tempfile.TemporaryDirectory() as td:
    cfg, w, v0, v2 = small_model_files(td, "dbg", cfgs["fusedA"])
    e = engine_from_v0(v0, B)
    m = orc.open(v0)
    seqs = [m.new_seq() for _ in range(B)]
    rng = np.random.default_rng(0)
    for step in range(12):
        tok = rng.integers(0, cfg.vocab_size, size=B).astype(np.int32)
        pos = np.full(B, step, np.int32)
        lg = e.forward_batch(tok, pos).numpy()
        for b in range(B):
            want = seqs[b].forward(tok[b], pos[b])
            ok, mad, mrd = close(lg[b], want, 1e-4)
            print(f"step {step} slot {b}: ok={ok} max abs {mad:.2e} launches {e.last_step_launches()}", flush=True)
print("done")
