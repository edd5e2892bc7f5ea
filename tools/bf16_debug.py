"""Development aid: bf16 path vs the oracle, buffer by buffer (one decode step at position 0)."""
import os, sys, tempfile
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
from pkgload import load_pkg
from oracle_lib import Oracle
from gpu_util import small_model_files
ck, eng = load_pkg("checkpoint"), load_pkg("engine")
td = tempfile.mkdtemp()
cfg, w, v0, v2 = small_model_files(td, "x", dict(dim=256, hidden_dim=704, n_layers=3, n_heads=8, n_kv_heads=8, vocab_size=512, seq_len=48))
wb = ck.bf16_weights(w)
p = os.path.join(td, "b.bin"); ck.write_v0(p, cfg, wb)
orc = Oracle()
for B in (1, 16):
    e = eng.Engine(eng.DeviceWeightsBF16(cfg, wb), B)
    m = orc.open(p, bf16=True)
    mf = orc.open(p)
    s = m.new_seq(); sf = mf.new_seq()
    T = int(sys.argv[1]) if len(sys.argv) > 1 else 5
    tok = np.full(B, T, np.int32); pos = np.zeros(B, np.int32)
    lg = e.forward_batch(tok, pos).numpy()[0]
    want = s.forward(T, 0); wantf = sf.forward(T, 0)
    def rel(a, b): return float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-30))
    print(f"B={B} logits vs bf16 oracle {rel(lg, want):.3e}   vs fp32-activation oracle {rel(lg, wantf):.3e}")
    kc = e.state.key_cache[0].cpu().numpy(); vc = e.state.value_cache[0].cpu().numpy()
    ok = s.buf("key_cache").reshape(cfg.n_layers, cfg.seq_len, -1); ov = s.buf("value_cache").reshape(cfg.n_layers, cfg.seq_len, -1)
    okf = sf.buf("key_cache").reshape(cfg.n_layers, cfg.seq_len, -1)
    for l in range(cfg.n_layers):
        print(f"   layer {l}: K row {rel(kc[l, 0], ok[l, 0]):.3e} (vs fp32-act {rel(kc[l, 0], okf[l, 0]):.3e})  V row {rel(vc[l, 0], ov[l, 0]):.3e}")
    print("   q", rel(e.state.q[0].cpu().numpy(), s.buf("q")), " x(final, un-normalised on GPU) n/a")
