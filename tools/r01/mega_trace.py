"""Development aid: per-phase time stamps of the persistent step kernel (THA_MEGA_TRACE=1)."""
import ctypes as C, os, sys
os.environ["THA_MEGA_TRACE"] = "1"
import numpy as np, torch
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests"))
from pkgload import load_pkg
ck = load_pkg("checkpoint"); eng = load_pkg("engine"); cabi = load_pkg("cabi")
B = int(sys.argv[1]) if len(sys.argv) > 1 else 1
pos0 = int(sys.argv[2]) if len(sys.argv) > 2 else 64
cfg = ck.STORIES110M
dw = eng.DeviceWeights(cfg, ck.random_weights(cfg, seed=1, std=0.02))
e = eng.Engine(dw, B, with_att=False)
lib = cabi.lib()
s = torch.cuda.Stream()
with torch.cuda.stream(s):
    tok = torch.randint(3, cfg.vocab_size, (B,), dtype=torch.int32, device="cuda")
    pos = torch.full((B,), pos0, dtype=torch.int32, device="cuda")
    nxt = torch.zeros(B, dtype=torch.int32, device="cuda")
    for _ in range(5):
        e.decode_greedy(tok, pos, nxt)
    s.synchronize()
nph = cfg.n_layers * 5 + 1
buf = np.zeros(148 * 1024 * 4, np.uint64)
grid = C.c_int(0)
lib.c.tha_mega_trace.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_int)]
assert lib.c.tha_mega_trace(buf.ctypes.data_as(C.c_void_p), buf.size, C.byref(grid)) == 0
g = grid.value
t = buf[: g * nph * 4].reshape(g, nph, 4).astype(np.int64)
t0 = t[:, 0, 0].min()
names = ["QKV", "ATTN", "WO", "W13", "W2"]
print(f"B={B} pos={pos0 + 4}: step span {(t[:, -1, 3].max() - t0) / 1e3:.1f} us")
tot = {}
for ph in range(nph):
    kind = "CLS" if ph == nph - 1 else names[ph % 5]
    wait = (t[:, ph, 1] - t[:, ph, 0]); stage = (t[:, ph, 2] - t[:, ph, 1]) if kind != "ATTN" else 0 * wait
    comp = (t[:, ph, 3] - (t[:, ph, 2] if kind != "ATTN" else t[:, ph, 1]))
    d = tot.setdefault(kind, [0, 0, 0, 0, 0])
    d[0] += 1; d[1] += wait.mean(); d[2] += np.mean(stage); d[3] += comp.mean(); d[4] += comp.max()
    if ph < 11 or ph == nph - 1:
        print(f"ph {ph:2d} {kind:4s} start {(t[:, ph, 0].min() - t0) / 1e3:7.2f} us  wait mean {wait.mean() / 1e3:5.2f} max {wait.max() / 1e3:5.2f}  "
              f"stage {np.mean(stage) / 1e3:5.2f}  compute mean {comp.mean() / 1e3:5.2f} max {comp.max() / 1e3:5.2f}")
for k, d in tot.items():
    print(f"{k:4s} x{d[0]:2d}: wait {d[1] / d[0] / 1e3:5.2f}  stage {d[2] / d[0] / 1e3:5.2f}  compute {d[3] / d[0] / 1e3:5.2f} (max {d[4] / d[0] / 1e3:5.2f}) us per phase")
