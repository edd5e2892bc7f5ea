#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
python -m pytest tests/test_gpu_bf16.py -q -s 2>&1 | grep -E "rows|per sequence|passed|failed|Error" | cut -c1-300
python - <<'PY'
import sys, os, time
sys.path.insert(0, "tests"); sys.path.insert(0, ".")
import numpy as np, torch, dataclasses
from pkgload import load_pkg
import bench
eng, ck = load_pkg("engine"), load_pkg("checkpoint")
dev = "cuda"
cfg = dataclasses.replace(ck.LLAMA2_7B, seq_len=2304)
dw = eng.DeviceWeightsBF16.random(cfg)
e = eng.Engine(dw, 1, with_att=False)
rng = np.random.default_rng(0)
T = 2048
seqs = [(0, 0, rng.integers(3, cfg.vocab_size, size=T).astype(np.int32))]
st = torch.cuda.Stream()
lg = torch.zeros(1, cfg.vocab_size, device=dev); nx = torch.zeros(1, dtype=torch.int32, device=dev)
with torch.cuda.stream(st):
    e.prefill(seqs, lg, nx, stream=st); st.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(st)
    for _ in range(3): e.prefill(seqs, lg, nx, stream=st)
    b.record(st); st.synchronize()
ms = a.elapsed_time(b) / 3
d, h, L, kvd = cfg.dim, cfg.hidden_dim, cfg.n_layers, cfg.kv_dim
flops = 2.0 * T * L * (2 * d * d + 2 * d * kvd + 3 * d * h) + 2.0 * L * T * T * d + 2.0 * cfg.vocab_size * d
print(f"7B bf16 prefill 2048: {ms:.2f} ms, {flops / ms / 1e9:.1f} useful TFLOP/s, launches {e.last_step_launches()}, err {e.lib.c.tha_gemm_check_error()}")
for B in (1, 8):
    e2 = eng.Engine(dw, B, with_att=False)
    v, msd = bench.resident_tok_s(e2, cfg, B, 3, 10, dev, pos0=0)
    nb = bench.step_bytes(cfg, dw.nbytes(), 6, B)
    print(f"7B bf16 decode B={B}: {v:.1f} tok/s, {msd:.3f} ms/step, {nb / msd / 1e6:.0f} GB/s = {nb / msd / 1e6 / 6532:.3f} of HBM, launches {e2.last_step_launches()}")
    del e2
c64 = dataclasses.replace(ck.LLAMA2_7B, seq_len=512); dw.cfg = c64
e3 = eng.Engine(dw, 64, with_att=False)
v, msd = bench.resident_tok_s(e3, c64, 64, 3, 10, dev)
print(f"7B bf16 decode B=64: {v:.1f} tok/s, {msd:.3f} ms/step")
PY
