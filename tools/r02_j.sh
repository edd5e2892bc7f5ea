#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_bf16.py -m gpu -q -s -k "prefill" 2>&1 | grep -E "per sequence|passed|failed|Error|error" | cut -c1-300
python - <<'PY'
import sys, os
sys.path.insert(0, "tests"); sys.path.insert(0, ".")
import numpy as np, torch, dataclasses
from pkgload import load_pkg
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
print(f"7B bf16 prefill 2048: {ms:.2f} ms, {flops / ms / 1e9:.1f} useful TFLOP/s = {flops / ms / 1e9 / 1381.3:.3f} of sustained bf16, launches {e.last_step_launches()}, err {e.lib.c.tha_gemm_check_error()}")
PY
