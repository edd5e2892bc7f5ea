#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
THA_TRACE=1 python - <<'PY'
import sys, os
sys.path.insert(0, "tests"); sys.path.insert(0, ".")
import numpy as np, torch, dataclasses
from pkgload import load_pkg
eng, ck = load_pkg("engine"), load_pkg("checkpoint")
cfg = dataclasses.replace(ck.LLAMA2_7B, seq_len=2304)
dw = eng.DeviceWeightsBF16.random(cfg)
e = eng.Engine(dw, 1, with_att=False)
rng = np.random.default_rng(0)
T = 2048
seqs = [(0, 0, rng.integers(3, cfg.vocab_size, size=T).astype(np.int32))]
lg = torch.zeros(1, cfg.vocab_size, device="cuda"); nx = torch.zeros(1, dtype=torch.int32, device="cuda")
e.prefill(seqs, lg, nx); torch.cuda.synchronize()
e.lib.c.tha_trace_dump(0)
e.prefill(seqs, lg, nx); torch.cuda.synchronize()
e.lib.c.tha_trace_dump(0)
PY
