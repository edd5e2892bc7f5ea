"""Development aid: time tha_prefill_f32 (CUDA events) and report prompt tok/s and useful TFLOP/s.
python tools/bench_prefill.py --model 7b --seqs 1 --len 2048"""
import argparse, os, sys, dataclasses
import numpy as np, torch
ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
sys.path.insert(0, os.path.join(ROOT, "tests")); sys.path.insert(0, ROOT)
from pkgload import load_pkg
import bench
ap = argparse.ArgumentParser()
ap.add_argument("--model", default="7b"); ap.add_argument("--seqs", type=int, default=1)
ap.add_argument("--len", type=int, default=2048); ap.add_argument("--reps", type=int, default=3)
ap.add_argument("--logits", type=int, default=1)
ap.add_argument("--bf16", type=int, default=0, help="1: the bf16 weight path (tha_prefill_bf16)")
a = ap.parse_args()
eng = load_pkg("engine")
args = argparse.Namespace(model=a.model, quant="f32", seq_len=0)
if a.bf16:
    cfg, _ck = bench.model_config(a.model)
    dw = eng.DeviceWeightsBF16.random(cfg, seed=1337, device="cuda")
else:
    cfg, dw, _ = bench.build_weights(args, "cuda")
e = eng.Engine(dw, a.seqs, device="cuda", with_att=False)
rng = np.random.default_rng(0)
seqs = [(i, 0, rng.integers(3, cfg.vocab_size, size=a.len).astype(np.int32)) for i in range(a.seqs)]
st = torch.cuda.Stream()
logits = torch.zeros(a.seqs, cfg.vocab_size, device="cuda") if a.logits else None
nxt = torch.zeros(a.seqs, dtype=torch.int32, device="cuda") if a.logits else None
with torch.cuda.stream(st):
    e.prefill(seqs, logits, nxt, stream=st); st.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(st)
    for _ in range(a.reps): e.prefill(seqs, logits, nxt, stream=st)
    e1.record(st); st.synchronize()
ms = e0.elapsed_time(e1) / a.reps
err = load_pkg("cabi").lib().c.tha_gemm_check_error()
T = a.seqs * a.len
d, h, L, kvd = cfg.dim, cfg.hidden_dim, cfg.n_layers, cfg.dim // cfg.n_heads * cfg.n_kv_heads
pmm = L * (d * d * 2 + 2 * d * kvd + 3 * d * h)
flops = 2 * T * pmm + 2 * L * a.seqs * a.len * a.len * d + 2 * a.seqs * cfg.vocab_size * d
print(f"{a.model} {'bf16' if a.bf16 else 'f32'} prefill {a.seqs} x {a.len}: {ms:.2f} ms  {T / ms * 1e3:.0f} prompt tok/s  {flops / ms / 1e9:.1f} useful TFLOP/s  launches {e.last_step_launches()} abort={err}")
