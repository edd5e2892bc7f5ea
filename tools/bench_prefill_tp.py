"""Prompt ingestion of ONE replica split tensor-parallel over the GPUs of a box (tha_tp_prefill_f32), CUDA events, max over ranks:
python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 tools/bench_prefill_tp.py --model 7b --len 2048"""
import argparse, os, sys
import numpy as np, torch
import torch.distributed as dist
ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
sys.path.insert(0, os.path.join(ROOT, "tests")); sys.path.insert(0, ROOT)
from pkgload import load_pkg
import bench
ap = argparse.ArgumentParser()
ap.add_argument("--model", default="7b"); ap.add_argument("--seqs", type=int, default=1)
ap.add_argument("--len", type=int, default=2048); ap.add_argument("--reps", type=int, default=3)
ap.add_argument("--quant", default="f32")
a = ap.parse_args()
rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr)
dist.init_process_group("nccl", device_id=torch.device(f"cuda:{lr}"))
tpm = load_pkg("tp")
cfg, _ = bench.model_config(a.model)
T = a.seqs * a.len
dw = tpm.device_random_shard(cfg, world, rank, a.quant == "q8", seed=1337, device=f"cuda:{lr}")
r = tpm.TpRank(cfg, dw, world, rank, a.seqs, device=f"cuda:{lr}", max_rows=T)
r.connect_distributed()
rng = np.random.default_rng(0)
seqs = [(i, 0, rng.integers(3, cfg.vocab_size, size=a.len).astype(np.int32)) for i in range(a.seqs)]
st = torch.cuda.Stream()
logits = torch.zeros(a.seqs, cfg.vocab_size, device="cuda")
nxt = torch.zeros(a.seqs, dtype=torch.int32, device="cuda")
with torch.cuda.stream(st):
    torch.cuda.synchronize(); dist.barrier()
    r.prefill(seqs, logits, nxt, stream=st); st.synchronize()
    torch.cuda.synchronize(); dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(st)
    for _ in range(a.reps): r.prefill(seqs, logits, nxt, stream=st)
    e1.record(st); st.synchronize()
ms = load_pkg("dp").max_over_ranks(e0.elapsed_time(e1) / a.reps, "cuda")
bad = r.check_error() or load_pkg("cabi").lib().c.tha_gemm_check_error()
tok = [torch.zeros_like(nxt) for _ in range(world)]
dist.all_gather(tok, nxt)
same = all(torch.equal(tok[0], t) for t in tok)
d, h, L, kvd = cfg.dim, cfg.hidden_dim, cfg.n_layers, cfg.dim // cfg.n_heads * cfg.n_kv_heads
flops = 2 * T * L * (d * d * 2 + 2 * d * kvd + 3 * d * h) + 2 * L * a.seqs * a.len * a.len * d + 2 * a.seqs * cfg.vocab_size * d
if rank == 0:
    print(f"tp{world} {a.model} {a.quant} prefill {a.seqs} x {a.len}: {ms:.2f} ms  {T / ms * 1e3:.0f} prompt tok/s  {flops / ms / 1e9:.1f} useful TFLOP/s (whole replica)  "
          f"launches {r.last_step_launches()}  error={int(bool(bad))}  ranks agree on the next token: {same}", flush=True)
torch.cuda.synchronize(); dist.barrier()
r.close()
dist.destroy_process_group()
