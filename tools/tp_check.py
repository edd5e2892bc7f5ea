#!/usr/bin/env python
"""Tensor-parallel parity on REAL multiple GPUs (one process per GPU, IPC-mapped exchange blocks, PDL on):

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tools/tp_check.py

Every rank builds the same seeded small model, keeps its shard, and runs ragged decode steps; rank 0 compares the
logits with the oracle of the UNSPLIT model (fp32 1e-4, Q8_0 norm-wise 5e-2) and all ranks must agree bit for bit.
tests/test_gpu_tp.py is the single-GPU form of the same check (all ranks on cuda:0)."""
import os
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    import torch
    import torch.distributed as dist
    from pkgload import load_pkg
    from gpu_util import assert_close, assert_close_scaled, small_model_files
    from oracle_lib import Oracle
    rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(lr)
    dist.init_process_group("nccl", device_id=torch.device(f"cuda:{lr}"))
    ck, eng, tpm = load_pkg("checkpoint"), load_pkg("engine"), load_pkg("tp")
    tp = world
    models = {
        2: dict(dim=256, hidden_dim=704, n_layers=3, n_heads=8, n_kv_heads=8, vocab_size=512, seq_len=48),
        4: dict(dim=512, hidden_dim=1408, n_layers=2, n_heads=8, n_kv_heads=8, vocab_size=384, seq_len=40),
        8: dict(dim=1024, hidden_dim=2752, n_layers=2, n_heads=16, n_kv_heads=8, vocab_size=384, seq_len=40),
    }
    ok = True
    with tempfile.TemporaryDirectory() as td:
        cfg, w, v0, v2 = small_model_files(td, f"tp{tp}_{rank}", models[tp])
        for q8 in (False, True):
            for B in (3, 20):
                if q8:
                    gs = ck.v2_group_size(cfg)
                    qt = {}
                    for n in ck.V2_QORDER:
                        t = w[n]
                        if n == "tok_emb":
                            q, s = ck.quantize_q80(t, gs)
                            qt[n] = (q.reshape(t.shape), s)
                        else:
                            qs, ss = zip(*[ck.quantize_q80(t[l], gs) for l in range(cfg.n_layers)])
                            qt[n] = (np.stack(qs).reshape(t.shape), np.stack(ss))
                    qt["wcls"] = qt["tok_emb"]
                    lcfg, lqt = tpm.shard_q8(cfg, gs, qt, tp, rank)
                    dw = eng.DeviceWeightsQ8(lcfg, gs, {k: w[k] for k in ("rms_att", "rms_ffn", "rms_final")}, lqt)
                else:
                    lcfg, lw = tpm.shard_f32(cfg, w, tp, rank)
                    dw = eng.DeviceWeights(lcfg, lw)
                r = tpm.TpRank(cfg, dw, tp, rank, B)
                r.connect_distributed()
                m = Oracle().open(v2 if q8 else v0, quantized=q8) if rank == 0 else None
                seqs = [m.new_seq() for _ in range(B)] if rank == 0 else None
                rng = np.random.default_rng(B)
                logits = torch.zeros(B, cfg.vocab_size, device="cuda")
                stream = torch.cuda.Stream()
                for step in range(10):
                    tok = rng.integers(0, cfg.vocab_size, size=B).astype(np.int32)
                    pos = np.array([step - (b % 3) for b in range(B)], np.int32)
                    pos[pos < 0] = -1
                    tok_d, pos_d = torch.from_numpy(tok).cuda(), torch.from_numpy(pos).cuda()
                    torch.cuda.synchronize()
                    dist.barrier()
                    r.decode_step(tok_d, pos_d, logits, None, stream=stream)
                    stream.synchronize()
                    if r.check_error():
                        print(f"rank {rank}: exchange timed out (q8={q8} B={B} step {step})", flush=True)
                        ok = False
                        break
                    got = logits.cpu().numpy()
                    act = pos >= 0
                    digest = torch.tensor([float(np.abs(got[act]).sum()), float(got[act].sum())], dtype=torch.float64, device="cuda")
                    all_d = [torch.zeros_like(digest) for _ in range(world)]
                    dist.all_gather(all_d, digest)
                    if any(not torch.equal(all_d[0], d) for d in all_d):
                        print(f"rank {rank}: ranks disagree (q8={q8} B={B} step {step})", flush=True)
                        ok = False
                    if rank == 0:
                        for b in range(B):
                            if pos[b] < 0:
                                continue
                            want = seqs[b].forward(int(tok[b]), int(pos[b]))
                            try:
                                (assert_close_scaled if q8 else assert_close)(got[b], want, 5e-2 if q8 else 1e-4, f"tp{tp} q8={q8} B={B} step {step} slot {b}")
                            except AssertionError as ex:
                                print("MISMATCH", ex, flush=True)
                                ok = False
                torch.cuda.synchronize()
                dist.barrier()   # peers may still be reading this rank's exchange block
                r.close()
                if rank == 0:
                    print(f"tp{tp} q8={q8} B={B}: {'ok' if ok else 'FAILED'}", flush=True)
    flag = torch.tensor([0 if ok else 1], device="cuda")
    dist.all_reduce(flag)
    dist.barrier()
    dist.destroy_process_group()
    if rank == 0:
        print("TP_CHECK", "PASS" if int(flag.item()) == 0 else "FAIL", flush=True)
    return 0 if int(flag.item()) == 0 else 1


if __name__ == "__main__":
    sys.exit(main())
