#!/usr/bin/env python
"""Llama-2-7B bf16 decode at small slot counts: the fused GEMV step (default for 1-2 slots) against the tensor-core row pipeline
(tha_set_option("tc", 1)).   python tools/bf16_small.py"""
import dataclasses, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests")); sys.path.insert(0, ROOT)
from pkgload import load_pkg
import bench
eng = load_pkg("engine")
lib = load_pkg("cabi").lib()
cfg, _ = bench.model_config("7b")
cfg = dataclasses.replace(cfg, seq_len=512)
dw = eng.DeviceWeightsBF16.random(cfg, seed=1, device="cuda")
W = 2.0 * (cfg.n_layers * (2 * cfg.dim * cfg.dim + 2 * cfg.dim * cfg.kv_dim + 3 * cfg.dim * cfg.hidden_dim) + cfg.vocab_size * cfg.dim)
for B in (1, 2, 3, 4, 8):
    for tc in (-1, 1):
        lib.check(lib.c.tha_set_option(b"tc", tc))
        e = eng.Engine(dw, B, device="cuda", with_att=False)
        v, ms = bench.resident_tok_s(e, cfg, B, 4, 12, "cuda", pos0=5)
        print(f"7B bf16 B={B} tc={tc:2d}: {v:8.1f} tok/s  {ms:.3f} ms/step  {W / ms / 1e9 / 6531.9 * 1e3 / 1e3:.3f} of HBM  launches {e.last_step_launches()}")
        del e
