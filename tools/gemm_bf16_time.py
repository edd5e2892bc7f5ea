#!/usr/bin/env python
"""Time tha_gemm_bf16 alone at the Llama-2-7B prompt shapes: python tools/gemm_bf16_time.py [N]   (THA_BF16_MT=1|2 forces the item height)"""
import ctypes as C, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
from pkgload import load_pkg
import torch
lib = load_pkg("cabi").lib()
N = int(sys.argv[1]) if len(sys.argv) > 1 else 2048
st = torch.cuda.current_stream().cuda_stream
P = lambda t: C.c_void_p(t.data_ptr())
for name, M, K in (("wo", 4096, 4096), ("wq|wk|wv", 12288, 4096), ("w1|w3", 22016, 4096), ("w2", 4096, 11008)):
    W = (torch.randn(M, K, device="cuda") * 0.05).to(torch.bfloat16)
    X = torch.randn(N, K, device="cuda").to(torch.bfloat16)
    out = torch.zeros(N, M, device="cuda")
    for _ in range(2):
        lib.check(lib.c.tha_gemm_bf16(C.c_void_p(st), P(out), M, P(W), M, K, P(X), N, K))
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        lib.check(lib.c.tha_gemm_bf16(C.c_void_p(st), P(out), M, P(W), M, K, P(X), N, K))
    e1.record(); torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / 5
    print(f"{name:9s} {M:6d} x {K:6d} x {N}: {us:8.1f} us  {2.0 * M * K * N / us / 1e6:7.1f} TFLOP/s  MT={os.environ.get('THA_BF16_MT', 'auto')}")
