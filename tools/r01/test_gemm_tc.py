"""Development aid: tcgen05 3xTF32 weight matmul against fp64, and its streaming rate."""
import ctypes as C, os, sys
import numpy as np, torch
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests"))
from pkgload import load_pkg
cabi = load_pkg("cabi")
lib = cabi.lib().c
lib.tha_gemm_tf32x3.argtypes = [C.c_void_p, C.c_void_p, C.c_longlong, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_longlong]
lib.tha_gemm_check_error.restype = C.c_int
torch.manual_seed(0)
st = torch.cuda.Stream()
def run(M, K, N, reps=0):
    W = torch.randn(M, K, device="cuda") * 0.05
    X = torch.randn(N, K, device="cuda")
    out = torch.full((N, M), float("nan"), device="cuda")
    torch.cuda.synchronize()
    with torch.cuda.stream(st):
        rc = lib.tha_gemm_tf32x3(st.cuda_stream, out.data_ptr(), M, W.data_ptr(), M, K, X.data_ptr(), N, K)
        st.synchronize()
    err = lib.tha_gemm_check_error()
    want = (X.double() @ W.double().T)
    ref32 = (X @ W.T)
    e = (out.double() - want).abs().max().item() / want.abs().max().item()
    e32 = (ref32.double() - want).abs().max().item() / want.abs().max().item()
    msg = f"M={M} K={K} N={N}: rc={rc} abort={err} relerr {e:.2e} (torch fp32 matmul {e32:.2e})"
    if reps:
        with torch.cuda.stream(st):
            for _ in range(3): lib.tha_gemm_tf32x3(st.cuda_stream, out.data_ptr(), M, W.data_ptr(), M, K, X.data_ptr(), N, K)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(st)
            for _ in range(reps): lib.tha_gemm_tf32x3(st.cuda_stream, out.data_ptr(), M, W.data_ptr(), M, K, X.data_ptr(), N, K)
            e1.record(st); st.synchronize()
        us = e0.elapsed_time(e1) / reps * 1e3
        msg += f"  {us:8.1f} us  W {M*K*4/us/1e3:7.0f} GB/s  {2*M*K*N/us/1e6:7.1f} TFLOP/s"
    print(msg, flush=True)
    return e
for shape in [(128, 32, 16), (128, 64, 16), (256, 768, 8), (768, 768, 64), (2304, 768, 33), (4096, 768, 64), (768, 2048, 128),
              (32000, 768, 64), (100, 132, 5), (4096, 4096, 256), (4096, 4096, 300)]:
    run(*shape)
for shape in [(32000, 768, 64), (4096, 4096, 16), (4096, 4096, 64), (11008, 4096, 64), (4096, 11008, 64), (32000, 4096, 64),
              (4096, 4096, 256), (11008, 4096, 2048), (4096, 4096, 2048)]:
    run(*shape, reps=20)
