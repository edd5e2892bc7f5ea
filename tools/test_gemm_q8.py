"""Development aid: streaming rate of the tcgen05 matmul with fp32 and Q8_0 weights."""
import ctypes as C, os, sys
import numpy as np, torch
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests"))
from pkgload import load_pkg
lib = load_pkg("cabi").lib().c
st = torch.cuda.Stream()
def run(M, K, N, q8, reps=20):
    X = torch.randn(N, K, device="cuda")
    out = torch.zeros(N, M, device="cuda")
    if q8:
        Wq = torch.randint(-127, 128, (M, K), dtype=torch.int8, device="cuda")
        Ws = torch.rand(M, K // 64, device="cuda") * 0.01
        call = lambda: lib.tha_gemm_q8_tf32x3(st.cuda_stream, out.data_ptr(), M, Wq.data_ptr(), Ws.data_ptr(), 64, M, K, X.data_ptr(), N, K)
        wbytes = M * K * 1.0625
    else:
        W = torch.randn(M, K, device="cuda") * 0.05
        call = lambda: lib.tha_gemm_tf32x3(st.cuda_stream, out.data_ptr(), M, W.data_ptr(), M, K, X.data_ptr(), N, K)
        wbytes = M * K * 4
    torch.cuda.synchronize()
    with torch.cuda.stream(st):
        for _ in range(3): rc = call()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(st)
        for _ in range(reps): call()
        e1.record(st); st.synchronize()
    us = e0.elapsed_time(e1) / reps * 1e3
    nkb = M / 128 * K / 32 * ((N + 127) // 128 if N > 64 else 1)
    print(f"{'q8 ' if q8 else 'f32'} M={M} K={K} N={N}: rc={rc} abort={lib.tha_gemm_check_error()} {us:8.1f} us  W {wbytes/us/1e3:7.0f} GB/s  "
          f"{2*M*K*N/us/1e6:7.1f} TFLOP/s  {us*1e-6*1.965e9*148/nkb:7.0f} SM-cycles per tile-kblock", flush=True)
for shape in [(4096, 4096, 64), (11008, 4096, 64), (4096, 11008, 64), (32000, 4096, 64), (22016, 4096, 16), (22016, 4096, 128), (22016, 4096, 2048)]:
    for q8 in (False, True):
        run(*shape, q8)
