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
    if True:
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
if not os.environ.get("Q8I_SHORT"):
    for shape in [(4096, 4096, 64), (11008, 4096, 64), (4096, 11008, 64), (32000, 4096, 64), (22016, 4096, 16), (22016, 4096, 128), (22016, 4096, 2048)]:
        run(*shape, False)

print("---- int8 tensor cores (tha_matmul_q8_tc)")
def run_i8(M, K, N, reps=20):
    Wq = torch.randint(-127, 128, (M, K), dtype=torch.int8, device="cuda")
    Ws = torch.rand(M, K // 64, device="cuda") * 0.01
    Xq = torch.randint(-127, 128, (N, K), dtype=torch.int8, device="cuda")
    ldxs = (N + 63) // 64 * 64
    XsT = torch.rand(K // 64, ldxs, device="cuda")
    out = torch.zeros(N, M, device="cuda")
    call = lambda: lib.tha_matmul_q8_tc(st.cuda_stream, out.data_ptr(), M, Wq.data_ptr(), Ws.data_ptr(), 64, M, K, Xq.data_ptr(), XsT.data_ptr(), ldxs, N)
    torch.cuda.synchronize()
    with torch.cuda.stream(st):
        for _ in range(3): rc = call()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(st)
        for _ in range(reps): call()
        e1.record(st); st.synchronize()
    us = e0.elapsed_time(e1) / reps * 1e3
    print(f"i8  M={M} K={K} N={N}: rc={rc} abort={lib.tha_gemm_check_error()} {us:8.1f} us  W {M*K*1.0625/us/1e3:7.0f} GB/s  {2*M*K*N/us/1e6:7.1f} TOP/s", flush=True)
shapes = [(4096, 4096, 1), (4096, 4096, 8), (4096, 4096, 64), (22016, 4096, 1), (22016, 4096, 8), (22016, 4096, 16), (22016, 4096, 32), (22016, 4096, 64),
          (22016, 4096, 256), (4096, 11008, 8), (32000, 4096, 8), (32000, 4096, 64), (32000, 768, 64), (2304, 768, 8)]
if os.environ.get("Q8I_SHORT"): shapes = [(151552, 4096, 8), (151552, 4096, 16), (151552, 4096, 32), (151552, 4096, 64), (56832, 11008, 8)]
for shape in shapes:
    run_i8(*shape)
