"""Development aid: per-slot timeline of CTA 0 of the int8 tensor-core matmul."""
import ctypes as C, os, sys
import numpy as np, torch
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests"))
from pkgload import load_pkg
lib = load_pkg("cabi").lib().c
lib.tha_gemm_trace.argtypes = [C.c_void_p, C.c_int]
M, K, N = (int(a) for a in sys.argv[1:4]) if len(sys.argv) > 3 else (18944, 4096, 8)
Wq = torch.randint(-127, 128, (M, K), dtype=torch.int8, device="cuda"); Ws = torch.rand(M, K // 64, device="cuda")
Xq = torch.randint(-127, 128, (N, K), dtype=torch.int8, device="cuda"); ldxs = (N + 63) // 64 * 64
XsT = torch.rand(K // 64, ldxs, device="cuda"); out = torch.zeros(N, M, device="cuda")
torch.cuda.synchronize()
st = torch.cuda.Stream()
lib.tha_gemm_trace(None, 1)
for _ in range(3):
    lib.tha_matmul_q8_tc(st.cuda_stream, out.data_ptr(), M, Wq.data_ptr(), Ws.data_ptr(), 64, M, K, Xq.data_ptr(), XsT.data_ptr(), ldxs, N)
st.synchronize()
buf = np.zeros((10, 64), np.int64)
lib.tha_gemm_trace(buf.ctypes.data, 1)
t0 = buf[buf > 0].min()
names = ["tmaIss", "mmaFull", "commit(st)", "epiSaw(st)", "epiRel(st)", "prodWoke", "mmaTop", "mma0", "mma1", "mma3"]
print("i  " + " ".join(f"{n:>10s}" for n in names))
for i in range(min(34, K // 128)):
    print(f"{i:2d} " + " ".join(f"{(buf[s, i] - t0) if buf[s, i] else -1:10d}" for s in range(10)))
