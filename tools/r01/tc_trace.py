"""Development aid: per-k-block timeline of one CTA of the tcgen05 matmul."""
import ctypes as C, os, sys
import numpy as np, torch
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests"))
from pkgload import load_pkg
lib = load_pkg("cabi").lib().c
lib.tha_gemm_trace.argtypes = [C.c_void_p, C.c_int]
M, K, N = (int(a) for a in sys.argv[1:4]) if len(sys.argv) > 3 else (4096, 4096, 64)
W = torch.randn(M, K, device="cuda"); X = torch.randn(N, K, device="cuda"); out = torch.zeros(N, M, device="cuda")
torch.cuda.synchronize()
st = torch.cuda.Stream()
lib.tha_gemm_trace(None, 1)
for _ in range(3):
    lib.tha_gemm_tf32x3(st.cuda_stream, out.data_ptr(), M, W.data_ptr(), M, K, X.data_ptr(), N, K)
st.synchronize()
buf = np.zeros((10, 64), np.int64)
lib.tha_gemm_trace(buf.ctypes.data, 1)
t0 = buf[buf > 0].min()
names = ["Wissue", "Xissue", "convW", "convSlot", "convDone", "mmaSaw", "mmaIss", "commit", "loopTop", "waited"]
print("kb " + " ".join(f"{n:>9s}" for n in names))
for i in range(min(40, K // 32)):
    print(f"{i:2d} " + " ".join(f"{(buf[s, i] - t0) if buf[s, i] else -1:9d}" for s in range(10)))
