"""Development aid: live per-kernel times of decode steps (events after every launch, graphs and PDL off).
   THA_TRACE=1 THA_GRAPH=0 THA_PDL=0 python tools/live_trace.py [bench.py args]"""
import os, sys
os.environ.setdefault("THA_TRACE", "1"); os.environ.setdefault("THA_GRAPH", "0"); os.environ.setdefault("THA_PDL", "0")
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import bench, torch, numpy as np
from pkgload import load_pkg
sys.argv = [sys.argv[0]] + sys.argv[1:]
args = bench.parse_args()
torch.cuda.set_device(0)
cfg, dw, _ = bench.build_weights(args, "cuda:0")
eng = load_pkg("engine")
e = eng.Engine(dw, args.batch, with_att=False)
B = args.batch
st = torch.cuda.Stream()
with torch.cuda.stream(st):
    tok = torch.randint(3, cfg.vocab_size, (B,), dtype=torch.int32, device="cuda")
    pos = torch.full((B,), args.pos0, dtype=torch.int32, device="cuda")
    nxt = torch.zeros(B, dtype=torch.int32, device="cuda")
    lg = torch.zeros(B, cfg.vocab_size, device="cuda")
    for _ in range(3): e.decode_greedy(tok, pos, nxt, lg, stream=st)
    st.synchronize()
    e.lib.c.tha_trace_dump(1 << 30)   # drop warm-up
    for _ in range(8): e.decode_greedy(tok, pos, nxt, lg, stream=st)
    st.synchronize()
    e.lib.c.tha_trace_dump(1)
