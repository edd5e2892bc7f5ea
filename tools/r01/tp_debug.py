"""Development aid: one tensor-parallel configuration with all ranks on cuda:0; on a timed-out exchange print every rank's flags / epoch."""
import os, sys, tempfile, ctypes as C
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
from pkgload import load_pkg
from gpu_util import small_model_files
sys.path.insert(0, os.path.join(ROOT, "tests"))
import test_gpu_tp as T

name, tp, q8, B = sys.argv[1], int(sys.argv[2]), sys.argv[3] == "q8", int(sys.argv[4])
pdl = int(sys.argv[5]) if len(sys.argv) > 5 else 0
lib = load_pkg("cabi").lib()
lib.c.tha_set_option(b"pdl", pdl)
with tempfile.TemporaryDirectory() as td:
    cfg, w, v0, v2 = small_model_files(td, name, T.TP_MODELS[name])
    ranks = T.make_ranks(cfg, w, tp, B, q8)
    streams = [torch.cuda.Stream() for _ in range(tp)]
    logits = [torch.zeros(B, cfg.vocab_size, device="cuda") for _ in range(tp)]
    rng = np.random.default_rng(0)
    for step in range(int(os.environ.get('STEPS', '6'))):
        tok = rng.integers(0, cfg.vocab_size, size=B).astype(np.int32)
        pos = np.full(B, step % (cfg.seq_len - 1), np.int32)
        tok_d = [torch.from_numpy(tok).cuda() for _ in range(tp)]
        pos_d = [torch.from_numpy(pos).cuda() for _ in range(tp)]
        torch.cuda.synchronize()
        for r in range(tp):
            ranks[r].decode_step(tok_d[r], pos_d[r], logits[r], None, stream=streams[r])
        torch.cuda.synchronize()
        errs = [ranks[r].check_error() for r in range(tp)]
        ge = lib.c.tha_gemm_check_error()
        if any(errs) or ge or step % 50 == 0:
            print("step", step, "errs", errs, "gemm_err", ge, "launches", ranks[0].last_step_launches(), flush=True)
        if any(errs):
            for r in range(tp):
                hdr = (C.c_uint * 16)()
                lib.c.tha_memcpy_d2h.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t]
                lib.c.tha_memcpy_d2h(hdr, C.c_void_p(ranks[r].local_ptr()), 64)
                print(" rank", r, "flags", list(hdr)[:8], "epoch", hdr[8], "err", hdr[9], flush=True)
            break
        same = all(torch.equal(logits[0], logits[r]) for r in range(1, tp))
        if not same:
            print("   ranks differ at step", step, flush=True)
