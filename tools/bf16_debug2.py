import os, sys, tempfile
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
from pkgload import load_pkg
from oracle_lib import Oracle
from gpu_util import small_model_files
import pathlib
sys.argv = sys.argv[:1]
import test_gpu_bf16 as T
orc = Oracle()
td = pathlib.Path(tempfile.mkdtemp())
eng = load_pkg("engine")
cfg, wb, path = T.bf16_model(td, "fusedA", T.MODELS["fusedA"])
for tokv in (242, 5):
    e = eng.Engine(eng.DeviceWeightsBF16(cfg, wb), 1)
    m = orc.open(path, bf16=True); mf = orc.open(path)
    s = m.new_seq(); sf = mf.new_seq()
    lg = e.forward_batch(np.array([tokv], np.int32), np.zeros(1, np.int32)).numpy()[0]
    a = s.forward(tokv, 0); b = sf.forward(tokv, 0)
    print(tokv, "vs bf16 oracle", np.abs(lg - a).max() / np.abs(a).max(), "vs fp32-act oracle", np.abs(lg - b).max() / np.abs(b).max())
