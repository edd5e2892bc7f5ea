"""hip_llama.cpp_b200/checkpoint.py writes the reference's two file formats byte for byte.
The *_v0.bin / *_v2.bin fixtures were written by the reference's own train/export.py."""
import os

import numpy as np
import pytest

from pkgload import load_pkg

ck = load_pkg("checkpoint")


@pytest.mark.parametrize("name", ["tinyA", "tinyB_gqa"])
def test_v0_and_v2_roundtrip_byte_identical(golden_dir, tmp_path, name):
    v0 = os.path.join(golden_dir, f"{name}_v0.bin")
    v2 = os.path.join(golden_dir, f"{name}_v2.bin")
    cfg, w, shared = ck.read_v0(v0)
    assert shared
    out0 = str(tmp_path / "o_v0.bin")
    ck.write_v0(out0, cfg, w)
    assert open(out0, "rb").read() == open(v0, "rb").read()
    out2 = str(tmp_path / "o_v2.bin")
    gs = ck.write_v2(out2, cfg, w, group_size=64)
    assert open(out2, "rb").read() == open(v2, "rb").read()
    c2, gs2, norms, qt, shared2 = ck.read_v2(v2)
    assert (c2, gs2, shared2) == (cfg, gs, True)
    assert np.array_equal(norms["rms_att"], w["rms_att"])
    q, s = ck.quantize_q80(w["wq"][1], gs)
    assert np.array_equal(qt["wq"][0][1].reshape(-1), q) and np.array_equal(qt["wq"][1][1], s)


def test_named_shapes():
    assert ck.hidden_dim_for(768, 32) == 2048
    assert ck.hidden_dim_for(4096, 256) == 11008
    n = sum(int(np.prod(s)) for s in ck.tensor_shapes(ck.STORIES110M).values())
    assert n == 109_529_856
    assert 28 + 4 * n + 2 * 4 * (1024 * 64 // 2) == 438_381_596


def test_random_weights_seeded():
    c = ck.Config(64, 192, 2, 4, 2, 128, 16)
    a = ck.random_weights(c, seed=5)
    b = ck.random_weights(c, seed=5)
    assert all(np.array_equal(a[k], b[k]) for k in a)
    assert a["wk"].shape == (2, 32, 64)
    assert abs(float(a["wq"].std()) - 0.02) < 2e-3 and abs(float(a["wo"].std()) - 0.01) < 1e-3
