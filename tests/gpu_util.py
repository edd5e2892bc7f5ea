"""Shared helpers for the -m gpu parity tests (CUDA path vs the oracle)."""
import ctypes as C
import os

import numpy as np

from pkgload import load_pkg


def close(got, want, eps):
    """The reference comparers' rule: pass if abs <= eps OR rel <= eps (scripts/test/*.test.cpp)."""
    got = np.asarray(got, np.float64); want = np.asarray(want, np.float64)
    ad = np.abs(got - want)
    rd = ad / np.maximum(np.abs(want), 1e-30)
    ok = (ad <= eps) | (rd <= eps)
    return bool(ok.all()), float(ad.max(initial=0.0)), float(np.where(ok, 0, rd).max(initial=0.0))


def assert_close(got, want, eps, what=""):
    ok, mad, mrd = close(got, want, eps)
    assert ok, f"{what}: max abs err {mad:.3e}, worst failing rel err {mrd:.3e}, eps {eps}"


def assert_close_scaled(got, want, eps, what=""):
    """Norm-wise check: max |got - want| <= eps * max |want| (used where the comparison is
    dominated by reordered fp32 sums or by Q8_0 rounding flips; the element-wise rule above is
    meaningless for outputs that cancel to ~0)."""
    got = np.asarray(got, np.float64); want = np.asarray(want, np.float64)
    err = float(np.abs(got - want).max(initial=0.0)); scale = float(np.abs(want).max(initial=0.0))
    assert err <= eps * max(scale, 1e-30), f"{what}: max abs err {err:.3e} vs scale {scale:.3e} (ratio {err / max(scale, 1e-30):.3e}, eps {eps})"


def tdev(a, dtype=None):
    import torch
    t = torch.from_numpy(np.ascontiguousarray(a))
    if dtype is not None:
        t = t.to(dtype)
    return t.cuda()


def P(t):
    return C.c_void_p(t.data_ptr()) if t is not None else None


def stream_ptr():
    import torch
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def small_model_files(tmp_dir, name, cfg_kwargs, seed=3, std=0.08):
    """Random-init tiny model written in both file formats (byte layout == reference exporter)."""
    ck = load_pkg("checkpoint")
    cfg = ck.Config(**cfg_kwargs)
    w = ck.random_weights(cfg, seed=seed, std=std)
    rng = np.random.default_rng(seed + 1)
    for k in ("rms_att", "rms_ffn", "rms_final"):
        w[k] = (1.0 + 0.1 * rng.standard_normal(w[k].shape)).astype(np.float32)
    v0 = os.path.join(tmp_dir, f"{name}_v0.bin")
    v2 = os.path.join(tmp_dir, f"{name}_v2.bin")
    ck.write_v0(v0, cfg, w)
    ck.write_v2(v2, cfg, w)
    return cfg, w, v0, v2


def engine_from_v0(path, batch):
    ck = load_pkg("checkpoint"); eng = load_pkg("engine")
    cfg, w, shared = ck.read_v0(path)
    return eng.Engine(eng.DeviceWeights(cfg, w, shared_classifier=shared), batch)


def engine_from_v2(path, batch):
    ck = load_pkg("checkpoint"); eng = load_pkg("engine")
    cfg, gs, norms, qt, shared = ck.read_v2(path)
    return eng.Engine(eng.DeviceWeightsQ8(cfg, gs, norms, qt), batch)
