import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))

GOLDEN = os.path.join(ROOT, "tests", "golden")

# tests/test_gpu_tp.py runs the ranks of a tensor-parallel replica as concurrent streams of ONE device; they hand-shake
# inside their kernels, so two ranks' streams must never share a hardware work queue (the default 8 queues are dealt to
# streams round-robin, and an aliased pair serialises: the first rank then waits for a peer queued behind it).  Must be
# set before the CUDA context exists.
os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def pytest_collection_modifyitems(config, items):
    """`pytest tests` on a box without a CUDA device (or without the built extension) skips the gpu-marked tests
    instead of failing inside torch's CUDA initialisation."""
    if not any("gpu" in it.keywords for it in items):
        return
    why = None
    if not os.path.exists(os.path.join(ROOT, "hip_llama.cpp_b200", "lib", "libtha_b200.so")):
        why = "libtha_b200.so is not built (python __graft_entry__.py)"
    else:
        try:
            import torch
            if not torch.cuda.is_available():
                why = "no CUDA device"
        except ImportError:
            why = "torch not importable"
    if why:
        skip = pytest.mark.skip(reason=why)
        for it in items:
            if "gpu" in it.keywords:
                it.add_marker(skip)


@pytest.fixture(scope="session")
def oracle():
    from oracle_lib import Oracle
    return Oracle()


@pytest.fixture(scope="session")
def golden_dir():
    return GOLDEN
