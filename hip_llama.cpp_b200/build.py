"""In-tree build recipes: nvcc for sm_100a only, g++ for the C++ host layer.

    python hip_llama.cpp_b200/build.py            # everything that is out of date
    python hip_llama.cpp_b200/build.py --force

Outputs (git-ignored, but they travel to the GPU box with the snapshot):
    hip_llama.cpp_b200/lib/libtha_b200.so    CUDA kernels + the extern "C" boundary (include/tha_b200.h)
    hip_llama.cpp_b200/lib/libthablas.so     C++ host layer with the reference's names (include/thaBLAS.hpp ...)
    build/apps/llama                         the batching driver (reference: Makefile:9-26 -> build/apps/llama)
    build/tests/*                            C++ API tests
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

PKG = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(PKG)
CSRC = os.path.join(PKG, "csrc")
HOST = os.path.join(PKG, "host")
LIB = os.path.join(PKG, "lib")
OBJ = os.path.join(ROOT, "build", "obj")
INCLUDE = os.path.join(ROOT, "include")

NVCC = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
CXX = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else "g++"
CUDA_HOME = os.path.dirname(os.path.dirname(os.path.realpath(NVCC)))

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
    "-ccbin", CXX, "-Xcompiler", "-fPIC", "-I", INCLUDE, "-I", CSRC,
]


def _newer(target: str, deps: list[str]) -> bool:
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps if os.path.exists(d))


def _run(cmd: list[str]) -> None:
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if r.returncode != 0:
        sys.stderr.write(" ".join(cmd) + "\n" + r.stdout + "\n")
        raise RuntimeError(f"build step failed: {os.path.basename(cmd[-1])}")


def cuda_sources() -> list[str]:
    return sorted(os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith(".cu"))


def headers() -> list[str]:
    hs = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cuh", ".h"))]
    for d, _, fs in os.walk(INCLUDE):
        hs += [os.path.join(d, f) for f in fs]
    return hs


def build_cabi(force: bool = False) -> str:
    """nvcc every csrc/*.cu for sm_100a and link libtha_b200.so."""
    os.makedirs(LIB, exist_ok=True)
    os.makedirs(OBJ, exist_ok=True)
    so = os.path.join(LIB, "libtha_b200.so")
    hdrs = headers()
    jobs = []
    objs = []
    for src in cuda_sources():
        obj = os.path.join(OBJ, os.path.basename(src)[:-3] + ".o")
        objs.append(obj)
        if force or _newer(obj, [src] + hdrs):
            jobs.append([NVCC] + NVCC_FLAGS + ["-c", src, "-o", obj])
    with ThreadPoolExecutor(max_workers=min(8, max(1, len(jobs)))) as ex:
        list(ex.map(_run, jobs))
    if force or jobs or _newer(so, objs):
        _run([NVCC, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-ccbin", CXX, "-o", so] + objs +
             ["-cudart", "static"])
    return so


def build_host(force: bool = False) -> str | None:
    """g++ the C++ host layer (reference-named API) against libtha_b200.so."""
    if not os.path.isdir(HOST):
        return None
    srcs = sorted(os.path.join(HOST, f) for f in os.listdir(HOST) if f.endswith(".cpp"))
    if not srcs:
        return None
    so = os.path.join(LIB, "libthablas.so")
    if force or _newer(so, srcs + headers() + [os.path.join(LIB, "libtha_b200.so")]):
        # no CUDA headers, no cudart: the host layer reaches the GPU only through the C-ABI
        _run([CXX, "-O2", "-std=c++17", "-fPIC", "-shared", "-fopenmp", "-I", INCLUDE, "-o", so] + srcs +
             ["-L", LIB, "-ltha_b200", "-Wl,-rpath,$ORIGIN"])
    return so


def build_programs(force: bool = False) -> list[str]:
    """The driver app and the C++ API tests (sources under apps/ and tests/cpp/)."""
    out = []
    targets = []
    apps = os.path.join(PKG, "apps")
    if os.path.isdir(apps):
        for f in sorted(os.listdir(apps)):
            if f.endswith(".cpp"):
                targets.append((os.path.join(apps, f), os.path.join(ROOT, "build", "apps", f[:-4])))
    tcpp = os.path.join(ROOT, "tests", "cpp")
    if os.path.isdir(tcpp):
        for f in sorted(os.listdir(tcpp)):
            if f.endswith(".cpp"):
                targets.append((os.path.join(tcpp, f), os.path.join(ROOT, "build", "tests", f[:-4])))
    for src, exe in targets:
        os.makedirs(os.path.dirname(exe), exist_ok=True)
        if force or _newer(exe, [src] + headers() + [os.path.join(LIB, "libthablas.so")]):
            _run([CXX, "-O2", "-std=c++17", "-fopenmp", "-I", INCLUDE, "-o", exe, src, "-L", LIB, "-lthablas",
                  "-ltha_b200", "-Wl,-rpath," + LIB])
        out.append(exe)
    return out


def build_all(force: bool = False) -> dict:
    res = {"cabi": build_cabi(force)}
    res["host"] = build_host(force)
    res["programs"] = build_programs(force) if res["host"] else []
    return res


if __name__ == "__main__":
    print(build_all(force="--force" in sys.argv))
