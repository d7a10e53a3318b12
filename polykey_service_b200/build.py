"""In-tree nvcc build of libpkroute.so (sm_100a only; no JIT cache, no other arch).

    python -m polykey_service_b200.build [--force]

The .so is git-ignored but travels to the GPU box with the gpurun snapshot.
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libpkroute.so")
SOURCES = ["pk_kernels.cu", "pk_abi.cu", "pk_noalloc.cu", "pk_gcm.cu"]
HEADERS = ["pk_kernels.cuh", "pk_spec.cuh", "pk_gcm.cuh", os.path.join("..", "..", "include", "pk_route.h"),
           os.path.join("..", "..", "include", "pk_gcm.h")]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
              "-Xcompiler", "-fPIC", "-Xcompiler", "-fvisibility=hidden", "-shared"]


def nvcc() -> str:
    for cand in (shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found: libpkroute.so cannot be built (there is no CPU fallback)")


def stale() -> bool:
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    deps = [os.path.join(CSRC, f) for f in SOURCES + HEADERS] + [os.path.abspath(__file__)]
    return any(os.path.getmtime(d) > t for d in deps if os.path.exists(d))


def build(force: bool = False, verbose: bool = False, defines=(), out: str | None = None) -> str:
    """defines/out: build a tuning variant (e.g. defines=["PK_PIPE=0"]) to another path."""
    if out is None and not force and not stale():
        return LIB
    out = out or LIB
    cmd = [nvcc(), *NVCC_FLAGS, *[f"-D{d}" for d in defines], "-o", out] + [os.path.join(CSRC, f) for f in SOURCES]
    if verbose:
        cmd.insert(1, "-Xptxas=-v")
    subprocess.check_call(cmd)
    return out


def build_latency_bench(force: bool = False) -> str:
    """g++ build of host/latency_bench.cpp (the load generator that drives the C++ batching
    front-end; bench.py's `latency` / `e2e.with_pack` legs run it) into build/latency_bench."""
    root = os.path.dirname(HERE)
    host = os.path.join(HERE, "host")
    exe = os.path.join(root, "build", "latency_bench")
    srcs = [os.path.join(host, f) for f in ("latency_bench.cpp", "batching_service.cpp", "service.cpp")]
    deps = srcs + [os.path.join(host, f) for f in ("batching_service.hpp", "service.hpp")] + [LIB]
    if not force and os.path.exists(exe) and all(os.path.getmtime(d) <= os.path.getmtime(exe) for d in deps if os.path.exists(d)):
        return exe
    os.makedirs(os.path.dirname(exe), exist_ok=True)
    subprocess.check_call(["g++", "-std=c++20", "-O2", "-pthread", "-o", exe, *srcs, "-L" + HERE, "-lpkroute",
                           "-Wl,-rpath," + HERE])
    return exe


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
