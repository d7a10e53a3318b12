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
ROOT = os.path.dirname(HERE)
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libpkroute.so")
# The nvcc command line lives in ONE place, deploy/Makefile.pkroute — the target a maintainer adds
# to the reference's Makefile (north_star: "the Makefile/Dockerfile (nvcc sm_100a build)").  This
# module only drives it, so what __graft_entry__.build() compiles is what a deployment compiles.
MAKEFILE = os.path.join(ROOT, "deploy", "Makefile.pkroute")


def _make(*args: str) -> None:
    if not shutil.which("make"):
        raise RuntimeError("make not found: libpkroute.so cannot be built (there is no CPU fallback)")
    subprocess.check_call(["make", "--no-print-directory", "-f", MAKEFILE, *args], cwd=ROOT)


def build(force: bool = False, verbose: bool = False, defines=(), out: str | None = None) -> str:
    """defines/out: build a tuning variant (e.g. defines=["PK_MIN_CTAS=4"]) to another path."""
    target = os.path.abspath(out) if out else LIB
    args = ["libpkroute-force" if (force or out) else "libpkroute", f"OUT={target}"]
    if defines:
        args.append("DEFINES=" + " ".join(defines))
    if verbose:
        args.append("PK_VERBOSE=1")
    _make(*args)
    return target


def build_latency_bench(force: bool = False) -> str:
    """g++ build of host/latency_bench.cpp (the load generator that drives the C++ batching
    front-end; bench.py's `latency` / `e2e.with_pack` legs run it) into build/latency_bench."""
    exe = os.path.join(ROOT, "build", "latency_bench")
    if force and os.path.exists(exe):
        os.remove(exe)
    _make("latency-bench")
    return exe


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
