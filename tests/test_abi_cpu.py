"""CPU: libpkroute.so loads, exports every symbol include/pk_route.h declares, agrees with the
header on struct sizes, and FAILS LOUDLY without a GPU (no CPU fallback in the product)."""
import ctypes as C
import os
import re
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    names = set()
    for hdr in ("pk_route.h", "pk_gcm.h"):
        src = open(os.path.join(ROOT, "include", hdr)).read()
        names |= set(re.findall(r"^(?:PK_API|PK_GCM_API)[^;(]*?\b(pk_\w+)\s*\(", src, flags=re.M))
    return sorted(names)


def test_every_declared_symbol_is_exported(pk):
    names = declared_symbols()
    assert len(names) >= 19 and "pk_route_batch" in names and "pk_init" in names
    L = pk.lib()
    missing = [n for n in names if not hasattr(L, n)]
    assert not missing, missing
    out = subprocess.check_output(["nm", "-D", "--defined-only", pk.LIB_PATH], text=True)
    exported = sorted(l.split()[-1] for l in out.splitlines() if " T " in l)
    assert [e for e in exported if e.startswith("pk_")] == names    # nothing undeclared leaks out


def test_struct_sizes_match_header(pk, tmp_path):
    prog = tmp_path / "sz.c"
    prog.write_text('#include <stdio.h>\n#include "pk_route.h"\nint main(){printf("%zu %zu %zu %zu %zu\\n",'
                    'sizeof(pk_tool),sizeof(pk_provider),sizeof(pk_table),sizeof(pk_decision),sizeof(pk_config));return 0;}')
    exe = tmp_path / "sz"
    subprocess.check_call(["gcc", "-I", os.path.join(ROOT, "include"), "-o", str(exe), str(prog)])
    sizes = [int(x) for x in subprocess.check_output([str(exe)], text=True).split()]
    assert sizes == [C.sizeof(pk.PkTool), C.sizeof(pk.PkProvider), C.sizeof(pk.PkTable),
                     pk.DECISION_DTYPE.itemsize, C.sizeof(pk.PkConfig)]
    assert sizes[3] == 24


def test_default_table_is_the_reference_switch(pk):
    t = pk.default_table()                      # mock.go:33,37,52 and default (:60)
    got = [(bytes(t.tools[k].name[:t.tools[k].name_len]), t.tools[k].class_id) for k in range(t.n_tools)]
    assert got == [(b"example_tool", 0), (b"struct_tool", 1), (b"file_tool", 2)] and t.default_class == 3
    assert t.n_providers == 8


def test_product_table_equals_oracle_table(pk, oracle):
    a, b = pk.default_table(), oracle.default_table()
    assert bytes(C.string_at(C.addressof(a), C.sizeof(a))) == bytes(C.string_at(C.addressof(b), C.sizeof(b)))


def test_strerror_and_version(pk):
    L = pk.lib()
    assert L.pk_abi_version() == 1
    assert L.pk_strerror(0) == b"ok" and b"16" in L.pk_strerror(-2)
    assert all(L.pk_strerror(c) for c in range(-9, 1))


def test_fails_loudly_without_gpu(pk, has_cuda):
    if has_cuda:
        pytest.skip("a CUDA device is present")
    with pytest.raises(pk.PkError) as e:
        pk.Router(device=0, max_requests=16, max_payload_bytes=1 << 16)
    assert e.value.code == -6                   # PK_ERR_NO_DEVICE: no silent CPU path


def test_null_and_bad_arguments(pk):
    L = pk.lib()
    assert L.pk_init(0, None, None, None) == -1
    h = C.c_void_p()
    cfg = pk.PkConfig(0, 1, 1 << 16, 1 << 10)
    t = pk.default_table()
    assert L.pk_init(0, C.byref(t), C.byref(cfg), C.byref(h)) == -1       # max_requests == 0
    t.n_tools = 99
    cfg = pk.PkConfig(16, 1, 1 << 16, 1 << 10)
    assert L.pk_init(0, C.byref(t), C.byref(cfg), C.byref(h)) == -1       # bad table
    assert L.pk_sync(None) == -1 and L.pk_wait(None, 0) == -1


def test_product_never_touches_the_oracle():
    """The shipped path must not import, link or name anything under oracle/."""
    pkg = os.path.join(ROOT, "polykey_service_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".hpp", ".h", ".go")):
                src = open(os.path.join(dirpath, f), errors="replace").read()
                assert "pk_oracle" not in src and "libpkoracle" not in src and "pk_ref_" not in src, f
    out = subprocess.check_output(["ldd", os.path.join(pkg, "libpkroute.so")], text=True)
    assert "oracle" not in out


def test_the_deploy_makefile_is_the_build(pk):
    """The nvcc command of deploy/Makefile.pkroute — the target a maintainer adds to the reference's
    Makefile (north_star: "the Makefile/Dockerfile (nvcc sm_100a build)") — is what builds
    libpkroute.so here: sm_100a only, every CUDA source, and the library is up to date with it."""
    import subprocess
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    dry = subprocess.run(["make", "-n", "-B", "-f", os.path.join(root, "deploy", "Makefile.pkroute"), "libpkroute"],
                         capture_output=True, text=True, cwd=root)
    assert dry.returncode == 0, dry.stderr
    cmd = dry.stdout
    assert "arch=compute_100a,code=sm_100a" in cmd and cmd.count("-gencode") == 1 and "-lineinfo" in cmd
    for src in ("pk_kernels.cu", "pk_abi.cu", "pk_noalloc.cu", "pk_gcm.cu"):
        assert src in cmd
    up_to_date = subprocess.run(["make", "-q", "-f", os.path.join(root, "deploy", "Makefile.pkroute"), "libpkroute"], cwd=root)
    assert up_to_date.returncode == 0                                 # build.py went through this very target
    from polykey_service_b200 import build
    assert build.MAKEFILE.endswith("deploy/Makefile.pkroute")


def test_deploy_files_are_well_formed():
    """deploy/compose.gpu.yml parses and asks for what the reference's compose.yml (compose.yml:24-31:
    0.5 CPU, 512 MiB, no device) does not: a GPU reservation; deploy/Dockerfile.gpu keeps the
    reference's stage names and builds with cgo on through the same Makefile target."""
    import yaml
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    comp = yaml.safe_load(open(os.path.join(root, "deploy", "compose.gpu.yml")))
    svc = comp["services"]["polykey-server"]
    dev = svc["deploy"]["resources"]["reservations"]["devices"][0]
    assert dev["driver"] == "nvidia" and "gpu" in dev["capabilities"]
    assert float(svc["deploy"]["resources"]["limits"]["cpus"]) >= 2
    assert svc["build"]["dockerfile"] == "deploy/Dockerfile.gpu"
    df = open(os.path.join(root, "deploy", "Dockerfile.gpu")).read()
    for stage in ("AS builder", "AS tester", "AS production", "AS server"):
        assert stage in df
    assert "make -f deploy/Makefile.pkroute libpkroute" in df and "make -f deploy/Makefile.pkroute build-gpu" in df
    mk = open(os.path.join(root, "deploy", "Makefile.pkroute")).read()
    assert "CGO_ENABLED=1" in mk and "CGO_ENABLED=0" not in mk.split("build-gpu:")[1]
