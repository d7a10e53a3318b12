"""The C++ host mirror of the reference's seam (polykey_service_b200/host/): MockService and
Server against the literals of internal/service/mock.go:22-67 and internal/server/server.go:27-43,
and the BatchingService (MPSC batcher, one flusher per GPU, shard = request id % n_gpus).

CPU: the batcher is driven with a stand-in backend that answers from the oracle, so the host
logic is covered without a GPU.  GPU: the same checks through libpkroute.so (CudaBackend)."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def host_test_exe(tmp_path_factory, oracle, pk):
    exe = str(tmp_path_factory.mktemp("host") / "host_service_test")
    host = os.path.join(ROOT, "polykey_service_b200", "host")
    subprocess.check_call([
        "g++", "-std=c++20", "-O2", "-Wall", "-Wextra", "-pthread", "-o", exe,
        os.path.join(ROOT, "tests", "host_service_test.cpp"),
        os.path.join(host, "service.cpp"), os.path.join(host, "batching_service.cpp"),
        "-L" + os.path.join(ROOT, "oracle"), "-lpkoracle",
        "-L" + os.path.join(ROOT, "polykey_service_b200"), "-lpkroute",
        "-Wl,-rpath," + os.path.join(ROOT, "oracle"),
        "-Wl,-rpath," + os.path.join(ROOT, "polykey_service_b200")])
    return exe


def test_host_mirror_cpu(host_test_exe):
    out = subprocess.run([host_test_exe], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0 and "HOST TESTS OK" in out.stdout, out.stdout + out.stderr


@pytest.mark.gpu
def test_host_mirror_cuda_backend(host_test_exe):
    out = subprocess.run([host_test_exe, "cuda"], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0 and "HOST CUDA TESTS OK" in out.stdout, out.stdout + out.stderr


@pytest.fixture(scope="module")
def security_test_exe(tmp_path_factory, pk):
    exe = str(tmp_path_factory.mktemp("sec") / "host_security_test")
    host = os.path.join(ROOT, "polykey_service_b200", "host")
    subprocess.check_call([
        "g++", "-std=c++20", "-O2", "-Wall", "-Wextra", "-o", exe,
        os.path.join(ROOT, "tests", "host_security_test.cpp"), os.path.join(host, "security.cpp"),
        "-L" + os.path.join(ROOT, "polykey_service_b200"), "-lpkroute",
        "-Wl,-rpath," + os.path.join(ROOT, "polykey_service_b200")])
    return exe


def test_security_mirror_cpu(security_test_exe):
    """Key validation and its message (cipher.go:15-19) need no GPU."""
    out = subprocess.run([security_test_exe], capture_output=True, text=True, timeout=120)
    assert out.returncode == 0 and "SECURITY TESTS OK" in out.stdout, out.stdout + out.stderr


@pytest.mark.gpu
def test_security_mirror_cuda(security_test_exe):
    """Encrypt / Decrypt / BatchEncrypt / BatchDecrypt (cipher.go:92-141) on the device."""
    out = subprocess.run([security_test_exe, "cuda"], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0 and "SECURITY TESTS OK" in out.stdout, out.stdout + out.stderr
