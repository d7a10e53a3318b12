import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def _has_cuda() -> bool:
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


@pytest.fixture(scope="session")
def oracle():
    """The CPU oracle (test infrastructure); compiles oracle/libpkoracle.so if needed."""
    from oracle import pk_oracle
    pk_oracle.build()
    pk_oracle.lib()
    return pk_oracle


@pytest.fixture(scope="session")
def pk():
    """The product binding; building libpkroute.so needs nvcc but no GPU."""
    import polykey_service_b200 as pk
    if not os.path.exists(pk.LIB_PATH):
        from polykey_service_b200 import build
        build.build()
    pk.lib()
    return pk


@pytest.fixture(scope="session")
def has_cuda():
    return _has_cuda()


def to_pk_table(pk, oracle_table):
    """Oracle and product declare the same C struct; copy bytes across."""
    import ctypes as C
    return pk.PkTable.from_buffer_copy(bytes(C.string_at(C.addressof(oracle_table), C.sizeof(oracle_table))))
