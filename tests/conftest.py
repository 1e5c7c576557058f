import os
import sys

import pytest

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")
    # the product library and the checker are built in-tree by __graft_entry__.build(); build them if a fresh
    # checkout has neither (nvcc cross-compiles sm_100a without a GPU)
    import subprocess
    if not os.path.exists(os.path.join(ROOT, "halo_b200", "libhalo_b200.so")):
        subprocess.check_call(["make", "-s", "-j4", "-C", os.path.join(ROOT, "halo_b200", "csrc")])
    if not os.path.exists(os.path.join(ROOT, "oracle", "liboracle.so")):
        subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "oracle")])


@pytest.fixture(scope="session")
def oracle():
    import oracle_lib
    oracle_lib.set_threads(min(16, os.cpu_count() or 1))
    return oracle_lib


@pytest.fixture(scope="session")
def golden_dir():
    return os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
