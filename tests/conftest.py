import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session")
def built_libs():
    """Build (or reuse) the in-tree native libraries; CPU-only cross-compile is fine."""
    from starpu_inference_server_b200 import build

    return build.build()


@pytest.fixture(scope="session")
def oracle_gs():
    """ctypes handle on the C oracle for gather/scatter (test infrastructure)."""
    import ctypes
    import subprocess

    odir = os.path.join(ROOT, "oracle")
    so = os.path.join(odir, "_build", "liboracle_gs.so")
    if not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(os.path.join(odir, "gather_scatter.c")):
        subprocess.check_call(["make", "-C", odir, "-s"])
    return ctypes.CDLL(so)
