import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "rust-raytracing_b200")
for p in (ROOT, PKG):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session", autouse=True)
def _built():
    """Native artefacts (CUDA library, oracle) are built once per session; up-to-date files are reused."""
    import __graft_entry__ as g
    g.build()


@pytest.fixture(scope="session")
def oracle():
    from oracle import oracle as o
    o.load()
    return o


@pytest.fixture(scope="session")
def rtx():
    import rtx_b200
    return rtx_b200


@pytest.fixture(scope="session")
def gpu(rtx):
    if rtx.device_count() < 1:
        pytest.fail("no CUDA device: -m gpu tests must run on the GPU box (there is no CPU fallback)")
    return 0


@pytest.fixture(scope="session")
def libm_check():
    """Builds and runs tests/native/sincos_check.cpp: (exit code, output).  Exit code 0 = the host build of
    rtx_sincos.cuh returns the bits of this machine's libm sin/cos on every sampled argument."""
    import subprocess
    src = os.path.join(ROOT, "tests", "native", "sincos_check.cpp")
    exe = os.path.join(PKG, "lib", "sincos_check")
    os.makedirs(os.path.dirname(exe), exist_ok=True)
    subprocess.check_call([os.environ.get("CXX", "g++"), "-O2", "-std=c++17", "-ffp-contract=off", "-mfma",
                           "-I", os.path.join(PKG, "csrc"), src, "-o", exe, "-lm"])
    out = subprocess.run([exe, "2000000"], capture_output=True, text=True, timeout=600)
    return out.returncode, out.stdout + out.stderr
