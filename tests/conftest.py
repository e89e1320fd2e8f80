import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session", autouse=True)
def _built():
    """Make sure the CPU-side libraries exist (cheap no-op when they are already built)."""
    import subprocess
    need = [os.path.join(ROOT, "oracle", "libm17port.so"), os.path.join(ROOT, "m17_cxx_demod_b200", "libm17gen.so")]
    if not all(os.path.exists(p) for p in need):
        subprocess.run(["make", "port"], cwd=os.path.join(ROOT, "oracle"), check=True)
        subprocess.run(["make", "libm17gen.so"], cwd=os.path.join(ROOT, "m17_cxx_demod_b200"), check=True)
    yield
