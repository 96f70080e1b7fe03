import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with `-m gpu`)")


@pytest.fixture(scope="session", autouse=True)
def _build_checkers():
    """Build the C restatement and the host emulation of the kernel source if they are stale."""
    import subprocess
    if not os.path.exists(os.path.join(ROOT, "oracle", "libminicfd_oracle.so")):
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle"), "port"])
    yield
