import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "oracle")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box via gpurun)")


@pytest.fixture(scope="session")
def orc():
    import oracle  # oracle/oracle.py -- the checker; tests are allowed to import it

    oracle.build()
    return oracle


@pytest.fixture(scope="session")
def synth():
    import k8sgpu_b200  # noqa: F401  (registers the package)
    from k8sgpu_b200 import synth as s

    return s
