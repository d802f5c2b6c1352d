import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200); run with -m gpu")


@pytest.fixture(scope="session")
def port():
    """The C restatement of the reference algorithm (oracle/nns_oracle.c)."""
    from oracle import oracle as O
    return O.port()


@pytest.fixture(scope="session")
def ref():
    """The reference's own code (oracle/_ref/libnns_ref.so); skipped when not built."""
    from oracle import oracle as O
    if not O.ref_available():
        pytest.skip("oracle/_ref/libnns_ref.so not built and /root/reference absent")
    return O.ref()


@pytest.fixture(scope="session")
def dev():
    """One GPU context + stream for the whole session (fails loudly without a GPU)."""
    from nnstreamer_b200 import Device
    d = Device(0)
    yield d
    d.close()
