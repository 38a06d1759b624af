import json
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLD = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def golden(name):
    with open(os.path.join(GOLD, name)) as f:
        return json.load(f)


@pytest.fixture(scope="session")
def oracle():
    """CPU oracle (oracle/nx_oracle.c).  Test infrastructure: the checker, never the thing under test."""
    from oracle.oracle import Oracle
    return Oracle()


@pytest.fixture(scope="session")
def ref():
    """The reference's own compiled code (oracle/_ref/libnxref.so) when it was built."""
    from oracle.oracle import Ref
    if not Ref.available():
        pytest.skip("oracle/_ref/libnxref.so not built (needs /root/reference)")
    return Ref()


@pytest.fixture(scope="session")
def nx():
    import nodexa_chain_core_b200 as m
    return m


def _has_gpu():
    try:
        import nodexa_chain_core_b200 as m
        return m.lib.nx_device_count() > 0
    except Exception:
        return False


@pytest.fixture(scope="session")
def gpu(nx):
    if nx.lib.nx_device_count() < 1:
        pytest.fail("no CUDA device visible: -m gpu tests must run on a GPU box")
    return 0


_CTX = {}


@pytest.fixture(scope="session")
def ctx_of(nx, gpu):
    """Epoch contexts on cuda:0, built once per session."""
    def get(epoch):
        if epoch not in _CTX:
            _CTX[epoch] = nx.create_epoch_context_full(epoch, gpu)
        return _CTX[epoch]
    yield get
    for c in _CTX.values():
        c.close()
    _CTX.clear()
