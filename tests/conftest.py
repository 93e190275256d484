import os
import sys

import pytest

# Virtual ranks (CFD_B200_VIRTUAL_RANKS=1: several slabs of a gang on ONE GPU, tests/test_gpu_runner.py) wait for each
# other inside kernels; two of their streams must never share a hardware work queue, or the waiting kernel blocks
# the one it waits for.  Must be set before the CUDA context exists.  (One rank per GPU needs nothing of the kind.)
os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")
# ... and no kernel may be lazily LOADED while another rank's kernel on the same GPU waits for it: loading a
# kernel on first launch can synchronise the context (CUDA programming guide, "Lazy Loading: concurrent execution").
os.environ.setdefault("CUDA_MODULE_LOADING", "EAGER")

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def _cuda_ok():
    try:
        from cfd_b200 import lib
        return bool(lib.load().gpu_is_available())
    except Exception:
        return False


def pytest_collection_modifyitems(config, items):
    if _cuda_ok():
        return
    skip = pytest.mark.skip(reason="no CUDA device in this container")
    for it in items:
        if "gpu" in it.keywords:
            it.add_marker(skip)


@pytest.fixture(scope="session")
def orc():
    import oracle
    return oracle.Oracle()


@pytest.fixture(scope="session")
def ref():
    import oracle
    try:
        return oracle.Ref()
    except FileNotFoundError as e:
        pytest.skip(str(e))
