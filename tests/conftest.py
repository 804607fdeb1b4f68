import os
import sys
import glob
import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with `-m gpu`)")


@pytest.fixture(scope="session", autouse=True)
def _built():
    """Make sure the C oracle and libgvss.so exist (compiling needs no GPU)."""
    import oracle
    from duckdb_vss_b200 import capi
    if not oracle.have_port() or (os.path.isdir("/root/reference") and not oracle.have_ref()):
        oracle.build()
    if not os.path.exists(capi.LIB_PATH):
        import __graft_entry__
        __graft_entry__.build()
    yield


def golden_cases():
    return sorted(os.path.basename(p)[:-4] for p in glob.glob(os.path.join(GOLDEN, "*.npz")))


def load_golden(name):
    z = np.load(os.path.join(GOLDEN, name + ".npz"), allow_pickle=False)
    blob = open(os.path.join(GOLDEN, name + ".usearch"), "rb").read()
    return z, blob


@pytest.fixture(scope="session")
def checker():
    """The CPU checker for GPU parity tests: the compiled reference when present, else the pinned C port."""
    import oracle
    if oracle.have_ref():
        from oracle import ref
        return ref.RefIndex, "reference"
    from oracle import port
    return port.PortIndex, "port"
