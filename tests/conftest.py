import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def golden_dir():
    return GOLDEN


@pytest.fixture(scope="session")
def built_lib():
    """The in-tree library; built on demand so a fresh checkout can run the CPU suite."""
    import b200ddp
    if not os.path.exists(b200ddp._lib.LIB_PATH):
        import __graft_entry__
        __graft_entry__.build()
    return b200ddp._lib.lib()


@pytest.fixture()
def ctx1(built_lib):
    """A one-rank context on cuda:0 with a small staging area."""
    import torch
    import b200ddp
    torch.cuda.set_device(0)
    c = b200ddp.Context(0, 1, 0, staging_bytes=64 << 20)
    yield c
    torch.cuda.synchronize()
    c.close()
