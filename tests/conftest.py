import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def lib_built():
    """The C-ABI library must exist for every test that touches the product; build it if the tree is fresh."""
    from fcaf3d_b200 import build
    return build.build()


@pytest.fixture(scope="session")
def oracle_built():
    from oracle import build_oracle
    return build_oracle.build()


@pytest.fixture(scope="session")
def dev(lib_built):
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    return torch.device("cuda", 0)


GOLDEN = os.path.join(ROOT, "tests", "golden")
