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


@pytest.fixture(scope="session")
def ref_models(dev, oracle_built):
    """The reference's own model classes (me_resnet.py, fcaf3d_neck_with_head.py, single_stage_sparse.py, unmodified),
    imported by path on the product's operator surfaces.  The files come from the reference checkout where it exists
    (build container) or from the copies oracle/build_oracle.py staged under oracle/_ref/py (GPU box)."""
    from fcaf3d_b200 import compat
    for root in (os.environ.get("FCAF3D_REFERENCE", "/root/reference"), oracle_built.get("ref_py", "")):
        if root and compat.reference_files_present(root):
            return compat.load_reference_models(root)
    pytest.fail("reference model files neither under /root/reference nor staged under oracle/_ref/py: run "
                "__graft_entry__.build() in the build container before shipping the tree")


GOLDEN = os.path.join(ROOT, "tests", "golden")
