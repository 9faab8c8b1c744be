import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def O():
    """The CPU oracle (test infrastructure)."""
    from oracle import oracle

    oracle.build()
    return oracle


@pytest.fixture(scope="session")
def hsv_golden():
    return np.load(os.path.join(GOLDEN, "hsv_cv2.npz"))


@pytest.fixture(scope="session")
def ref_fixtures():
    return np.load(os.path.join(GOLDEN, "reference_fixtures.npz"))


@pytest.fixture(scope="session")
def pcs():
    import point_cloud_segmentation_b200 as p

    return p


@pytest.fixture(scope="session")
def ctx(pcs):
    """One pcs_ctx on cuda:0 through the C ABI.  No CUDA device -> the test fails loudly (there is no fallback)."""
    c = pcs.Context(0)
    yield c
    c.destroy()
