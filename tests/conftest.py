import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden", "reference_outputs.npz")
REFERENCE = "/root/reference"


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (sm_100a) GPU; run with -m gpu on the GPU box")


@pytest.fixture(scope="session")
def golden():
    return np.load(GOLDEN, allow_pickle=False)


@pytest.fixture(scope="session")
def ref_modules():
    """The live reference, only available in the build container."""
    if not os.path.isdir(os.path.join(REFERENCE, "binary")):
        pytest.skip("/root/reference not present (GPU box): live-reference checks run in the build container")
    sys.path.insert(0, REFERENCE)
    import binary.modules as rm
    import binary.utils as ru
    return rm, ru


@pytest.fixture(scope="session")
def cuda():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    major, _ = torch.cuda.get_device_capability(0)
    assert major == 10, "these tests expect a B200 (sm_100)"
    return torch.device("cuda:0")
