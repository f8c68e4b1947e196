import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")
REFERENCE = "/root/reference"


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run on the GPU box with -m gpu)")


def pytest_collection_modifyitems(config, items):
    # GPU tests must never silently pass without the CUDA path: fail loudly if selected on a box with no GPU
    pass


@pytest.fixture(scope="session")
def golden_dir():
    return GOLDEN


@pytest.fixture(scope="session")
def ae30():
    d = np.load(os.path.join(GOLDEN, "ae30_5epochs.npz"))
    return {k: d[k] for k in d.files}


@pytest.fixture(scope="session")
def ae30_full():
    d = np.load(os.path.join(GOLDEN, "ae30_100epochs.npz"))
    return {k: d[k] for k in d.files}


@pytest.fixture(scope="session")
def have_reference():
    return os.path.isdir(REFERENCE)


@pytest.fixture(scope="session")
def cuda_device():
    import torch

    assert torch.cuda.is_available(), "GPU test selected but no CUDA device is visible"
    import b200_sensor as bs

    bs._ffi.lib()  # raises if libb2sensor.so is missing: no fallback
    return torch.device("cuda", 0)


@pytest.fixture
def lstm_path(cuda_device):
    """Select the LSTM kernel for one test ("tcgen05" = the tensor-core kernel with one tile per CTA, "tcgen05x2" = its two-tile
    form forced on whatever the batch size, "fp32" = the FMA one) through the context option; the default is restored afterwards."""
    import b200_sensor as bs

    def pick(path):
        bs.keras.set_option("lstm_tc", {"tcgen05": 1, "tcgen05x2": 3, "fp32": 0}[path])

    yield pick
    bs.keras.set_option("lstm_tc", 2)
