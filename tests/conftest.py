import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def keys():
    """Seeded key set in TFHEpp's raw layouts, from the oracle's generator."""
    from oracle import orc
    return orc.Keys(20261104)


@pytest.fixture(scope="session")
def engine(keys):
    """The CUDA engine initialised on cuda:0 with `keys` (GPU tests only)."""
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    import cufhe_b200 as cf
    cf.SetGPUNum(1)
    cf.Initialize(keys.bk, keys.ksk)
    yield cf
    cf.CleanUp()
