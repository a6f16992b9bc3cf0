import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")
    config.addinivalue_line("markers", "multigpu: needs >= 2 CUDA devices")


def pytest_collection_modifyitems(config, items):
    try:
        import torch
        have = torch.cuda.is_available()
        ndev = torch.cuda.device_count() if have else 0
    except Exception:
        have, ndev = False, 0
    for item in items:
        if "gpu" in item.keywords and not have:
            item.add_marker(pytest.mark.skip(reason="no CUDA device"))
        if "multigpu" in item.keywords and ndev < 2:
            item.add_marker(pytest.mark.skip(reason="needs >= 2 CUDA devices"))


@pytest.fixture(scope="session")
def income():
    from fedb200 import prepare
    return prepare(with_mean=True)


@pytest.fixture(scope="session")
def income_nomean():
    from fedb200 import prepare
    return prepare(with_mean=False)
