import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with `-m gpu` under gpurun)")


@pytest.fixture(scope="session")
def pkg():
    import adapt_b200

    return adapt_b200


@pytest.fixture(scope="session")
def model50(pkg):
    return pkg.ResNet50(weights=None)
