import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def pytest_collection_modifyitems(config, items):
    """A stalled kernel (a barrier that never completes) would otherwise hold the GPU box until the caller's limit and
    leave no output: with pytest-timeout present every test gets 300 s, enforced from a watchdog thread (a signal
    cannot interrupt a thread blocked inside a CUDA call), which prints the stacks and ends the run."""
    if not config.pluginmanager.hasplugin("timeout"):
        return
    for item in items:
        if item.get_closest_marker("timeout") is None:
            item.add_marker(pytest.mark.timeout(300, method="thread"))


@pytest.fixture(scope="session")
def golden_dir():
    return GOLDEN
