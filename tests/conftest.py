import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with `pytest -m gpu` under gpurun)")
    config.addinivalue_line("markers", "multigpu(n): needs at least n GPUs")


def pytest_collection_modifyitems(config, items):
    import torch
    ngpu = torch.cuda.device_count() if torch.cuda.is_available() else 0
    for it in items:
        m = it.get_closest_marker("multigpu")
        if m is not None:
            it.add_marker(pytest.mark.gpu)
            need = m.args[0] if m.args else 2
            if ngpu < need:
                it.add_marker(pytest.mark.skip(reason=f"needs {need} GPUs, have {ngpu}"))
        elif it.get_closest_marker("gpu") is not None and ngpu == 0:
            it.add_marker(pytest.mark.skip(reason="no GPU visible"))


@pytest.fixture
def repo_root():
    return ROOT


@pytest.fixture(autouse=True)
def _fresh_world():
    yield
    import b200dp
    from b200dp.utils import logging as L
    if b200dp.is_initialized() and b200dp.size() == 1:
        b200dp.shutdown()
    L.reset_logger()
