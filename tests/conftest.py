import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.dirname(os.path.abspath(__file__))):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs at least one CUDA device (run on the B200 box)")
    config.addinivalue_line("markers", "multigpu: needs two or more CUDA devices")


def pytest_collection_modifyitems(config, items):
    try:
        import torch

        ndev = torch.cuda.device_count() if torch.cuda.is_available() else 0
    except Exception:
        ndev = 0
    skip_gpu = pytest.mark.skip(reason="no CUDA device")
    skip_multi = pytest.mark.skip(reason="needs >= 2 CUDA devices")
    for item in items:
        if "gpu" in item.keywords and ndev == 0:
            item.add_marker(skip_gpu)
        if "multigpu" in item.keywords and ndev < 2:
            item.add_marker(skip_multi)
