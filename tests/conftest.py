import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")
    config.addinivalue_line("markers", "multigpu: needs >= 2 CUDA devices")


def pytest_collection_modifyitems(config, items):
    import torch
    if torch.cuda.is_available():
        ngpu = torch.cuda.device_count()
        skip_multi = pytest.mark.skip(reason="needs >= 2 GPUs")
        for it in items:
            if "multigpu" in it.keywords and ngpu < 2:
                it.add_marker(skip_multi)
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for it in items:
        if "gpu" in it.keywords or "multigpu" in it.keywords:
            it.add_marker(skip)
