"""pytest configuration: the ``gpu`` marker and collection rules.

``python -m pytest tests -m "not gpu"`` must pass on a CPU-only box; ``-m gpu`` runs on a B200.
GPU tests are skipped (not failed) when no CUDA device exists.
"""
import os
import sys
import pathlib

import pytest
import torch

ROOT = pathlib.Path(__file__).resolve().parent.parent
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))

os.environ.setdefault("MASTER_ADDR", "127.0.0.1")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200); run with -m gpu")
    config.addinivalue_line("markers", "multigpu: needs >= 2 CUDA devices")


def pytest_collection_modifyitems(config, items):
    has_cuda = torch.cuda.is_available()
    n = torch.cuda.device_count() if has_cuda else 0
    for item in items:
        if "gpu" in item.keywords and not has_cuda:
            item.add_marker(pytest.mark.skip(reason="no CUDA device"))
        if "multigpu" in item.keywords and n < 2:
            item.add_marker(pytest.mark.skip(reason="needs >= 2 GPUs"))
