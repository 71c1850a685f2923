import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def dev():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    return torch.device("cuda:0")


def pytest_report_header(config):
    try:
        from yolo_dl_b200 import _lib
        return f"yolohot library: {_lib.LIB_PATH} — {_lib.load().yh_version().decode()}"
    except Exception as e:  # the CPU suite still reports why
        return f"yolohot library not loaded: {e}"
