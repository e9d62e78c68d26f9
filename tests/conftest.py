import importlib.util
import os
import socket
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

REFERENCE_TRAIN = "/root/reference/train.py"
REF_COPY = os.path.join(ROOT, "baseline", "_ref", "train.py")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def pytest_collection_modifyitems(config, items):
    import torch
    if torch.cuda.is_available():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


def free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


@pytest.fixture(scope="session")
def reference_module():
    """The reference's train.py imported as a module (oracle for parity tests)."""
    path = REFERENCE_TRAIN if os.path.exists(REFERENCE_TRAIN) else REF_COPY
    if not os.path.exists(path):
        pytest.skip("reference train.py not available")
    spec = importlib.util.spec_from_file_location("reference_train", path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod
