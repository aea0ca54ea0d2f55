import sys
from pathlib import Path

import pytest
import torch

ROOT = Path(__file__).resolve().parent.parent
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))

GOLDEN = ROOT / "tests" / "golden"


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def load_golden(name: str):
    return torch.load(GOLDEN / name, weights_only=False)


@pytest.fixture(scope="session")
def cuda_lib():
    """The C-ABI library, loaded once.  GPU tests fail loudly when it is missing."""
    from vllm_b200 import _lib

    return _lib.load()
