import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def golden_ops():
    import torch
    return torch.load(os.path.join(ROOT, "tests", "golden", "ops.pt"))


@pytest.fixture(scope="session")
def golden_models():
    import torch
    return torch.load(os.path.join(ROOT, "tests", "golden", "models.pt"))


@pytest.fixture(scope="session")
def native():
    """The ctypes binding; building is `__graft_entry__.build()`'s job, loading must work."""
    import nano_qwen3tts_vllm_b200._native as nat
    nat.lib()
    return nat
