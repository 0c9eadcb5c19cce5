import os
import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))
os.environ.setdefault("PYTHONDONTWRITEBYTECODE", "1")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (sm_100a) — run with `-m gpu`")


@pytest.fixture(scope="session")
def cuda_lib():
    """The product's CUDA library; GPU tests must fail loudly (not skip) if it cannot be used."""
    import torch

    from s2t_b200 import lib

    assert torch.cuda.is_available(), "GPU test selected but no CUDA device is visible"
    handle = lib.load()
    arch = handle.s2t_device_arch()
    assert arch == 100, f"s2t_b200 kernels are sm_100a-only; device reports arch {arch}"
    return handle
