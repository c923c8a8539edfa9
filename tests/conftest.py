import os
import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session")
def oracle():
    import oracle_lib
    return oracle_lib.load()


@pytest.fixture(scope="session")
def vbctx():
    import torch
    if not torch.cuda.is_available():
        pytest.fail("GPU test selected but no CUDA device is present (there is no CPU fallback)")
    from vision_b200 import Context
    ctx = Context(0)
    yield ctx
    ctx.close()
