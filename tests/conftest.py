import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200); run with -m gpu on the GPU box")
    config.addinivalue_line("markers", "multigpu: needs at least 2 CUDA devices")


def pytest_collection_modifyitems(config, items):
    import torch

    ngpu = torch.cuda.device_count() if torch.cuda.is_available() else 0
    for item in items:
        if "gpu" in item.keywords and ngpu == 0:
            item.add_marker(pytest.mark.skip(reason="no CUDA device"))
        if "multigpu" in item.keywords and ngpu < 2:
            item.add_marker(pytest.mark.skip(reason="needs >= 2 CUDA devices"))


@pytest.fixture(scope="session")
def golden():
    def load(name):
        return np.load(os.path.join(GOLDEN, name))
    return load


@pytest.fixture(scope="session", autouse=True)
def _built_library():
    """Compile libb200dp.so and the two host shims when one of them is MISSING (a fresh checkout: built
    files are git-ignored; nvcc cross-compiles without a GPU).  Files that exist are used as they are —
    bringing them up to date is __graft_entry__.build()'s job, and a test run on the GPU box must exercise
    exactly the binaries that were shipped to it."""
    from dbx_distributed_pytorch_examples_b200 import _lib
    if not (os.path.exists(_lib.LIB_PATH) and os.path.exists(_lib._fast_path()) and os.path.exists(_lib._autograd_path())):
        _lib.build()
