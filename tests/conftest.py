"""
pytest configuration.

Markers
  gpu        needs a CUDA device (B200); these are the parity tests proper and call the product
             through its C-ABI.  The driver runs `-m gpu` on the GPU box and `-m "not gpu"` on CPU.
  reference  needs /root/reference (build container only); skipped elsewhere.
"""
import pathlib
import sys

import pytest

ROOT = pathlib.Path(__file__).resolve().parent.parent
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))

GOLDEN = ROOT / "tests" / "golden"


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device; parity tests through the C-ABI")
    config.addinivalue_line("markers", "reference: needs /root/reference (build container only)")


def _has_gpu():
    try:
        import torch

        return torch.cuda.is_available()
    except Exception:
        return False


def pytest_collection_modifyitems(config, items):
    has_gpu = None
    ref_ok = pathlib.Path("/root/reference/swarmrl").is_dir()
    for item in items:
        if "gpu" in item.keywords:
            if has_gpu is None:
                has_gpu = _has_gpu()
            if not has_gpu:
                item.add_marker(pytest.mark.skip(reason="no CUDA device"))
        if "reference" in item.keywords and not ref_ok:
            item.add_marker(pytest.mark.skip(reason="/root/reference not present"))


@pytest.fixture(scope="session")
def golden():
    import numpy as np

    def load(name):
        return np.load(GOLDEN / name, allow_pickle=False)

    return load
