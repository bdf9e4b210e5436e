import sys
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parent.parent
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))

GOLDEN = Path(__file__).resolve().parent / "golden"


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (sm_100a); run on the B200 box")


def pytest_collection_modifyitems(config, items):
    """`pytest tests` on a box without a CUDA device: the gpu-marked tests are skipped (with the reason), not
    failed; `-m gpu` on the B200 box runs them all."""
    try:
        import torch
        have = torch.cuda.is_available()
    except Exception:
        have = False
    if have:
        return
    skip = pytest.mark.skip(reason="needs a CUDA device (sm_100a); aircraft_b200 has no CPU path")
    for it in items:
        if "gpu" in it.keywords:
            it.add_marker(skip)


@pytest.fixture(scope="session")
def built():
    """The in-tree CUDA library and the CPU checkers exist (build once per session)."""
    import __graft_entry__ as g
    g.build_cuda()
    g.build_oracle()
    return True


@pytest.fixture(scope="session")
def trim_com():
    from aircraft_b200.config import TRIM_COM
    return TRIM_COM


@pytest.fixture(scope="session", params=["trim_com", "json_com"])
def com(request):
    from aircraft_b200.config import TRIM_COM
    return TRIM_COM if request.param == "trim_com" else (0.133, 0.0, 0.003)


@pytest.fixture(scope="session")
def golden_dir():
    return GOLDEN
