import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def _has_gpu():
    try:
        import torch

        return torch.cuda.is_available()
    except Exception:
        return False


@pytest.fixture(scope="session")
def ctx():
    """One libmvstereo context on cuda:0.  No fallback: without a GPU the gpu-marked tests must not be selected."""
    from multiview_b200 import api

    c = api.Context(0)
    yield c
    c.close()


@pytest.fixture(scope="session", autouse=True)
def _build_oracle():
    import oracle

    oracle.build()
