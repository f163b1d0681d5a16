import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line('markers', 'gpu: needs a CUDA device (run on the B200 box)')


@pytest.fixture(scope='session', autouse=True)
def _built_libraries():
    """Build the CPU oracle (and, where /root/reference exists, the reference host build) plus the C-ABI
    library if they are missing.  On the GPU box the prebuilt files travel with the snapshot."""
    from oracle import sampler
    if not os.path.exists(os.path.join(ROOT, 'oracle', '_build', 'liboracle.so')):
        sampler.build(ref=True)
    from arcnerf_b200 import build as b
    if not os.path.exists(b.LIB_PATH):
        b.build()
    yield


def has_gpu():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


@pytest.fixture(scope='session')
def cuda():
    import torch
    if not torch.cuda.is_available():
        pytest.skip('no CUDA device')
    return torch.device('cuda:0')
