import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, 'tests', 'golden')


def pytest_configure(config):
    config.addinivalue_line('markers', 'gpu: needs a CUDA device (run on the B200 box with -m gpu)')


@pytest.fixture(scope='session')
def oracle():
    """The CPU oracle (test infrastructure only)."""
    sys.path.insert(0, os.path.join(ROOT, 'oracle'))
    import sfm_oracle
    return sfm_oracle


@pytest.fixture(scope='session')
def engine():
    import torch
    if not torch.cuda.is_available():
        pytest.skip('no CUDA device')
    import sfm_b200
    eng = sfm_b200.Engine(device=0)
    yield eng
    eng.close()


def golden(name):
    import numpy as np
    return np.load(os.path.join(GOLDEN, name))
