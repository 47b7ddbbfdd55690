import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN_DIR = os.path.join(ROOT, 'tests', 'golden')


def pytest_configure(config):
    config.addinivalue_line('markers', 'gpu: needs a CUDA device (run on the B200 box)')


@pytest.fixture(scope='session')
def golden():
    """Outputs of the reference's own Python code (tests/golden/make_golden.py)."""
    return dict(np.load(os.path.join(GOLDEN_DIR, 'reference_run.npz')))


@pytest.fixture(scope='session')
def device():
    """Reference fixture tests/fixtures/torch.py:4-11."""
    import torch
    return torch.device('cuda:0') if torch.cuda.is_available() else torch.device('cpu')
