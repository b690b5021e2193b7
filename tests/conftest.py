import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, 'pytorch-kaldi-cgs_b200')
for p in (PKG, ROOT):
    if p not in sys.path:
        sys.path.insert(0, p)
GOLDEN = os.path.join(ROOT, 'tests', 'golden')


def pytest_configure(config):
    config.addinivalue_line('markers', 'gpu: needs a CUDA device (run on the B200 box)')
    # the reference's own modules (tests/test_reference_loop.py) carry old-style regex literals
    config.addinivalue_line('filterwarnings', 'ignore::SyntaxWarning')
    config.addinivalue_line('filterwarnings', 'ignore::DeprecationWarning')


@pytest.fixture(scope='session')
def golden_dir():
    return GOLDEN
