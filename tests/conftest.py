import os
import sys
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, 'tests', 'golden')


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope='session')
def orc():
    """The CPU oracle (test infrastructure): oracle/oracle.py over oracle/_build/libbsgo.so."""
    from oracle import oracle
    oracle.lib()
    return oracle


def load_golden(name):
    import numpy as np
    return np.load(os.path.join(GOLDEN, f'golden_{name}.npz'))
