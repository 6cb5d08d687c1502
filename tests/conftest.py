import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, 'oracle')):
    if p not in sys.path:
        sys.path.insert(0, p)
GOLDEN = os.path.join(ROOT, 'tests', 'golden')


def pytest_configure(config):
    config.addinivalue_line('markers', 'gpu: needs a CUDA device (run on the B200 box)')


@pytest.fixture(scope='session')
def built_lib():
    """The in-tree CUDA library; built on demand (nvcc cross-compiles without a GPU)."""
    import __graft_entry__ as g
    g.build()
    from contact_zones_b200 import _lib
    return _lib.load()


def golden_cases():
    """The fixed-state cases `A_...` .. `G_...` (oracle/make_golden.py); run_*, results_*, ops, swaps are other fixtures."""
    return sorted(f[:-4] for f in os.listdir(GOLDEN) if f.endswith('.npz') and f[0].isupper() and f[1] == '_')
