import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def pytest_collection_modifyitems(config, items):
    # keep the no-alloc entry point's test file (test_zz_*) last, whatever the collection order of the rest
    items.sort(key=lambda it: (os.path.basename(str(it.fspath)).startswith("test_zz_"),))
