import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with -m gpu on the GPU box)")
    config.addinivalue_line("markers", "slow: long-running (Llama-2-7B sized) case")


def pytest_collection_modifyitems(config, items):
    # `-m gpu` on a machine without a GPU is a usage error, not a skip: the product has no CPU path.
    pass
