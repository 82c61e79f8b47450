import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
os.environ.setdefault("CIB_CLUSTER_DIR", os.path.join("/tmp", "cib200_test_clusters_%d" % os.getpid()))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 GPU and the built libcib200.so (run with -m gpu)")


def pytest_collection_modifyitems(config, items):
    try:
        import torch
        has_gpu = torch.cuda.is_available()
    except Exception:
        has_gpu = False
    if has_gpu:
        return
    skip = pytest.mark.skip(reason="no GPU on this machine")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


@pytest.fixture
def fresh_names():
    from cori_intml_examples_b200.nn import layers
    layers.reset_name_counters()
    yield
