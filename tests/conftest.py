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


@pytest.fixture(autouse=True)
def _gpu_tests_start_on_device_0(request):
    """Every GPU test starts with device 0 current, in PyTorch's own bookkeeping as well as in the CUDA runtime: the multi-GPU tests
    walk over the devices, the library's cib_init() calls cudaSetDevice behind PyTorch's back, and a later test that creates
    tensors on plain "cuda" must not end up with operands on two different GPUs."""
    if "gpu" in request.keywords:
        try:
            import torch
            if torch.cuda.is_available():
                torch.cuda.set_device(0)
        except Exception:
            pass
    yield
