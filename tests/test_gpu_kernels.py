"""sm_100a kernels vs plain PyTorch fp32 references (tf32 tolerance).  Needs a B200: `pytest -m gpu`."""
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "scripts"))

pytestmark = pytest.mark.gpu


def _cases():
    import gpu_kernel_checks as K
    return K.CASES


def test_library_is_loaded_and_device_is_blackwell():
    import torch
    from cori_intml_examples_b200.ops import _lib
    assert _lib.available(), "libcib200.so must be built in-tree (scripts/build.sh)"
    L = _lib.lib()
    _lib.check(L.cib_init(0))
    assert torch.cuda.get_device_capability(0)[0] == 10
    with open("/proc/self/maps") as f:
        assert "libcib200.so" in f.read()          # the native library really is mapped into this process


@pytest.mark.parametrize("name", ["gemm_kk_128", "gemm_kk_k32", "gemm_kmn", "gemm_mnk", "gemm_mnmn", "gemm_splitk", "gemm_m64", "gemm_n256"])
def test_tcgen05_gemm(name):
    assert _cases()[name]()


@pytest.mark.parametrize("name", ["dense_128_4096", "dense_64_2304"])
def test_fused_dense_relu_forward(name):
    assert _cases()[name]()


@pytest.mark.parametrize("name", ["conv_fwd_16_32_pool", "conv_fwd_32_64_pool", "conv_fwd_16_32_valid", "conv_fwd_32_32_nopool", "conv_fwd_64_128", "conv_fwd_128_256_wide"])
def test_tcgen05_conv_forward(name):
    assert _cases()[name]()


@pytest.mark.parametrize("name", ["dgrad_16_32_mask", "dgrad_32_64_unpool", "dgrad_16_16_valid", "dgrad_128_256_wide"])
def test_tcgen05_conv_dgrad(name):
    assert _cases()[name]()


@pytest.mark.parametrize("name", ["wgrad_16_32", "wgrad_32_64", "wgrad_32_32_valid", "wgrad_16_32_valid", "wgrad_64_128", "wgrad_128_256_wide"])
def test_tcgen05_conv_wgrad(name):
    assert _cases()[name]()


@pytest.mark.parametrize("name", ["conv1_fwd_3_16_pool", "conv1_fwd_1_16_pool", "conv1_fwd_1_32_valid", "conv1_fwd_3_64_pool",
                                  "conv1_wgrad_3_16_pool", "conv1_wgrad_1_16_pool", "conv1_wgrad_1_32_valid", "conv1_wgrad_3_64_pool"])
def test_first_layer_tensor_core_kernels(name):
    assert _cases()[name]()


@pytest.mark.parametrize("name", ["conv_direct_3_16", "conv_direct_1_8_valid"])
def test_cuda_core_conv(name):
    assert _cases()[name]()


def test_tma_per_tap_kernels_still_agree():
    """The first-generation (TMA per tap) producers stay as the fall-back for filters that do not fit in shared memory."""
    from cori_intml_examples_b200.ops import _lib
    L = _lib.lib()
    _lib.check(L.cib_set_conv_version(1))
    try:
        c = _cases()
        assert c["conv_fwd_32_64_pool"]() and c["dgrad_32_64_unpool"]() and c["wgrad_32_64"]()
    finally:
        _lib.check(L.cib_set_conv_version(2))
