"""Allocation-free dense-layer forward, called through the C ABI (this file is named so that pytest collects it last).

Who needs it: an HPO trial that shares its GPU with other trials must not call cudaMalloc/cudaFree while running (both
synchronise the whole device and would stall the neighbours' graph replays), and a trial confined to the HBM slice the
scheduler gave it cannot ask for more.  `cib_dense_relu_fwd_noalloc` therefore keeps its split-K scratch in an array
built into the library instead of taking a buffer from the caller.

The test calls the existing entry point (caller-provided scratch) and the allocation-free one on the same inputs, twice in
a row -- the built-in scratch is shared by all calls and keeps the previous call's contents, so the second call would
expose a missing clear -- and also checks that a problem larger than the built-in scratch is refused with a clear message.
"""
import ctypes as C

import pytest

pytestmark = pytest.mark.gpu

TOL = 6e-3      # same relative tolerance as the other tf32 kernel tests


def _run(M, N, K, seed):
    import torch
    from cori_intml_examples_b200.ops import _lib
    L = _lib.lib()
    _lib.check(L.cib_init(0))
    torch.cuda.set_device(0)
    g = torch.Generator(device="cuda").manual_seed(seed)
    A = torch.randn(M, K, device="cuda", generator=g)
    W = torch.randn(K, N, device="cuda", generator=g) * 0.05
    b = torch.randn(N, device="cuda", generator=g)
    ref_out = torch.empty(M, N, device="cuda")
    z = torch.zeros(M, N, device="cuda")
    _lib.check(L.cib_dense_relu_fwd(_lib.ptr(A), _lib.ptr(W), _lib.ptr(b), _lib.ptr(ref_out), _lib.ptr(z), M, N, K, None))
    outs = []
    for _ in range(2):                                    # two calls in a row on the same (default) stream
        out = torch.full((M, N), float("nan"), device="cuda")
        rc = L.cib_dense_relu_fwd_noalloc(C.c_void_p(A.data_ptr()), C.c_void_p(W.data_ptr()), C.c_void_p(b.data_ptr()),
                                          C.c_void_p(out.data_ptr()), M, N, K, None)
        assert rc == 0, L.cib_last_error().decode()
        outs.append(out)
    torch.cuda.synchronize()
    fp32 = torch.relu(A @ W + b)
    scale = float(ref_out.abs().max()) + 1e-12
    for out in outs:
        assert torch.isfinite(out).all()
        assert float((out - ref_out).abs().max()) / scale < TOL          # agrees with the existing entry point
        assert float((out - fp32).abs().max()) / (float(fp32.abs().max()) + 1e-12) < TOL
    return outs


def test_noalloc_matches_existing_entry_point_twice_in_a_row():
    _run(128, 128, 4096, seed=11)


def test_noalloc_smaller_problem_after_a_larger_one():
    # a different shape right after: the shared scratch still holds the previous call's layout and must be cleared
    _run(128, 128, 4096, seed=12)
    _run(64, 32, 2304, seed=13)


def test_noalloc_rejects_problems_beyond_the_built_in_scratch():
    import torch
    from cori_intml_examples_b200.ops import _lib
    L = _lib.lib()
    _lib.check(L.cib_init(0))
    M, N, K = 512, 256, 64                               # 131072 floats > 65536
    A = torch.zeros(M, K, device="cuda"); W = torch.zeros(K, N, device="cuda"); b = torch.zeros(N, device="cuda")
    out = torch.zeros(M, N, device="cuda")
    rc = L.cib_dense_relu_fwd_noalloc(_lib.ptr(A), _lib.ptr(W), _lib.ptr(b), _lib.ptr(out), M, N, K, None)
    assert rc != 0
    msg = L.cib_last_error().decode()
    assert "exceeds the built-in scratch" in msg and "65536" in msg
