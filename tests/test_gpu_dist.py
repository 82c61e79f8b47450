"""Multi-GPU checks (need >= 2 GPUs in one box; skipped on a single-GPU box): fused peer-to-peer all-reduce against a PyTorch sum,
broadcast, data-parallel engines staying bit-identical, and a barrier stress test (many back-to-back collectives so the
epoch counters of the flag barrier advance thousands of times, alternating one-shot and two-shot)."""
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "scripts"))

pytestmark = pytest.mark.gpu


def _ngpu():
    try:
        import torch
        return torch.cuda.device_count()
    except Exception:
        return 0


needs2 = pytest.mark.skipif(_ngpu() < 2, reason="needs at least 2 GPUs")


@pytest.fixture(autouse=True)
def _back_to_device_0():
    """The single-process multi-GPU checks walk torch.cuda.set_device over all GPUs; later tests in the same process create their
    tensors on "the current device" and must find device 0 again."""
    yield
    try:
        import torch
        if torch.cuda.is_available():
            torch.cuda.set_device(0)
    except Exception:
        pass


@needs2
def test_fused_allreduce_and_broadcast_match_pytorch():
    import gpu_dist_checks as D
    res = D.check_allreduce(min(_ngpu(), 8))
    assert res and all(r["ok"] for r in res), [r for r in res if not r["ok"]]


@needs2
def test_data_parallel_engines_stay_bit_identical():
    import gpu_dist_checks as D
    res = D.check_dp_engines(min(_ngpu(), 8))
    for mode in ("mode1", "mode2"):
        assert res[mode]["replicas_identical"] and res[mode]["ok"], res[mode]


@needs2
def test_barrier_stress_many_back_to_back_collectives():
    import torch
    import gpu_dist_checks as D
    from cori_intml_examples_b200.ops import _lib
    world = min(_ngpu(), 8)
    n = 4096
    L, hs, ptrs = D.symbufs(world, n)
    views = [D.as_tensor(ptrs[r], n, r) for r in range(world)]
    for r in range(world):
        views[r].fill_(1.0)
    for r in range(world):
        torch.cuda.synchronize(r)
    iters = 3000
    for it in range(iters):                      # averaging keeps the value at exactly 1.0 whatever the order
        for r in range(world):
            _lib.check(L.cib_symbuf_allreduce(hs[r], n, 1, 1 + (it & 1), None))
    for r in range(world):
        torch.cuda.synchronize(r)
    err = (_lib.C.c_uint * 4)()
    assert all(float(v.min()) == 1.0 == float(v.max()) for v in views)
    _lib.check(L.cib_device_error(err))
    assert list(err) == [0, 0, 0, 0], "a bounded barrier wait tripped: %s" % [hex(x) for x in err]


@needs2
def test_nvls_multicast_allreduce_matches_sum_or_is_refused_cleanly():
    """In-switch reduction (multimem.ld_reduce / multimem.st).  Hosts that do not expose NVSwitch multicast must refuse with a
    message, not crash; where it is available the result equals the sum on every replica."""
    import ctypes as C
    import torch
    import gpu_dist_checks as D
    from cori_intml_examples_b200.ops import _lib
    world = min(_ngpu(), 8)
    L = _lib.lib()
    for r in range(world):
        _lib.check(L.cib_init(r))
    n = 1 << 18
    h = C.c_void_p()
    if L.cib_nvls_create_local(C.byref(h), world, n):
        msg = L.cib_last_error().decode()
        assert "nvls" in msg
        pytest.skip("multicast not available here: " + msg)
    views = []
    for r in range(world):
        p = C.c_void_p(); _lib.check(L.cib_nvls_ptr(h, r, C.byref(p)))
        views.append(D.as_tensor(p.value, n, r))
    for rep in range(3):
        data = [torch.randn(n, device="cuda:%d" % r) for r in range(world)]
        ref = sum(d.cpu().double() for d in data)
        ref = (ref / world if rep == 1 else ref).float()
        for r in range(world):
            views[r].copy_(data[r]); torch.cuda.synchronize(r)
        for r in range(world):
            _lib.check(L.cib_nvls_allreduce(h, r, n, int(rep == 1), None))
        for r in range(world):
            torch.cuda.synchronize(r)
        got = [v.cpu() for v in views]
        assert max(float((g - ref).abs().max()) for g in got) < 1e-4
        assert all(torch.equal(got[0], g) for g in got[1:])
    _lib.check(L.cib_nvls_destroy(h))


def _torchrun(nproc, script_args, env_extra=None, timeout=600):
    import json
    import socket
    import subprocess
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    env = dict(os.environ); env.update(env_extra or {})
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(nproc), "--master-addr", "127.0.0.1",
           "--master-port", str(port)] + script_args
    r = subprocess.run(cmd, cwd=ROOT, env=env, capture_output=True, text=True, timeout=timeout)
    lines = [json.loads(l) for l in r.stdout.splitlines() if l.startswith("{")]
    return r, lines


@needs2
def test_multiprocess_dp_gradient_replicas_and_slots_over_cuda_ipc():
    """The path bench.py / train_rpv.py really use: one PROCESS per GPU, arenas mapped through CUDA IPC.  Reduced gradient ==
    single-GPU gradient of the concatenated batch (tf32 tolerance, via an SGD lr=1 step), replicas bit-identical after 50 steps of
    Adam / Nadam / Adadelta in one-shot and two-shot mode, optimizer moments complete on every rank after the slot gather."""
    r, lines = _torchrun(2, [os.path.join(ROOT, "scripts", "gpu_dp_mp_check.py")])
    assert lines, (r.returncode, r.stdout[-2000:], r.stderr[-3000:])
    res = lines[-1]
    assert res["ok"] and r.returncode == 0, (r.returncode, res, r.stderr[-1500:])
    assert all(c["ok"] for c in res["gradient"].values()), res["gradient"]
    assert all(c["ok"] for c in res["replicas"].values()), res["replicas"]


@needs2
def test_dead_rank_becomes_a_device_error_not_a_hang():
    """Rank 1 exits before its first collective step: rank 0's fused all-reduce must give up after CIB_PEER_TIMEOUT_S with an error
    code (cib_device_error / a failed CUDA call) instead of spinning for ever."""
    r, lines = _torchrun(2, [os.path.join(ROOT, "scripts", "gpu_dp_mp_check.py"), "--die-rank", "1"], {"CIB_PEER_TIMEOUT_S": "3"}, timeout=240)
    wd = [l for l in lines if l.get("watchdog")]
    assert wd and wd[0]["rc"] != 0, (r.returncode, r.stdout[-2000:], r.stderr[-2000:])


@needs2
def test_multiprocess_nvls_training_path_matches_the_single_gpu_gradient():
    """CIB_NVLS=1: every rank binds its gradient buffer into one NVSwitch multicast object (the handle travels between the processes
    as a POSIX file descriptor) and the large bucket is reduced inside the switch by the fused kernel.  Same oracle as the
    peer-to-peer path: reduced gradient == single-GPU gradient, replicas bit-identical, moments complete.  Where the platform does
    not expose multicast the set-up must fall back to the peer-to-peer kernels on ALL ranks and still pass."""
    r, lines = _torchrun(2, [os.path.join(ROOT, "scripts", "gpu_dp_mp_check.py")], {"CIB_NVLS": "1"})
    assert lines, (r.returncode, r.stdout[-2000:], r.stderr[-3000:])
    res = lines[-1]
    assert res["ok"] and r.returncode == 0, (r.returncode, res, r.stderr[-1500:])
    flags = {c["nvls"] for c in res["gradient"].values()}
    assert len(flags) == 1                                   # every engine of the run made the same choice
