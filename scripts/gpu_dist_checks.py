#!/usr/bin/env python
"""Multi-GPU checks in ONE process over same-process peer access: fused all-reduce vs a PyTorch sum, broadcast,
data-parallel engines (bit-identical replicas, agreement with a single large-batch oracle).
Reference invariant checked: all ranks end with identical weights / test loss (DistTrain_mnist.ipynb:526-548)."""
import ctypes as C
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch

from cori_intml_examples_b200.ops import _lib


def symbufs(world, nfloats):
    L = _lib.lib()
    hs = (C.c_void_p * world)()
    for r in range(world):
        _lib.check(L.cib_init(r))
        h = C.c_void_p()
        _lib.check(L.cib_symbuf_create(C.byref(h), r, nfloats))
        hs[r] = h
    _lib.check(L.cib_symbuf_attach_local(hs, world))
    ptrs = []
    for r in range(world):
        p = C.c_void_p(); _lib.check(L.cib_symbuf_ptr(hs[r], C.byref(p))); ptrs.append(p.value)
    return L, hs, ptrs


def as_tensor(ptr, n, dev):
    """float32 tensor view on raw device memory (through the CUDA array interface)."""
    class _W(object):
        pass
    w = _W()
    w.__cuda_array_interface__ = {"shape": (n,), "typestr": "<f4", "data": (ptr, False), "version": 3}
    with torch.cuda.device(dev):
        return torch.as_tensor(w, device="cuda:%d" % dev)


def check_allreduce(world):
    out = []
    nmax = 16 * 1024 * 1024
    L, hs, ptrs = symbufs(world, nmax)
    views = [as_tensor(ptrs[r], nmax, r) for r in range(world)]
    for n in (1024, 4096, 65536, 1 << 20, nmax):
        for mode in (1, 2):
            if mode == 1 and n * 4 > (1 << 20):
                continue
            for rep in range(3):
                data = [torch.randn(n, device="cuda:%d" % r, generator=torch.Generator(device="cuda:%d" % r).manual_seed(100 * rep + r)) for r in range(world)]
                ref = sum(d.cpu().double() for d in data)
                for r in range(world):
                    views[r][:n].copy_(data[r])
                for r in range(world):
                    torch.cuda.synchronize(r)
                for r in range(world):
                    _lib.check(L.cib_symbuf_allreduce(hs[r], n, rep % 2, mode, None))
                for r in range(world):
                    torch.cuda.synchronize(r)
                exp = (ref / world if rep % 2 else ref).float()
                got = [views[r][:n].cpu() for r in range(world)]
                err = max(float((g - exp).abs().max()) for g in got)
                same = all(torch.equal(got[0], g) for g in got[1:])
                out.append({"n": n, "mode": mode, "rep": rep, "err": err, "replicas_identical": same, "ok": err < 1e-4 and same})
    # broadcast
    for r in range(world):
        views[r][:4096].fill_(float(r + 1))
    for r in range(world):
        torch.cuda.synchronize(r)
    for r in range(world):
        _lib.check(L.cib_symbuf_broadcast(hs[r], 4096, world - 1, None))
    for r in range(world):
        torch.cuda.synchronize(r)
    out.append({"broadcast_ok": all(float(views[r][:4096].min()) == float(world) == float(views[r][:4096].max()) for r in range(world)),
                "ok": all(float(views[r][:4096].min()) == float(world) for r in range(world))})
    return out


def check_dp_engines(world, steps=4, B=16):
    from cori_intml_examples_b200.models import rpv
    from cori_intml_examples_b200.utils import synthetic
    x, y, _ = synthetic.rpv_like(B * world, seed=3, channels=3)
    y = y.reshape(-1, 1)
    models = []
    for r in range(world):
        torch.cuda.set_device(r)
        m = rpv.build_model((64, 64, 3), [16, 32, 64], [128], dropout=0.0, optimizer="Adam", lr=1e-3, device="cuda:%d" % r)
        models.append(m)
    w0 = models[0].get_weights()
    L = _lib.lib()
    engs = (C.c_void_p * world)()
    for r, m in enumerate(models):
        m.set_weights(w0)
        m.backend._ensure(B)
        engs[r] = m.backend.engine
    comm = C.c_void_p()
    _lib.check(L.cib_comm_create_local(C.byref(comm), engs, world))
    res = {}
    for mode in (2, 1):
        for r, m in enumerate(models):
            m.set_weights(w0); m.backend.reset_optimizer(); m.optimizer.iterations = 0; m.backend.push_hyper()
            m.backend.attach_comm(comm, r, mode)
            m.backend._bind(x[r * B:(r + 1) * B], y[r * B:(r + 1) * B], None)
        for s in range(steps):
            for r, m in enumerate(models):                      # async launches; the kernels meet on the GPUs
                torch.cuda.set_device(r)
                _lib.check(L.cib_engine_train_steps(m.backend.engine, 1))
            for r, m in enumerate(models):
                m.backend._bind(x[r * B:(r + 1) * B], y[r * B:(r + 1) * B], None)   # rewind the cursor (synchronises)
        ws = [m.get_weights() for m in models]
        identical = all(all(np.array_equal(a, b) for a, b in zip(ws[0], w)) for w in ws[1:])
        iters = models[0].backend.metrics().iterations
        res["mode%d" % mode] = {"replicas_identical": identical, "iterations": int(iters), "ok": bool(identical and iters == steps)}
    # Reduced gradient against a single-GPU engine fed the concatenated world x B batch (SURVEY section 4 tier 3: DP training equals
    # single-GPU training with an N x batch).  One SGD(lr = 1, momentum = 0) step moves the weights by exactly minus the gradient, so
    # w0 - w1 IS the (reduced) gradient: compared at tf32 tolerance, before any adaptive optimizer can amplify sign noise.
    torch.cuda.set_device(0)
    single = rpv.build_model((64, 64, 3), [16, 32, 64], [128], dropout=None, optimizer="SGD", lr=1.0, device="cuda:0")
    single.set_weights(w0)
    single.backend._ensure(B * world); single.backend._bind(x, y, None)
    _lib.check(L.cib_engine_train_steps(single.backend.engine, 1))
    g_ref = [a - b for a, b in zip(w0, single.get_weights())]
    single.backend.close()
    sgd = []
    for r in range(world):
        torch.cuda.set_device(r)
        m = rpv.build_model((64, 64, 3), [16, 32, 64], [128], dropout=None, optimizer="SGD", lr=1.0, device="cuda:%d" % r)
        m.set_weights(w0); m.backend._ensure(B); sgd.append(m)
    engs2 = (C.c_void_p * world)(*[m.backend.engine for m in sgd])
    comm2 = C.c_void_p()
    _lib.check(L.cib_comm_create_local(C.byref(comm2), engs2, world))
    for mode in (2, 1):
        for r, m in enumerate(sgd):
            m.set_weights(w0); m.backend.attach_comm(comm2, r, mode)
            m.backend._bind(x[r * B:(r + 1) * B], y[r * B:(r + 1) * B], None)
        for r, m in enumerate(sgd):
            torch.cuda.set_device(r)
            _lib.check(L.cib_engine_train_steps(m.backend.engine, 1))
        g = [a - b for a, b in zip(w0, sgd[0].get_weights())]
        rel = [float(np.abs(u - v).max() / (np.abs(v).max() + 1e-12)) for u, v in zip(g, g_ref)]
        res["mode%d" % mode]["reduced_gradient_max_rel_err_vs_single_gpu"] = max(rel)
        res["mode%d" % mode]["ok"] = bool(res["mode%d" % mode]["ok"] and max(rel) < 2e-2)
    return res


if __name__ == "__main__":
    world = min(torch.cuda.device_count(), int(sys.argv[1]) if len(sys.argv) > 1 else 8)
    out = {"world": world, "allreduce": check_allreduce(world)}
    print("allreduce cases:", len(out["allreduce"]), "bad:", [c for c in out["allreduce"] if not c["ok"]][:5], flush=True)
    out["dp"] = check_dp_engines(world)
    print("dp:", json.dumps(out["dp"])[:900], flush=True)
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    json.dump(out, open(os.path.join(ROOT, "gpurun_out", "dist_checks_%dgpu.json" % world), "w"), indent=1)
