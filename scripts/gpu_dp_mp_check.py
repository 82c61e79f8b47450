#!/usr/bin/env python
"""Multi-PROCESS data-parallel correctness on real GPUs (one rank per GPU over the CUDA-IPC path that bench.py / train_rpv.py
use).  Run under torchrun:

    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 scripts/gpu_dp_mp_check.py

Checks (SURVEY section 4 tier 3; reference invariant DistTrain_mnist.ipynb:526-548: every rank prints the same test loss):
  1. gradient  -- one SGD(lr=1, momentum=0) step moves the weights by exactly minus the REDUCED gradient, so w0 - w1 of the
                  N-rank job is compared with w0 - w1 of a single-GPU engine fed the concatenated N x B batch, at tf32 tolerance,
                  before any adaptive optimizer can amplify sign noise; one-shot and two-shot, early bucket on and off.
  2. replicas  -- 50 steps of Adam / Nadam / Adadelta with dropout: parameter buffers hashed on every rank must be bit-identical.
  3. slots     -- after two-shot steps + cib_engine_gather_slots every rank holds the same, complete optimizer moments, close to
                  the one-shot run's (not bit-equal: the wgrad kernels accumulate with atomics).
With CIB_NVLS=1 the two-shot cases run the in-switch (multimem.ld_reduce) kernel on a multicast-bound gradient buffer; each case
reports whether that path was really active ("nvls").
Rank 0 prints ONE JSON line with the verdicts; exit code 0 only if everything passed.
`--die-rank R` makes rank R exit before its first collective step instead (watchdog test: the survivors must see a device
error within CIB_PEER_TIMEOUT_S, not hang).
"""
import argparse
import hashlib
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def digest(arrays):
    h = hashlib.sha256()
    for a in arrays:
        h.update(np.ascontiguousarray(a).tobytes())
    return h.hexdigest()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--die-rank", type=int, default=-1)
    ap.add_argument("--steps", type=int, default=50)
    a = ap.parse_args()
    import torch
    from cori_intml_examples_b200.models import rpv
    from cori_intml_examples_b200.parallel import dist as hvd
    from cori_intml_examples_b200.utils import synthetic
    from cori_intml_examples_b200.ops import _lib

    hvd.init()
    rank, world = hvd.rank(), hvd.size()
    dev = hvd.local_rank() % torch.cuda.device_count()
    torch.cuda.set_device(dev)
    L = _lib.lib()
    B = 16
    x, y, _ = synthetic.rpv_like(B * world, seed=11, channels=3)
    y = y.reshape(-1, 1)
    out = {"world": world}

    def dp_model(opt, lr, mode, dropout, early):
        _lib.check(L.cib_set_early_update(early))
        from cori_intml_examples_b200.nn import optimizers
        o = hvd.DistributedOptimizer(optimizers.get(opt, lr=lr), mode=mode)
        inputs, outputs = rpv._rpv_stack((64, 64, 3), [(c, 1, True) for c in (16, 32, 64)], [128], dropout)
        from cori_intml_examples_b200 import nn
        m = nn.Model(inputs=inputs, outputs=outputs, name="RPVClassifier")
        m.compile(optimizer=o, loss="binary_crossentropy", metrics=["accuracy"], device="cuda:%d" % dev)
        return m

    if a.die_rank >= 0:
        m = dp_model("Adam", 1e-3, 0, 0.0, 1)
        m.set_weights([hvd.broadcast(w, 0) for w in m.get_weights()])
        m.backend._ensure(B)
        m.backend._bind(x[rank * B:(rank + 1) * B], y[rank * B:(rank + 1) * B], None)
        hvd.barrier()
        if rank == a.die_rank:
            os._exit(0)
        rc = L.cib_engine_train_steps(m.backend.engine, 1)
        rc = rc or L.cib_engine_sync(m.backend.engine)
        err = (_lib.C.c_uint * 4)()
        L.cib_device_error(err)
        print(json.dumps({"watchdog": True, "rank": rank, "rc": int(rc), "device_error": [hex(v) for v in err],
                          "message": L.cib_last_error().decode("utf-8", "replace")[:200]}), flush=True)
        os._exit(0 if rc != 0 else 3)

    # ---- single-GPU oracle of the reduced gradient: SGD lr=1 on the concatenated batch
    _lib.check(L.cib_set_early_update(0))
    single = rpv.build_model((64, 64, 3), [16, 32, 64], [128], dropout=None, optimizer="SGD", lr=1.0, device="cuda:%d" % dev)
    w0 = [hvd.broadcast(w, 0) for w in single.get_weights()]
    single.set_weights(w0)
    single.backend._ensure(B * world)
    single.backend._bind(x, y, None)
    _lib.check(L.cib_engine_train_steps(single.backend.engine, 1))
    g_ref = [a0 - b for a0, b in zip(w0, single.get_weights())]
    single.backend.close()

    grad_cases = {}
    for mode in (1, 2):
        for early in (0, 1):
            m = dp_model("SGD", 1.0, mode, None, early)
            m.set_weights(w0)
            m.backend._ensure(B)
            m.backend._bind(x[rank * B:(rank + 1) * B], y[rank * B:(rank + 1) * B], None)
            hvd.barrier()
            _lib.check(L.cib_engine_train_steps(m.backend.engine, 1))
            g = [a0 - b for a0, b in zip(w0, m.get_weights())]
            # tf32 tolerance relative to each tensor's gradient scale
            rel = [float(np.abs(u - v).max() / (np.abs(v).max() + 1e-12)) for u, v in zip(g, g_ref)]
            same = len(set(hvd.allgather_object(digest(m.get_weights())))) == 1
            grad_cases["mode%d_early%d" % (mode, early)] = {"max_rel_err": max(rel), "replicas_identical": same, "ok": bool(max(rel) < 2e-2 and same),
                                                             "nvls": bool(L.cib_engine_nvls_active(m.backend.engine))}
            m.backend.close()
            hvd.barrier()
    out["gradient"] = grad_cases

    rep = {}
    slots_ref = {}
    for opt, lr in (("Adam", 1e-3), ("Nadam", 1e-3), ("Adadelta", 1.0)):
        for mode in (1, 2):
            m = dp_model(opt, lr * world, mode, 0.2, 1)
            m.seed = 7                         # same dropout masks in the one-shot and the two-shot run (and on every rank)
            m.set_weights(w0)
            m.backend._ensure(B)
            m.backend._bind(x[rank * B:(rank + 1) * B], y[rank * B:(rank + 1) * B], None)
            _lib.check(L.cib_engine_set_wrap(m.backend.engine, 1))
            hvd.barrier()
            # stage 1: three steps, then the moments of the two modes are compared while the trajectories still coincide (the wgrad
            # kernels accumulate with atomics, so runs are not bit-equal and drift apart later); a sharded, un-gathered buffer
            # would be (W-1)/W zeros, i.e. a relative distance of ~0.7
            _lib.check(L.cib_engine_train_steps(m.backend.engine, 3))
            m.backend._gather_slots()
            _lib.check(L.cib_engine_sync(m.backend.engine))
            early_slots = [np.concatenate([t.ravel() for t in group]) for group in m.backend.get_slots()]
            _lib.check(L.cib_engine_train_steps(m.backend.engine, a.steps - 3))
            m.backend._gather_slots()
            _lib.check(L.cib_engine_sync(m.backend.engine))
            dw = digest(m.get_weights())
            ds = digest([t for group in m.backend.get_slots() for t in group])
            same_w = len(set(hvd.allgather_object(dw))) == 1
            same_s = len(set(hvd.allgather_object(ds))) == 1
            finite = all(np.isfinite(w).all() for w in m.get_weights())
            it = int(m.backend.metrics().iterations)
            key = "%s_mode%d" % (opt, mode)
            rep[key] = {"weights_identical": same_w, "slots_identical": same_s, "finite": finite, "iterations": it,
                        "ok": bool(same_w and same_s and finite and it == a.steps)}
            if mode == 1:
                slots_ref[opt] = early_slots
            else:
                d = [float(np.linalg.norm(u - v) / (np.linalg.norm(v) + 1e-30)) for u, v in zip(early_slots, slots_ref[opt])]
                rep[key]["slots_rel_dist_to_one_shot_after_3_steps"] = d
                rep[key]["ok"] = bool(rep[key]["ok"] and max(d) < 0.05)
            m.backend.close()
            hvd.barrier()
    out["replicas"] = rep
    err = (_lib.C.c_uint * 4)()
    _lib.check(L.cib_device_error(err))
    out["device_error"] = [int(v) for v in err]
    ok = all(c["ok"] for c in grad_cases.values()) and all(c["ok"] for c in rep.values()) and not any(err)
    out["ok"] = bool(ok)
    oks = hvd.allgather_object(bool(ok))
    if rank == 0:
        out["ok"] = bool(all(oks))
        print(json.dumps(out), flush=True)
    hvd.shutdown()
    return 0 if all(oks) else 1


if __name__ == "__main__":
    sys.exit(main())
