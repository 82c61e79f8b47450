#!/usr/bin/env python
"""In-kernel timeline of the fused gradient all-reduce inside a real training step (torchrun, one rank per GPU).

Enables cib_engine_ar_trace (globaltimer stamps written by CTA 0 of the push all-reduce kernels and by five one-thread marker
kernels in the step graph), replays a few steps of the bench configuration and prints, per rank and relative to the step start:
when the early (dense) bucket became eligible, when its kernel started / saw all peers' flags / finished reducing + pushing
parameters / saw the parameter flags, when the backward joined, and the same for the last (convolution) bucket.
The roofline the all-reduce is compared against: bytes that must cross NVLink per rank / 900 GB/s.
"""
import ctypes as C
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch
    from cori_intml_examples_b200.models import rpv
    from cori_intml_examples_b200.parallel import dist as hvd
    from cori_intml_examples_b200.utils import synthetic
    from cori_intml_examples_b200.ops import _lib
    hvd.init()
    rank, world = hvd.rank(), hvd.size()
    dev = hvd.local_rank() % torch.cuda.device_count()
    torch.cuda.set_device(dev)
    B = 128
    model = rpv.build_model((64, 64, 3), conv_sizes=[16, 32, 64], fc_sizes=[128], dropout=0.2, optimizer="Adam", lr=0.001 * world,
                            use_horovod=world > 1, device="cuda:%d" % dev)
    if world > 1:
        model.set_weights([hvd.broadcast(w, 0) for w in model.get_weights()])
    x, y, _ = synthetic.rpv_like(2048, seed=rank, channels=3)
    be = model.backend
    be._ensure(B)
    be._bind(x, y.reshape(-1, 1), np.arange(len(x), dtype=np.int32))
    L = be.lib
    _lib.check(L.cib_engine_set_wrap(be.engine, 1))
    _lib.check(L.cib_engine_ar_trace(be.engine, 1, None))
    hvd.barrier()
    _lib.check(L.cib_engine_train_steps(be.engine, 10))
    rows = []
    for _ in range(5):
        _lib.check(L.cib_engine_train_steps(be.engine, 1))
        buf = (C.c_ulonglong * 32)()
        _lib.check(L.cib_engine_ar_trace(be.engine, 0, buf))
        t = np.array(list(buf), dtype=np.float64)
        t0 = t[16]
        rel = lambda v: None if v == 0 else round((v - t0) / 1e3, 2)
        rows.append({"early_eligible": rel(t[17]), "early_start": rel(t[8]), "early_flags_seen": rel(t[9]), "early_reduced": rel(t[10]),
                     "early_param_flags_seen": rel(t[11]), "early_done_marker": rel(t[18]), "backward_joined": rel(t[19]),
                     "last_start": rel(t[0]), "last_flags_seen": rel(t[1]), "last_reduced": rel(t[2]), "last_param_flags_seen": rel(t[3]),
                     "step_done": rel(t[20])})
    nparam = be.plan.nparam
    out = {"rank": rank, "world": world, "unit": "us after step start (globaltimer)", "steps": rows,
           "bytes_per_rank_over_nvlink": int(2 * (world - 1) / world * nparam * 4),
           "roofline_us_at_900GBs": round(2 * (world - 1) / world * nparam * 4 / 900e9 * 1e6, 2)}
    allr = hvd.allgather_object(out)
    if rank == 0:
        print(json.dumps(allr))
    hvd.shutdown()


if __name__ == "__main__":
    main()
