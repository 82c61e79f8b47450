#!/usr/bin/env python
"""Diagnostics for trials on a GPU engine: direct in-process timing, then the same through a 1-engine cluster."""
import faulthandler
import functools
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "scripts"))
faulthandler.dump_traceback_later(int(os.environ.get("DIAG_DUMP_AFTER", "100")), exit=True)


def log(*a):
    print("[%7.2f]" % (time.time() - T0), *a, flush=True)


T0 = time.time()
stage = sys.argv[1] if len(sys.argv) > 1 else "direct"
if stage == "direct":
    from cori_intml_examples_b200.models import mnist
    x, y, xt, yt = mnist.load_data(60000, 10000)
    log("data", x.shape)
    for cfg in (dict(h1=4, h2=8, h3=32, dropout=0.5, optimizer="Adadelta"), dict(h1=64, h2=8, h3=128, dropout=0.02, optimizer="Nadam"),
                dict(h1=32, h2=64, h3=128, dropout=0.25, optimizer="Adam")):
        m = mnist.build_model(device="gpu", **cfg)
        log("built", cfg)
        print(m.backend.plan.describe() if hasattr(m, "backend") and hasattr(m.backend, "plan") else "", flush=True)
        t = time.time()
        h = m.fit(x, y, batch_size=128, epochs=2, validation_split=0.17, verbose=2)
        log("fit 2 epochs: %.2f s" % (time.time() - t), {k: round(v[-1], 4) for k, v in h.history.items()})
else:
    import hpo_bench
    from cori_intml_examples_b200.cluster import LocalCluster
    with LocalCluster(1, mode="process", slots=int(os.environ.get("DIAG_SLOTS", "1"))) as cl:
        log("cluster started, gpus", cl.gpus)
        c = cl.client()
        log("client connected", c.ids)
        log("trivial apply ->", c[0].apply_sync(lambda: __import__("os").environ.get("CUDA_VISIBLE_DEVICES")))
        lv = c.load_balanced_view()
        f = functools.partial(hpo_bench.mnist_trial, 6000, 1000, 2, 128, "gpu")
        ars = [lv.apply(f, h1=4, h2=8, h3=32, dropout=0.5, optimizer="Adadelta") for _ in range(int(os.environ.get("DIAG_TASKS", "1")))]
        for ar in ars:
            try:
                r = ar.get(timeout=60)
                log("trial done", {k: (round(v[-1], 4) if isinstance(v, list) else round(v, 3)) for k, v in r.items()})
            except Exception as e:
                log("trial failed:", repr(e)[:2000])
                log("stdout:", ar.stdout[-2000:] if ar.stdout else None)
                log("stderr:", ar.stderr[-2000:] if ar.stderr else None)
        c.close()
    log("done")
