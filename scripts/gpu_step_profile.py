#!/usr/bin/env python
"""Runs a few training steps of one model so that `ncu --metrics gpu__time_duration.sum` can list its kernels.
   python scripts/gpu_step_profile.py mnist 32 64 128   |   rpv 16 32 64 128   |   legacy"""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np

kind = sys.argv[1] if len(sys.argv) > 1 else "mnist"
sizes = [int(v) for v in sys.argv[2:]]
B = int(os.environ.get("BATCH", "128"))
steps = int(os.environ.get("STEPS", "10"))
if kind == "mnist":
    from cori_intml_examples_b200.models import mnist
    h1, h2, h3 = sizes or (32, 64, 128)
    x, y, _, _ = mnist.load_data(B * steps, 16)
    m = mnist.build_model(h1=h1, h2=h2, h3=h3, dropout=0.25, optimizer="Adam", device="gpu")
elif kind == "rpv":
    from cori_intml_examples_b200.models import rpv
    from cori_intml_examples_b200.utils import synthetic
    c = sizes or (16, 32, 64, 128)
    x, y, _ = synthetic.rpv_like(B * steps, seed=0, channels=int(os.environ.get("CH", "1")))
    m = rpv.build_model(x.shape[1:], conv_sizes=list(c[:-1]), fc_sizes=[c[-1]], dropout=0.2, optimizer="Adam", device="gpu")
else:
    from cori_intml_examples_b200.models import rpv
    from cori_intml_examples_b200.utils import synthetic
    x, y, _ = synthetic.rpv_like(B * steps, seed=0)
    m = rpv.build_legacy_model(x.shape[1:], device="gpu")
m._ensure_backend() if hasattr(m, "_ensure_backend") else None
for ep in range(2):
    t = time.time()
    h = m.fit(x, y, batch_size=B, epochs=1, verbose=0, shuffle=True)
    print("epoch %d: %d steps in %.2f ms -> %.3f ms/step  loss %.4f" % (ep, steps, 1e3 * (time.time() - t), 1e3 * (time.time() - t) / steps, h.history["loss"][-1]), flush=True)
try:
    print(m.backend.plan.describe())
except Exception as e:
    print("no plan:", e)
