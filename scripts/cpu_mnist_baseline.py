#!/usr/bin/env python
"""BASELINE.json config 1: mnist.build_model (defaults, 37,562 parameters) trained in a single process on the CPU, batch 128,
1 epoch of 60,000 synthetic 28x28 images -- our CPU oracle backend (PyTorch CPU ops).  Context number from the reference:
10-11 s/epoch on a Cori Haswell node with 3 TF threads (GridSearchCV_mnist.ipynb:199-229).  Also times the same epoch on the
GPU engine when one is present."""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch

from cori_intml_examples_b200.models import mnist

threads = int(os.environ.get("CPU_THREADS", "3"))
torch.set_num_threads(threads)
x, y, xt, yt = mnist.load_data(60000, 10000)
res = {"config": "mnist.build_model() defaults, batch 128, 1 epoch x 60000 synthetic 28x28", "cpu_threads": threads}
m = mnist.build_model(device="cpu")
t = time.time(); h = m.fit(x, y, batch_size=128, epochs=1, verbose=0); res["cpu_s_per_epoch"] = time.time() - t
res["cpu_loss"] = h.history["loss"][-1]
if torch.cuda.is_available():
    g = mnist.build_model(device="gpu")
    g.fit(x, y, batch_size=128, epochs=1, verbose=0)            # warm-up: graph capture, dataset upload
    t = time.time(); h = g.fit(x, y, batch_size=128, epochs=1, verbose=0); res["gpu_s_per_epoch"] = time.time() - t
    res["gpu_loss"] = h.history["loss"][-1]
print(json.dumps(res))
out = os.path.join(ROOT, "gpurun_out", "cpu_mnist_baseline.json")
os.makedirs(os.path.dirname(out), exist_ok=True)
json.dump(res, open(out, "w"), indent=1)
