#!/usr/bin/env python
"""Throughput of the reference's old single-GPU model (Train_rpv.ipynb:204-218,281: four 3x3 convolutions with strides 1,2,1,2 at
64/128/256/256 channels + Dense 512, 34,515,201 parameters, batch 128, Adam) -- the one FLOP-bound workload of the reference, which
records 51-56 s per epoch of 64,000 samples = 1,140-1,255 samples/s on its (unnamed) GPU.  Device-timed graph replay, CUDA events."""
import ctypes as C
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np

from cori_intml_examples_b200.models import rpv
from cori_intml_examples_b200.ops import _lib
from cori_intml_examples_b200.utils import synthetic

B, steps = 128, int(os.environ.get("STEPS", "40"))
x, y, _ = synthetic.rpv_like(B * 16, seed=0, channels=1)
m = rpv.build_legacy_model(x.shape[1:], device="gpu")
be = m.backend
be._ensure(B)
be._bind(x, y.reshape(-1, 1), np.arange(len(x), dtype=np.int32))
L = be.lib
_lib.check(L.cib_engine_set_wrap(be.engine, 1))
_lib.check(L.cib_engine_train_steps(be.engine, 5))
_lib.check(L.cib_engine_sync(be.engine))
best = 1e30
for _ in range(3):
    ms = C.c_float(0)
    _lib.check(L.cib_engine_time_steps_aligned(be.engine, 1, steps, C.byref(ms)))
    best = min(best, ms.value / steps)
fwd = 0.0
for b in be.plan.blocks:
    fwd += 2.0 * B * (b.Ho * b.Wo * b.ksize * b.ksize * b.ci_logical * b.co_logical if b.kind == 0 else b.k_logical * b.n_logical)
first = 2.0 * B * be.plan.blocks[0].Ho * be.plan.blocks[0].Wo * 9 * be.plan.blocks[0].ci_logical * be.plan.blocks[0].co_logical
flops = 3 * fwd - first
peaks = {}
try:
    peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
except Exception:
    pass
tf32 = float(peaks.get("bf16_tflops_sustained", 2250.0)) / 2
out = {"model": "build_legacy_model (64,64,1): 34,515,201 params", "batch": B, "ms_per_step": best, "samples_per_s": B / best * 1e3,
       "reference_samples_per_s": [1140, 1255], "step_gflop": flops / 1e9, "achieved_tflops": flops / best / 1e9,
       "fraction_of_tf32_peak": flops / best / 1e9 / tf32, "plan": be.plan.describe().split("\n"), "loss": float(be.metrics().step_loss)}
print(json.dumps(out))
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
json.dump(out, open(os.path.join(ROOT, "gpurun_out", "legacy_bench.json"), "w"), indent=1)
