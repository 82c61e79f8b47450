#!/usr/bin/env python
"""Times the first-layer tensor-core kernels (conv1_tc.cuh) on the shapes of the headline models, CUDA events, L2 flushed by
rotating over input copies larger than L2.  Prints one JSON line per shape."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch

from cori_intml_examples_b200.ops import _lib

L = _lib.lib(); _lib.check(L.cib_init(0))
SHAPES = [("rpv_3_16", 128, 64, 64, 3, 16, 1, 1), ("rpv_1_16", 128, 64, 64, 1, 16, 1, 1), ("mnist_1_32", 128, 28, 28, 1, 32, 0, 0),
          ("rpv_3_64", 128, 64, 64, 3, 64, 1, 1)]
only = sys.argv[1] if len(sys.argv) > 1 else ""
for name, n, H, W, Ci, Co, pad, pool in SHAPES:
    if only and only not in name:
        continue
    Ho, Wo = H + 2 * pad - 2, W + 2 * pad - 2
    Hs, Ws = (Ho // 2, Wo // 2) if pool else (Ho, Wo)
    copies = max(2, int(200e6 // (n * H * W * Ci * 4)) + 1)
    xs = [torch.randn(n, H, W, Ci, device="cuda") for _ in range(min(copies, 40))]
    w = torch.randn(3, 3, Ci, Co, device="cuda") * 0.2
    b = torch.zeros(Co, device="cuda")
    out = torch.empty(n, Hs, Ws, Co, device="cuda")
    g = torch.randn(n, Hs, Ws, Co, device="cuda")
    dW = torch.zeros(3, 3, Ci, Co, device="cuda"); db = torch.zeros(Co, device="cuda")
    res = {"shape": name}
    for what in ("fwd", "wgrad"):
        def run(i):
            x = xs[i % len(xs)]
            if what == "fwd":
                _lib.check(L.cib_conv1_fwd_tc(_lib.ptr(x), _lib.ptr(w), _lib.ptr(b), _lib.ptr(out), n, H, W, Ci, Co, pad, pool, None))
            else:
                _lib.check(L.cib_conv1_wgrad_tc(_lib.ptr(x), _lib.ptr(g), _lib.ptr(out), _lib.ptr(dW), _lib.ptr(db), n, H, W, Ci, Co, pad, pool, None))
        for i in range(5):
            run(i)
        torch.cuda.synchronize()
        a, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for i in range(50):
            run(i)
        e.record(); torch.cuda.synchronize()
        res[what + "_us"] = round(a.elapsed_time(e) * 1e3 / 50, 2)
    print(json.dumps(res), flush=True)
